"""debug/measure the narrow tensor-core stage kernel (kernel_path 6) against the register-resident
one (kernel_path 2) and the oracle: python scripts/ntc_debug.py [B_big]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import O, args_of, cnf, make_icnf, oracle_cfg, relerr, wl  # noqa: E402


def run(cid, B, kp):
    inp = wl.make_inputs(cid, B=B)
    icnf = make_icnf(cid, kernel_path=kp)
    t = time.time()
    lp, regs = cnf.inference(icnf, cnf.TestMode(), inp["xs"].T, *args_of(inp), eps=inp["eps"].T)
    st = cnf.last_stats(icnf)
    return inp, np.asarray(lp), st, time.time() - t


SMALL = () if os.environ.get("NTC_BIG_ONLY") else (("C2", 128), ("C2", 4096), ("C2", 1000), ("C2", 1 << 17))
for cid, B in SMALL:
    inp, lp6, st6, _ = run(cid, B, 0)
    _, lp2, st2, _ = run(cid, B, 2)
    print(f"{cid} B={B}: path {st6['kernel_path']} vs {st2['kernel_path']}: naccept {st6['naccept']}/{st2['naccept']} "
          f"nreject {st6['nreject']}/{st2['nreject']} max rel diff {relerr(lp6, lp2):.3e}", flush=True)
    if B <= 4096:
        c = wl.CONFIGS[cid]
        rlp, _, _, rst = O.logdensity(oracle_cfg(cid), inp["theta"], 0, inp["xs"], None, inp["eps"], *c["tspan"])
        print(f"   oracle naccept {rst['naccept']}: path6 err {relerr(lp6, rlp):.3e}  path2 err {relerr(lp2, rlp):.3e}", flush=True)

Bbig = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
for kp in ((0,) if os.environ.get("NTC_BIG_ONLY") else (0, 2)):
    inp = wl.make_inputs("C2", B=Bbig)
    icnf = make_icnf("C2", kernel_path=kp)
    cnf.set_option(icnf, "time_stages", 1)
    for rep in range(3):
        lp, _ = cnf.inference(icnf, cnf.TestMode(), inp["xs"].T, *args_of(inp), eps=inp["eps"].T)
        st = cnf.last_stats(icnf)
    print(f"B={Bbig} kernel_path {st['kernel_path']}: ms_solve {st['ms_solve']:.3f} stage ms {st['ms_stage'] / max(st['n_stage'], 1):.4f} "
          f"x{st['n_stage']} naccept {st['naccept']} nreject {st['nreject']}", flush=True)

