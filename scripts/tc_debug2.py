"""debug: repeated evaluations / whole solves with the ring on vs off"""
import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import O, args_of, cnf, make_icnf, oracle_cfg, wl

cid = "C3"
c = wl.CONFIGS[cid]; D = c["nvars"] + c["naug"]
for B in (300, 20000):
    inp = wl.make_inputs(cid, B=B)
    res = {}
    for ring in (0, 1):
        icnf = make_icnf(cid)
        cnf.set_option(icnf, "tc_ring", ring)
        outs = []
        for rep in range(3):   # same handle, different inputs: stale data would show from the 2nd call on
            u = np.random.default_rng(5 + rep).normal(0, 1, (B, D + 3)).astype(np.float32)
            outs.append(np.asarray(cnf.augmented_f(icnf, cnf.TrainMode(), u.T, *args_of(inp), eps=inp["eps"].T)).T.copy())
        lp, regs = cnf.inference(icnf, cnf.TrainMode(), inp["xs"].T, *args_of(inp), eps=inp["eps"].T)
        st = cnf.last_stats(icnf)
        res[ring] = (outs, np.asarray(lp).copy(), st["naccept"], st["nreject"])
    for rep in range(3):
        a, b = res[0][0][rep], res[1][0][rep]
        print(f"B={B} aug_f rep {rep}: max |ring - noring| = {np.abs(a - b).max():.3e}")
    print(f"B={B} solve: naccept {res[0][2]}/{res[1][2]} nreject {res[0][3]}/{res[1][3]} max|dlogp| {np.abs(res[0][1]-res[1][1]).max():.3e}", flush=True)
