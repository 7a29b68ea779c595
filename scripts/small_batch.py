"""small-batch (launch-bound) breakdown: one loss+gradient call at B=32 on the README shape"""
import sys, os, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import args_of, cnf, make_icnf, wl
for cid, B in (("C1", 32), ("C1", 1024), ("C0", 64), ("C3", 128)):
    inp = wl.make_inputs(cid, B=B)
    icnf = make_icnf(cid)
    for it in range(3):
        t = time.perf_counter()
        val, g = cnf.loss_and_grad(icnf, cnf.TrainMode(), inp["xs"].T, *args_of(inp), eps=inp["eps"].T)
        dt = time.perf_counter() - t
    st = cnf.last_stats(icnf)
    print(cid, B, f"wall {1e3*dt:.2f} ms  ms_total {st['ms_total']:.2f} ms_solve {st['ms_solve']:.2f} naccept {st['naccept']} nreject {st['nreject']} launches {st['gpu_launches']} path {st['kernel_path']}", flush=True)
    t = time.perf_counter()
    lp = cnf.inference(icnf, cnf.TestMode(), inp["xs"].T, *args_of(inp))
    dt = time.perf_counter() - t
    st = cnf.last_stats(icnf)
    print("   inference TestMode", f"wall {1e3*dt:.2f} ms  ms_total {st['ms_total']:.2f} ms_solve {st['ms_solve']:.2f} naccept {st['naccept']} launches {st['gpu_launches']} path {st['kernel_path']}", flush=True)
