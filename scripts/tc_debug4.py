"""debug: ring reverse sweep at multi-tile sizes; gradx is deterministic (no atomics) -> exact comparison"""
import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import O, args_of, cnf, make_icnf, oracle_cfg, wl, relnorm

def run(cid, B, opts):
    inp = wl.make_inputs(cid, B=B)
    icnf = make_icnf(cid)
    for k, v in opts.items():
        cnf.set_option(icnf, k, v)
    val, g, gx = cnf.loss_and_grad(icnf, cnf.TrainMode(), inp["xs"].T, *args_of(inp), eps=inp["eps"].T, want_grad_x=True)
    return val, np.asarray(g), np.asarray(gx).T.copy(), cnf.last_stats(icnf)

for cid, B in (("C3", 128), ("C3", 20000), ("C4", 60000)):
    v0, g0, x0, st0 = run(cid, B, {"tc_ring": 0})
    for opts in ({"tc_dbg": 32 + 4}, {"tc_dbg": 32}, {}, {}, {"tc_dbg": 4}):
        v1, g1, x1, st1 = run(cid, B, opts)
        d = np.abs(x1 - x0)
        rows = np.unique(np.argwhere(d > 1e-6 * np.abs(x0).max())[:, 0])
        print(cid, B, opts, f"naccept {st1['naccept']} grad rel {relnorm(g1, g0):.2e} gradx rel {relnorm(x1, x0):.2e} bad rows {len(rows)}",
              (rows[:8].tolist(), (rows[:8] % 128).tolist()) if len(rows) else "", flush=True)
