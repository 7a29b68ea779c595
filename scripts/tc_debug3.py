"""debug: tensor reverse sweep with the ring, variants, vs generic reverse sweep"""
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
    return val, np.asarray(g), np.asarray(gx), cnf.last_stats(icnf)

for cid, B in (("C3", 256), ("C4", 1024)):
    v0, g0, x0, st0 = run(cid, B, {"tc_backward": 0, "tc_ring": 0})
    for opts in ({"tc_ring": 0}, {}, {"tc_dbg": 1}, {"tc_dbg": 2}, {"tc_dbg": 8}, {"tc_dbg": 16}, {"tc_dbg": 32}):
        try:
            v1, g1, x1, st1 = run(cid, B, opts)
            print(cid, B, opts, f"naccept {st0['naccept']}/{st1['naccept']} grad rel {relnorm(g1, g0):.2e} gradx rel {relnorm(x1, x0):.2e}", flush=True)
        except Exception as e:
            print(cid, B, opts, "FAILED", e, flush=True)
            sys.exit(1)
