"""debug helper: stage-level parity of the tensor path vs the oracle, with debug switches"""
import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import O, args_of, cnf, make_icnf, oracle_cfg, wl

cases = (("C3", 300), ("C3", 40000))
for opts in ({"tc_ring": 0}, {}, {"tc_dbg": 1}, {"tc_dbg": 2}, {"tc_dbg": 4}, {"tc_dbg": 5}):
    for cid, B in cases:
        tm = 0
        inp = wl.make_inputs(cid, B=B)
        c = wl.CONFIGS[cid]
        D = c["nvars"] + c["naug"]
        u = np.random.default_rng(5).normal(0, 1, (B, D + 3)).astype(np.float32)
        icnf = make_icnf(cid, trace_mode=tm)
        for k, v in opts.items():
            cnf.set_option(icnf, k, v)
        du = np.asarray(cnf.augmented_f(icnf, cnf.TrainMode(), u.T, *args_of(inp), eps=inp["eps"].T)).T
        ref = O.augmented_f(oracle_cfg(cid, trace_mode=tm), inp["theta"], 1, u, inp["ys"], inp["eps"])
        scale = np.maximum(np.abs(ref).max(axis=0, keepdims=True), 1e-6)
        err = np.abs(du - ref) / scale
        bad = np.argwhere(err > 1e-3)
        rows = np.unique(bad[:, 0]) if len(bad) else []
        print(opts, cid, B, "max err zdot %.2e ldot %.2e n %.2e" % (err[:, :D].max(), err[:, D].max(), err[:, D + 2].max()),
              "bad rows", len(rows), (rows[:6].tolist(), rows[-3:].tolist()) if len(rows) else "", flush=True)
