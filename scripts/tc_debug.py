"""debug helper: stage-level parity of the tensor path vs the oracle"""
import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import O, args_of, cnf, make_icnf, oracle_cfg, wl

for cid, B in (("C4", 512), ("C3", 300), ("C4", 40000)):
    for tm in (0, 1):
        inp = wl.make_inputs(cid, B=B)
        c = wl.CONFIGS[cid]
        D = c["nvars"] + c["naug"]
        u = np.random.default_rng(5).normal(0, 1, (B, D + 3)).astype(np.float32)
        icnf = make_icnf(cid, trace_mode=tm)
        du = np.asarray(cnf.augmented_f(icnf, cnf.TrainMode(), u.T, *args_of(inp), eps=inp["eps"].T)).T
        ref = O.augmented_f(oracle_cfg(cid, trace_mode=tm), inp["theta"], 1, u, inp["ys"], inp["eps"])
        scale = np.maximum(np.abs(ref).max(axis=0, keepdims=True), 1e-6)
        err = np.abs(du - ref) / scale
        print(cid, B, "tm", tm, "max err zdot %.2e ldot %.2e E %.2e n %.2e" % (err[:, :D].max(), err[:, D].max(), err[:, D + 1].max(), err[:, D + 2].max()),
              "nan" if np.isnan(du).any() else "", flush=True)
        if err.max() > 1e-3:
            bad = np.argwhere(err > 1e-3)
            print("  first bad (row, col):", bad[:8].tolist(), "du", du[bad[0][0], :4], "ref", ref[bad[0][0], :4])
