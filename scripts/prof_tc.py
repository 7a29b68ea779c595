"""profiling driver for the tensor path: `eval` = a few augmented_f evaluations (TrainMode),
`grad` = one loss+gradient call.  usage: prof_tc.py eval|grad CID B"""
import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import args_of, cnf, make_icnf, wl

what, cid, B = sys.argv[1], sys.argv[2], int(sys.argv[3])
inp = wl.make_inputs(cid, B=B)
c = wl.CONFIGS[cid]
D = c["nvars"] + c["naug"]
icnf = make_icnf(cid)
if what == "eval":
    u = np.random.default_rng(5).normal(0, 1, (B, D + 3)).astype(np.float32)
    for _ in range(3):
        cnf.augmented_f(icnf, cnf.TrainMode(), u.T, *args_of(inp), eps=inp["eps"].T)
else:
    val, g = cnf.loss_and_grad(icnf, cnf.TrainMode(), inp["xs"].T, *args_of(inp), eps=inp["eps"].T)
    print(val, cnf.last_stats(icnf))
