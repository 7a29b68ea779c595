import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import args_of, cnf, make_icnf, wl
cid, B = sys.argv[1], int(sys.argv[2])
inp = wl.make_inputs(cid, B=B)
icnf = make_icnf(cid)
val, g = cnf.loss_and_grad(icnf, cnf.TrainMode(), inp["xs"].T, *args_of(inp), eps=inp["eps"].T)
print(val, cnf.last_stats(icnf))
