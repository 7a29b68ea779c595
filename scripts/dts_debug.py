"""accepted step sizes of the library vs the oracle on one case (CNFB200_DEBUG_DTS=1 prints the library's)"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
os.environ["CNFB200_DEBUG_DTS"] = "1"
from helpers import O, args_of, cnf, make_icnf, oracle_cfg, wl
cid, B, tm, mode = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
kp = int(sys.argv[5]) if len(sys.argv) > 5 else 0
inp = wl.make_inputs(cid, B=B)
c = wl.CONFIGS[cid]
icnf = make_icnf(cid, trace_mode=tm, kernel_path=kp)
m = cnf.TrainMode() if mode else cnf.TestMode()
lp, regs = cnf.inference(icnf, m, inp["xs"].T, *args_of(inp), eps=inp["eps"].T)
st = cnf.last_stats(icnf)
rlp, _, _, rst = O.logdensity(oracle_cfg(cid, trace_mode=tm), inp["theta"], mode, inp["xs"], inp["ys"], inp["eps"], *c["tspan"])
print("path", st["kernel_path"], "naccept", st["naccept"], "oracle", rst["naccept"], "nreject", st["nreject"], rst["nreject"])
print("oracle dts:", " ".join(f"{d:.7g}" for d in O.last_dts()))
print("max rel logp err", float(np.max(np.abs(np.asarray(lp) - rlp) / np.abs(rlp))))
