"""debug helper: the tensor-path reverse sweep (k_tc_gemm<2,true> + k_tc_wgrad) against the generic
SIMT reverse sweep on the same forward solve; per-layer gradient errors; if the weight gradient is
off, tries the other plausible MN-major descriptor strides."""
import sys, os, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import O, args_of, cnf, make_icnf, oracle_cfg, wl, relnorm


def layer_slices(dims):
    o, out = 0, []
    for l in range(len(dims) - 1):
        nw = dims[l] * dims[l + 1]
        out.append(("W%d" % (l + 1), slice(o, o + nw))); o += nw
        out.append(("b%d" % (l + 1), slice(o, o + dims[l + 1]))); o += dims[l + 1]
    return out


def run(cid, B, opts):
    inp = wl.make_inputs(cid, B=B)
    icnf = make_icnf(cid)
    for k, v in opts.items():
        cnf.set_option(icnf, k, v)
    t = time.time()
    val, g, gx = cnf.loss_and_grad(icnf, cnf.TrainMode(), inp["xs"].T, *args_of(inp), eps=inp["eps"].T, want_grad_x=True)
    return val, np.asarray(g), np.asarray(gx), cnf.last_stats(icnf), time.time() - t


cases = [("C4", 1024), ("C3", 256), ("C3", 20000)] if len(sys.argv) < 2 else [(sys.argv[1], int(sys.argv[2]))]
for cid, B in cases:
    dims = wl.CONFIGS[cid]["dims"]
    v0, g0, x0, st0, t0 = run(cid, B, {"tc_backward": 0})
    v1, g1, x1, st1, t1 = run(cid, B, {})
    print(f"{cid} B={B}: loss {v0:.6f}/{v1:.6f} naccept {st0['naccept']}/{st1['naccept']} "
          f"grad rel {relnorm(g1, g0):.2e} gradx rel {relnorm(x1, x0):.2e}  "
          f"launches {st0['gpu_launches']}/{st1['gpu_launches']} wall {t0:.2f}s/{t1:.2f}s nan={np.isnan(g1).any()}", flush=True)
    for name, sl in layer_slices(dims):
        print(f"   {name}: rel {relnorm(g1[sl], g0[sl]):.2e}  |ref| {np.linalg.norm(g0[sl]):.3e} |tc| {np.linalg.norm(g1[sl]):.3e}")
    if relnorm(g1, g0) > 1e-3 and relnorm(x1, x0) < 1e-3:
        for lbo, sbo in ((1024, 8192), (8192, 2048), (128, 1024), (1024, 128), (16, 1024), (4096, 1024)):
            try:
                v2, g2, x2, st2, _ = run(cid, B, {"wg_lbo": lbo, "wg_sbo": sbo})
                print(f"   variant lbo={lbo} sbo={sbo}: grad rel {relnorm(g2, g0):.2e}", flush=True)
            except Exception as e:  # noqa
                print("   variant", lbo, sbo, "failed:", e)
                break
