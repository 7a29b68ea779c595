"""debug driver of the fused multi-layer kernel (cnf_stage_fused.cu): one augmented_f evaluation on C4-shaped inputs,
fused vs per-layer tensor path vs the CPU oracle, error per output row group"""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cnfb200_loader
cnf = cnfb200_loader.load_package()
wl = cnf.workloads
from oracle import oracle as O
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from helpers import make_icnf, oracle_cfg, args_of

cid = sys.argv[1] if len(sys.argv) > 1 else "C4"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
c = wl.CONFIGS[cid]
D = c["nvars"] + c["naug"]
R = D + 3
inp = wl.make_inputs(cid, B=B)
u = np.random.default_rng(5).normal(0, 1, (B, R)).astype(np.float32)
ref = O.augmented_f(oracle_cfg(cid), inp["theta"], 1, u, inp["ys"], inp["eps"])
scale = np.maximum(np.abs(ref).max(axis=0, keepdims=True), 1e-6)
for fused in (0, 1):
    icnf = make_icnf(cid)
    cnf.set_option(icnf, "fused", fused)
    du = np.asarray(cnf.augmented_f(icnf, cnf.TrainMode(), u.T, *args_of(inp), eps=inp["eps"].T)).T
    e = np.abs(du - ref) / scale
    print(f"fused={fused}: zdot {e[:, :D].max():.2e}  ldot {e[:, D].max():.2e}  Edot {e[:, D + 1].max():.2e}  ndot {e[:, D + 2].max():.2e}"
          f"  worst rows {np.argsort(-e.max(axis=1))[:4]}", flush=True)
    if fused and e.max() > 1e-4:
        b = int(np.argmax(e.max(axis=1)))
        print(" row", b, "got", du[b], "\n want", ref[b])
# whole solve + timing
for fused in (0, 1):
    icnf = make_icnf(cid)
    cnf.set_option(icnf, "fused", fused)
    cnf.set_option(icnf, "time_stages", 1)
    Bt = int(sys.argv[3]) if len(sys.argv) > 3 else (1 << 17)
    big = wl.make_inputs(cid, B=Bt)
    for _ in range(2):
        lp, regs = cnf.inference(icnf, cnf.TrainMode(), big["xs"].T, *args_of(big), eps=big["eps"].T)
    st = cnf.last_stats(icnf)
    print(f"fused={fused}: B={Bt} naccept={st['naccept']} nf={st['nf']} path={st['kernel_path']} ms_total={st['ms_total']:.2f} "
          f"stage_ms={st['ms_stage'] / max(st['n_stage'], 1):.4f} launches={st['gpu_launches']}", flush=True)
    if fused == 0:
        lp0 = np.asarray(lp).copy()
    else:
        print("  max rel logp diff fused vs per-layer:", float(np.max(np.abs(np.asarray(lp) - lp0) / np.abs(lp0))))

# timing experiments (wrong results by design for dbg != 0)
if len(sys.argv) > 4:
    Bt = 1 << 20
    big = wl.make_inputs(cid, B=Bt)
    for dbg in [int(x) for x in sys.argv[4].split(",")]:
        icnf = make_icnf(cid)
        cnf.set_option(icnf, "time_stages", 1)
        cnf.set_option(icnf, "tc_dbg", dbg)
        try:
            for _ in range(2):
                cnf.inference(icnf, cnf.TrainMode(), big["xs"].T, *args_of(big), eps=big["eps"].T)
            st = cnf.last_stats(icnf)
            print(f"dbg={dbg}: B={Bt} nf={st['nf']} retcode={st['retcode']} stage_ms={st['ms_stage'] / max(st['n_stage'], 1):.4f}", flush=True)
        except Exception as e:
            print(f"dbg={dbg}: {e}")

if len(sys.argv) > 5:   # phase trace of CTA 0 (tc_dbg & 16): one evaluation
    icnf = make_icnf(cid)
    cnf.set_option(icnf, "tc_dbg", 16)
    Bt = 1 << 20
    big = wl.make_inputs(cid, B=Bt)
    uu = np.zeros((Bt, R), np.float32); uu[:, :big["xs"].shape[1]] = big["xs"]
    cnf.augmented_f(icnf, cnf.TrainMode(), uu.T, *args_of(big), eps=big["eps"].T)
