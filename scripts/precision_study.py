"""Precision study for the tensor-core path: emulate operand rounding of every
layer GEMM (forward and VJP chain) and measure the effect on per-sample log-density / loss against
the FP64 restatement on the same frozen step sequence.  CPU only (torch float64 carrier)."""
import sys, os
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import O, oracle_cfg, wl
from oracle import ref_f64 as F

def rnd(x, mant):
    """round-to-nearest to `mant` explicit mantissa bits (tf32: 10, bf16: 7), float32 exponent range"""
    x32 = x.to(torch.float32)
    i = x32.view(torch.int32)
    drop = 23 - mant
    half = 1 << (drop - 1)
    i = ((i + half) >> drop) << drop
    return i.view(torch.float32).to(torch.float64)

def split(x, mant, parts):
    out, r = [], x.to(torch.float32).to(torch.float64)
    for _ in range(parts):
        p = rnd(r, mant); out.append(p); r = r - p
    return out

SCHEMES = {
    "fp32": None,
    "tf32x1": (10, [(0, 0)]),
    "tf32x2_A": (10, [(0, 0), (1, 0)]),
    "tf32x3": (10, [(0, 0), (1, 0), (0, 1)]),
    "bf16x1": (7, [(0, 0)]),
    "bf16x3": (7, [(0, 0), (1, 0), (0, 1)]),
    "bf16x6": (7, [(0, 0), (1, 0), (0, 1), (1, 1), (2, 0), (0, 2)]),
}

class RMM(torch.autograd.Function):
    scheme = None
    @staticmethod
    def mm(a, b):
        s = RMM.scheme
        if s is None:
            return (a.to(torch.float32) @ b.to(torch.float32)).to(torch.float64)
        mant, terms = s
        n = max(max(t) for t in terms) + 1
        A, B = split(a, mant, n), split(b, mant, n)
        return sum(A[i] @ B[j] for i, j in terms)
    @staticmethod
    def forward(ctx, a, b):
        ctx.save_for_backward(a, b)
        return RMM.mm(a, b)
    @staticmethod
    def backward(ctx, g):
        a, b = ctx.saved_tensors
        return RMM.mm(g, b.T), RMM.mm(a.T, g)

def mlp_r(layers, h):
    for W, b in layers:
        h = torch.tanh(RMM.apply(h, W.T.contiguous()) + b)
    return h

def main():
    F.mlp = mlp_r
    for cid, B in (("C3", 48), ("C4", 64)):
        c = wl.CONFIGS[cid]; inp = wl.make_inputs(cid, B=B); cfg = oracle_cfg(cid)
        O.logdensity(cfg, inp["theta"], 1, inp["xs"], inp["ys"], inp["eps"], *c["tspan"])
        dts = [float(d) for d in O.last_dts()]
        F.mlp = F.__dict__.get("_orig_mlp", None) or mlp_r
        RMM.scheme = "exact"
        res = {}
        for name, sch in SCHEMES.items():
            RMM.scheme = sch
            r = F.run(cfg, inp["theta"], 1, inp["xs"], inp["ys"], inp["eps"], *c["tspan"], frozen_dts=dts)
            res[name] = r
        # float64 exact reference
        class Exact:  # exact matmul
            pass
        RMM.mm_backup = RMM.mm
        RMM.mm = staticmethod(lambda a, b: a @ b)
        ref = F.run(cfg, inp["theta"], 1, inp["xs"], inp["ys"], inp["eps"], *c["tspan"], frozen_dts=dts)
        RMM.mm = RMM.mm_backup
        print(f"== {cid} B={B} steps={len(dts)}  |logp| ~ {np.abs(ref['logp']).mean():.1f}")
        for name, r in res.items():
            e = np.max(np.abs(r["logp"] - ref["logp"]) / np.abs(ref["logp"]))
            el = abs(r["loss"] - ref["loss"]) / abs(ref["loss"])
            ez = np.max(np.abs(r["z"] - ref["z"]))
            print(f"  {name:10s} max rel logp err {e:.2e}   loss rel {el:.2e}   max |dz| {ez:.2e}")

main()
