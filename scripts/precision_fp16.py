"""one augmented_f (forward chain + eps chain) with emulated operand splitting: noise of f vs float64.
fp16x2: x = f1 + f2 in IEEE half (22-bit significand for |x| <~ 1); bf16xN as in the tensor path."""
import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cnfb200_loader
wl = cnfb200_loader.load_package().workloads

def split_bf16(x, parts):
    out, r = [], x.astype(np.float32)
    for _ in range(parts):
        i = r.view(np.int32)
        p = (((i + 0x8000) >> 16) << 16).view(np.float32)
        out.append(p.astype(np.float64)); r = (r - p).astype(np.float32)
    return out
def split_fp16(x, parts):
    out, r = [], x.astype(np.float32)
    for _ in range(parts):
        p = r.astype(np.float16).astype(np.float32)
        out.append(p.astype(np.float64)); r = (r - p).astype(np.float32)
    return out
SCH = {"fp32": None,
       "bf16x3": (split_bf16, [(0,0),(1,0),(0,1)]),
       "bf16x6": (split_bf16, [(0,0),(1,0),(0,1),(1,1),(2,0),(0,2)]),
       "fp16x3": (split_fp16, [(0,0),(1,0),(0,1)]),
       "fp16x4": (split_fp16, [(0,0),(1,0),(0,1),(1,1)])}
def mm(a, b, sch):
    if sch == "f64": return a @ b
    if sch is None: return (a.astype(np.float32) @ b.astype(np.float32)).astype(np.float64)
    sp, terms = sch
    n = max(max(t) for t in terms) + 1
    A, B = sp(a, n), sp(b, n)
    acc = np.zeros((a.shape[0], b.shape[1]), np.float32)
    for i, j in terms: acc = (acc + (A[i] @ B[j]).astype(np.float32)).astype(np.float32)
    return acc.astype(np.float64)
def f(Ws, bs, z, eps, sch):
    hs, h = [], z
    for W, b in zip(Ws, bs):
        h = np.tanh(mm(h, W.T, sch) + b); hs.append(h)
    if sch not in ("f64",): hs = [x.astype(np.float32).astype(np.float64) for x in hs]
    g = eps * (1 - hs[-1] ** 2)
    for l in range(len(Ws) - 1, 0, -1):
        g = mm(g, Ws[l], sch) * (1 - hs[l - 1] ** 2)
    v = mm(g, Ws[0], sch)[:, :eps.shape[1]]
    return np.concatenate([hs[-1], -(v * eps).sum(1, keepdims=True), np.linalg.norm(hs[-1], axis=1, keepdims=True), np.linalg.norm(v, axis=1, keepdims=True)], 1)
for dims in ([32,256,256,32], [16,128,128,8], [128,1024,1024,128]):
    rng = np.random.default_rng(0); B = 256
    Ws = [rng.uniform(-1, 1, (dims[i+1], dims[i])) * np.sqrt(6/(dims[i]+dims[i+1])) for i in range(len(dims)-1)]
    Ws = [w.astype(np.float32).astype(np.float64) for w in Ws]
    bs = [np.zeros(d) for d in dims[1:]]
    D = dims[-1]
    z = rng.normal(0, 2, (B, dims[0])).astype(np.float32).astype(np.float64); eps = rng.normal(0, 1, (B, D)).astype(np.float32).astype(np.float64)
    Ws0 = [w.copy() for w in Ws]; Ws0[0] = Ws[0]
    ref = f(Ws, bs, z, eps, "f64")
    sc = np.abs(ref).max(0, keepdims=True)
    print(dims)
    for name, sch in SCH.items():
        r = f(Ws, bs, z, eps, sch)
        e = np.abs(r - ref) / sc
        print(f"  {name:8s} zdot {e[:, :D].max():.2e}  ldot {e[:, D].max():.2e}  E {e[:, D+1].max():.2e}  n {e[:, D+2].max():.2e}   rms {np.sqrt((e**2).mean()):.2e}")
