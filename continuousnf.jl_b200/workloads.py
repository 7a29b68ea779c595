"""Synthetic workloads C0..C5 of SURVEY.md 8.0 / 8.4 (BASELINE.json `configs`).

Host-generated with numpy `default_rng(seed)`, float32; the same bytes go to the GPU path and
to the CPU oracle.  numpy only: no dependency on the CUDA library nor on oracle/.
"""
import numpy as np

TRACE_VJP, TRACE_JVP = 0, 1

# id -> model / solver description.  `mode` is the headline mode of the config.
CONFIGS = {
    # reference's own micro-benchmark shape: benchmark/benchmarks.jl:24-46
    "C0": dict(tag="RNODE", nvars=8, naug=8, C=0, dims=[16, 16], B=64, tspan=(0.0, 13.0),
               trace_mode=TRACE_VJP, lambda1=1e-2, lambda2=1e-2, lambda3=1e-2, mode="train",
               reltol=0.0, abstol=0.0, data="uniform"),
    # README.md:39-68
    "C1": dict(tag="RNODE", nvars=1, naug=1, C=0, dims=[2, 6, 2], B=1024, tspan=(0.0, 13.0),
               trace_mode=TRACE_VJP, lambda1=1e-2, lambda2=1e-2, lambda3=1e-2, mode="train",
               reltol=float(np.sqrt(np.finfo(np.float32).eps)),
               abstol=float(np.finfo(np.float32).eps), data="beta"),
    # BASELINE.json configs[1]: the N=1 bench workload
    "C2": dict(tag="ICNF", nvars=2, naug=0, C=0, dims=[2, 64, 64, 2], B=1 << 20,
               tspan=(0.0, 1.0), trace_mode=TRACE_JVP, lambda1=0.0, lambda2=0.0, lambda3=0.0,
               mode="test", reltol=0.0, abstol=0.0, data="mixture"),
    "C3": dict(tag="RNODE", nvars=16, naug=16, C=0, dims=[32, 256, 256, 32], B=1 << 18,
               tspan=(0.0, 1.0), trace_mode=TRACE_VJP, lambda1=1e-2, lambda2=1e-2, lambda3=0.0,
               mode="train", reltol=0.0, abstol=0.0, data="mixture"),
    "C4": dict(tag="CondRNODE", nvars=8, naug=0, C=8, dims=[16, 128, 128, 8], B=1 << 20,
               tspan=(0.0, 1.0), trace_mode=TRACE_VJP, lambda1=1e-2, lambda2=1e-2, lambda3=0.0,
               mode="train", reltol=0.0, abstol=0.0, data="mixture"),
    "C5": dict(tag="RNODE", nvars=64, naug=64, C=0, dims=[128, 1024, 1024, 128], B=1 << 22,
               tspan=(0.0, 1.0), trace_mode=TRACE_VJP, lambda1=1e-2, lambda2=1e-2, lambda3=0.0,
               mode="train", reltol=0.0, abstol=0.0, data="mixture"),
}
_IDS = list(CONFIGS)


def seed_of(cid):
    return 20261104 + _IDS.index(cid)


def nparams(dims):
    return sum(dims[l] * dims[l + 1] + dims[l + 1] for l in range(len(dims) - 1))


def glorot_theta(dims, rng):
    """Glorot-uniform weights (column-major d_{l+1} x d_l), zero biases, Lux flat layout."""
    parts = []
    for l in range(len(dims) - 1):
        lim = np.sqrt(6.0 / (dims[l] + dims[l + 1]))
        parts.append(rng.uniform(-lim, lim, dims[l] * dims[l + 1]).astype(np.float32))
        parts.append(np.zeros(dims[l + 1], np.float32))
    return np.concatenate(parts)


def make_inputs(cid, B=None, seed=None, offset=0):
    """returns dict(xs (B,nvars), ys (B,C)|None, eps (B,D), theta).  Arrays are sample-major,
    i.e. the memory image of Julia's column-major `rows x B` matrices.  `offset` selects a
    different shard of the same synthetic stream (rank r uses offset=r)."""
    c = CONFIGS[cid]
    B = c["B"] if B is None else B
    seed = seed_of(cid) if seed is None else seed
    nv, D, Cc = c["nvars"], c["nvars"] + c["naug"], c["C"]
    theta = glorot_theta(c["dims"], np.random.default_rng(seed))
    rng = np.random.default_rng([seed, 7, offset])
    if c["data"] == "mixture":
        K = 8
        mu = np.random.default_rng([seed, 3]).normal(0.0, 2.0, (K, nv))
        comp = rng.integers(0, K, B)
        clean = mu[comp] + 0.5 * rng.normal(0.0, 1.0, (B, nv))
    elif c["data"] == "beta":
        clean = rng.beta(2.0, 4.0, (B, nv))
    else:
        clean = rng.uniform(0.0, 1.0, (B, nv))
    xs = clean.astype(np.float32)
    ys = None
    if Cc:
        ys = (clean[:, :Cc] + 0.1 * rng.normal(0.0, 1.0, (B, Cc))).astype(np.float32)
    eps = np.random.default_rng([seed + 1, offset]).normal(0.0, 1.0, (B, D)).astype(np.float32)
    return dict(xs=xs, ys=ys, eps=eps, theta=theta)


def flops_per_eval(cid, mode=None):
    """algorithmic flops per sample per augmented_f evaluation:
    Hutchinson F_H = 2P + 2(P - C d1); exact trace F_E = 2P(1+D); backward F_B = 8P."""
    c = CONFIGS[cid]
    dims = c["dims"]
    P = sum(dims[l] * dims[l + 1] for l in range(len(dims) - 1))
    D = c["nvars"] + c["naug"]
    mode = c["mode"] if mode is None else mode
    if mode == "test":
        return 2 * P * (1 + D)
    return 2 * P + 2 * (P - c["C"] * dims[1])


def tc_bytes_per_eval(cid):
    """algorithmic HBM bytes per sample of one TrainMode/VecJac augmented_f evaluation on the tensor
    path: every layer product is one GEMM launch whose operands and results are
    hi/lo 16-bit planes (4 bytes per element, widths padded to 64); the epsilon chain re-reads h_l."""
    c = CONFIGS[cid]
    dims = c["dims"]
    L = len(dims) - 1
    D = c["nvars"] + c["naug"]
    pad = lambda n: (n + 63) // 64 * 64
    e = D + pad(dims[0])                                              # prep: z rows -> planes of h_0
    for l in range(1, L):
        e += pad(dims[l - 1]) + pad(dims[l])                          # h_l = tanh(W h_{l-1} + b)
    e += pad(dims[L - 1]) + D + pad(D)                                # last layer: zdot rows + g_L planes
    for l in range(L - 1, 0, -1):
        e += pad(dims[l + 1] if l + 1 < L else D) + 2 * pad(dims[l])  # g_l: read g_{l+1}, h_l; write g_l
    e += pad(dims[1]) + 3 * D                                         # v = W_1^T g_1 against eps; 3 output rows
    return 4 * e


def tc_bytes_per_eval_exact(cid):
    """same for one TestMode (exact trace) evaluation on the tensor path: the forward chain, then D one-hot
    tangent chains (first layer elementwise from the h_1 planes, then one GEMM per layer)."""
    c = CONFIGS[cid]
    dims = c["dims"]
    L = len(dims) - 1
    D = c["nvars"] + c["naug"]
    pad = lambda n: (n + 63) // 64 * 64
    e = D + pad(dims[0])                                              # prep
    for l in range(1, L):
        e += pad(dims[l - 1]) + pad(dims[l])                          # forward hidden layers
    e += pad(dims[L - 1]) + D                                         # last layer: zdot rows
    chain = 2 * pad(dims[1])                                          # tau_1 = W_1[:, i] .* (1 - h_1^2)
    for l in range(2, L):
        chain += pad(dims[l - 1]) + 2 * pad(dims[l])                  # tau_l: read tau_{l-1}, h_l; write tau_l
    chain += pad(dims[L - 1]) + D + 2                                 # last layer: read tau_{L-1}, y rows; trace row r/w
    return 4 * (e + D * chain)


def flops_backward_per_eval(cid):
    dims = CONFIGS[cid]["dims"]
    return 8 * sum(dims[l] * dims[l + 1] for l in range(len(dims) - 1))
