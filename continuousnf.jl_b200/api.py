"""Host-side mirror of the reference's operator interface for the hot path, over the C ABI.

Same names and argument meaning as ContinuousNormalizingFlows.jl (Julia is not installed here, so
this Python layer plays the role of the `ext/` Julia shim; INTEGRATION.md shows the `ccall` form):
  construct            src/base_icnf.jl:1-77
  inference / generate / loss   src/base_icnf.jl:407-508, src/icnf.jl:481-502
  TestMode / TrainMode, DIVecJacMatrixMode / DIJacVecMatrixMode   src/types.jl:1-35
  RNODE / FFJORD / CondRNODE / CondFFJORD tags                    src/icnf.jl:1-56
Matrices follow the Julia convention `rows x B`; they are handed to the library as column-major
memory (sample-major), which is what Julia would pass.
"""
import ctypes as C
import math

import weakref

import numpy as np

from . import _cabi


# ---------------------------------------------------------------- tags, modes, resources
class Mode:
    pass


class TestMode(Mode):
    __test__ = False
    code = _cabi.MODE_TEST


class TrainMode(Mode):
    code = _cabi.MODE_TRAIN


class ComputeMode:
    def __init__(self, adback=None):
        self.adback = adback  # accepted for signature parity; the kernels are hand-written


class DIVecJacMatrixMode(ComputeMode):
    trace_mode = _cabi.TRACE_VJP


class DIJacVecMatrixMode(ComputeMode):
    trace_mode = _cabi.TRACE_JVP


class DIVecJacVectorMode(ComputeMode):
    """VectorMode: every sample its own adaptive solve (own dt, own error norm).  TestMode on
    the narrow 3-layer shapes runs as one register-resident launch (kernel_path 4); every other net, width and mode
    (TrainMode Hutchinson, conditional flows, the tensor-core paths) runs the MatrixMode stage kernels under a per-sample
    controller (kernel_path 8).  The reverse sweep is MatrixMode-only."""
    trace_mode = _cabi.TRACE_VJP
    vector = True


class DIJacVecVectorMode(ComputeMode):
    trace_mode = _cabi.TRACE_JVP
    vector = True


class AbstractICNF:
    cond, regularised = False, False


class RNODE(AbstractICNF):
    regularised = True


class FFJORD(AbstractICNF):
    pass


class CondRNODE(AbstractICNF):
    cond, regularised = True, True


class CondFFJORD(AbstractICNF):
    cond = True


class Planar(AbstractICNF):
    """the tag whose nn is a PlanarLayer (src/icnf.jl:1-18); same augmented_f as FFJORD"""


class CondPlanar(AbstractICNF):
    cond = True


class B200Libs:
    """The new resource type (the `resource = ...` keyword of construct selects the path)."""

    def __init__(self, device=0, kernel_path=0):
        self.device, self.kernel_path = device, kernel_path


# ---------------------------------------------------------------- nn description (Lux mirror)
class Dense:
    def __init__(self, in_dims, out_dims, activation="tanh"):
        name = getattr(activation, "__name__", activation)
        if name != "tanh":
            raise ValueError("B200Libs path supports Dense(in => out, tanh) layers only")
        self.in_dims, self.out_dims = int(in_dims), int(out_dims)


class PlanarLayer:
    """PlanarLayer(nvars, tanh; n_cond) — src/layers/planar_layer.jl:1-88: z -> u * tanh(w'[z; ys] + b),
    parameters (u[nvars], w[nvars + n_cond], b[1]).  Row N3: runs as Dense(d0 => 1, tanh) followed
    by a bias-free linear product (CNF_FLAG_LAST_LINEAR) on the generic kernels."""

    def __init__(self, nvars, activation="tanh", *, n_cond=0, use_bias=True):
        name = getattr(activation, "__name__", activation)
        if name != "tanh" or not use_bias:
            raise ValueError("B200Libs path supports PlanarLayer(nvars, tanh; use_bias = true) only")
        self.nvars, self.n_cond = int(nvars), int(n_cond)
        self.in_dims, self.out_dims = self.nvars + self.n_cond, self.nvars


class Chain:
    planar = False

    def __init__(self, *layers):
        if len(layers) == 1 and isinstance(layers[0], (list, tuple)):
            layers = tuple(layers[0])
        if any(isinstance(l, PlanarLayer) for l in layers):
            if len(layers) != 1:
                raise ValueError("B200Libs path: a PlanarLayer must be the whole network")
            self.planar = True
        for a, b in zip(layers[:-1], layers[1:]):
            if a.out_dims != b.in_dims:
                raise ValueError("Chain: layer widths do not match")
        self.layers = layers

    @property
    def dims(self):
        if self.planar:
            return [self.layers[0].in_dims, 1, self.layers[0].out_dims]
        return [self.layers[0].in_dims] + [l.out_dims for l in self.layers]


def param_offsets(dims, planar=False):
    """Lux/ComponentArray flat layout: layer_k = (weight[d_{k+1} x d_k col-major], bias);
    PlanarLayer: (u, w, b) (src/layers/planar_layer.jl:38-60)."""
    if planar:
        d0, D = dims[0], dims[2]
        return np.array([D, 0], np.int64), np.array([D + d0, 0], np.int64), D + d0 + 1
    w, b, o = [], [], 0
    for l in range(len(dims) - 1):
        w.append(o)
        o += dims[l] * dims[l + 1]
        b.append(o)
        o += dims[l + 1]
    return np.array(w, np.int64), np.array(b, np.int64), o


class ICNF:
    """mirror of `struct ICNF` (src/icnf.jl:69-104); owns the device handle."""

    def __init__(self, **kw):
        self.__dict__.update(kw)
        self._handle = C.c_void_p()
        self._cfg = None
        self._last_stats = None

    def __del__(self):
        try:
            if self._handle:
                _cabi.load().cnfb200_destroy(self._handle)
        except Exception:
            pass

    # -- handle management
    def _config(self):
        c = _cabi.CnfConfig()
        dims = self.nn.dims if isinstance(self.nn, Chain) else Chain(self.nn).dims
        c.D, c.nvars = self.nvars + self.naugmented, self.nvars
        c.C = dims[0] - c.D
        c.n_layers = len(dims) - 1
        for i, d in enumerate(dims):
            c.dims[i] = d
        c.trace_mode = self.compute_mode.trace_mode
        fl = 0
        if self.cond:
            fl |= _cabi.FLAG_COND
        if self.naugmented != 0:
            fl |= _cabi.FLAG_AUGMENTED
        if self.lambda1 != 0:
            fl |= _cabi.FLAG_NORM_Z
        if self.lambda2 != 0:
            fl |= _cabi.FLAG_NORM_J
        if self.lambda3 != 0:
            fl |= _cabi.FLAG_NORM_Z_AUG
        if self.nn.planar:
            fl |= _cabi.FLAG_LAST_LINEAR
        c.flags = fl
        c.lambda1, c.lambda2, c.lambda3 = self.lambda1, self.lambda2, self.lambda3
        sk = dict(self.sol_kwargs)
        alg = sk.pop("alg", None)
        if alg is not None and str(alg).lower() not in ("tsit5", "tsit5()"):
            raise ValueError("B200Libs path implements the Tsit5 branch only")
        for ignored in ("progress", "save_everystep", "save_start", "save_end", "alg_hints"):
            sk.pop(ignored, None)
        c.reltol = sk.pop("reltol", 0.0)
        c.abstol = sk.pop("abstol", 0.0)
        c.maxiters = int(min(sk.pop("maxiters", 0), 2 ** 62))
        c.dt0 = sk.pop("dt", 0.0)
        c.dtmax = sk.pop("dtmax", 0.0)
        c.dtmin = sk.pop("dtmin", 0.0)
        if sk:
            raise ValueError(f"unsupported sol_kwargs: {sorted(sk)}")
        c.device = self.resource.device
        c.kernel_path = self.resource.kernel_path
        return c

    def handle(self):
        if not self._handle:
            lib = _cabi.load()
            self._cfg = self._config()
            rc = lib.cnfb200_create(C.byref(self._cfg), C.byref(self._handle))
            if rc != 0:
                msg = lib.cnfb200_last_error(self._handle).decode() if self._handle else "create failed"
                if self._handle:
                    lib.cnfb200_destroy(self._handle)
                    self._handle = C.c_void_p()
                raise _cabi.CnfError(f"cnfb200_create: {msg} (rc={rc})")
            if getattr(self.compute_mode, "vector", False):
                lib.cnfb200_set_option(self._handle, b"vector_mode", 1)
        return self._handle

    def _check(self, rc, what):
        if rc < 0:
            msg = _cabi.load().cnfb200_last_error(self._handle).decode()
            raise _cabi.CnfError(f"{what}: {msg} (rc={rc})")
        return rc

    @property
    def D(self):
        return self.nvars + self.naugmented

    @property
    def dims(self):
        return self.nn.dims


def construct(aicnf, nn, nvars, naugmented=0, *, data_type=np.float32, compute_mode=None,
              inplace=False, cond=None, resource=None, basedist=None, tspan=(0.0, 1.0),
              steer_rate=0.0, epsdist=None, sol_kwargs=None, rng=None, lambda1=None, lambda2=None,
              lambda3=0.0):
    """construct(RNODE, nn, nvars, naugmented; ...) — src/base_icnf.jl:1-77.  Unsupported
    combinations raise (there is no CPU fallback in this path)."""
    if np.dtype(data_type) != np.float32:
        raise ValueError("B200Libs path is Float32 only")
    compute_mode = DIVecJacMatrixMode() if compute_mode is None else compute_mode
    if getattr(compute_mode, "trace_mode", None) is None:
        raise ValueError("B200Libs path: compute_mode must be one of the DI*MatrixMode / DI*VectorMode classes")
    if basedist is not None or epsdist is not None:
        raise ValueError("B200Libs path: basedist/epsdist must be the default MvNormal(0, I)")
    resource = B200Libs() if resource is None else resource
    if not isinstance(resource, B200Libs):
        raise ValueError("this package implements resource = B200Libs() only")
    if not isinstance(nn, Chain):
        nn = Chain(nn)
    cond = aicnf.cond if cond is None else cond
    reg = 1e-2 if aicnf.regularised else 0.0
    lambda1 = reg if lambda1 is None else lambda1
    lambda2 = reg if lambda2 is None else lambda2
    D = nvars + naugmented
    dims = nn.dims
    if dims[-1] != D or dims[0] < D or (not cond and dims[0] != D) or (cond and dims[0] == D):
        raise ValueError(f"nn dims {dims} do not fit nvars+naugmented={D}, cond={cond}")
    return ICNF(nn=nn, nvars=int(nvars), naugmented=int(naugmented), compute_mode=compute_mode,
                inplace=inplace, cond=bool(cond), resource=resource, tspan=tuple(map(float, tspan)),
                steer_rate=float(steer_rate), sol_kwargs=dict(sol_kwargs or {}),
                rng=np.random.default_rng() if rng is None else rng,
                lambda1=float(lambda1), lambda2=float(lambda2), lambda3=float(lambda3))


def setup(rng, icnf):
    """LuxCore.setup(rng, icnf) -> (ps, st): Glorot-uniform weights, zero biases, flat Float32
    vector in ComponentArray order (src/exts/mlj_ext/core_icnf.jl:37-38)."""
    parts = []
    dims = icnf.dims
    if icnf.nn.planar:   # glorot_uniform on vectors (nfan = (1, n)), zero bias
        for n in (dims[2], dims[0]):
            lim = math.sqrt(6.0 / (1 + n))
            parts.append(rng.uniform(-lim, lim, n).astype(np.float32))
        parts.append(np.zeros(1, np.float32))
        return np.concatenate(parts), {}
    for l in range(len(dims) - 1):
        lim = math.sqrt(6.0 / (dims[l] + dims[l + 1]))
        parts.append(rng.uniform(-lim, lim, dims[l] * dims[l + 1]).astype(np.float32))
        parts.append(np.zeros(dims[l + 1], np.float32))
    return np.concatenate(parts), {}


# ---------------------------------------------------------------- array marshalling
def _is_torch(a):
    return type(a).__module__.startswith("torch")


class _Mat:
    """a `rows x B` matrix as column-major (sample-major) float32 memory, host or device"""

    def __init__(self, a, rows, name):
        self.torch = _is_torch(a)
        if a.ndim == 1:
            a = a.reshape(rows, -1) if not self.torch else a.reshape(rows, -1)
        if a.shape[0] != rows:
            raise ValueError(f"{name}: expected {rows} rows, got shape {tuple(a.shape)}")
        self.B = int(a.shape[1])
        if self.torch:
            import torch
            t = a.t().contiguous().to(torch.float32)  # (B, rows) row-major == rows x B col-major
            self.keep, self.ptr, self.device = t, t.data_ptr(), t.device
        else:
            t = np.ascontiguousarray(np.asarray(a).T, dtype=np.float32)
            self.keep, self.ptr, self.device = t, t.ctypes.data, None


class _PinnedPool:
    """page-locked host buffers for output matrices (cnfb200_host_alloc): the device->host copy of a result
    runs at PCIe rate into them, ~6x faster than into pageable numpy memory.  A buffer goes back to the pool
    when the array that wraps it is garbage-collected; small outputs stay ordinary numpy arrays.  Sizes are
    rounded up to powers of two (so varying batch sizes share buffers) and the pool never keeps more than
    MAX_FREE_BYTES of idle page-locked memory: beyond that a returned buffer is freed."""
    MIN_BYTES = 1 << 16
    MAX_KEEP = 8                 # free buffers kept per size class
    MAX_FREE_BYTES = 1 << 30     # idle page-locked bytes kept in total

    def __init__(self):
        self.free = {}
        self.free_bytes = 0

    @staticmethod
    def _cls(nbytes):
        return 1 << max(int(nbytes) - 1, 0).bit_length()

    def take(self, size):
        lst = self.free.get(size)
        if lst:
            self.free_bytes -= size
            return lst.pop()
        p = C.c_void_p()
        if _cabi.load().cnfb200_host_alloc(size, C.byref(p)) != 0 or not p.value:
            return None
        return p.value

    def give(self, size, addr):
        lst = self.free.setdefault(size, [])
        if len(lst) < self.MAX_KEEP and self.free_bytes + size <= self.MAX_FREE_BYTES:
            lst.append(addr)
            self.free_bytes += size
        else:
            _cabi.load().cnfb200_host_free(addr)

    def empty(self, shape):
        nbytes = int(np.prod(shape)) * 4
        size = self._cls(nbytes)
        addr = self.take(size) if nbytes >= self.MIN_BYTES else None
        if addr is None:
            return np.empty(shape, np.float32)
        buf = (C.c_float * (nbytes // 4)).from_address(addr)
        weakref.finalize(buf, self.give, size, addr)     # the numpy array keeps `buf` alive through .base
        return np.frombuffer(buf, dtype=np.float32).reshape(shape)


_pinned = _PinnedPool()


def _out(rows, B, like):
    """allocate a `rows x B` output in sample-major memory; returns (holder, ptr)"""
    if like is not None and like.torch and like.device.type == "cuda":
        import torch
        t = torch.empty((B, rows), dtype=torch.float32, device=like.device)
        return t, t.data_ptr()
    t = _pinned.empty((B, rows))
    return t, t.ctypes.data


def _rows_view(holder):
    """(B, rows) holder -> `rows x B` view (Julia orientation)"""
    return holder.t() if _is_torch(holder) else holder.T


def _upload_params(icnf, ps):
    lib = _cabi.load()
    h = icnf.handle()
    w, b, n = param_offsets(icnf.dims, icnf.nn.planar)
    if _is_torch(ps):
        import torch
        p = ps.detach().to(torch.float32).contiguous()
        ptr, cnt = p.data_ptr(), p.numel()
    else:
        p = np.ascontiguousarray(ps, dtype=np.float32)
        ptr, cnt = p.ctypes.data, p.size
        # the same host parameter vector as in the previous call (pdf / logpdf / rand of a fitted model):
        # every derived form is already on the device, skip the upload and its synchronisations
        last = getattr(icnf, "_theta_uploaded", None)
        if last is not None and last[0] == h and last[1].shape == p.reshape(-1).shape and np.array_equal(last[1], p.reshape(-1)):
            return h
    if cnt != n:
        raise ValueError(f"ps has {cnt} parameters, the network needs {n}")
    icnf._check(lib.cnfb200_set_params(h, ptr, n, w.ctypes.data_as(_cabi._I64P),
                                       b.ctypes.data_as(_cabi._I64P)), "cnfb200_set_params")
    icnf._theta_uploaded = None if _is_torch(ps) else (h, p.reshape(-1).copy())
    return h


def steer_tspan(icnf, mode):
    """src/base_icnf.jl:108-121: TrainMode with steer_rate != 0 re-randomises t1 per call"""
    t0, t1 = icnf.tspan
    if isinstance(mode, TrainMode) and icnf.steer_rate != 0:
        r = np.float32(icnf.rng.uniform(-icnf.steer_rate, icnf.steer_rate))
        t1 = float(np.float32(abs(t1 - t0)) * r + np.float32(t1))
    return t0, t1


def _draw_eps(icnf, mode, B, like, eps):
    """eps ~ MvNormal(0, I), D x B, drawn once per call on the host (src/base_icnf.jl:277-278).
    TestMode never reads eps, so no draw is made there."""
    if eps is not None:
        return _Mat(eps, icnf.D, "eps")
    if not isinstance(mode, TrainMode):
        return None
    if getattr(icnf, "_device_rng", False):
        return None   # drawn inside the library (seed_device_rng)
    e = icnf.rng.standard_normal((B, icnf.D), dtype=np.float32)
    m = _Mat.__new__(_Mat)
    m.torch, m.B, m.keep, m.ptr, m.device = False, B, e, e.ctypes.data, None
    if like is not None and like.torch and like.device.type == "cuda":
        import torch
        t = torch.from_numpy(e).to(like.device)
        m.torch, m.keep, m.ptr, m.device = True, t, t.data_ptr(), t.device
    return m


def seed_device_rng(icnf, seed, sample_offset=0):
    """from now on eps (TrainMode) and, in generate, z0 are drawn on the device
    by the library's Philox stream instead of on the host from icnf.rng (src/base_icnf.jl:277-278,
    365-372); `sample_offset` = first global column of this rank's shard."""
    lib = _cabi.load()
    rc = lib.cnfb200_seed_rng(icnf.handle(), int(seed), int(sample_offset))
    icnf._check(rc, "cnfb200_seed_rng")
    icnf._device_rng = True


def last_eps(icnf, B):
    """the D x B eps of the last TrainMode call (drawn on the device or passed in)"""
    lib = _cabi.load()
    out = np.empty((int(B), icnf.D), np.float32)
    rc = lib.cnfb200_last_eps(icnf.handle(), out.ctypes.data, int(B))
    icnf._check(rc, "cnfb200_last_eps")
    return out.T


def fit_epochs(icnf, X, Y, ps, *, batch_size, n_epochs, perm=None, eta=1e-3, beta=(0.9, 0.999)):
    """`n_epochs` of the reference's fit loop (src/exts/mlj_ext/core_icnf.jl:30-90:
    shuffled DataLoader, loss + gradient, Optimisers.Lion step per mini-batch) in one library call with
    data, parameters and optimiser state resident on the device.  X: nvars x N (Y: C x N); perm:
    (n_epochs, N) column indices or None (shuffled inside).  Needs seed_device_rng (eps is drawn on
    the device).  Returns (ps, loss_history)."""
    lib = _cabi.load()
    h = _upload_params(icnf, ps)
    x = _Mat(X, icnf.nvars, "X")
    y = _Mat(Y, icnf._cfg.C, "Y") if icnf.cond else None
    p = None
    if perm is not None:
        p = np.ascontiguousarray(perm, dtype=np.int64)
        if p.shape != (n_epochs, x.B):
            raise ValueError(f"perm must have shape ({n_epochs}, {x.B})")
    # the reference re-draws the steered t1 in every loss call of the fit (src/base_icnf.jl:108-121, :283): the
    # library draws it per mini-batch from its own documented stream (cnfb200_set_steer, `steer_sequence` below)
    t0, t1 = icnf.tspan
    icnf._check(lib.cnfb200_set_steer(h, float(icnf.steer_rate)), "cnfb200_set_steer")
    out = np.empty(param_offsets(icnf.dims, icnf.nn.planar)[2], np.float32)
    hist = np.empty(n_epochs, np.float32)
    st_ = _cabi.CnfStats()
    rc = lib.cnfb200_fit_epochs(h, x.ptr, y.ptr if y else None, x.B, int(batch_size), int(n_epochs),
                                p.ctypes.data if p is not None else None, t0, t1, float(eta), float(beta[0]),
                                float(beta[1]), out.ctypes.data, hist.ctypes.data, C.byref(st_))
    icnf._theta_uploaded = None      # the device copy of theta has been trained: never skip the next upload
    icnf._check(rc, "cnfb200_fit_epochs")
    icnf._last_stats = st_.as_dict()
    return out, hist


def steer_sequence(seed, n, steer_rate, tspan):
    """the t1 values cnfb200_fit_epochs uses for its first `n` mini-batches after seed_device_rng(icnf, seed)
    (include/cnfb200.h, cnfb200_set_steer): host-side restatement for callers that want to replay a fit"""
    m64 = (1 << 64) - 1
    s = (int(seed) ^ 0xA5A5A5A55A5A5A5A) & m64
    t0, t1 = np.float32(tspan[0]), np.float32(tspan[1])
    out = []
    for _ in range(n):
        s = (s + 0x9E3779B97F4A7C15) & m64
        z = s
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & m64
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & m64
        z ^= z >> 31
        u = (z >> 40) / 16777216.0
        r = np.float32((2.0 * u - 1.0) * float(np.float32(steer_rate)))
        # fmaf(|t1 - t0|, r, t1): one rounding
        out.append(float(np.float32(float(np.float32(abs(t1 - t0))) * float(r) + float(t1))))
    return out


def _split(args, cond, what):
    """(xs?, [ys,] ps, st) argument tails, like the Julia method pairs"""
    if cond:
        if len(args) != 3:
            raise TypeError(f"{what}: conditional flow expects (..., ys, ps, st)")
        return args[0], args[1], args[2]
    if len(args) != 2:
        raise TypeError(f"{what}: expects (..., ps, st)")
    return None, args[0], args[1]


# ---------------------------------------------------------------- the operator interface
def inference(icnf, mode, xs, *args, eps=None, return_z=False):
    """inference(icnf, mode, xs[, ys], ps, st) -> (logp̂x, (Ė, ṅ, Ȧ)) in TrainMode,
    (logp̂x, (Ȧ,)) in TestMode — src/base_icnf.jl:407-426, :167-189."""
    ys, ps, st = _split(args, icnf.cond, "inference")
    lib = _cabi.load()
    h = _upload_params(icnf, ps)
    x = _Mat(xs, icnf.nvars, "xs")
    y = _Mat(ys, icnf._cfg.C, "ys") if icnf.cond else None
    e = _draw_eps(icnf, mode, x.B, x, eps)
    t0, t1 = steer_tspan(icnf, mode)
    nreg = 3 if isinstance(mode, TrainMode) else 1
    lp, lp_ptr = _out(1, x.B, x)
    rg, rg_ptr = _out(nreg, x.B, x)
    z, z_ptr = _out(icnf.D, x.B, x) if return_z else (None, None)
    st_ = _cabi.CnfStats()
    rc = lib.cnfb200_logdensity(h, mode.code, x.ptr, y.ptr if y else None, e.ptr if e else None,
                                x.B, t0, t1, lp_ptr, rg_ptr, z_ptr, C.byref(st_))
    icnf._check(rc, "cnfb200_logdensity")
    icnf._last_stats = st_.as_dict()
    regs = tuple(_rows_view(rg)[i] for i in range(nreg))
    out = (lp.reshape(-1), regs)
    return out + (_rows_view(z),) if return_z else out


def generate(icnf, mode, *args, eps=None, z0=None):
    """generate(icnf, mode, [ys,] ps, st, n) -> nvars x n — src/base_icnf.jl:447-466, :358-405.
    z0 ~ basedist is drawn on the host from icnf.rng unless given."""
    n = args[-1]
    ys, ps, st = _split(args[:-1], icnf.cond, "generate")
    lib = _cabi.load()
    h = _upload_params(icnf, ps)
    y = _Mat(ys, icnf._cfg.C, "ys") if icnf.cond else None
    if z0 is None and getattr(icnf, "_device_rng", False):
        z, nB, like = None, int(n), y
    else:
        if z0 is None:
            z0 = icnf.rng.standard_normal((n, icnf.D), dtype=np.float32).T
        z = _Mat(z0, icnf.D, "z0")
        nB, like = z.B, z
    e = _draw_eps(icnf, mode, nB, like, eps)
    t0, t1 = steer_tspan(icnf, mode)
    xo, xo_ptr = _out(icnf.nvars, nB, like)
    st_ = _cabi.CnfStats()
    rc = lib.cnfb200_generate(h, mode.code, z.ptr if z else None, y.ptr if y else None, e.ptr if e else None,
                              nB, t1, t0, xo_ptr, C.byref(st_))
    icnf._check(rc, "cnfb200_generate")
    icnf._last_stats = st_.as_dict()
    return _rows_view(xo)


def loss(icnf, mode, xs, *args, eps=None):
    """loss(icnf, mode, xs[, ys], ps, st): TrainMode mean(-logp̂x + λ₁Ė + λ₂ṅ + λ₃Ȧ)
    (src/icnf.jl:481-502); TestMode -mean(logp̂x) (src/base_icnf.jl:489-508)."""
    if isinstance(mode, TrainMode):
        return loss_and_grad(icnf, mode, xs, *args, eps=eps, want_grad=False)[0]
    lp, _ = inference(icnf, mode, xs, *args, eps=eps)
    return float(-(lp.double().mean() if _is_torch(lp) else lp.astype(np.float64).mean()))


def loss_and_grad(icnf, mode, xs, *args, eps=None, want_grad=True, want_grad_x=False):
    """loss and its gradient w.r.t. ps (and xs): what Enzyme/SciMLSensitivity provide to the
    reference's `fit` (src/exts/mlj_ext/core_icnf.jl:59-63); here a discrete adjoint."""
    ys, ps, st = _split(args, icnf.cond, "loss")
    lib = _cabi.load()
    h = _upload_params(icnf, ps)
    x = _Mat(xs, icnf.nvars, "xs")
    y = _Mat(ys, icnf._cfg.C, "ys") if icnf.cond else None
    e = _draw_eps(icnf, mode, x.B, x, eps)
    t0, t1 = steer_tspan(icnf, mode)
    _, _, n = param_offsets(icnf.dims, icnf.nn.planar)
    val = C.c_float(0.0)
    g = np.zeros(n, np.float32) if want_grad else None
    gx, gx_ptr = _out(icnf.nvars, x.B, x) if want_grad_x else (None, None)
    st_ = _cabi.CnfStats()
    rc = lib.cnfb200_loss_grad_mode(h, mode.code, x.ptr, y.ptr if y else None, e.ptr if e else None, x.B, t0, t1,
                                    C.byref(val), g.ctypes.data if want_grad else None, gx_ptr, C.byref(st_))
    icnf._check(rc, "cnfb200_loss_grad_mode")
    icnf._last_stats = st_.as_dict()
    out = (val.value, g)
    return out + (_rows_view(gx),) if want_grad_x else out


def augmented_f(icnf, mode, u, *args, eps=None):
    """one evaluation of augmented_f(u, p, t, icnf, mode, nn, st, eps): src/icnf.jl:148-164,
    :318-350, :384-420.  u: R x B."""
    ys, ps, st = _split(args, icnf.cond, "augmented_f")
    lib = _cabi.load()
    h = _upload_params(icnf, ps)
    R = icnf.D + (3 if isinstance(mode, TrainMode) else 1)
    um = _Mat(u, R, "u")
    y = _Mat(ys, icnf._cfg.C, "ys") if icnf.cond else None
    e = _Mat(eps, icnf.D, "eps") if eps is not None else None
    du, du_ptr = _out(R, um.B, um)
    rc = lib.cnfb200_augmented_f(h, mode.code, um.ptr, y.ptr if y else None, e.ptr if e else None,
                                 um.B, du_ptr)
    icnf._check(rc, "cnfb200_augmented_f")
    return _rows_view(du)


def augmented_f_static(icnf, mode, u, *args, eps=None, out):
    """cnfb200_augmented_f_static: the same evaluation with no device allocation inside the call (no reference
    counterpart; csrc/cnf_noalloc.cu).  u, ys, eps and `out` are CUDA torch tensors that already hold the
    sample-major memory the ABI reads: shape (B, rows), float32, contiguous.  Nothing is copied or allocated here
    either; `out` (B, R) is filled and returned.  (2D + C + R) * round_up(B, 256) floats must fit
    _cabi.STATIC_SCRATCH_BYTES."""
    ys, ps, st = _split(args, icnf.cond, "augmented_f_static")
    lib = _cabi.load()
    h = _upload_params(icnf, ps)
    R = icnf.D + (3 if isinstance(mode, TrainMode) else 1)
    for name, t, rows in (("u", u, R), ("ys", ys if icnf.cond else None, icnf._cfg.C), ("eps", eps, icnf.D), ("out", out, R)):
        if t is None:
            continue
        if not (_is_torch(t) and t.device.type == "cuda" and t.is_contiguous() and t.dtype.is_floating_point
                and t.element_size() == 4 and t.ndim == 2 and t.shape[1] == rows and t.shape[0] == u.shape[0]):
            raise ValueError(f"{name}: expected a contiguous float32 CUDA tensor of shape (B, {rows})")
    rc = lib.cnfb200_augmented_f_static(h, mode.code, u.data_ptr(), ys.data_ptr() if icnf.cond else None,
                                        eps.data_ptr() if eps is not None else None, int(u.shape[0]), out.data_ptr())
    icnf._check(rc, "cnfb200_augmented_f_static")
    return out


def set_option(icnf, key, value):
    """runtime options of the handle: 'time_stages', 'kernel_path'"""
    icnf._check(_cabi.load().cnfb200_set_option(icnf.handle(), key.encode(), int(value)), "cnfb200_set_option")


def last_stats(icnf):
    return icnf._last_stats
