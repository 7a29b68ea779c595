"""ctypes binding of the CPU oracle (oracle/cnf_oracle.c).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module; the product package (continuousnf.jl_b200/) never does.

PARITY UNPINNED: the reference is Julia and cannot run here; see cnf_oracle.c's header.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

MAX_LAYERS = 8
MODE_TEST, MODE_TRAIN = 0, 1
TRACE_VJP, TRACE_JVP = 0, 1
FLAG_COND, FLAG_AUGMENTED, FLAG_NORM_Z, FLAG_NORM_J, FLAG_NORM_Z_AUG, FLAG_LAST_LINEAR = 1, 2, 4, 8, 16, 32


class Config(C.Structure):
    """mirror of include/cnfb200.h cnf_config"""
    _fields_ = [
        ("D", C.c_int32), ("C", C.c_int32), ("nvars", C.c_int32), ("n_layers", C.c_int32),
        ("dims", C.c_int32 * (MAX_LAYERS + 1)), ("trace_mode", C.c_int32), ("flags", C.c_uint32),
        ("lambda1", C.c_float), ("lambda2", C.c_float), ("lambda3", C.c_float),
        ("reltol", C.c_float), ("abstol", C.c_float), ("maxiters", C.c_int64),
        ("dt0", C.c_float), ("dtmax", C.c_float), ("dtmin", C.c_float),
        ("device", C.c_int32), ("kernel_path", C.c_int32),
    ]


class Stats(C.Structure):
    """mirror of include/cnfb200.h cnf_stats"""
    _fields_ = [
        ("naccept", C.c_int64), ("nreject", C.c_int64), ("nf", C.c_int64),
        ("retcode", C.c_int32), ("kernel_path", C.c_int32),
        ("t_final", C.c_float), ("dt_last", C.c_float),
        ("ms_total", C.c_double), ("ms_solve", C.c_double), ("gpu_launches", C.c_int64),
        ("ms_stage", C.c_double), ("n_stage", C.c_int64),
        ("stiff_max", C.c_float), ("stiff_run", C.c_int32),
    ]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


def build(force=False):
    so = os.path.join(_HERE, "libcnf_oracle.so")
    src = os.path.join(_HERE, "cnf_oracle.c")
    if force or not os.path.exists(so) or (
            os.path.exists(src) and os.path.getmtime(so) < os.path.getmtime(src)):
        subprocess.check_call(["make", "-C", _HERE, "-B"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
    return _LIB


def make_config(nvars, naug, dims, *, C_=0, trace_mode=TRACE_VJP, lambda1=0.0, lambda2=0.0,
                lambda3=0.0, reltol=0.0, abstol=0.0, maxiters=0, dt0=0.0, dtmax=0.0, dtmin=0.0, planar=False):
    cfg = Config()
    D = nvars + naug
    assert dims[0] == D + C_ and dims[-1] == D, (dims, D, C_)
    cfg.D, cfg.C, cfg.nvars, cfg.n_layers = D, C_, nvars, len(dims) - 1
    for i, d in enumerate(dims):
        cfg.dims[i] = d
    cfg.trace_mode = trace_mode
    fl = 0
    if C_ > 0:
        fl |= FLAG_COND
    if naug != 0:
        fl |= FLAG_AUGMENTED
    if lambda1 != 0:
        fl |= FLAG_NORM_Z
    if lambda2 != 0:
        fl |= FLAG_NORM_J
    if lambda3 != 0:
        fl |= FLAG_NORM_Z_AUG
    if planar:   # PlanarLayer(nvars, tanh; n_cond): u * tanh(w'[z; ys] + b) = Dense(d0 => 1, tanh) then a bias-free product
        assert len(dims) == 3 and dims[1] == 1
        fl |= FLAG_LAST_LINEAR
    cfg.flags = fl
    cfg.lambda1, cfg.lambda2, cfg.lambda3 = lambda1, lambda2, lambda3
    cfg.reltol, cfg.abstol, cfg.maxiters = reltol, abstol, maxiters
    cfg.dt0, cfg.dtmax, cfg.dtmin = dt0, dtmax, dtmin
    return cfg


def offsets(dims, planar=False):
    """Lux field order: layer_k = (weight[d_{k+1} x d_k col-major], bias[d_{k+1}]); PlanarLayer
    (src/layers/planar_layer.jl:38-60): (u[nvars], w[nvars + n_cond], b[1])"""
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


def _f(a):
    return None if a is None else a.ctypes.data_as(C.POINTER(C.c_float))


def _i64(a):
    return a.ctypes.data_as(C.POINTER(C.c_int64))


def _prep(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def R_of(cfg, mode):
    return cfg.D + (3 if mode == MODE_TRAIN else 1)


def augmented_f(cfg, theta, mode, u, ys=None, eps=None):
    """u: (B, R) sample-major. returns du (B, R)"""
    w, b, _ = offsets(list(cfg.dims[:cfg.n_layers + 1]), bool(cfg.flags & FLAG_LAST_LINEAR))
    u, ys, eps, theta = _prep(u), _prep(ys), _prep(eps), _prep(theta)
    B = u.shape[0]
    du = np.empty_like(u)
    lib().cnf_oracle_augmented_f(C.byref(cfg), _f(theta), _i64(w), _i64(b), C.c_int32(mode),
                                 _f(u), _f(ys), _f(eps), C.c_int64(B), _f(du))
    return du


def logdensity(cfg, theta, mode, xs, ys=None, eps=None, t0=0.0, t1=1.0):
    """xs: (B, nvars). returns logp (B,), regs (B, 3|1), z (B, D), stats dict"""
    w, b, _ = offsets(list(cfg.dims[:cfg.n_layers + 1]), bool(cfg.flags & FLAG_LAST_LINEAR))
    xs, ys, eps, theta = _prep(xs), _prep(ys), _prep(eps), _prep(theta)
    B = xs.shape[0]
    if eps is None:
        eps = np.zeros((B, cfg.D), np.float32)
    nreg = 3 if mode == MODE_TRAIN else 1
    logp = np.empty(B, np.float32)
    regs = np.empty((B, nreg), np.float32)
    z = np.empty((B, cfg.D), np.float32)
    st = Stats()
    lib().cnf_oracle_logdensity(C.byref(cfg), _f(theta), _i64(w), _i64(b), C.c_int32(mode),
                                _f(xs), _f(ys), _f(eps), C.c_int64(B), C.c_float(t0),
                                C.c_float(t1), _f(logp), _f(regs), _f(z), C.byref(st))
    return logp, regs, z, st.as_dict()


def generate(cfg, theta, mode, z0, ys=None, eps=None, t_from=1.0, t_to=0.0):
    w, b, _ = offsets(list(cfg.dims[:cfg.n_layers + 1]), bool(cfg.flags & FLAG_LAST_LINEAR))
    z0, ys, eps, theta = _prep(z0), _prep(ys), _prep(eps), _prep(theta)
    n = z0.shape[0]
    if eps is None:
        eps = np.zeros((n, cfg.D), np.float32)
    xs = np.empty((n, cfg.nvars), np.float32)
    st = Stats()
    lib().cnf_oracle_generate(C.byref(cfg), _f(theta), _i64(w), _i64(b), C.c_int32(mode),
                              _f(z0), _f(ys), _f(eps), C.c_int64(n), C.c_float(t_from),
                              C.c_float(t_to), _f(xs), C.byref(st))
    return xs, st.as_dict()


def loss_grad(cfg, theta, xs, ys=None, eps=None, t0=0.0, t1=1.0, want_grad=True, mode=1):
    w, b, n = offsets(list(cfg.dims[:cfg.n_layers + 1]), bool(cfg.flags & FLAG_LAST_LINEAR))
    xs, ys, eps, theta = _prep(xs), _prep(ys), _prep(eps), _prep(theta)
    B = xs.shape[0]
    loss = C.c_float(0)
    gth = np.zeros(n, np.float32) if want_grad else None
    gx = np.zeros_like(xs) if want_grad else None
    st = Stats()
    lib().cnf_oracle_loss_grad_mode(C.byref(cfg), _f(theta), C.c_int64(n), _i64(w), _i64(b), C.c_int32(mode),
                                    _f(xs), _f(ys), _f(eps), C.c_int64(B), C.c_float(t0), C.c_float(t1),
                                    C.byref(loss), _f(gth), _f(gx), C.byref(st))
    return loss.value, gth, gx, st.as_dict()


def philox4x32_10(ctr, key):
    """one Philox4x32-10 block (known-answer tests)"""
    c = (C.c_uint32 * 4)(*ctr); k = (C.c_uint32 * 2)(*key); o = (C.c_uint32 * 4)()
    lib().cnf_oracle_philox4x32_10(c, k, o)
    return [int(v) for v in o]


def philox_normal(seed, sample_offset, draw, rows, B):
    """the matrix cnfb200_seed_rng's draw number `draw` produces: rows x B (column = sample)"""
    out = np.empty((B, rows), np.float32)
    f = lib().cnf_oracle_philox_normal
    f.argtypes = [C.c_uint64, C.c_uint64, C.c_uint32, C.c_int32, C.c_int64, C.c_void_p]
    f.restype = None
    f(seed, sample_offset, draw, rows, B, out.ctypes.data)
    return out.T


def num_threads():
    return int(lib().cnf_oracle_num_threads())


def set_num_threads(n=None):
    """use `n` OpenMP threads (default: every host core; torchrun pins OMP_NUM_THREADS=1)"""
    n = (os.cpu_count() or 1) if n is None else int(n)
    lib().cnf_oracle_set_num_threads(C.c_int32(n))
    return num_threads()


def last_dts():
    """accepted dt sequence of the most recent oracle solve in this process"""
    buf = np.zeros(4096, np.float32)
    lib().cnf_oracle_last_dts.restype = C.c_int64
    n = lib().cnf_oracle_last_dts(_f(buf), C.c_int64(4096))
    return buf[:min(n, 4096)].copy()


def set_perturb(on):
    """rounding-sensitivity probe: evaluate tanh with libm tanhf instead of the tanh_fast rational
    (a ~3e-7 perturbation of every activation).  The spread between the two oracle runs is the
    conditioning floor a second FP32 implementation cannot be expected to beat."""
    lib().cnf_oracle_set_perturb(C.c_int(1 if on else 0))
