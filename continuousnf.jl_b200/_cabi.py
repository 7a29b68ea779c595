"""ctypes binding of libcnfb200.so (include/cnfb200.h).  This is the same boundary the Julia
extension binds with `ccall`.  There is no CPU fallback: if the CUDA library is
missing or no B200 is visible, calls fail loudly."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# CNFB200_LIB: debugging aid (load a differently built library, e.g. to reproduce a fixed bug); the product is the in-tree file
LIB_PATH = os.environ.get("CNFB200_LIB") or os.path.join(_HERE, "libcnfb200.so")

MAX_LAYERS = 8
MODE_TEST, MODE_TRAIN = 0, 1
TRACE_VJP, TRACE_JVP = 0, 1
FLAG_COND, FLAG_AUGMENTED, FLAG_NORM_Z, FLAG_NORM_J, FLAG_NORM_Z_AUG, FLAG_LAST_LINEAR = 1, 2, 4, 8, 16, 32
STATIC_SCRATCH_BYTES = 16 << 20   # CNFB200_STATIC_SCRATCH_BYTES
RETCODES = {0: "Success", 1: "MaxIters", 2: "DtLessThanMin", 3: "Unstable"}

EXPORTS = [
    "cnfb200_version", "cnfb200_create", "cnfb200_destroy", "cnfb200_set_option", "cnfb200_last_error",
    "cnfb200_set_params", "cnfb200_comm_unique_id", "cnfb200_comm_init", "cnfb200_comm_p2p", "cnfb200_augmented_f",
    "cnfb200_logdensity", "cnfb200_loss_grad", "cnfb200_loss_grad_mode", "cnfb200_generate", "cnfb200_measure_fp32_peak",
    "cnfb200_seed_rng", "cnfb200_last_eps", "cnfb200_fit_epochs", "cnfb200_host_alloc", "cnfb200_host_free", "cnfb200_set_steer",
    "cnfb200_augmented_f_static",
]


class CnfConfig(C.Structure):
    _fields_ = [
        ("D", C.c_int32), ("C", C.c_int32), ("nvars", C.c_int32), ("n_layers", C.c_int32),
        ("dims", C.c_int32 * (MAX_LAYERS + 1)), ("trace_mode", C.c_int32), ("flags", C.c_uint32),
        ("lambda1", C.c_float), ("lambda2", C.c_float), ("lambda3", C.c_float),
        ("reltol", C.c_float), ("abstol", C.c_float), ("maxiters", C.c_int64),
        ("dt0", C.c_float), ("dtmax", C.c_float), ("dtmin", C.c_float),
        ("device", C.c_int32), ("kernel_path", C.c_int32),
    ]


class CnfStats(C.Structure):
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


class CnfError(RuntimeError):
    pass


_lib = None
_FP = C.POINTER(C.c_float)
_I64P = C.POINTER(C.c_int64)


def load():
    """dlopen libcnfb200.so; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise CnfError(f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; "
                       "g.build()'` (nvcc, sm_100a). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    lib.cnfb200_version.restype = C.c_char_p
    lib.cnfb200_last_error.restype = C.c_char_p
    lib.cnfb200_last_error.argtypes = [C.c_void_p]
    lib.cnfb200_create.argtypes = [C.POINTER(CnfConfig), C.POINTER(C.c_void_p)]
    lib.cnfb200_destroy.argtypes = [C.c_void_p]
    lib.cnfb200_set_option.argtypes = [C.c_void_p, C.c_char_p, C.c_int64]
    lib.cnfb200_set_params.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, _I64P, _I64P]
    lib.cnfb200_comm_unique_id.argtypes = [C.c_char_p]
    lib.cnfb200_comm_init.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_char_p]
    lib.cnfb200_comm_p2p.argtypes = [C.c_void_p]
    lib.cnfb200_comm_p2p.restype = C.c_int32
    lib.cnfb200_augmented_f.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                        C.c_int64, C.c_void_p]
    lib.cnfb200_augmented_f_static.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                               C.c_int64, C.c_void_p]
    lib.cnfb200_logdensity.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                       C.c_int64, C.c_float, C.c_float, C.c_void_p, C.c_void_p,
                                       C.c_void_p, C.POINTER(CnfStats)]
    lib.cnfb200_loss_grad.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                      C.c_float, C.c_float, _FP, C.c_void_p, C.c_void_p,
                                      C.POINTER(CnfStats)]
    lib.cnfb200_loss_grad_mode.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                      C.c_float, C.c_float, _FP, C.c_void_p, C.c_void_p,
                                      C.POINTER(CnfStats)]
    lib.cnfb200_generate.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_int64, C.c_float, C.c_float, C.c_void_p,
                                     C.POINTER(CnfStats)]
    lib.cnfb200_measure_fp32_peak.argtypes = [C.c_int32, C.POINTER(C.c_double)]
    lib.cnfb200_host_alloc.argtypes = [C.c_size_t, C.POINTER(C.c_void_p)]
    lib.cnfb200_host_free.argtypes = [C.c_void_p]
    lib.cnfb200_set_steer.argtypes = [C.c_void_p, C.c_float]
    lib.cnfb200_seed_rng.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64]
    lib.cnfb200_last_eps.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
    lib.cnfb200_fit_epochs.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_void_p,
                                       C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, C.c_void_p, C.c_void_p,
                                       C.POINTER(CnfStats)]
    for name in EXPORTS:
        if name not in ("cnfb200_version", "cnfb200_last_error"):
            getattr(lib, name).restype = C.c_int32
    _lib = lib
    return lib
