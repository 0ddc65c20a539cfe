"""ctypes binding of libssvi_b200.so (include/ssvi_b200.h).

There is NO CPU fallback: if the shared library is missing, importing the compute path raises.
Build it with `python __graft_entry__.py` (nvcc, sm_100a; the driver's build() entry point).
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SSVI_B200_LIB", os.path.join(_HERE, "libssvi_b200.so"))   # override: kernel A/B builds

MAX_ORDER = 8
MAX_RANK = 64

COV_FULL, COV_DIAG = 0, 1
MODEL_D, MODEL_SIMPLE, MODEL_ROBUST = 0, 1, 2
LIK_NORMAL, LIK_BERNOULLI, LIK_POISSON = 0, 1, 2
UPDATE_S, UPDATE_N = 0, 1
NOISE_F64, NOISE_F32 = 0, 1
TABLE_F64, TABLE_F32, TABLE_F32_TC = 0, 1, 2

TUNE_L2_PREFETCH, TUNE_LIN, TUNE_PF2, TUNE_PF3 = 1, 2, 4, 8

FLAG_NONFINITE_ACC, FLAG_NONFINITE_COV, FLAG_PHI_ZERO, FLAG_CHOL_FAIL, FLAG_EIG_NOCONV, FLAG_TC_RANGE = 1, 2, 4, 8, 16, 32


class SsviConfig(ctypes.Structure):
    _fields_ = [
        ("order", ctypes.c_int32), ("rank", ctypes.c_int32), ("cov_kind", ctypes.c_int32),
        ("model", ctypes.c_int32), ("likelihood", ctypes.c_int32), ("k1", ctypes.c_int32),
        ("k2", ctypes.c_int32), ("mean_update", ctypes.c_int32), ("cov_update", ctypes.c_int32),
        ("noise_dtype", ctypes.c_int32), ("kernel_path", ctypes.c_int32), ("tuning", ctypes.c_int32), ("table_dtype", ctypes.c_int32), ("tc_scale_log2", ctypes.c_int32),
        ("eta", ctypes.c_double), ("cov_eta", ctypes.c_double), ("sigma_eta", ctypes.c_double),
        ("poisson_eta", ctypes.c_double), ("min_eigenval", ctypes.c_double), ("cov_floor", ctypes.c_double),
        ("likelihood_param", ctypes.c_double), ("dims", ctypes.c_int64 * MAX_ORDER),
    ]


class SsviLoop(ctypes.Structure):
    """include/ssvi_b200.h: SsviLoop (argument block of ssvi_factorize_iterations)."""
    _fields_ = [
        ("rowptr", ctypes.c_void_p * MAX_ORDER), ("perm", ctypes.c_void_p * MAX_ORDER),
        ("rowptr_host", ctypes.c_void_p * MAX_ORDER),
        ("z_other", ctypes.c_void_p), ("z_own", ctypes.c_void_p), ("r_w", ctypes.c_void_p), ("z_f", ctypes.c_void_p),
        ("pSigma", ctypes.c_void_p), ("w_state", ctypes.c_void_p), ("hyper_out", ctypes.c_void_p),
        ("stop", ctypes.c_void_p), ("workspace", ctypes.c_void_p), ("workspace_bytes", ctypes.c_int64),
        ("epsilon", ctypes.c_double),
        ("noise_seed", ctypes.c_uint64), ("noise_draws", ctypes.c_uint64),
        ("minibatch_seed", ctypes.c_uint64), ("minibatch_draws", ctypes.c_uint64),
        ("col", ctypes.c_int64 * MAX_ORDER), ("launches", ctypes.c_int64),
        ("time_step", ctypes.c_int32 * MAX_ORDER),
        ("pool_slots", ctypes.c_int32), ("batch_size", ctypes.c_int32), ("chunk_entries", ctypes.c_int32),
        ("fixed_covariance", ctypes.c_int32),
    ]


# name -> (restype, argtypes); must list every symbol include/ssvi_b200.h declares
_P = ctypes.c_void_p
_I32, _I64 = ctypes.c_int32, ctypes.c_int64
SIGNATURES = {
    "ssvi_version": (ctypes.c_char_p, []),
    "ssvi_acc_len": (_I64, [ctypes.POINTER(SsviConfig)]),
    "ssvi_sampler_row_len": (_I64, [ctypes.POINTER(SsviConfig)]),
    "ssvi_noise_block_len": (_I64, [ctypes.POINTER(SsviConfig)]),
    "ssvi_noise_offset": (_I64, [ctypes.POINTER(SsviConfig), _I32, _I32]),
    "ssvi_create": (ctypes.c_int, [ctypes.POINTER(SsviConfig), ctypes.POINTER(_P)]),
    "ssvi_destroy": (None, [_P]),
    "ssvi_set_tuning": (ctypes.c_int, [_P, _I32]),
    "ssvi_set_row_windows": (ctypes.c_int, [_P, _P, _P]),
    "ssvi_last_error": (ctypes.c_char_p, [_P]),
    "ssvi_bind_data": (ctypes.c_int, [_P, _P, _P, _I64]),
    "ssvi_bind_state": (ctypes.c_int, [_P] + [ctypes.POINTER(_P)] * 6),
    "ssvi_bind_noise": (ctypes.c_int, [_P, _P, _P, _P, _P, _I32]),
    "ssvi_bind_quadrature": (ctypes.c_int, [_P, _P, _I32]),
    "ssvi_refresh_sampler": (ctypes.c_int, [_P, _I32, _P, _I64, _P]),
    "ssvi_accumulate": (ctypes.c_int, [_P, _I32, _I64, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "ssvi_sweep_supported": (ctypes.c_int, [_P]),
    "ssvi_sweep_table_bytes": (_I64, [_P, _I32, _I32]),
    "ssvi_sweep_accumulate": (ctypes.c_int, [_P, _I32, _I32, _I32, _I64, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I64, _P, _I64, _I64, _I64, _P, _P]),
    "ssvi_sweep_work_bytes": (_I64, [_P, _I64]),
    "ssvi_reduce_chunks": (ctypes.c_int, [_P, _I64, _P, _P, _P, _P]),
    "ssvi_reduce_chunks_into": (ctypes.c_int, [_P, _I64, _P, _P, _P, _I32, _P]),
    "ssvi_row_update": (ctypes.c_int, [_P, _I32, _I64, _P, _P, _P, _P, _P, _I32, _I32, _P, _P]),
    "ssvi_sigma_update": (ctypes.c_int, [_P, _I64, _P, _P, _P, _P]),
    "ssvi_hyper_reduce": (ctypes.c_int, [_P, _I32, _P, _P]),
    "ssvi_estimate_entries": (ctypes.c_int, [_P, _I32, _P, _P, _I64, _I32, _P, _P, _P]),
    "ssvi_predict_mean": (ctypes.c_int, [_P, _P, _I64, _P, _P]),
    "ssvi_draw_minibatches": (ctypes.c_int, [_P, _I64, _P, _P, _P, _I32, ctypes.c_uint64, _P, _P, _P]),
    "ssvi_fill_noise": (ctypes.c_int, [_P, _P, _P, _P, _P, _I32, ctypes.c_uint64, ctypes.c_uint64, _P]),
    "ssvi_loop_workspace_bytes": (_I64, [_P, _I32]),
    "ssvi_factorize_iterations": (ctypes.c_int, [_P, ctypes.POINTER(SsviLoop), _I64, _I64, _P]),
    "ssvi_static_scratch_bytes": (_I64, []),
    "ssvi_accumulate_rows_static": (ctypes.c_int, [_P, _I32, _I64, _P, _P, _P, _P, _P, _P, _P, _I64, _P, _P, _P]),
    "ssvi_check_flags": (ctypes.c_int, [_P, ctypes.POINTER(_I32), _P]),
}

_lib = None


def load():
    """Load the shared library once; raise loudly if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "libssvi_b200.so not found at %s: build the CUDA extension first "
            "(python -c 'import __graft_entry__ as g; g.build()'). There is no CPU fallback." % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


class SsviError(RuntimeError):
    pass


def check(lib, handle, rc):
    if rc != 0:
        msg = lib.ssvi_last_error(handle)
        raise SsviError("libssvi_b200 error %d: %s" % (rc, msg.decode() if msg else "?"))
