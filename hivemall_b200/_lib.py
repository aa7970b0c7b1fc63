"""ctypes binding of libhivemall_b200.so (the C-ABI in include/hivemall_b200.h).

Loading fails loudly when the library has not been built; nothing here falls back to a
CPU implementation.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libhivemall_b200.so")

HM_OK = 0
STATUS_NAMES = {0: "HM_OK", 1: "HM_ERR_INVALID_ARGUMENT", 2: "HM_ERR_CUDA", 3: "HM_ERR_STATE",
                4: "HM_ERR_OOM", 5: "HM_ERR_UNSUPPORTED"}


class HivemallError(RuntimeError):
    """Raised for every non-zero hm_status; .status holds the code."""

    def __init__(self, status: int, message: str):
        super().__init__(f"{STATUS_NAMES.get(status, status)}: {message}")
        self.status = status
        self.message = message


class UDFArgumentError(HivemallError, ValueError):
    """HM_ERR_INVALID_ARGUMENT: the role UDFArgumentException plays for a Hive UDTF."""


class HmConfig(C.Structure):
    _fields_ = [
        ("task", C.c_int32), ("loss", C.c_int32), ("opt", C.c_int32), ("eta_sched", C.c_int32),
        ("reg", C.c_int32), ("grad_norm", C.c_int32), ("mini_batch", C.c_int32), ("iters", C.c_int32),
        ("dims", C.c_int64), ("total_steps", C.c_double),
        ("eta0", C.c_float), ("power_t", C.c_float), ("lambda_", C.c_float), ("eps", C.c_float),
        ("rho", C.c_float), ("beta1", C.c_float), ("beta2", C.c_float),
        ("alpha", C.c_float), ("beta", C.c_float), ("lambda1", C.c_float), ("lambda2", C.c_float),
        ("factors", C.c_int32), ("fm_classification", C.c_int32), ("sigma", C.c_float),
        ("lambda_w0", C.c_float), ("lambda_w", C.c_float), ("lambda_v", C.c_float),
        ("seed", C.c_uint32), ("mix_interval", C.c_int32), ("hash_seed", C.c_uint32),
        ("pa_c", C.c_float), ("pa_epsilon", C.c_float), ("cov_r", C.c_float), ("cov_phi", C.c_float), ("loss_param", C.c_float),
        ("reserved", C.c_int32 * 3),
    ]


class HmStats(C.Structure):
    _fields_ = [("rows", C.c_int64), ("nnz", C.c_int64), ("mini_batches", C.c_int64),
                ("kernel_launches", C.c_int64), ("h2d_bytes", C.c_int64), ("d2h_bytes", C.c_int64),
                ("reserved0", C.c_double)]


# name -> (restype, argtypes).  tests/test_abi.py checks this table against the header.
_P = C.c_void_p
SIGNATURES = {
    "hm_last_error": (C.c_char_p, []),
    "hm_abi_version": (C.c_int, []),
    "hm_device_count": (C.c_int, []),
    "hm_host_alloc": (C.c_int, [C.POINTER(_P), C.c_int64]),
    "hm_host_free": (C.c_int, [_P]),
    "hm_mhash": (C.c_int, [_P, _P, C.c_int64, C.c_uint32, C.c_int32, _P, C.c_int]),
    "hm_mhash_device": (C.c_int, [_P, _P, C.c_int64, C.c_uint32, C.c_int32, _P, _P]),
    "hm_feature_hashing": (C.c_int, [_P, _P, C.c_int64, C.c_uint32, C.c_int32, _P, _P, C.c_int]),
    "hm_parse_options": (C.c_int, [C.c_char_p, C.c_char_p, C.POINTER(HmConfig)]),
    "hm_udtf_create": (C.c_int, [C.c_char_p, C.c_char_p, C.c_int, C.POINTER(_P)]),
    "hm_udtf_create_from_config": (C.c_int, [C.POINTER(HmConfig), C.c_int, C.POINTER(_P)]),
    "hm_udtf_process": (C.c_int, [_P, C.c_int64, _P, _P, _P, _P]),
    "hm_udtf_process_split": (C.c_int, [_P, C.c_int64, _P, _P, _P, _P, _P, _P]),
    "hm_udtf_process_columns": (C.c_int, [_P, C.c_int64, C.c_int32, _P, _P, C.c_int32, _P, _P]),
    "hm_udtf_process_strings": (C.c_int, [_P, C.c_int64, _P, _P, _P, _P]),
    "hm_udtf_process_device": (C.c_int, [_P, C.c_int64, C.c_int64, _P, _P, _P, _P]),
    "hm_udtf_close": (C.c_int, [_P, C.POINTER(C.c_int64)]),
    "hm_udtf_forward": (C.c_int, [_P, _P, _P, _P, C.c_int64, C.POINTER(C.c_int64)]),
    "hm_udtf_destroy": (C.c_int, [_P]),
    "hm_udtf_num_arrays": (C.c_int, [_P]),
    "hm_udtf_array": (C.c_int, [_P, C.c_int, C.POINTER(C.c_char_p), C.POINTER(_P), C.POINTER(C.c_int64),
                                C.POINTER(C.c_int32)]),
    "hm_udtf_read_array": (C.c_int, [_P, C.c_int, _P, C.c_int64]),
    "hm_udtf_write_array": (C.c_int, [_P, C.c_int, _P, C.c_int64]),
    "hm_udtf_get_step": (C.c_int, [_P, C.POINTER(C.c_int64)]),
    "hm_udtf_set_step": (C.c_int, [_P, C.c_int64]),
    "hm_udtf_sync": (C.c_int, [_P]),
    "hm_udtf_stream": (_P, [_P]),
    "hm_udtf_scale_replica": (C.c_int, [_P, C.c_float]),
    "hm_udtf_predict": (C.c_int, [_P, C.c_int64, _P, _P, _P, _P]),
    "hm_udtf_predict_device": (C.c_int, [_P, C.c_int64, _P, _P, _P, _P]),
    "hm_udtf_last_margins": (C.c_int, [_P, _P, C.c_int64, C.POINTER(C.c_int64)]),
    "hm_udtf_save_model_table": (C.c_int, [_P, C.c_char_p, C.c_int, C.c_char, C.POINTER(C.c_int64)]),
    "hm_udtf_load_model_table": (C.c_int, [_P, C.c_char_p, C.c_int, C.c_char, C.POINTER(C.c_int64)]),
    "hm_udtf_config": (C.c_int, [_P, C.POINTER(HmConfig)]),
    "hm_udtf_hot_features": (C.c_int, [_P, _P, C.c_int32, C.POINTER(C.c_int32)]),
    "hm_model_table_noalloc": (C.c_int, [_P, C.c_int64, _P, _P, C.c_int64, _P, _P]),
    "hm_udtf_stats": (C.c_int, [_P, C.POINTER(HmStats)]),
    "hm_udtf_set_profiling": (C.c_int, [_P, C.c_int]),
    "hm_udtf_num_kernels": (C.c_int, [_P]),
    "hm_udtf_kernel_time": (C.c_int, [_P, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_int64),
                                      C.POINTER(C.c_double), C.POINTER(C.c_double)]),
}

_lib = None


def load() -> C.CDLL:
    """Load the shared library (once).  Raises if it is missing: build it with
    `python -m hivemall_b200.build` (needs nvcc); there is no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing. hivemall_b200 is CUDA-only: build it with "
            "`python -m hivemall_b200.build` (nvcc, sm_100a). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here means header and library disagree
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(status: int) -> None:
    if status == HM_OK:
        return
    msg = load().hm_last_error().decode("utf-8", "replace")
    if status == 1:
        raise UDFArgumentError(status, msg)
    raise HivemallError(status, msg)
