"""ctypes binding of libdla_b200.so (include/dla_b200.h).  Loading fails loudly:
there is no CPU fallback anywhere in this package."""
from __future__ import annotations

import ctypes as C
import os
import threading

import numpy as np

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG_DIR, "libdla_b200.so")

DLA_OK = 0
STATUS_NAMES = {0: "DLA_OK", -1: "DLA_ERR_BAD_ARG", -2: "DLA_ERR_CUDA", -3: "DLA_ERR_NO_DEVICE",
                -4: "DLA_ERR_OOM", -5: "DLA_ERR_STATE"}

NOALLOC_MAX_POINTS = 16384   # DLA_NOALLOC_MAX_POINTS of include/dla_b200.h

# every symbol include/dla_b200.h declares
EXPORTED_SYMBOLS = (
    "dla_abi_version", "dla_device_count", "dla_create", "dla_destroy", "dla_last_error",
    "dla_set_samples", "dla_eval_batch", "dla_stage_batch", "dla_run_staged", "dla_fetch_staged",
    "dla_get_timing", "dla_voigt", "dla_voigt_noalloc", "dla_log_mvnpdf_low_rank", "dla_set_model", "dla_eval_spectra",
    "dla_train_set_data", "dla_train_objective",
)


class DlaError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"{STATUS_NAMES.get(status, status)}: {message}")
        self.status = status


class dla_params(C.Structure):
    _fields_ = [("k", C.c_int32), ("num_lines", C.c_int32), ("num_samples", C.c_int64),
                ("store_sample_ll", C.c_int32), ("reserved", C.c_int32), ("workspace_bytes", C.c_int64)]


class dla_prep_params(C.Structure):
    _fields_ = [("min_lambda", C.c_double), ("max_lambda", C.c_double), ("lya_wavelength", C.c_double),
                ("lyman_limit", C.c_double), ("max_z_cut", C.c_double), ("min_z_cut", C.c_double),
                ("pixel_spacing", C.c_double), ("width", C.c_int32), ("reserved", C.c_int32),
                ("log_c_0", C.c_double), ("log_tau_0", C.c_double), ("log_beta", C.c_double)]


class dla_timing(C.Structure):
    _fields_ = [("ms_total", C.c_double), ("ms_phi", C.c_double), ("ms_voigt", C.c_double),
                ("ms_contract", C.c_double), ("ms_reduce", C.c_double), ("ms_h2d", C.c_double),
                ("ms_d2h", C.c_double), ("launches", C.c_int64), ("h2d_bytes", C.c_int64),
                ("d2h_bytes", C.c_int64), ("chunks", C.c_int64), ("ms_chol", C.c_double),
                ("contraction", C.c_int64)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


_pd = C.POINTER(C.c_double)
_pi64 = C.POINTER(C.c_int64)
_pi32 = C.POINTER(C.c_int32)
_ctx = C.c_void_p

_INPUTS = [_pi64, _pi64, _pd, _pi32, _pd, _pd, _pd, _pd, _pd, _pd, _pd, _pd, _pd]
_OUTPUTS = [_pd, _pd, _pd, _pd, _pd, _pd, _pi32]

_lib = None
_load_lock = threading.Lock()


def load() -> C.CDLL:
    """Load (building first if the sources are newer and nvcc exists) and type the library."""
    global _lib
    if _lib is not None:
        return _lib
    with _load_lock:
        return _load_locked()


def _load_locked() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    from . import build
    if build.is_stale():
        # missing, or older than a kernel source: rebuild when nvcc is here (the GPU box always has it); a stale
        # binary without nvcc is an error, never a silent run of old kernels
        try:
            build.build_library()
        except RuntimeError as e:
            if os.path.exists(LIB_PATH) and "nvcc not found" in str(e):
                raise ImportError(f"{LIB_PATH} is older than its sources and nvcc is not available to rebuild it") from e
            raise
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it with `python -m gp_dla_detection_b200.build` "
                          "(nvcc, sm_100a). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    lib.dla_abi_version.restype = C.c_int
    lib.dla_device_count.restype = C.c_int
    lib.dla_create.restype = C.c_int
    lib.dla_create.argtypes = [C.POINTER(_ctx), C.c_int, C.POINTER(dla_params)]
    lib.dla_destroy.restype = None
    lib.dla_destroy.argtypes = [_ctx]
    lib.dla_last_error.restype = C.c_char_p
    lib.dla_last_error.argtypes = [_ctx]
    lib.dla_set_samples.restype = C.c_int
    lib.dla_set_samples.argtypes = [_ctx, _pd, _pd]
    lib.dla_eval_batch.restype = C.c_int
    lib.dla_eval_batch.argtypes = [_ctx, C.c_int64] + _INPUTS + _OUTPUTS
    lib.dla_stage_batch.restype = C.c_int
    lib.dla_stage_batch.argtypes = [_ctx, C.c_int64] + _INPUTS
    lib.dla_run_staged.restype = C.c_int
    lib.dla_run_staged.argtypes = [_ctx]
    lib.dla_fetch_staged.restype = C.c_int
    lib.dla_fetch_staged.argtypes = [_ctx] + _OUTPUTS
    lib.dla_get_timing.restype = C.c_int
    lib.dla_get_timing.argtypes = [_ctx, C.POINTER(dla_timing)]
    lib.dla_voigt.restype = C.c_int
    lib.dla_voigt.argtypes = [_ctx, _pd, C.c_int64, C.c_double, C.c_double, C.c_int32, _pd]
    lib.dla_voigt_noalloc.restype = C.c_int
    lib.dla_voigt_noalloc.argtypes = [_ctx, _pd, C.c_int64, C.c_double, C.c_double, C.c_int32, _pd]
    lib.dla_log_mvnpdf_low_rank.restype = C.c_int
    lib.dla_log_mvnpdf_low_rank.argtypes = [_ctx, _pd, _pd, _pd, _pd, C.c_int64, C.c_int32, _pd]
    lib.dla_set_model.restype = C.c_int
    lib.dla_set_model.argtypes = [_ctx, C.c_int64, _pd, _pd, _pd, _pd, C.POINTER(dla_prep_params)]
    lib.dla_eval_spectra.restype = C.c_int
    lib.dla_eval_spectra.argtypes = ([_ctx, C.c_int64, _pi64, _pd, _pd, _pd, C.POINTER(C.c_uint8), _pd, _pd, _pd, C.c_int32]
                                     + _OUTPUTS + [_pd, _pd])
    lib.dla_train_set_data.restype = C.c_int
    lib.dla_train_set_data.argtypes = [_ctx, C.c_int64, C.c_int64, _pd, _pd, _pd]
    lib.dla_train_objective.restype = C.c_int
    lib.dla_train_objective.argtypes = [_ctx, _pd, C.c_int32, _pd, _pd]
    _lib = lib
    return lib


def ptr(a: np.ndarray | None, ctype):
    if a is None:
        return None
    return a.ctypes.data_as(C.POINTER(ctype))
