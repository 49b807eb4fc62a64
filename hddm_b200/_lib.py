"""ctypes loader for libhddm_b200.so (the C ABI in include/hddm_b200.h).

Fails loudly: if the shared library is missing it is built with nvcc; if it cannot be built or
no CUDA device is usable every call raises RuntimeError.  There is no CPU fallback.
"""
import ctypes as C
import os

import numpy as np

from . import build as _build

OK, ERR_CUDA, ERR_ARG, ERR_SUPPORT, ERR_RT_RANGE = 0, 1, 2, 3, 4
MAX_DEPTH = 24

c_dp = C.POINTER(C.c_double)
c_lp = C.POINTER(C.c_int64)
c_ip = C.POINTER(C.c_int32)
c_up = C.POINTER(C.c_ulonglong)

_lib = None

# name -> (restype, argtypes); mirrors include/hddm_b200.h one to one (tests/test_boundary.py
# checks that every symbol declared in the header is listed here and exported).
_d, _i, _l, _v = C.c_double, C.c_int, C.c_int64, C.c_void_p
_NUM = [_d, _i, _i, _i, _d]  # err, n_st, n_sz, use_adaptive, simps_err
SIGNATURES = {
    "hddm_b200_version": (_i, []),
    "hddm_b200_last_error": (C.c_char_p, []),
    "hddm_b200_init": (_i, [_i]),
    "hddm_b200_device_count": (_i, [C.POINTER(_i)]),
    "hddm_b200_synchronize": (_i, []),
    "hddm_b200_fp64_peak": (_i, [_i, c_dp, C.POINTER(C.c_float)]),
    "hddm_b200_set_kernel": (_i, [_i]),
    "hddm_b200_set_sort": (_i, [_i]),
    "hddm_sort_trials_scratch": (_l, [_l, _l]),
    "hddm_sort_trials_dev": (_i, [_v, _l, _v, _l, _v, _v, _l, _v]),
    "hddm_series_table": (_i, [_d, c_dp, c_ip, C.POINTER(_i)]),
    "hddm_full_pdf": (_i, [_d] * 8 + _NUM + [c_dp]),
    "hddm_pdf_array": (_i, [c_dp, _l] + [_d] * 7 + [_d, _i, _i, _i, _i, _d, _d, _d, c_dp]),
    "hddm_wiener_like": (_i, [c_dp, _l] + [_d] * 7 + _NUM + [_d, _d, c_dp]),
    "hddm_wiener_like_multi": (_i, [c_dp, _l, C.POINTER(c_dp), c_lp] + _NUM + [_d, _d, c_dp]),
    "hddm_wiener_like_contaminant": (_i, [c_dp, c_ip, _l] + [_d] * 9 + _NUM + [c_dp]),
    "hddm_gen_cdf_using_pdf": (_i, [c_dp, _l] + [_d] * 7 + _NUM + [_d, _d, c_dp]),
    "hddm_gen_rts_from_cdf": (_i, [c_dp, _l] + [_d] * 7 + [c_dp, c_dp, _l, c_dp, c_dp]),
    "hddm_dmat_cdf_array": (_i, [c_dp, _l] + [_d] * 9 + [c_dp]),
    "hddm_wiener_like_batch": (_i, [c_dp, _l, c_lp, c_dp, _l] + _NUM + [c_dp, c_up]),
    "hddm_wiener_like_multi_batch": (_i, [c_dp, _l, c_dp, c_dp, _l] + _NUM + [c_dp, c_up]),
    "hddm_gen_rts_from_cdf_batch": (_i, [c_dp, _l, c_dp, _l, c_dp, c_dp, _l, c_dp, c_dp]),
    "hddm_gen_rts_from_cdf_batch_seeded": (_i, [c_dp, _l, c_dp, _l, C.c_ulonglong, _l, _l, c_dp, c_dp]),
    "hddm_ppc_stats": (_i, [c_dp, _l, _l, c_dp]),
    "hddm_dmat_cdf_batch": (_i, [c_dp, _l, c_dp, _l, c_dp, c_ip]),
    "hddm_wiener_like_regression_batch": (_i, [c_dp, _l, c_dp, _l, c_dp, c_dp, _l] + _NUM + [c_dp, c_up]),
    "hddm_regdata_upload": (_i, [c_dp, _l, _i, c_lp, C.POINTER(c_dp), C.POINTER(_v)]),
    "hddm_regdata_free": (_i, [_v]),
    "hddm_regdata_like_batch": (_i, [_v, C.POINTER(_i), C.POINTER(_i), C.POINTER(c_dp), c_dp, _l] + _NUM + [c_dp, c_up]),
    "hddm_simulate_drift_batch": (_i, [c_dp, _l, _l, _d, _d, C.c_ulonglong, c_dp, c_dp, _l, c_dp]),
    "hddm_simulate_drift_switch_batch": (_i, [c_dp, c_dp, _l, _l, _d, _d, C.c_ulonglong, c_dp, c_dp, _l, c_dp]),
    "hddm_wiener_like_rlddm": (_i, [c_dp, c_lp, c_dp, c_lp, _l] + [_d] * 10 + _NUM + [_d, _d, c_dp]),
    "hddm_wiener_like_rlddm_batch": (_i, [c_dp, c_lp, c_dp, c_lp, _l, c_dp, _l] + _NUM + [c_dp, c_up]),
    "hddm_wiener_like_rl": (_i, [c_lp, c_dp, c_lp, _l] + [_d] * 7 + [c_dp]),
    "hddm_wiener_like_rl_batch": (_i, [c_lp, c_dp, c_lp, _l, c_dp, _l, c_dp]),
    "hddm_wiener_like_multi_rlddm": (_i, [c_dp, c_lp, c_dp, c_lp, _l, _d, C.POINTER(c_dp), c_lp] + _NUM +
                                     [_d, _d, c_dp]),
    "hddm_trials_upload": (_i, [c_dp, _l, C.POINTER(_v)]),
    "hddm_trials_free": (_i, [_v]),
    "hddm_trials_wiener_like": (_i, [_v] + [_d] * 7 + _NUM + [_d, _d, c_dp]),
    "hddm_trials_wiener_like_batch": (_i, [_v, c_lp, c_dp, _l] + _NUM + [c_dp, c_up]),
    "hddm_wiener_like_nodes": (_i, [c_dp, _l, c_lp, _l, c_lp, c_dp, _l] + _NUM + [c_dp, c_up]),
    "hddm_trials_wiener_like_nodes": (_i, [_v, c_lp, _l, c_lp, c_dp, _l] + _NUM + [c_dp, c_up]),
    "hddm_hier_ddm_sample": (_i, [_v, c_lp, _l, _l, c_dp, c_dp, _d, _d, _d] + _NUM +
                             [C.POINTER(C.c_uint64), _l, c_dp, c_dp, c_lp]),
    "hddm_wiener_like_batch_scratch": (_l, [_l, _l]),
    "hddm_wiener_like_batch_dev": (_i, [_v, _l, _v, _l, _v, _l] + _NUM + [_v, _v, _v, _v]),
    "hddm_wiener_like_batch_dev_static": (_i, [_v, _l, _v, _l, _v, _l] + _NUM + [_v, _v, _v]),
    "hddm_pdf_array_dev": (_i, [_v, _l] + [_d] * 7 + [_d, _i, _i, _i, _i, _d, _d, _d, _v, _v, _v]),
    "hddm_dmat_cdf_array_dev": (_i, [_v, _l] + [_d] * 9 + [_v, _v]),
}


def library_path():
    return _build.LIB


def lib():
    """Load (building if needed) the native library.  Raises RuntimeError when impossible."""
    global _lib
    if _lib is None:
        path = _build.LIB
        if not os.path.exists(path):
            path = _build.build_library()
        L = C.CDLL(path)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc):
    """Map a C-ABI status to the exception the reference raises in the same situation."""
    if rc == OK:
        return
    msg = lib().hddm_b200_last_error().decode("utf-8", "replace")
    if rc == ERR_SUPPORT:
        raise ValueError(msg)          # src/wfpt.pyx:397, src/cdfdif_wrapper.pyx:25
    if rc == ERR_RT_RANGE:
        raise AssertionError(msg)      # src/cdfdif_wrapper.pyx:21
    if rc == ERR_ARG:
        raise ValueError(msg)
    raise RuntimeError("hddm_b200: " + msg)


def as_rt_buffer(x, name="x"):
    """Reproduce the Cython buffer contract `np.ndarray[double, ndim=1]` (src/wfpt.pyx:32,54,...):
    only float64 1-D ndarrays are accepted (any stride); lists raise TypeError."""
    if not isinstance(x, np.ndarray):
        raise TypeError("Argument '%s' has incorrect type (expected numpy.ndarray, got %s)"
                        % (name, type(x).__name__))
    if x.ndim != 1:
        raise ValueError("Buffer has wrong number of dimensions (expected 1, got %d)" % x.ndim)
    if x.dtype != np.float64:
        raise ValueError("Buffer dtype mismatch, expected 'double' but got '%s'" % x.dtype)
    return np.ascontiguousarray(x)


def dptr(a):
    return a.ctypes.data_as(c_dp)
