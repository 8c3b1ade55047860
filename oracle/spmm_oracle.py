"""ctypes wrapper around oracle/spmm_oracle.c (TEST INFRASTRUCTURE; see oracle/__init__.py).

Restates /root/reference/utils.py:92-97 (``sgc_precompute``) with torch.spmm's CPU arithmetic
(sequential ascending-column fp32 FMA) written out in C.
"""
import ctypes
import os
import subprocess
import time

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle_spmm.so")
_lib = None


def build(force=False):
    """Compile oracle/spmm_oracle.c with the committed Makefile (gcc)."""
    src = os.path.join(_HERE, "spmm_oracle.c")
    if (not force and os.path.exists(_LIB_PATH)
            and os.path.getmtime(_LIB_PATH) >= os.path.getmtime(src)):
        return _LIB_PATH
    subprocess.check_call(["make", "-s", "-C", _HERE, "-B", "liboracle_spmm.so"])
    return _LIB_PATH


def _load():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        lib = ctypes.CDLL(_LIB_PATH)
        i64, p = ctypes.c_int64, ctypes.c_void_p
        lib.sgc_oracle_spmm_f32.argtypes = [i64, p, p, p, p, i64, p, i64, i64]
        lib.sgc_oracle_spmm_f32.restype = None
        lib.sgc_oracle_spmm_standardized_f32.argtypes = [i64, p, p, p, p, i64, p, p, p, i64, i64]
        lib.sgc_oracle_spmm_standardized_f32.restype = None
        lib.sgc_oracle_rsqrt_refined_array.argtypes = [i64, p, p]
        lib.sgc_oracle_rsqrt_refined_array.restype = None
        lib.sgc_oracle_col_stats.argtypes = [i64, i64, p, i64, p, p]
        lib.sgc_oracle_col_stats.restype = None
        _lib = lib
    return _lib


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _csr_args(indptr, indices, data):
    indptr = np.ascontiguousarray(indptr, dtype=np.int64)
    indices = np.ascontiguousarray(indices, dtype=np.int32)
    data = np.ascontiguousarray(data, dtype=np.float32)
    return indptr, indices, data


def spmm_hop(indptr, indices, data, x, mean=None, std=None):
    """One hop ``S @ x`` (fp32, sequential-FMA order).  With ``mean``/``std`` the gathered operand
    is standardised first, as utils.py:119 does before the reference's first hop."""
    lib = _load()
    indptr, indices, data = _csr_args(indptr, indices, data)
    x = np.ascontiguousarray(x, dtype=np.float32)
    n_rows = indptr.shape[0] - 1
    f = x.shape[1]
    y = np.empty((n_rows, f), dtype=np.float32)
    if mean is None:
        lib.sgc_oracle_spmm_f32(n_rows, _ptr(indptr), _ptr(indices), _ptr(data), _ptr(x), f,
                                _ptr(y), f, f)
    else:
        mean = np.ascontiguousarray(mean, dtype=np.float32)
        std = np.ascontiguousarray(std, dtype=np.float32)
        lib.sgc_oracle_spmm_standardized_f32(n_rows, _ptr(indptr), _ptr(indices), _ptr(data),
                                             _ptr(x), f, _ptr(mean), _ptr(std), _ptr(y), f, f)
    return y


def sgc_precompute_oracle(x, indptr, indices, data, degree):
    """utils.py:92-97: ``for i in range(degree): features = spmm(adj, features)``; returns
    ``(features, seconds)`` like the reference.  degree == 0 returns the input unchanged."""
    t = time.perf_counter()
    for _ in range(degree):
        x = spmm_hop(indptr, indices, data, x)
    return x, time.perf_counter() - t


def rsqrt_refined(r):
    lib = _load()
    r = np.ascontiguousarray(r, dtype=np.float64)
    out = np.empty_like(r)
    lib.sgc_oracle_rsqrt_refined_array(r.size, _ptr(r), _ptr(out))
    return out


def col_stats(x):
    lib = _load()
    x = np.ascontiguousarray(x, dtype=np.float32)
    n, f = x.shape
    mean = np.empty(f, dtype=np.float64)
    std = np.empty(f, dtype=np.float64)
    lib.sgc_oracle_col_stats(n, f, _ptr(x), f, _ptr(mean), _ptr(std))
    return mean, std
