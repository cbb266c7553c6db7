"""ORACLE — TEST INFRASTRUCTURE ONLY.

Rook-pivoted Bunch-Kaufman (LAPACK dsytrf_rook / dsytrs_rook / dsytri_rook), the
factorization behind the reference's `bunchkaufman!(A, true)` (src/linearalgebra/dense.jl:199-201).
scipy ships no Python wrapper for the rook routines, but the OpenBLAS it bundles exports them
(`scipy_dsytrf_rook_` ...): they are bound here with ctypes (Fortran calling convention:
every argument by reference, hidden string length last).
"""
from __future__ import annotations

import ctypes as C
import glob
import os

import numpy as np
import scipy

_lib = None


def _load():
    global _lib
    if _lib is None:
        pat = os.path.join(os.path.dirname(os.path.dirname(scipy.__file__)), "scipy.libs", "libscipy_openblas*.so")
        paths = sorted(glob.glob(pat))
        if not paths:
            raise RuntimeError("scipy's bundled OpenBLAS was not found: " + pat)
        _lib = C.CDLL(paths[0])
    return _lib


def _i(v):
    return C.byref(C.c_int(int(v)))


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def dsytrf_rook(A: np.ndarray, uplo: bytes = b"U"):
    """Rook Bunch-Kaufman of symmetric A from its `uplo` triangle.  Returns (ldu, ipiv, info), LAPACK
    conventions (ipiv 1-based, negative entries mark 2 x 2 pivots)."""
    lib = _load()
    a = np.array(A, dtype=np.float64, order="F")
    n = a.shape[0]
    ipiv = np.zeros(max(n, 1), dtype=np.int32)
    info = C.c_int(0)
    lwork = max(1, 64 * n)
    work = np.zeros(lwork)
    lib.scipy_dsytrf_rook_(C.c_char_p(uplo), _i(n), _dp(a), _i(max(n, 1)), ipiv.ctypes.data_as(C.POINTER(C.c_int)), _dp(work),
                           _i(lwork), C.byref(info), C.c_size_t(1))
    return a, ipiv[:n], int(info.value)


def dsytrs_rook(ldu: np.ndarray, ipiv: np.ndarray, B: np.ndarray, uplo: bytes = b"U") -> np.ndarray:
    lib = _load()
    b = np.array(B, dtype=np.float64, order="F")
    one_d = b.ndim == 1
    if one_d:
        b = b.reshape(-1, 1, order="F")
    n, nrhs = b.shape
    info = C.c_int(0)
    ip = np.ascontiguousarray(ipiv, dtype=np.int32)
    lib.scipy_dsytrs_rook_(C.c_char_p(uplo), _i(n), _i(nrhs), _dp(ldu), _i(max(n, 1)), ip.ctypes.data_as(C.POINTER(C.c_int)),
                           _dp(b), _i(max(n, 1)), C.byref(info), C.c_size_t(1))
    assert info.value == 0
    return b[:, 0] if one_d else b


def dsytri_rook(ldu: np.ndarray, ipiv: np.ndarray) -> np.ndarray:
    """Inverse from the factorization; returns the full symmetric matrix."""
    lib = _load()
    a = np.array(ldu, dtype=np.float64, order="F")
    n = a.shape[0]
    work = np.zeros(max(n, 1))
    info = C.c_int(0)
    ip = np.ascontiguousarray(ipiv, dtype=np.int32)
    lib.scipy_dsytri_rook_(C.c_char_p(b"U"), _i(n), _dp(a), _i(max(n, 1)), ip.ctypes.data_as(C.POINTER(C.c_int)), _dp(work),
                           C.byref(info), C.c_size_t(1))
    assert info.value == 0
    return np.triu(a) + np.triu(a, 1).T
