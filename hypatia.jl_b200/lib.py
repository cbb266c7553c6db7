"""ctypes binding of the C-ABI declared in include/hypatia_b200.h.

There is NO CPU fallback: if the shared library is missing or no CUDA device
is present, loading / context creation raises.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# HYB_LIB: an alternative build of the same library (profiling builds of tools/); never a different backend
LIB_PATH = os.environ.get("HYB_LIB") or os.path.join(_HERE, "libhypatia_b200.so")

c_double_p = C.POINTER(C.c_double)
c_int_p = C.POINTER(C.c_int)


class ConeDesc(C.Structure):
    """struct hyb_cone_desc"""

    _fields_ = [
        ("type", C.c_int32),
        ("use_dual_barrier", C.c_int32),
        ("offset", C.c_int64),
        ("dim", C.c_int64),
        ("side_or_R", C.c_int64),
        ("U", C.c_int64),
        ("nPs", C.c_int32),
        ("_pad", C.c_int32),
        ("Ps", C.POINTER(c_double_p)),
        ("Ls", C.POINTER(C.c_int64)),
    ]


# name -> (restype, argtypes); kept in sync with include/hypatia_b200.h (tests check the header)
SIGNATURES = {
    "hyb_ctx_create": (C.c_void_p, [C.c_int]),
    "hyb_ctx_destroy": (None, [C.c_void_p]),
    "hyb_last_error": (C.c_char_p, [C.c_void_p]),
    "hyb_launch_count": (C.c_int64, []),
    "hyb_model_load": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, c_double_p, C.c_int64, C.POINTER(ConeDesc)]),
    "hyb_cones_load_point": (C.c_int, [C.c_void_p, c_double_p, c_double_p, C.c_double]),
    "hyb_cones_is_feas": (C.c_int, [C.c_void_p, c_int_p, c_int_p]),
    "hyb_cones_check": (C.c_int, [C.c_void_p, C.c_double, C.c_int, C.c_double, C.c_double, c_int_p, c_double_p]),
    "hyb_cones_grad": (C.c_int, [C.c_void_p, c_double_p]),
    "hyb_cones_hess_prod": (C.c_int, [C.c_void_p, C.c_int, c_double_p, c_double_p, C.c_int64]),
    "hyb_cones_dder3": (C.c_int, [C.c_void_p, c_double_p, c_double_p]),
    "hyb_schur_update": (C.c_int, [C.c_void_p, c_double_p, c_double_p, c_int_p]),
    "hyb_schur_assemble": (C.c_int, [C.c_void_p, c_double_p]),
    "hyb_schur_factor": (C.c_int, [C.c_void_p, c_double_p, c_int_p]),
    "hyb_device_buffer": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "hyb_peer_root_create": (C.c_int, [C.c_void_p, C.c_int, C.c_char_p]),
    "hyb_peer_attach": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int, C.c_int]),
    "hyb_peer_set_slot": (C.c_int, [C.c_void_p, C.c_void_p]),
    "hyb_schur_assemble_peer": (C.c_int, [C.c_void_p, c_double_p]),
    "hyb_peer_root_sum": (C.c_int, [C.c_void_p, c_double_p]),
    "hyb_schur_solve": (C.c_int, [C.c_void_p, c_double_p, c_double_p, C.c_int64]),
    "hyb_schur_get": (C.c_int, [C.c_void_p, c_double_p]),
    "hyb_G_mul": (C.c_int, [C.c_void_p, C.c_char, c_double_p, c_double_p]),
    "hyb_solve_subsystem3": (C.c_int, [C.c_void_p, c_double_p, c_double_p, c_double_p, c_double_p]),
    "hyb_dirs_setup": (C.c_int, [C.c_void_p, c_double_p, c_double_p]),
    "hyb_get_directions": (
        C.c_int,
        [C.c_void_p, C.c_int, c_double_p, c_double_p, C.c_int64, C.c_int, C.c_double, C.c_double, C.c_double, C.c_int,
         C.c_double, c_int_p, c_double_p, c_double_p],
    ),
    "hyb_search_setup": (C.c_int, [C.c_void_p, c_double_p]),
    "hyb_dirs_keep": (C.c_int, [C.c_void_p, C.c_int, C.c_int]),
    "hyb_search_load_point": (C.c_int, [C.c_void_p, c_double_p, c_double_p]),
    "hyb_search_gates": (C.c_int, [C.c_void_p, c_double_p, c_double_p, c_double_p]),
    "hyb_search_check": (C.c_int, [C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_int, C.c_double, c_int_p, c_double_p]),
    "hyb_search_alpha_supported": (C.c_int, [C.c_void_p]),
    "hyb_search_alpha": (
        C.c_int,
        [C.c_void_p, C.c_int, c_double_p, c_double_p, C.c_double, C.c_double, C.c_double, C.c_int, c_int_p, c_double_p, c_double_p],
    ),
    "hyb_step_point_set": (C.c_int, [C.c_void_p, c_double_p, c_double_p, c_double_p]),
    "hyb_step_point_get": (C.c_int, [C.c_void_p, c_double_p, c_double_p, c_double_p]),
    "hyb_step_convergence": (C.c_int, [C.c_void_p, C.c_double, c_double_p]),
    "hyb_step_rhs": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double]),
    "hyb_step_directions": (
        C.c_int,
        [C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, C.c_int, C.c_double, c_int_p, c_double_p, c_double_p],
    ),
    "hyb_step_point_update": (C.c_int, [C.c_void_p, c_double_p, c_double_p]),
    "hyb_geqp3": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, c_double_p, C.c_int64, c_int_p, c_double_p]),
    "hyb_ormqr_right": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, c_double_p, C.c_int64, c_double_p, c_double_p, C.c_int64]),
    "hyb_outer_prod": (
        C.c_int,
        [C.c_void_p, C.c_int64, C.c_int64, C.c_double, c_double_p, C.c_int64, C.c_double, c_double_p, C.c_int64],
    ),
    "hyb_gemm_tn_host": (
        C.c_int,
        [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_double, c_double_p, C.c_int64, c_double_p, C.c_int64,
         C.c_double, c_double_p, C.c_int64],
    ),
    "hyb_posdef_solve": (
        C.c_int,
        [C.c_void_p, C.c_int64, c_double_p, C.c_int64, c_double_p, C.c_int64, C.c_int64, c_int_p],
    ),
    "hyb_multi_create": (C.c_void_p, [C.c_int, c_int_p]),
    "hyb_multi_destroy": (None, [C.c_void_p]),
    "hyb_multi_root": (C.c_void_p, [C.c_void_p]),
    "hyb_multi_model_load": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, c_double_p, C.c_int64, C.POINTER(ConeDesc)]),
    "hyb_multi_plan": (C.c_int, [C.c_int64, C.c_int64, C.POINTER(ConeDesc), C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_int64),
                                 C.POINTER(C.c_int64)]),
    "hyb_multi_shard_rows": (C.c_int, [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "hyb_multi_schur_update": (C.c_int, [C.c_void_p, c_double_p, c_double_p, c_double_p, c_int_p]),
    "hyb_ctx_set_assemble_only": (C.c_int, [C.c_void_p, C.c_int]),
    "hyb_symm_solve": (
        C.c_int,
        [C.c_void_p, C.c_int64, c_double_p, C.c_int64, c_int_p, c_double_p, C.c_int64, C.c_int64, c_int_p],
    ),
    "hyb_schur_set": (C.c_int, [C.c_void_p, c_double_p]),
    "hyb_symindef_load": (C.c_int, [C.c_void_p]),
    "hyb_symindef_update": (C.c_int, [C.c_void_p, c_double_p, c_int_p]),
    "hyb_symindef_solve": (C.c_int, [C.c_void_p, c_double_p, c_double_p, C.c_int64]),
    "hyb_symindef_get": (C.c_int, [C.c_void_p, c_double_p]),
    "hyb_gemv_n_noalloc": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "hyb_stats": (C.c_int, [C.c_void_p, C.c_int, c_double_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "hyb_last_schur_timings": (C.c_int, [C.c_void_p, c_double_p, c_double_p, c_double_p]),
    "hyb_bench_gemm_tn": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int, C.c_int, c_double_p]),
    "hyb_bench_potrf": (C.c_int, [C.c_void_p, C.c_int64, C.c_int, c_double_p, c_double_p]),
}

_lib: Optional[C.CDLL] = None


def load() -> C.CDLL:
    """dlopen the CUDA library; raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with __graft_entry__.build() (make -C hypatia.jl_b200/csrc). "
            "hypatia_b200 has no CPU fallback."
        )
    lib = C.CDLL(LIB_PATH)
    for name, (restype, argtypes) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    return lib


class HybError(RuntimeError):
    pass


def dptr(a: np.ndarray):
    return a.ctypes.data_as(c_double_p)


def f64(a, order="F") -> np.ndarray:
    return np.require(a, dtype=np.float64, requirements=["F_CONTIGUOUS" if order == "F" else "C_CONTIGUOUS", "ALIGNED"])


class Context:
    """Owns one hyb_ctx (one GPU)."""

    def __init__(self, device: int = 0):
        self.lib = load()
        self.ptr = self.lib.hyb_ctx_create(int(device))
        if not self.ptr:
            msg = self.lib.hyb_last_error(None)
            raise HybError("hyb_ctx_create failed: " + (msg.decode() if msg else "unknown error"))

    def check(self, rc: int, what: str) -> int:
        if rc < 0:
            msg = self.lib.hyb_last_error(self.ptr)
            raise HybError(f"{what} failed ({rc}): " + (msg.decode() if msg else ""))
        return rc

    def close(self):
        if getattr(self, "ptr", None):
            if not getattr(self, "_borrowed", False):  # a MultiContext's root belongs to the hyb_multi
                self.lib.hyb_ctx_destroy(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def launch_count(self) -> int:
        return int(self.lib.hyb_launch_count())

    def stats(self, reset: bool = False):
        """(device_ms, h2d_bytes, d2h_bytes) accumulated since the last reset."""
        ms, a, b = C.c_double(0), C.c_int64(0), C.c_int64(0)
        self.check(self.lib.hyb_stats(self.ptr, int(reset), C.byref(ms), C.byref(a), C.byref(b)), "hyb_stats")
        return ms.value, a.value, b.value

    def last_schur_timings(self):
        a, b, c_ = C.c_double(0), C.c_double(0), C.c_double(0)
        self.check(self.lib.hyb_last_schur_timings(self.ptr, C.byref(a), C.byref(b), C.byref(c_)), "hyb_last_schur_timings")
        return a.value, b.value, c_.value

    # ---- measurement ----------------------------------------------------------
    def bench_gemm_tn(self, M, N, K, upper=False, reps=5) -> float:
        ms = C.c_double(0)
        self.check(self.lib.hyb_bench_gemm_tn(self.ptr, M, N, K, int(upper), reps, C.byref(ms)), "hyb_bench_gemm_tn")
        return ms.value

    def bench_potrf(self, N, reps=3):
        mf, ms = C.c_double(0), C.c_double(0)
        self.check(self.lib.hyb_bench_potrf(self.ptr, N, reps, C.byref(mf), C.byref(ms)), "hyb_bench_potrf")
        return mf.value, ms.value

    # ---- preprocessing -------------------------------------------------------
    def geqp3(self, M: np.ndarray, overwrite_a: bool = False):
        """Column-pivoted QR on the GPU; returns (qr, tau, piv) like
        scipy.linalg.qr(M, mode="raw", pivoting=True).  ``overwrite_a``: factor in M's own storage when M is a
        Fortran-ordered float64 array (no second host copy of a multi-gigabyte matrix)."""
        if overwrite_a and isinstance(M, np.ndarray) and M.dtype == np.float64 and M.flags.f_contiguous and M.flags.writeable:
            A = M
        else:
            A = np.array(M, dtype=np.float64, order="F")
        m, n = A.shape
        k = min(m, n)
        tau = np.zeros(k)
        piv = np.zeros(max(n, 1), dtype=np.int32)
        self.check(
            self.lib.hyb_geqp3(self.ptr, m, n, dptr(A), max(m, 1), piv.ctypes.data_as(c_int_p), dptr(tau) if k else None),
            "hyb_geqp3",
        )
        return A, tau, piv[:n].astype(np.int64)

    def ormqr_right(self, refl: np.ndarray, tau: np.ndarray, M: np.ndarray, overwrite: bool = False) -> np.ndarray:
        """M Q for the Q of a factored QR (reflectors below the diagonal of ``refl``, n x k): dormqr('R', 'N') on the GPU.
        ``overwrite``: work in M's own storage when it is a Fortran-ordered float64 array."""
        V, t = f64(refl), f64(tau)
        if overwrite and isinstance(M, np.ndarray) and M.dtype == np.float64 and M.flags.f_contiguous and M.flags.writeable:
            out = M
        else:
            out = np.array(M, dtype=np.float64, order="F")
        m, n = out.shape
        assert V.shape[0] == n and V.shape[1] >= t.size
        self.check(self.lib.hyb_ormqr_right(self.ptr, m, n, t.size, dptr(V), max(n, 1), dptr(t), dptr(out), max(m, 1)),
                   "hyb_ormqr_right")
        return out

    # ---- dense helpers -----------------------------------------------------
    def outer_prod(self, A: np.ndarray, C_: Optional[np.ndarray] = None, alpha=1.0, beta=0.0) -> np.ndarray:
        A = f64(A)
        K, N = A.shape
        out = np.zeros((N, N), order="F") if C_ is None else f64(C_).copy(order="F")
        self.check(self.lib.hyb_outer_prod(self.ptr, K, N, alpha, dptr(A), max(K, 1), beta, dptr(out), max(N, 1)), "hyb_outer_prod")
        return out

    def gemm_tn(self, A: np.ndarray, B: np.ndarray, C_: Optional[np.ndarray] = None, alpha=1.0, beta=0.0) -> np.ndarray:
        A, B = f64(A), f64(B)
        K, M = A.shape
        K2, N = B.shape
        assert K == K2
        out = np.zeros((M, N), order="F") if C_ is None else f64(C_).copy(order="F")
        self.check(
            self.lib.hyb_gemm_tn_host(self.ptr, M, N, K, alpha, dptr(A), max(K, 1), dptr(B), max(K, 1), beta, dptr(out), max(M, 1)),
            "hyb_gemm_tn_host",
        )
        return out

    def symm_solve(self, A: np.ndarray, B: Optional[np.ndarray] = None):
        """Rook-pivoted Bunch-Kaufman on the GPU: returns (ldu, ipiv, X, info) in LAPACK dsytrf_rook's layout."""
        A = f64(A).copy(order="F")
        N = A.shape[0]
        ipiv = np.zeros(max(N, 1), dtype=np.int32)
        info = C.c_int(0)
        if B is None:
            B2, nrhs, one_d = np.zeros((max(N, 1), 1), order="F"), 0, False
        else:
            B2 = f64(B).copy(order="F")
            one_d = B2.ndim == 1
            if one_d:
                B2 = B2.reshape(-1, 1, order="F")
            nrhs = B2.shape[1]
        self.check(
            self.lib.hyb_symm_solve(self.ptr, N, dptr(A), max(N, 1), ipiv.ctypes.data_as(c_int_p), dptr(B2), max(N, 1), nrhs,
                                    C.byref(info)),
            "hyb_symm_solve",
        )
        X = None if B is None else (B2[:, 0] if one_d else B2)
        return A, ipiv[:N], X, info.value

    def posdef_solve(self, A: np.ndarray, B: np.ndarray):
        """Returns (U, X, info): upper Cholesky factor, solution of A X = B, LAPACK info."""
        A = f64(A).copy(order="F")
        B2 = f64(B).copy(order="F")
        one_d = B2.ndim == 1
        if one_d:
            B2 = B2.reshape(-1, 1, order="F")
        N = A.shape[0]
        info = C.c_int(0)
        self.check(
            self.lib.hyb_posdef_solve(self.ptr, N, dptr(A), max(N, 1), dptr(B2), max(N, 1), B2.shape[1], C.byref(info)),
            "hyb_posdef_solve",
        )
        return np.triu(A), (B2[:, 0] if one_d else B2), info.value


class MultiContext:
    """Owns one hyb_multi: several GPUs driven from this process through the C-ABI alone.  ``root`` is the
    ordinary Context of devices[0] (the factorizing GPU)."""

    def __init__(self, devices):
        self.lib = load()
        self.devices = [int(d) for d in devices]
        arr = (C.c_int * len(self.devices))(*self.devices)
        self.ptr = self.lib.hyb_multi_create(len(self.devices), arr)
        if not self.ptr:
            msg = self.lib.hyb_last_error(None)
            raise HybError("hyb_multi_create failed: " + (msg.decode() if msg else "unknown error"))
        self.root = Context.__new__(Context)
        self.root.lib = self.lib
        self.root.ptr = self.lib.hyb_multi_root(self.ptr)
        self.root._borrowed = True

    def check(self, rc: int, what: str) -> int:
        return self.root.check(rc, what)

    def shard_rows(self):
        n = len(self.devices)
        lo, hi = (C.c_int64 * n)(), (C.c_int64 * n)()
        self.check(self.lib.hyb_multi_shard_rows(self.ptr, lo, hi), "hyb_multi_shard_rows")
        return [(int(a), int(b)) for a, b in zip(lo, hi)]

    def close(self):
        if getattr(self, "ptr", None):
            self.root.ptr = None
            self.lib.hyb_multi_destroy(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
