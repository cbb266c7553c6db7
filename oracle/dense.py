"""ORACLE — TEST INFRASTRUCTURE ONLY.

Dense factorization helpers, restating src/linearalgebra/dense.jl:
  posdef_fact_copy!  :226-249  Cholesky -> Bunch-Kaufman -> diagonal shift + BK
  increase_diag!     :100-107
  inv_fact!          :20-24    (potri)
  outer_prod!        :72-79    (syrk, upper)

The symmetric-indefinite fallback is rook-pivoted Bunch-Kaufman like the reference's
(`bunchkaufman!(A, true)` -> dsytrf_rook, dense.jl:199-201), through oracle/lapack_rook.py.
"""
from __future__ import annotations

import numpy as np
from scipy.linalg import lapack

from .lapack_rook import dsytrf_rook, dsytri_rook, dsytrs_rook

EPS = float(np.finfo(np.float64).eps)


def increase_diag(A: np.ndarray) -> np.ndarray:
    diag_pert = 1 + 1e-5
    diag_min = 1000 * EPS
    idx = np.diag_indices_from(A)
    A[idx] = diag_pert * np.maximum(A[idx], diag_min)
    return A


class Fact:
    """kind: 'chol' (upper factor in .factor) or 'bk' (dsytrf_rook output)."""

    def __init__(self, kind, factor, ipiv, success):
        self.kind, self.factor, self.ipiv, self.success = kind, factor, ipiv, success
        self.shifted = False

    @property
    def kind_code(self) -> int:
        """0 Cholesky, 1 Bunch-Kaufman, 2 diagonal shift + Bunch-Kaufman, -1 failed (the C-ABI's fact_kind)."""
        if not self.success:
            return -1
        return 0 if self.kind == "chol" else (2 if self.shifted else 1)

    def solve(self, arr: np.ndarray) -> np.ndarray:
        b = np.asfortranarray(arr, dtype=np.float64)
        if self.kind == "chol":
            x, info = lapack.dpotrs(self.factor, b, lower=0)
        else:
            x = dsytrs_rook(self.factor, self.ipiv, b)
        return x

    def inverse(self) -> np.ndarray:
        if self.kind == "chol":
            inv, info = lapack.dpotri(self.factor, lower=0)
        else:
            return dsytri_rook(self.factor, self.ipiv)
        return np.triu(inv) + np.triu(inv, 1).T


def _bk(mat: np.ndarray) -> Fact:
    ldu, ipiv, info = dsytrf_rook(mat)
    return Fact("bk", ldu, ipiv, info == 0)


def posdef_fact_copy(mat: np.ndarray, try_shift: bool = True) -> Fact:
    """mat: symmetric, upper triangle referenced (Symmetric(:U))."""
    c, info = lapack.dpotrf(np.array(mat, order="F"), lower=0, clean=0)
    if info == 0:
        return Fact("chol", c, None, True)
    fact = _bk(np.array(mat, order="F"))
    if try_shift and not fact.success:
        fact = _bk(increase_diag(np.array(mat, order="F")))
        fact.shifted = True
    return fact


def outer_prod_upper(A: np.ndarray) -> np.ndarray:
    """A'A, upper triangle valid (dsyrk 'U','T')."""
    from scipy.linalg import blas

    return blas.dsyrk(1.0, np.asfortranarray(A), trans=1, lower=0)
