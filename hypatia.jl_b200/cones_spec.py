"""Cone descriptors and the integer index maps of the hot path.

A descriptor says WHAT a cone is (type, dimension, parameters, barrier side,
initial point); the oracles live in a backend: ``oracle/cones.py`` (CPU
checker) or the CUDA library (``gpu.py``).  Index maps follow the reference
bit-exactly (0-based here, 1-based there):

  svec_length / svec_side / svec_idx / block_idxs   src/Cones/arrayutilities.jl:75,93-97,119-124,132
  smat_to_svec / svec_to_smat / scale_svec(_incr)   src/Cones/arrayutilities.jl:140-176,183-197,236-251
  build_cone_idxs                                   src/Models/Models.jl:63-73
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import List, Sequence

import numpy as np

# type codes shared with include/hypatia_b200.h (hyb_cone_type)
NONNEGATIVE = 0
EPINORMEUCL = 1
EPIPERSQUARE = 2
POSSEMIDEFTRI = 3
WSOSINTERPNONNEGATIVE = 4
WSOSINTERPPOSSEMIDEFTRI = 5
EPINORMINF = 6
DOUBLYNONNEGATIVETRI = 7
POSSEMIDEFTRISPARSE = 8
WSOSINTERPEPINORMEUCL = 9
WSOSINTERPEPINORMONE = 10

RT2 = math.sqrt(2.0)


# ----------------------------------------------------------------------------
# index maps (arrayutilities.jl)
# ----------------------------------------------------------------------------
def svec_length(side: int) -> int:
    """arrayutilities.jl:75"""
    return (side * (side + 1)) // 2


def svec_side(length: int) -> int:
    """arrayutilities.jl:93-97 (isqrt-based inverse, asserted exact)."""
    side = math.isqrt(1 + 8 * length) // 2
    assert side * (side + 1) == 2 * length, "not a triangular number"
    return side


def svec_idx(row: int, col: int) -> int:
    """1-based svec index of (row, col), arrayutilities.jl:119-124."""
    if row < col:
        row, col = col, row
    return ((row - 1) * row) // 2 + col


def block_idxs(incr: int, block: int) -> range:
    """1-based block -> 0-based python range, arrayutilities.jl:132."""
    return range(incr * (block - 1), incr * block)


def smat_to_svec(mat: np.ndarray, rt2: float = RT2) -> np.ndarray:
    """Upper triangle, column by column, off-diagonals * rt2 (arrayutilities.jl:183-197)."""
    m = mat.shape[0]
    out = np.empty(svec_length(m), dtype=mat.dtype)
    k = 0
    for j in range(m):
        out[k : k + j] = mat[:j, j] * rt2
        out[k + j] = mat[j, j]
        k += j + 1
    return out


def svec_to_smat(vec: np.ndarray, rt2: float = RT2, full: bool = True) -> np.ndarray:
    """Inverse of smat_to_svec (arrayutilities.jl:236-251); ``full`` mirrors copytri!."""
    m = svec_side(vec.shape[0])
    mat = np.zeros((m, m), dtype=vec.dtype)
    inv = 1.0 / rt2
    k = 0
    for j in range(m):
        mat[:j, j] = vec[k : k + j] * inv
        mat[j, j] = vec[k + j]
        k += j + 1
    if full:
        mat = mat + np.triu(mat, 1).T
    return mat


def scale_svec(arr: np.ndarray, scal: float) -> np.ndarray:
    """In place: off-diagonal rows *= scal (arrayutilities.jl:140-152)."""
    n = arr.shape[0]
    k = 0
    for i in range(1, svec_side(n) + 1):
        arr[k : k + i - 1] *= scal
        k += i
    assert k == n
    return arr


def scale_svec_incr(arr: np.ndarray, scal: float, incr: int) -> np.ndarray:
    """Blocked variant (arrayutilities.jl:160-176)."""
    n = arr.shape[0]
    d, r = divmod(n, incr)
    assert r == 0
    k = 0
    for i in range(1, svec_side(d) + 1):
        arr[k : k + incr * (i - 1)] *= scal
        k += incr * i
    assert k == n
    return arr


# ----------------------------------------------------------------------------
# descriptors
# ----------------------------------------------------------------------------
@dataclass
class ConeSpec:
    dim: int
    type_code: int = field(init=False, default=-1)
    use_dual_barrier: bool = field(init=False, default=False)

    @property
    def nu(self) -> float:
        raise NotImplementedError

    def initial_point(self) -> np.ndarray:
        raise NotImplementedError

    # scalar per cone for rescale_data unless per-row (process.jl:41-52)
    per_row_h_scale = False


@dataclass
class Nonnegative(ConeSpec):
    """src/Cones/nonnegative.jl:16-53"""

    per_row_h_scale = True

    def __post_init__(self):
        assert self.dim >= 1
        self.type_code = NONNEGATIVE

    @property
    def nu(self):
        return float(self.dim)

    def initial_point(self):
        return np.ones(self.dim)


@dataclass
class EpiNormEucl(ConeSpec):
    """src/Cones/epinormeucl.jl:16-59"""

    def __post_init__(self):
        assert self.dim >= 2
        self.type_code = EPINORMEUCL

    @property
    def nu(self):
        return 2.0

    def initial_point(self):
        arr = np.zeros(self.dim)
        arr[0] = math.sqrt(2.0)
        return arr


@dataclass
class EpiNormInf(ConeSpec):
    """src/Cones/epinorminf.jl:16-98 (real case): epigraph of the infinity norm; with ``use_dual`` the
    dual barrier is used, i.e. the cone is the epigraph of the 1-norm."""

    use_dual: bool = False

    def __post_init__(self):
        assert self.dim >= 2
        self.type_code = EPINORMINF
        self.use_dual_barrier = bool(self.use_dual)

    @property
    def nu(self):
        return float(self.dim)  # 1 + d with d = dim - 1 (epinorminf.jl:90)

    def initial_point(self):
        arr = np.zeros(self.dim)
        arr[0] = math.sqrt(self.nu)
        return arr


@dataclass
class EpiPerSquare(ConeSpec):
    """src/Cones/epipersquare.jl:17-70"""

    def __post_init__(self):
        assert self.dim >= 3
        self.type_code = EPIPERSQUARE

    @property
    def nu(self):
        return 2.0

    def initial_point(self):
        arr = np.zeros(self.dim)
        arr[:2] = 1.0
        return arr


@dataclass
class PosSemidefTri(ConeSpec):
    """Real svec PSD cone, src/Cones/possemideftri.jl:17-87"""

    side: int = field(init=False, default=0)

    def __post_init__(self):
        assert self.dim >= 1
        self.type_code = POSSEMIDEFTRI
        self.side = svec_side(self.dim)

    @property
    def nu(self):
        return float(self.side)

    def initial_point(self):
        arr = np.zeros(self.dim)
        k = 0
        for i in range(1, self.side + 1):  # possemideftri.jl:78-87
            arr[k] = 1.0
            k += i + 1
        return arr


@dataclass
class DoublyNonnegativeTri(ConeSpec):
    """Real svec doubly nonnegative cone (PSD and entrywise nonnegative), src/Cones/doublynonnegativetri.jl:17-59."""

    use_dual: bool = False
    side: int = field(init=False, default=0)

    def __post_init__(self):
        assert self.dim >= 1
        self.type_code = DOUBLYNONNEGATIVETRI
        self.side = svec_side(self.dim)
        self.use_dual_barrier = bool(self.use_dual)

    @property
    def nu(self):
        return float(self.dim)  # doublynonnegativetri.jl:79

    def offdiag_idxs(self) -> np.ndarray:
        """0-based svec positions of the strictly upper-triangular entries (doublynonnegativetri.jl:55)."""
        return np.array([svec_length(i - 1) + j for i in range(2, self.side + 1) for j in range(i - 1)], dtype=np.int64)

    def initial_point(self):  # doublynonnegativetri.jl:81-132
        side, d = self.side, float(self.dim)
        if side == 1:
            on_diag = off_diag = 1.0
        elif side == 2:
            on_diag, off_diag = math.sqrt(5.0) / 2.0, 1.0 / RT2
        else:
            n = float(side)
            # ascending coefficients in the reference; numpy.roots wants them descending
            p1 = [-n - 1, 0, n * n + n + 7, 0, -2 * n * n - 8, 0, n * n]
            on_diag, off_diag = n + 1, 1.0
            rtol = math.sqrt(np.finfo(np.float64).eps)
            close = lambda a, b: abs(a - b) <= rtol * max(abs(a), abs(b))  # noqa: E731  (isapprox)
            for r in np.roots(p1[::-1]):
                offd = float(np.real(r))
                if offd > 0:
                    temp = d - (d - n) * offd * offd
                    if temp > rtol:
                        ond = math.sqrt(temp / n)
                        denom = ond * ond + (n - 2) / RT2 * ond * offd - (n - 1) * offd * offd / 2
                        if close(ond * RT2 + (n - 2) * offd, ond * denom * RT2) and close(denom, offd * offd * (denom + 1)):
                            on_diag, off_diag = ond, offd
                            break
        arr = np.full(self.dim, off_diag)
        k = 0
        for i in range(1, side + 1):
            arr[k] = on_diag
            k += i + 1
        return arr


@dataclass
class PosSemidefTriSparse(ConeSpec):
    """Real sparse PSD cone, dense implementation: src/Cones/possemideftrisparse/possemideftrisparse.jl:17-117,
    denseimpl.jl.  ``rows`` / ``cols`` are the 1-based lower-triangle positions of the pattern (every diagonal entry
    present), in the order of the point's coordinates; off-diagonal coordinates carry the sqrt2 scaling."""

    side: int = 0
    rows: Sequence[int] = ()
    cols: Sequence[int] = ()
    use_dual: bool = False

    def __post_init__(self):
        self.type_code = POSSEMIDEFTRISPARSE
        self.rows0 = np.asarray(self.rows, dtype=np.int64) - 1
        self.cols0 = np.asarray(self.cols, dtype=np.int64) - 1
        assert self.rows0.size == self.cols0.size == self.dim >= 1
        assert np.all(self.cols0 <= self.rows0) and np.all(self.rows0 < self.side) and np.all(self.cols0 >= 0)
        diag = self.rows0[self.rows0 == self.cols0]
        assert diag.size == self.side == np.unique(diag).size, "every diagonal entry exactly once"
        self.use_dual_barrier = bool(self.use_dual)

    @property
    def nu(self):
        return float(self.side)

    def initial_point(self):  # possemideftrisparse.jl:103-117
        return (self.rows0 == self.cols0).astype(np.float64)


def psd_sparse(side: int, rows, cols, use_dual: bool = False) -> PosSemidefTriSparse:
    return PosSemidefTriSparse(dim=len(rows), side=side, rows=list(rows), cols=list(cols), use_dual=use_dual)


@dataclass
class WSOSInterpNonnegative(ConeSpec):
    """src/Cones/wsosinterpnonnegative.jl:17-97; dual barrier unless use_dual."""

    Ps: Sequence[np.ndarray] = ()
    use_dual: bool = False

    def __post_init__(self):
        self.type_code = WSOSINTERPNONNEGATIVE
        self.Ps = [np.asfortranarray(P, dtype=np.float64) for P in self.Ps]
        for P in self.Ps:
            assert P.shape[0] == self.dim
        self.use_dual_barrier = not self.use_dual

    @property
    def nu(self):
        return float(sum(P.shape[1] for P in self.Ps))

    def initial_point(self):
        return np.ones(self.dim)


@dataclass
class WSOSInterpPosSemidefTri(ConeSpec):
    """src/Cones/wsosinterppossemideftri.jl:17-123; dim = U*R(R+1)/2."""

    R: int = 1
    U: int = 0
    Ps: Sequence[np.ndarray] = ()
    use_dual: bool = False

    def __post_init__(self):
        self.type_code = WSOSINTERPPOSSEMIDEFTRI
        self.Ps = [np.asfortranarray(P, dtype=np.float64) for P in self.Ps]
        for P in self.Ps:
            assert P.shape[0] == self.U
        assert self.dim == self.U * svec_length(self.R)
        self.use_dual_barrier = not self.use_dual

    @property
    def nu(self):
        return float(self.R * sum(P.shape[1] for P in self.Ps))

    def initial_point(self):
        arr = np.zeros(self.dim)
        block = 1
        for i in range(1, self.R + 1):  # wsosinterppossemideftri.jl:115-123
            arr[block_idxs(self.U, block).start : block_idxs(self.U, block).stop] = 1.0
            block += i + 1
        return arr


@dataclass
class WSOSInterpEpiNormEucl(ConeSpec):
    """src/Cones/wsosinterpepinormeucl.jl:17-88: polynomial epigraph of the Euclidean norm of R - 1 polynomials in
    the interpolant basis; dim = U * R; dual barrier unless use_dual."""

    R: int = 2
    U: int = 0
    Ps: Sequence[np.ndarray] = ()
    use_dual: bool = False

    def __post_init__(self):
        self.type_code = WSOSINTERPEPINORMEUCL
        self.Ps = [np.asfortranarray(P, dtype=np.float64) for P in self.Ps]
        for P in self.Ps:
            assert P.shape[0] == self.U
        assert self.R >= 2 and self.dim == self.U * self.R
        self.use_dual_barrier = not self.use_dual

    @property
    def nu(self):
        return float(2 * sum(P.shape[1] for P in self.Ps))

    def initial_point(self):  # wsosinterpepinormeucl.jl:125-129
        arr = np.zeros(self.dim)
        arr[: self.U] = 1.0
        return arr


def wsos_epinormeucl(R: int, U: int, Ps, use_dual: bool = False) -> WSOSInterpEpiNormEucl:
    return WSOSInterpEpiNormEucl(dim=U * R, R=R, U=U, Ps=Ps, use_dual=use_dual)


@dataclass
class WSOSInterpEpiNormOne(ConeSpec):
    """src/Cones/wsosinterpepinormone.jl:17-99: polynomial epigraph of the 1-norm of R - 1 polynomials in the
    interpolant basis; dim = U * R; dual barrier unless use_dual."""

    R: int = 2
    U: int = 0
    Ps: Sequence[np.ndarray] = ()
    use_dual: bool = False

    def __post_init__(self):
        self.type_code = WSOSINTERPEPINORMONE
        self.Ps = [np.asfortranarray(P, dtype=np.float64) for P in self.Ps]
        for P in self.Ps:
            assert P.shape[0] == self.U
        assert self.R >= 2 and self.dim == self.U * self.R
        self.use_dual_barrier = not self.use_dual

    @property
    def nu(self):
        return float(self.R * sum(P.shape[1] for P in self.Ps))

    def initial_point(self):  # wsosinterpepinormone.jl:155-159
        arr = np.zeros(self.dim)
        arr[: self.U] = 1.0
        return arr


def wsos_epinormone(R: int, U: int, Ps, use_dual: bool = False) -> WSOSInterpEpiNormOne:
    return WSOSInterpEpiNormOne(dim=U * R, R=R, U=U, Ps=Ps, use_dual=use_dual)


def wsos_psd(R: int, U: int, Ps, use_dual: bool = False) -> WSOSInterpPosSemidefTri:
    return WSOSInterpPosSemidefTri(dim=U * svec_length(R), R=R, U=U, Ps=Ps, use_dual=use_dual)


def build_cone_idxs(q: int, cones: List[ConeSpec]) -> np.ndarray:
    """Offsets array of length K+1 (cone k owns [off[k], off[k+1])); Models.jl:63-73."""
    offs = np.zeros(len(cones) + 1, dtype=np.int64)
    for k, cone in enumerate(cones):
        offs[k + 1] = offs[k] + cone.dim
    assert offs[-1] == q, "sum of cone dimensions must equal q"
    return offs
