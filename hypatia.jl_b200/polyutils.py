"""WSOS interpolation data (setup-time only; host numpy).

Restates the part of src/PolyUtils/realinterp.jl the in-scope instances need:
  interpolate            :18-54   (box domains with n <= 2 take the deterministic
                                   Chebyshev / Padua route; free domains sample)
  interp_sample          :59-72
  interp_box             :74-126
  cheb2_pts / chebyshev  :128-165
  cheb2_data             :167-198 (points and P0 only)
  padua_data             :200-273 (points and P0 only)
  approxfekete_data      :277-305
  make_wsos_arrays / choose_interp_pts / vandermonde   :307-385
and realdomains.jl:41-63 (FreeDomain), :65-103 (BoxDomain).
Within one total degree the basis columns follow descending-lex exponent order
(the order of Combinatorics.multiexponents); any within-degree order spans the
same cone.
"""
from __future__ import annotations

import itertools
import math
from typing import List, Optional, Sequence, Tuple

import numpy as np
import scipy.linalg as sla


class FreeDomain:
    def __init__(self, n: int):
        self.n = n

    def dimension(self):
        return self.n

    def degree(self):
        return 0

    def weights(self, pts):
        return []

    def sample(self, npts, rng):
        return BoxDomain(-np.ones(self.n), np.ones(self.n)).sample(npts, rng)


class BoxDomain:
    def __init__(self, l: Sequence[float], u: Sequence[float]):
        self.l = np.asarray(l, dtype=np.float64)
        self.u = np.asarray(u, dtype=np.float64)
        assert self.l.size == self.u.size

    def dimension(self):
        return self.l.size

    def degree(self):
        return 2

    def weights(self, pts):
        return [(pts[:, i] - self.l[i]) * (self.u[i] - pts[:, i]) for i in range(pts.shape[1])]

    def sample(self, npts, rng):
        pts = rng.random((npts, self.dimension())) - 0.5
        return pts * (self.u - self.l)[None, :] + 0.5 * (self.u + self.l)[None, :]


def get_L(n: int, d: int) -> int:
    return math.comb(n + d, n)


def get_U(n: int, d: int) -> int:
    return math.comb(n + 2 * d, n)


def _multiexponents(n: int, t: int):
    """All exponent vectors of length n with sum t, descending lex."""
    if n == 1:
        yield (t,)
        return
    for first in range(t, -1, -1):
        for rest in _multiexponents(n - 1, t - first):
            yield (first,) + rest


def n_deg_exponents(n: int, deg: int) -> List[Tuple[int, ...]]:
    return [xp for t in range(deg + 1) for xp in _multiexponents(n, t)]


def cheb2_pts(k: int) -> np.ndarray:
    # -cospi(j / (k - 1)), j = 0..k-1
    return np.array([-math.cos(math.pi * j / (k - 1)) for j in range(k)])


def calc_univariate_chebyshev(pts_i: np.ndarray, d: int) -> np.ndarray:
    u = np.zeros((pts_i.size, d + 1))
    u[:, 0] = 1
    u[:, 1] = pts_i
    for t in range(2, d + 1):
        u[:, t] = 2 * pts_i * u[:, t - 1] - u[:, t - 2]
    return u


def make_chebyshev_vandermonde(pts: np.ndarray, deg: int) -> np.ndarray:
    n = pts.shape[1]
    expos = n_deg_exponents(n, deg)
    univ = [calc_univariate_chebyshev(pts[:, i], deg) for i in range(n)]
    V = np.zeros((pts.shape[0], len(expos)))
    for col, xp in enumerate(expos):
        v = univ[0][:, xp[0]].copy()
        for j in range(1, n):
            v *= univ[j][:, xp[j]]
        V[:, col] = v
    return V


def cheb2_data(d: int):
    U, L = get_U(1, d), get_L(1, d)
    pts = cheb2_pts(U).reshape(-1, 1)
    P0 = make_chebyshev_vandermonde(pts, d)
    return U, pts, P0, P0[:, : get_L(1, d - 1)]


def padua_data(d: int):
    U = get_U(2, d)
    cheba = cheb2_pts(2 * d + 1)
    chebb = cheb2_pts(2 * d + 2)
    pts = np.zeros((U, 2))
    j = 0
    for a in range(2 * d + 1):
        for b in range(2 * d + 2):
            if (a + b) % 2 == 0:
                pts[j, 0] = -cheba[a]
                pts[U - 1 - j, 1] = -chebb[2 * d + 1 - b]
                j += 1
    P0 = make_chebyshev_vandermonde(pts, d)
    return U, pts, P0, P0[:, : get_L(2, d - 1)]


def _prod_consec(n: int, d: int, j: int = 0) -> int:
    out = 1
    for v in range(2 * d + 1 + j, 2 * d + n + 1):
        out *= v
    return out


def choose_interp_pts(cand_pts: np.ndarray, d: int):
    n = cand_pts.shape[1]
    U = get_U(n, d)
    V = make_chebyshev_vandermonde(cand_pts, 2 * d)
    _, piv = sla.qr(V.T, mode="r", pivoting=True)
    keep = piv[:U]
    return V[keep, :], keep


def make_wsos_arrays(dom, cand_pts: np.ndarray, d: int):
    n = cand_pts.shape[1]
    V, keep = choose_interp_pts(cand_pts, d)
    pts = cand_pts[keep, :]
    P0 = V[:, : get_L(n, d)]
    Lsub = get_L(n, (2 * d - dom.degree()) // 2)
    return pts, P0, P0[:, :Lsub]


def approxfekete_data(n: int, d: int):
    npts = _prod_consec(n, d)
    cand = np.zeros((npts, n))
    for j in range(1, n + 1):
        ig = _prod_consec(n, d, j)
        cs = cheb2_pts(2 * d + j)
        i, l = 0, 0
        while True:
            cand[i : i + ig, j - 1] = cs[l]
            i += ig
            l += 1
            if l >= 2 * d + j:
                if i >= npts:
                    break
                l = 0
    dom = BoxDomain(-np.ones(n), np.ones(n))
    pts, P0, P0sub = make_wsos_arrays(dom, cand, d)
    return pts.shape[0], pts, P0, P0sub


def _interp_box_raw(n: int, d: int):
    if n == 1:
        return cheb2_data(d)
    if n == 2:
        return padua_data(d)
    return approxfekete_data(n, d)


def interpolate(dom, d: int, sample: Optional[bool] = None, sample_factor: int = 0, rng=None):
    """Returns (U, pts, Ps).  realinterp.jl:18-54."""
    assert d >= 1
    n = dom.dimension()
    U = get_U(n, d)
    if sample is None:
        sample = (not isinstance(dom, BoxDomain)) or n >= 7 or _prod_consec(n, d) > 35_000
    if sample:
        if sample_factor <= 0:
            sample_factor = 10 if U <= 12_000 else (5 if U <= 15_000 else (2 if U <= 22_000 else 1))
        rng = np.random.default_rng(1) if rng is None else rng
        cand = dom.sample(U * sample_factor, rng)
        pts, P0, P0sub = make_wsos_arrays(dom, cand, d)
        g = dom.weights(pts)
        Ps = [P0] + [np.sqrt(gi)[:, None] * P0sub for gi in g]
        return U, pts, [np.asfortranarray(P) for P in Ps]
    U, pts, P0, P0sub = _interp_box_raw(n, d)
    if isinstance(dom, FreeDomain):
        return U, pts, [np.asfortranarray(P0)]
    nd = dom.dimension()
    pscale = np.array([0.5 * (dom.u[j % nd] - dom.l[j % nd]) for j in range(n)])
    pshift = np.array([0.5 * (dom.u[j % nd] + dom.l[j % nd]) for j in range(n)])
    PWts = [(np.sqrt(1 - pts[:, j] ** 2) * pscale[j])[:, None] * P0sub for j in range(nd)]
    trpts = pts * pscale[None, :] + pshift[None, :]
    return U, trpts, [np.asfortranarray(P) for P in [P0] + PWts]
