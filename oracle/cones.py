"""ORACLE — TEST INFRASTRUCTURE ONLY (never on the product path).

CPU restatement (numpy / scipy -> OpenBLAS) of the reference's cone barrier
oracles for the hot path.  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this package.

Parity pinning: checked against the reference's own RNG-free known-answer
instances (test/nativeinstances.jl, see tests/test_oracle_instances.py) and the
oracle identities of test/cone.jl:31-129 (tests/test_oracle_cones.py).  The
reference itself (Julia) cannot run in this image, so iteration-level numbers
are pinned by this restatement only.

Each class follows one reference file:
  Cone base / generic oracles      src/Cones/Cones.jl:34-304
  Nonnegative                      src/Cones/nonnegative.jl
  EpiNormEucl                      src/Cones/epinormeucl.jl
  EpiPerSquare                     src/Cones/epipersquare.jl
  PosSemidefTri (real)             src/Cones/possemideftri.jl
  WSOSInterpNonnegative (real)     src/Cones/wsosinterpnonnegative.jl
  WSOSInterpPosSemidefTri          src/Cones/wsosinterppossemideftri.jl
"""
from __future__ import annotations

import math
from typing import List, Tuple

import numpy as np
import scipy.linalg as sla
from scipy.linalg import lapack

import hypatia_b200  # noqa: F401  (registers the package under its importable name)
from hypatia_b200 import cones_spec as cs
from hypatia_b200.host.coneset import ConeSet

from .dense import posdef_fact_copy

EPS = float(np.finfo(np.float64).eps)
RT2 = math.sqrt(2.0)


def _as2d(arr: np.ndarray) -> Tuple[np.ndarray, bool]:
    if arr.ndim == 1:
        return arr.reshape(-1, 1), True
    return arr, False


class Cone:
    """Generic plumbing of Cones.jl: lazy caches, generic Hessian-based products,
    check_numerics (:267-284), get_proxsqr (:288-304)."""

    uses_generic_hess = False

    def __init__(self, spec: cs.ConeSpec):
        self.spec = spec
        self.dim = spec.dim
        self.nu = spec.nu
        self.use_dual_barrier = spec.use_dual_barrier

    # -- setup_data! (Cones.jl:147-160)
    def setup_data(self):
        self.reset_data()
        d = self.dim
        self.point = np.zeros(d)
        self.dual_point = np.zeros(d)
        self.grad_vec = np.zeros(d)
        self.is_feas_flag = False
        self.hess_fact = None
        self.setup_extra_data()

    def setup_extra_data(self):
        pass

    def reset_data(self):  # Cones.jl:184-190
        self.feas_updated = self.grad_updated = self.hess_updated = False
        self.inv_hess_updated = self.hess_fact_updated = False

    def load_point(self, point, scal=None):  # Cones.jl:164-168
        if scal is None:
            self.point[:] = point
        else:
            self.point[:] = scal * point

    def load_dual_point(self, point):
        self.dual_point[:] = point

    def is_feas(self) -> bool:  # Cones.jl:63
        return self.is_feas_flag if self.feas_updated else self.update_feas()

    def is_dual_feas(self) -> bool:
        return True

    def grad(self) -> np.ndarray:  # Cones.jl:78
        return self.grad_vec if self.grad_updated else self.update_grad()

    def hess(self) -> np.ndarray:
        """Full symmetric Hessian (tests and generic-Hessian cones)."""
        if not self.hess_updated:
            self.update_hess()
        return self.hess_mat

    def inv_hess(self) -> np.ndarray:
        if not self.inv_hess_updated:
            self.update_inv_hess()
        return self.inv_hess_mat

    # generic products with stored H / chol(H) (Cones.jl:108-125)
    def hess_prod(self, arr):
        if not self.hess_updated:
            self.update_hess()
        return self.hess_mat @ arr

    def update_hess_fact(self) -> bool:  # Cones.jl:233-245
        if self.hess_fact_updated:
            return self.hess_fact.success
        if not self.hess_updated:
            self.update_hess()
        self.hess_fact = posdef_fact_copy(self.hess_mat, try_shift=False)
        self.hess_fact_updated = True
        return self.hess_fact.success

    def inv_hess_prod(self, arr):
        self.update_hess_fact()
        return self.hess_fact.solve(arr)

    def update_inv_hess(self):  # Cones.jl:246-252
        self.update_hess_fact()
        self.inv_hess_mat = self.hess_fact.inverse()
        self.inv_hess_updated = True

    def use_sqrt_hess_oracles(self, arr_dim: int) -> bool:  # Cones.jl:193-199
        if not self.hess_fact_updated:
            if arr_dim < self.dim:
                return False
            if not self.update_hess_fact():
                return False
        return self.hess_fact.kind == "chol"

    def sqrt_hess_prod(self, arr):  # Cones.jl:202-206  (U * arr)
        assert self.hess_fact_updated
        return np.triu(self.hess_fact.factor) @ arr

    def inv_sqrt_hess_prod(self, arr):  # Cones.jl:209-214  (U' \ arr)
        assert self.hess_fact_updated
        return sla.solve_triangular(self.hess_fact.factor, arr, trans="T", lower=False)

    def hess_prod_slow(self, arr):  # Cones.jl:229-231
        return self.hess_prod(arr)

    def use_dder3(self) -> bool:
        return True

    def check_numerics(self) -> bool:  # Cones.jl:267-284
        gtol = math.sqrt(math.sqrt(EPS))
        Htol = 10 * math.sqrt(gtol)
        g = self.grad()
        dim = g.size
        if abs(1 + float(g @ self.point) / self.nu) > gtol * dim:
            return False
        Hig = self.inv_hess_prod(g)
        if abs(1 - float(Hig @ g) / self.nu) > Htol * dim:
            return False
        return True

    def get_proxsqr(self, irtmu: float, use_max_prox: bool) -> float:  # Cones.jl:288-304
        g = self.grad()
        vec1 = irtmu * self.dual_point + g
        vec2 = self.inv_hess_prod(vec1)
        prox_sqr = float(vec2 @ vec1)
        if prox_sqr < -math.sqrt(EPS) * g.size:
            return math.inf
        return abs(prox_sqr)


# ---------------------------------------------------------------------------
class Nonnegative(Cone):
    def use_sqrt_hess_oracles(self, arr_dim):
        return True

    def update_feas(self):  # nonnegative.jl:55-60
        self.is_feas_flag = bool(np.all(self.point > EPS))
        self.feas_updated = True
        return self.is_feas_flag

    def is_dual_feas(self):  # :62
        return bool(np.all(self.dual_point > EPS))

    def update_grad(self):  # :64-69
        self.grad_vec = -1.0 / self.point
        self.grad_updated = True
        return self.grad_vec

    def update_hess(self):
        self.hess_mat = np.diag(self.grad() ** 2)
        self.hess_updated = True

    def update_inv_hess(self):
        self.inv_hess_mat = np.diag(self.point**2)
        self.inv_hess_updated = True

    def _col(self, arr):
        return self.point if arr.ndim == 1 else self.point[:, None]

    def hess_prod(self, arr):  # :93-97
        p = self._col(arr)
        return arr / p / p

    def inv_hess_prod(self, arr):  # :99-103
        p = self._col(arr)
        return arr * p * p

    def sqrt_hess_prod(self, arr):  # :105-109
        return arr / self._col(arr)

    def inv_sqrt_hess_prod(self, arr):  # :111-119
        return arr * self._col(arr)

    def dder3(self, direction):  # :121-124
        return (direction / self.point) ** 2 / self.point

    def get_proxsqr(self, irtmu, use_max_prox):  # :136-141
        v = (self.point * self.dual_point * irtmu - 1.0) ** 2
        return float(v.max()) if use_max_prox else float(v.sum())


# ---------------------------------------------------------------------------
class EpiNormEucl(Cone):
    def use_sqrt_hess_oracles(self, arr_dim):
        return True

    def update_feas(self):  # epinormeucl.jl:62-76
        u = self.point[0]
        if u > EPS:
            w = self.point[1:]
            self.dist = (u * u - float(w @ w)) / 2
            self.is_feas_flag = self.dist > EPS
        else:
            self.is_feas_flag = False
        self.feas_updated = True
        return self.is_feas_flag

    def is_dual_feas(self):  # :78-88
        u = self.dual_point[0]
        if u > EPS:
            w = self.dual_point[1:]
            return (u * u - float(w @ w)) > 2 * EPS
        return False

    def update_grad(self):  # :90-98
        g = self.point / self.dist
        g[0] *= -1
        self.grad_vec = g
        self.grad_updated = True
        return g

    def update_hess(self):  # :100-113
        g = self.grad()
        H = np.outer(g, g)
        inv_dist = 1.0 / self.dist
        H[np.diag_indices_from(H)] += inv_dist
        H[0, 0] -= inv_dist + inv_dist
        self.hess_mat = H
        self.hess_updated = True

    def update_inv_hess(self):  # :115-127
        Hi = np.outer(self.point, self.point)
        Hi[np.diag_indices_from(Hi)] += self.dist
        Hi[0, 0] -= self.dist + self.dist
        self.inv_hess_mat = Hi
        self.inv_hess_updated = True

    def hess_prod(self, arr):  # :129-144
        a, one_d = _as2d(arr)
        u, w = self.point[0], self.point[1:]
        uj, wj = a[0], a[1:]
        ga = (w @ wj - u * uj) / self.dist
        prod = np.empty_like(a)
        prod[0] = -ga * u - uj
        prod[1:] = ga[None, :] * w[:, None] + wj
        prod /= self.dist
        return prod[:, 0] if one_d else prod

    def inv_hess_prod(self, arr):  # :146-157
        a, one_d = _as2d(arr)
        pa = self.point @ a
        prod = pa[None, :] * self.point[:, None]
        prod[0] -= self.dist * a[0]
        prod[1:] += self.dist * a[1:]
        return prod[:, 0] if one_d else prod

    def sqrt_hess_prod(self, arr):  # :159-182
        a, one_d = _as2d(arr)
        u, w = self.point[0], self.point[1:]
        distrt2 = self.dist * RT2
        rtdist = math.sqrt(self.dist)
        urtdist = u + rtdist * RT2
        uj, wj = a[0], a[1:]
        dotwwj = w @ wj
        prod = np.empty_like(a)
        prod[0] = (u * uj - dotwwj) / distrt2
        wmulj = (dotwwj / urtdist - uj) / distrt2
        prod[1:] = w[:, None] * wmulj[None, :] + wj / rtdist
        return prod[:, 0] if one_d else prod

    def inv_sqrt_hess_prod(self, arr):  # :184-206
        a, one_d = _as2d(arr)
        u, w = self.point[0], self.point[1:]
        rtdist = math.sqrt(self.dist)
        urtdist = u + rtdist * RT2
        uj, wj = a[0], a[1:]
        dotwwj = w @ wj
        prod = np.empty_like(a)
        prod[0] = (u * uj + dotwwj) / RT2
        wmulj = (dotwwj / urtdist + uj) / RT2
        prod[1:] = w[:, None] * wmulj[None, :] + wj * rtdist
        return prod[:, 0] if one_d else prod

    def dder3(self, direction):  # :208-228
        point = self.point
        u, w = point[0], point[1:]
        u_dir, w_dir = direction[0], direction[1:]
        jdotpd = u * u_dir - float(w @ w_dir)
        d3 = self.hess_prod(direction)
        dotdHd = -float(direction @ d3)
        dotpHd = float(point @ d3)
        d3 = d3 * jdotpd
        d3[1:] += dotdHd * w + dotpHd * w_dir
        d3[0] += -dotdHd * u - dotpHd * u_dir
        d3 /= 2 * self.dist
        return d3


# ---------------------------------------------------------------------------
class EpiNormInf(Cone):
    """src/Cones/epinorminf.jl (real case, R == T): all oracles are O(d) closed forms; no sqrt oracles
    (use_sqrt_hess_oracles = false, :73), so this cone takes the gemm branch of the Schur assembly."""

    def use_sqrt_hess_oracles(self, arr_dim):  # :73
        return False

    def reset_data(self):  # :63-71
        super().reset_data()
        self.inv_hess_aux_updated = False

    def update_feas(self):  # :100-113
        u = self.point[0]
        if u > EPS:
            self.w = self.point[1:].copy()
            self.is_feas_flag = (u - float(np.max(np.abs(self.w)))) > EPS
        else:
            self.is_feas_flag = False
        self.feas_updated = True
        return self.is_feas_flag

    def is_dual_feas(self):  # :115-128
        u = self.dual_point[0]
        if u > EPS:
            return (u - float(np.sum(np.abs(self.dual_point[1:])))) > EPS
        return False

    def update_grad(self):  # :136-162
        u, w = self.point[0], self.w
        d = self.dim - 1
        self.mu = w / u
        self.zeta = 0.5 * (u - self.mu * w)
        self.cu = (d - 1) / u
        g = np.empty(self.dim)
        g[0] = self.cu - float(np.sum(1.0 / self.zeta))
        g[1:] = self.mu / self.zeta
        self.grad_vec = g
        self.grad_updated = True
        return g

    def update_hess(self):  # :164-203 (dense restatement of the arrow-shaped sparse Hessian)
        g = self.grad()
        u, zeta = self.point[0], self.zeta
        ui = 1.0 / u
        tzi2 = (1.0 / zeta - ui) / zeta
        H = np.zeros((self.dim, self.dim))
        H[0, 0] = float(np.sum(tzi2)) - self.cu / u
        off = -g[1:] / zeta
        H[0, 1:] = off
        H[1:, 0] = off
        H[np.arange(1, self.dim), np.arange(1, self.dim)] = tzi2
        self.hess_mat = H
        self.hess_updated = True

    def hess_prod(self, arr):  # :205-229
        self.grad()
        a, one_d = _as2d(arr)
        u, mu, zeta = self.point[0], self.mu[:, None], self.zeta[:, None]
        p, r = a[0], a[1:]
        pui = p / u
        s1 = (p[None, :] - mu * r) / zeta
        prod = np.empty_like(a)
        prod[0] = np.sum((s1 - pui[None, :]) / zeta, axis=0) - self.cu * pui
        prod[1:] = (r / u - s1 * mu) / zeta
        return prod[:, 0] if one_d else prod

    def _update_inv_hess_aux(self):  # :257-273
        self.grad()
        u, w = self.point[0], self.w
        umz = u - self.zeta
        self.Zu = -self.cu + float(np.sum(1.0 / umz))
        self.wumzi = w / umz
        self.zti = u - self.wumzi * w
        self.inv_hess_aux_updated = True

    def update_inv_hess(self):  # :275-312
        if not self.inv_hess_aux_updated:
            self._update_inv_hess_aux()
        u = self.point[0]
        Hi = np.empty((self.dim, self.dim))
        hiuu = Hi[0, 0] = u / self.Zu
        Hi[1:, 0] = self.wumzi * hiuu
        Hi[0, 1:] = Hi[1:, 0]
        Hi[1:, 1:] = np.outer(self.wumzi, self.wumzi * hiuu)
        Hi[np.arange(1, self.dim), np.arange(1, self.dim)] += self.zeta * self.zti
        self.inv_hess_mat = Hi
        self.inv_hess_updated = True

    def inv_hess_prod(self, arr):  # :314-335
        if not self.inv_hess_aux_updated:
            self._update_inv_hess_aux()
        a, one_d = _as2d(arr)
        u = self.point[0]
        p, r = a[0], a[1:]
        c1 = u * (p + self.wumzi @ r) / self.Zu
        prod = np.empty_like(a)
        prod[0] = c1
        prod[1:] = c1[None, :] * self.wumzi[:, None] + (self.zeta * self.zti)[:, None] * r
        return prod[:, 0] if one_d else prod

    def sqrt_hess_prod(self, arr):
        raise NotImplementedError("EpiNormInf has no sqrt Hessian oracles (epinorminf.jl:73)")

    inv_sqrt_hess_prod = sqrt_hess_prod

    def dder3(self, direction):  # :366-392
        self.grad()
        u, mu, zeta = self.point[0], self.mu, self.zeta
        p, r = direction[0], direction[1:]
        pui = p / u
        rui = r / u
        s1 = (p - mu * r) / zeta
        s2 = 0.5 * (p * pui - rui * r) / zeta - s1 * s1
        c1 = float(np.sum((s1 * pui + s2) / zeta))
        out = np.empty(self.dim)
        out[0] = -c1 - self.cu * pui * pui
        out[1:] = (s1 * rui + s2 * mu) / zeta
        return out


# ---------------------------------------------------------------------------
class EpiPerSquare(Cone):
    def use_sqrt_hess_oracles(self, arr_dim):
        return True

    def reset_data(self):  # epipersquare.jl:51-60
        super().reset_data()
        self.sqrt_hess_prod_updated = self.inv_sqrt_hess_prod_updated = False

    def update_feas(self):  # :73-88
        u, v = self.point[0], self.point[1]
        if u > EPS and v > EPS:
            w = self.point[2:]
            self.dist = u * v - float(w @ w) / 2
            self.is_feas_flag = self.dist > EPS
        else:
            self.is_feas_flag = False
        self.feas_updated = True
        return self.is_feas_flag

    def is_dual_feas(self):  # :90-100
        u, v = self.dual_point[0], self.dual_point[1]
        if u > EPS and v > EPS:
            w = self.dual_point[2:]
            return (u * v - float(w @ w) / 2) > EPS
        return False

    def update_grad(self):  # :102-112
        g = self.point / self.dist
        g2 = g[1]
        g[1] = -g[0]
        g[0] = -g2
        self.grad_vec = g
        self.grad_updated = True
        return g

    def update_hess(self):  # :114-128
        g = self.grad()
        H = np.outer(g, g)
        inv_dist = 1.0 / self.dist
        for j in range(2, self.dim):
            H[j, j] += inv_dist
        H[0, 1] -= inv_dist
        H[1, 0] = H[0, 1]
        self.hess_mat = H
        self.hess_updated = True

    def update_inv_hess(self):  # :130-143
        Hi = np.outer(self.point, self.point)
        for j in range(2, self.dim):
            Hi[j, j] += self.dist
        Hi[0, 1] -= self.dist
        Hi[1, 0] = Hi[0, 1]
        self.inv_hess_mat = Hi
        self.inv_hess_updated = True

    def hess_prod(self, arr):  # :145-162
        a, one_d = _as2d(arr)
        u, v, w = self.point[0], self.point[1], self.point[2:]
        uj, vj, wj = a[0], a[1], a[2:]
        ga = (w @ wj - v * uj - u * vj) / self.dist
        prod = np.empty_like(a)
        prod[0] = -ga * v - vj
        prod[1] = -ga * u - uj
        prod[2:] = ga[None, :] * w[:, None] + wj
        prod /= self.dist
        return prod[:, 0] if one_d else prod

    def inv_hess_prod(self, arr):  # :164-176
        a, one_d = _as2d(arr)
        pa = self.point @ a
        prod = pa[None, :] * self.point[:, None]
        prod[0] -= self.dist * a[1]
        prod[1] -= self.dist * a[0]
        prod[2:] += self.dist * a[2:]
        return prod[:, 0] if one_d else prod

    def _update_sqrt(self):  # :178-194
        rtdist = self.rtdist = math.sqrt(self.dist)
        self.denom = 2 * rtdist + self.point[0] + self.point[1]
        vec = np.empty(self.dim)
        vec[2:] = self.point[2:] / rtdist
        vec[0] = -self.point[1] / rtdist - 1
        vec[1] = -self.point[0] / rtdist - 1
        self.sqrt_hess_vec = vec
        self.sqrt_hess_prod_updated = True

    def _update_inv_sqrt(self):  # :196-211
        rtdist = self.rtdist = math.sqrt(self.dist)
        self.denom = 2 * rtdist + self.point[0] + self.point[1]
        vec = self.point.copy()
        vec[:2] += rtdist
        self.inv_sqrt_hess_vec = vec
        self.inv_sqrt_hess_prod_updated = True

    def sqrt_hess_prod(self, arr):  # :213-233
        if not self.sqrt_hess_prod_updated:
            self._update_sqrt()
        a, one_d = _as2d(arr)
        vec, rtdist = self.sqrt_hess_vec, self.rtdist
        dotj = (vec @ a) / self.denom
        prod = dotj[None, :] * vec[:, None]
        prod[0] -= a[1] / rtdist
        prod[1] -= a[0] / rtdist
        prod[2:] += a[2:] / rtdist
        return prod[:, 0] if one_d else prod

    def inv_sqrt_hess_prod(self, arr):  # :235-255
        if not self.inv_sqrt_hess_prod_updated:
            self._update_inv_sqrt()
        a, one_d = _as2d(arr)
        vec, rtdist = self.inv_sqrt_hess_vec, self.rtdist
        dotj = (vec @ a) / self.denom
        prod = dotj[None, :] * vec[:, None]
        prod[0] -= a[1] * rtdist
        prod[1] -= a[0] * rtdist
        prod[2:] += a[2:] * rtdist
        return prod[:, 0] if one_d else prod

    def dder3(self, direction):  # :257-280
        point = self.point
        u, v, w = point[0], point[1], point[2:]
        u_dir, v_dir, w_dir = direction[0], direction[1], direction[2:]
        jdotpd = u * v_dir + v * u_dir - float(w @ w_dir)
        d3 = self.hess_prod(direction)
        dotdHd = -float(direction @ d3)
        dotpHd = float(point @ d3)
        d3 = d3 * jdotpd
        d3[2:] += dotdHd * w + dotpHd * w_dir
        d3[0] += -dotdHd * v - dotpHd * v_dir
        d3[1] += -dotdHd * u - dotpHd * u_dir
        d3 /= 2 * self.dist
        return d3


# ---------------------------------------------------------------------------
def _symm_kron(mat: np.ndarray) -> np.ndarray:
    """Symmetric Kronecker product of a symmetric matrix with itself in svec
    coordinates (arrayutilities.jl:287-324) — only tests / non-QRChol paths."""
    side = mat.shape[0]
    dim = cs.svec_length(side)
    H = np.zeros((dim, dim))
    e = np.zeros(dim)
    for j in range(dim):
        e[:] = 0
        e[j] = 1
        E = cs.svec_to_smat(e)
        H[:, j] = cs.smat_to_svec(mat @ E @ mat)
    return H


class PosSemidefTri(Cone):
    def use_sqrt_hess_oracles(self, arr_dim):
        return True

    def setup_extra_data(self):
        self.side = self.spec.side

    def update_feas(self):  # possemideftri.jl:89-99
        self.mat = cs.svec_to_smat(self.point)  # full symmetric (copytri done eagerly)
        c, info = lapack.dpotrf(self.mat, lower=0, clean=1)
        self.fact_U = c
        self.is_feas_flag = info == 0
        self.feas_updated = True
        return self.is_feas_flag

    def is_dual_feas(self):  # :101-104
        mat3 = cs.svec_to_smat(self.dual_point)
        _, info = lapack.dpotrf(mat3, lower=0)
        return info == 0

    def update_grad(self):  # :106-116
        inv_mat, info = lapack.dpotri(self.fact_U, lower=0)
        inv_mat = np.triu(inv_mat) + np.triu(inv_mat, 1).T
        self.inv_mat = inv_mat
        self.grad_vec = -cs.smat_to_svec(inv_mat)
        self.grad_updated = True
        return self.grad_vec

    def update_hess(self):  # :118-125
        self.grad()
        self.hess_mat = _symm_kron(self.inv_mat)
        self.hess_updated = True

    def update_inv_hess(self):  # :127-133
        assert self.is_feas()
        self.inv_hess_mat = _symm_kron(self.mat)
        self.inv_hess_updated = True

    def _idx(self):
        """svec <-> smat index maps (arrayutilities.jl:183-251), cached."""
        if not hasattr(self, "_iu"):
            side = self.side
            cols = np.repeat(np.arange(side), np.arange(1, side + 1))
            rows = np.concatenate([np.arange(j + 1) for j in range(side)])
            self._iu = (rows, cols)
            self._scale = np.where(rows == cols, 1.0, 1.0 / RT2)
        return self._iu, self._scale

    def _cols(self, arr, fn):
        """Apply fn to smat of every column (the reference loops over columns through one
        scratch matrix, possemideftri.jl:138-144; batched here, same arithmetic per column)."""
        a, one_d = _as2d(arr)
        (rows, cols), scale = self._idx()
        m = a.shape[1]
        M = np.zeros((m, self.side, self.side))
        vals = a.T * scale[None, :]
        M[:, rows, cols] = vals
        M[:, cols, rows] = vals
        R = fn(M)
        prod = (R[:, rows, cols] / scale[None, :]).T
        prod = np.ascontiguousarray(prod) if not one_d else prod[:, 0]
        return prod

    def hess_prod(self, arr):  # :135-147   X^-1 A X^-1 (rdiv + ldiv with chol)
        U = self.fact_U

        def fn(M):
            m, s_, _ = M.shape
            # M / X then X \ M with the Cholesky factor, for all columns at once
            T = sla.cho_solve((U, False), M.transpose(1, 0, 2).reshape(s_, m * s_))  # X^-1 M_k
            T = T.reshape(s_, m, s_).transpose(1, 0, 2)  # (m, s, s) = X^-1 M_k
            T = T.transpose(0, 2, 1)  # (X^-1 M_k)' = M_k X^-1
            T2 = sla.cho_solve((U, False), T.transpose(1, 0, 2).reshape(s_, m * s_))
            return T2.reshape(s_, m, s_).transpose(1, 0, 2)

        return self._cols(arr, fn)

    def inv_hess_prod(self, arr):  # :149-160   X A X
        X = self.mat
        return self._cols(arr, lambda M: X @ (M @ X))

    def sqrt_hess_prod(self, arr):  # :162-174   U^-T A U^-1
        U = self.fact_U

        def fn(M):
            m, s_, _ = M.shape
            T = sla.solve_triangular(U, M.transpose(1, 0, 2).reshape(s_, m * s_), trans="T", lower=False)  # U^-T M_k
            T = T.reshape(s_, m, s_).transpose(1, 0, 2).transpose(0, 2, 1)  # M_k U^-1
            T2 = sla.solve_triangular(U, T.transpose(1, 0, 2).reshape(s_, m * s_), trans="T", lower=False)
            return T2.reshape(s_, m, s_).transpose(1, 0, 2)

        return self._cols(arr, fn)

    def inv_sqrt_hess_prod(self, arr):  # :176-192   U A U'
        U = np.triu(self.fact_U)
        return self._cols(arr, lambda M: U @ (M @ U.T))

    def dder3(self, direction):  # :194-204   S S', S = X^-1 D U^-1
        U = self.fact_U
        S = cs.svec_to_smat(direction)
        S = sla.cho_solve((U, False), S)
        S = sla.solve_triangular(U, S.T, trans="T", lower=False).T  # S / U
        return cs.smat_to_svec(S @ S.T)


class DoublyNonnegativeTri(PosSemidefTri):
    """src/Cones/doublynonnegativetri.jl: -logdet(S) - sum log(off-diagonal entries).  The products other than
    hess_prod! and the factorization-based oracles are the generic ones of Cones.jl (stored Hessian + its
    Cholesky), like for the WSOS cones."""

    uses_generic_hess = True

    def use_sqrt_hess_oracles(self, arr_dim):  # Cones.jl:193-199 (generic)
        return Cone.use_sqrt_hess_oracles(self, arr_dim)

    def setup_extra_data(self):
        self.side = self.spec.side
        self.offdiag = self.spec.offdiag_idxs()

    def update_feas(self):  # :134-147
        if np.all(self.point > EPS):
            return PosSemidefTri.update_feas(self)
        self.is_feas_flag = False
        self.feas_updated = True
        return False

    def is_dual_feas(self):  # Cones.jl:71
        return True

    def update_grad(self):  # :149-160
        PosSemidefTri.update_grad(self)
        self.inv_vec = 1.0 / self.point[self.offdiag]
        self.grad_vec[self.offdiag] -= self.inv_vec
        return self.grad_vec

    def update_hess(self):  # :162-175
        self.grad()
        H = _symm_kron(self.inv_mat)
        H[self.offdiag, self.offdiag] += self.inv_vec**2
        self.hess_mat = H
        self.hess_updated = True

    def update_inv_hess(self):  # Cones.jl:246-252
        Cone.update_inv_hess(self)

    def hess_prod(self, arr):  # :177-196
        prod = PosSemidefTri.hess_prod(self, arr)
        od = self.offdiag
        s_off = self.point[od]
        a = np.asarray(arr, dtype=np.float64)
        if a.ndim == 1:
            prod[od] += a[od] / s_off / s_off
        else:
            prod[od, :] += a[od, :] / s_off[:, None] / s_off[:, None]
        return prod

    def inv_hess_prod(self, arr):  # Cones.jl:115-119
        return Cone.inv_hess_prod(self, arr)

    def sqrt_hess_prod(self, arr):  # Cones.jl:202-206
        return Cone.sqrt_hess_prod(self, arr)

    def inv_sqrt_hess_prod(self, arr):  # Cones.jl:209-214
        return Cone.inv_sqrt_hess_prod(self, arr)

    def dder3(self, direction):  # :198-212
        out = PosSemidefTri.dder3(self, direction)
        od = self.offdiag
        s_off = self.point[od]
        out[od] += (direction[od] / s_off) ** 2 / s_off
        return out


# ---------------------------------------------------------------------------
class _GenericHessCone(Cone):
    """Cones whose products go through a stored Hessian and its Cholesky, with
    the hess_prod_slow switch (Cones.jl:218-227)."""

    uses_generic_hess = True

    def reset_data(self):
        super().reset_data()
        self.use_hess_prod_slow = False
        self.use_hess_prod_slow_updated = False

    def update_use_hess_prod_slow(self):  # Cones.jl:218-227
        if not self.hess_updated:
            self.update_hess()
        rel_viol = abs(1 - float(self.point @ (self.hess_mat @ self.point)) / self.nu)
        self.use_hess_prod_slow = rel_viol > self.dim * math.sqrt(EPS)
        self.use_hess_prod_slow_updated = True


class PosSemidefTriSparse(_GenericHessCone):
    """src/Cones/possemideftrisparse/denseimpl.jl (real case): -logdet of the matrix with the given sparsity
    pattern; gradient and Hessian restricted to the pattern, every other product generic (stored Hessian + its
    Cholesky), hess_prod_slow! through the Cholesky factor of the matrix."""

    def setup_extra_data(self):
        sp = self.spec
        self.side, self.ri, self.ci = sp.side, sp.rows0, sp.cols0
        self.offd = self.ri != self.ci
        self.scale = np.where(self.offd, 1.0 / RT2, 1.0)

    def _to_mat(self, vec):  # svec_to_smat_sparse! (:216-231) + copytri!
        M = np.zeros((self.side, self.side))
        x = vec * self.scale
        M[self.ri, self.ci] = x
        M[self.ci, self.ri] = x
        return M

    def _to_vec(self, M):  # smat_to_svec_sparse! (:251-267)
        return M[self.ri, self.ci] / self.scale

    def update_feas(self):  # :32-44
        self.mat = self._to_mat(self.point)
        c, info = lapack.dpotrf(self.mat, lower=0, clean=1)
        self.fact_U = c
        self.is_feas_flag = info == 0
        self.feas_updated = True
        return self.is_feas_flag

    def update_grad(self):  # :46-59
        inv_mat, _ = lapack.dpotri(self.fact_U, lower=0)
        self.inv_mat = np.triu(inv_mat) + np.triu(inv_mat, 1).T
        self.grad_vec = -self._to_vec(self.inv_mat)
        self.grad_updated = True
        return self.grad_vec

    def update_hess(self):  # :61-88
        self.grad()
        Li, ri, ci, od = self.inv_mat, self.ri, self.ci, self.offd
        i1, j1 = ri[:, None], ci[:, None]
        i2, j2 = ri[None, :], ci[None, :]
        both_diag = (~od)[:, None] & (~od)[None, :]
        one_diag = (~od)[:, None] ^ (~od)[None, :]
        H = Li[i1, i2] * Li[j1, j2] + Li[i1, j2] * Li[j1, i2]
        H = np.where(one_diag, RT2 * Li[i1, i2] * Li[j1, j2], H)
        H = np.where(both_diag, Li[i1, i2] ** 2, H)
        self.hess_mat = H
        self.hess_updated = True

    def hess_prod_slow(self, arr):  # :142-163
        if not self.use_hess_prod_slow_updated:
            self.update_use_hess_prod_slow()
        if not self.use_hess_prod_slow:
            return self.hess_prod(arr)
        a, one_d = _as2d(arr)
        out = np.empty_like(a)
        for j in range(a.shape[1]):
            L = self._to_mat(a[:, j])
            L = sla.cho_solve((self.fact_U, False), L)
            L = sla.cho_solve((self.fact_U, False), L.T).T
            out[:, j] = self._to_vec(L)
        return out[:, 0] if one_d else out

    def dder3(self, direction):  # :165-176: columns of L^-1 D Lambda^-1, pairwise dot products on the pattern
        U = self.fact_U
        D = self._to_mat(direction)
        M = sla.solve_triangular(U, D, trans="T", lower=False)  # L^-1 D  (L = U')
        M = sla.cho_solve((U, False), M.T).T  # (L^-1 D) Lambda^-1
        x = np.einsum("ki,ki->i", M[:, self.ri], M[:, self.ci])
        return x / self.scale


class WSOSInterpNonnegative(_GenericHessCone):
    def setup_extra_data(self):
        self.Ps = self.spec.Ps

    def update_feas(self):  # wsosinterpnonnegative.jl:99-128
        self.LamF = []
        self.is_feas_flag = True
        for Pk in self.Ps:
            LUk = Pk.T * self.point[None, :]
            LLk = LUk @ Pk
            c, info = lapack.dpotrf(LLk, lower=1, clean=1)
            if info != 0:
                self.is_feas_flag = False
                break
            self.LamF.append(c)
        self.feas_updated = True
        return self.is_feas_flag

    def is_dual_feas(self):  # :130-133
        return bool(np.all(self.dual_point > EPS))

    def update_grad(self):  # :135-149
        g = np.zeros(self.dim)
        self.LamFLP = []
        for Pk, Lk in zip(self.Ps, self.LamF):
            F = sla.solve_triangular(Lk, Pk.T, lower=True)
            self.LamFLP.append(F)
            g -= np.sum(F * F, axis=0)
        self.grad_vec = g
        self.grad_updated = True
        return g

    def update_hess(self):  # :151-166
        self.grad()
        H = np.zeros((self.dim, self.dim))
        for F in self.LamFLP:
            UU = F.T @ F
            H += UU * UU
        self.hess_mat = H
        self.hess_updated = True

    @staticmethod
    def _partial_lambda(direction, F):  # :206-211
        LU = F * direction[None, :]
        LL = LU @ F.T
        return LL @ F

    def hess_prod_slow(self, arr):  # :168-191
        if not self.use_hess_prod_slow_updated:
            self.update_use_hess_prod_slow()
        if not self.use_hess_prod_slow:
            return self.hess_prod(arr)
        a, one_d = _as2d(arr)
        prod = np.zeros_like(a)
        for F in self.LamFLP:
            for j in range(a.shape[1]):
                LU = self._partial_lambda(a[:, j], F)
                prod[:, j] += np.sum(F * LU, axis=0)
        return prod[:, 0] if one_d else prod

    def dder3(self, direction):  # :193-204
        self.grad()
        d3 = np.zeros(self.dim)
        for F in self.LamFLP:
            LU = self._partial_lambda(direction, F)
            d3 += np.sum(LU * LU, axis=0)
        return d3


class WSOSInterpPosSemidefTri(_GenericHessCone):
    def setup_extra_data(self):
        self.Ps = self.spec.Ps
        self.R = self.spec.R
        self.U = self.spec.U

    def _blk(self, vec, p, q):
        """block (p,q), 1-based, of an svec-of-blocks vector; off-diagonals / rt2."""
        b = cs.block_idxs(self.U, cs.svec_idx(p, q))
        v = vec[b.start : b.stop]
        return v / RT2 if p != q else v

    def _lambda(self, vec, Pk):
        """Block matrix [P' diag(vec_pq) P]_{pq}, full symmetric (update_feas :139-148)."""
        R, L = self.R, Pk.shape[1]
        Lam = np.zeros((R * L, R * L))
        for p in range(1, R + 1):
            for q in range(1, p + 1):
                B = (Pk.T * self._blk(vec, p, q)[None, :]) @ Pk
                Lam[(p - 1) * L : p * L, (q - 1) * L : q * L] = B
                if p != q:
                    Lam[(q - 1) * L : q * L, (p - 1) * L : p * L] = B.T
        return Lam

    def update_feas(self):  # wsosinterppossemideftri.jl:125-158
        self.LamFL = []
        self.is_feas_flag = True
        for Pk in self.Ps:
            Lam = self._lambda(self.point, Pk)
            c, info = lapack.dpotrf(Lam, lower=1, clean=1)
            if info != 0:
                self.is_feas_flag = False
                break
            self.LamFL.append(c)
        self.feas_updated = True
        return self.is_feas_flag

    def is_dual_feas(self):  # :160-169
        block = 1
        for i in range(1, self.R + 1):
            b = cs.block_idxs(self.U, block)
            if not np.all(self.dual_point[b.start : b.stop] > EPS):
                return False
            block += i + 1
        return True

    def _block_diag_prod(self, vect, mat1, mat2):  # :289-312
        U, R = self.U, self.R
        for j in range(R):
            for i in range(j + 1):
                blk = cs.svec_idx(j + 1, i + 1) - 1
                d = np.sum(mat1[:, i * U : (i + 1) * U] * mat2[:, j * U : (j + 1) * U], axis=0)
                vect[blk * U : (blk + 1) * U] += d * (RT2 if i != j else 1.0)

    def update_grad(self):  # :171-220:  LamFLP = LamFL \ kron(I_R, P')
        assert self.is_feas()
        g = np.zeros(self.dim)
        self.LamFLP = []
        R, U = self.R, self.U
        for Pk, Lc in zip(self.Ps, self.LamFL):
            L = Pk.shape[1]
            kron = np.zeros((R * L, R * U))
            for p in range(R):
                kron[p * L : (p + 1) * L, p * U : (p + 1) * U] = Pk.T
            F = sla.solve_triangular(Lc, kron, lower=True)
            self.LamFLP.append(F)
            self._block_diag_prod(g, F, F)
        self.grad_vec = -g
        self.grad_updated = True
        return self.grad_vec

    def update_hess(self):  # :222-270
        self.grad()
        R, U = self.R, self.U
        H = np.zeros((self.dim, self.dim))
        for F in self.LamFLP:
            PLP = F.T @ F  # (RU x RU) = kron(I,P) Lam^-1 kron(I,P')
            blk = lambda a, b: PLP[(a - 1) * U : a * U, (b - 1) * U : b * U]  # noqa: E731
            for p in range(1, R + 1):
                for q in range(1, p + 1):
                    block = cs.svec_idx(p, q)
                    for p2 in range(1, R + 1):
                        for q2 in range(1, p2 + 1):
                            block2 = cs.svec_idx(p2, q2)
                            if block2 < block:
                                continue
                            scal = RT2 if ((p == q) != (p2 == q2)) else 1.0
                            Hv = blk(p, p2) * blk(q, q2) * scal
                            if p != q and p2 != q2:
                                Hv = Hv + blk(p, q2) * blk(q, p2)
                            H[(block - 1) * U : block * U, (block2 - 1) * U : block2 * U] += Hv
        H = np.triu(H) + np.triu(H, 1).T
        self.hess_mat = H
        self.hess_updated = True

    def _partial_prod(self, arr, use_symm_prod):  # :314-358
        self.grad()
        a, one_d = _as2d(arr)
        prod = np.zeros_like(a)
        for Pk, Lc, F in zip(self.Ps, self.LamFL, self.LamFLP):
            for j in range(a.shape[1]):
                M = self._lambda(a[:, j], Pk)
                M = sla.solve_triangular(Lc, M, lower=True)
                M = sla.solve_triangular(Lc, M.T, lower=True).T  # M / L'
                LRUR = M @ F
                left = LRUR if use_symm_prod else F
                col = np.zeros(self.dim)
                self._block_diag_prod(col, left, LRUR)
                prod[:, j] += col
        return prod[:, 0] if one_d else prod

    def hess_prod_slow(self, arr):  # :272-281
        if not self.use_hess_prod_slow_updated:
            self.update_use_hess_prod_slow()
        if not self.use_hess_prod_slow:
            return self.hess_prod(arr)
        return self._partial_prod(arr, False)

    def dder3(self, direction):  # :283-286
        return self._partial_prod(direction, True)


class WSOSInterpEpiNormEucl(_GenericHessCone):
    """src/Cones/wsosinterpepinormeucl.jl: the arrow matrix [Lam(p_1) Lam(p_r)'; Lam(p_r) Lam(p_1) ...] is handled
    through the Cholesky of Lam(p_1) and of its Schur complement `mat`, per P_k.  Products go through the stored
    Hessian (generic oracles of Cones.jl)."""

    def setup_extra_data(self):  # :90-123
        self.R, self.U, self.Ps = self.spec.R, self.spec.U, self.spec.Ps

    def _views(self, vec):
        U = self.U
        return [vec[r * U : (r + 1) * U] for r in range(self.R)]

    def update_feas(self):  # :131-178
        pv = self._views(self.point)
        self.L11, self.Lm, self.LiL = [], [], []
        self.is_feas_flag = True
        for P in self.Ps:
            Lam11 = P.T @ (pv[0][:, None] * P)
            c, info = lapack.dpotrf(Lam11, lower=1, clean=1)
            if info != 0:
                self.is_feas_flag = False
                break
            mat = Lam11.copy()
            LiL_k = []
            for r in range(1, self.R):
                M = sla.solve_triangular(c, P.T @ (pv[r][:, None] * P), lower=True)  # L11 \ Lam(p_r)
                LiL_k.append(M)
                mat -= M.T @ M
            cm, info = lapack.dpotrf(mat, lower=1, clean=1)
            if info != 0:
                self.is_feas_flag = False
                break
            self.L11.append(c)
            self.Lm.append(cm)
            self.LiL.append(LiL_k)
        self.feas_updated = True
        return self.is_feas_flag

    def is_dual_feas(self):  # :180-183
        return bool(np.all(self.dual_point[: self.U] > EPS))

    def update_grad(self):  # :185-223
        U, R = self.U, self.R
        g = np.zeros(self.dim)
        self.L11iP, self.matLiP, self.edge = [], [], []
        for k, P in enumerate(self.Ps):
            A = sla.solve_triangular(self.L11[k], P.T, lower=True)  # Lam11 L^-1 P'
            B = sla.solve_triangular(self.Lm[k], P.T, lower=True)  # mat L^-1 P'
            E = [sla.solve_triangular(self.Lm[k], -(self.LiL[k][r].T @ A), lower=True) for r in range(R - 1)]
            self.L11iP.append(A)
            self.matLiP.append(B)
            self.edge.append(E)
            g[:U] -= np.sum(A * A, axis=0) + np.sum(B * B, axis=0)
            for r in range(1, R):
                g[r * U : (r + 1) * U] -= 2.0 * np.sum(E[r - 1] * B, axis=0)
                g[:U] -= np.sum(E[r - 1] * E[r - 1], axis=0)
        self.grad_vec = g
        self.grad_updated = True
        return g

    def update_hess(self):  # :225-296
        self.grad()
        U, R = self.U, self.R
        R2 = R - 2
        H = np.zeros((self.dim, self.dim))
        blk = lambda r: slice(r * U, (r + 1) * U)  # noqa: E731  (0-based block r)
        for k in range(len(self.Ps)):
            A, Bm, E = self.L11iP[k], self.matLiP[k], self.edge[k]
            P11 = A.T @ A
            Bk = [[None] * R for _ in range(R)]  # P Lam^-1 P blocks, (r, s)
            Bk[0][0] = Bm.T @ Bm
            for r in range(1, R):
                Bk[r][0] = E[r - 1].T @ Bm
                Bk[r][r] = E[r - 1].T @ E[r - 1] + P11
                for r2 in range(1, r):
                    Bk[r][r2] = E[r - 1].T @ E[r2 - 1]
            for r in range(R):
                for s_ in range(r + 1, R):
                    Bk[r][s_] = Bk[s_][r].T
            H[blk(0), blk(0)] -= np.triu(P11 * P11) * R2
            H[blk(0), blk(0)] += Bk[0][0] ** 2
            for r in range(1, R):
                for s_ in range(0, r):
                    UU = Bk[r][s_] ** 2
                    H[blk(0), blk(0)] += UU + UU.T
                    H[blk(0), blk(r)] += Bk[s_][0] * Bk[s_][r]
                H[blk(0), blk(0)] += Bk[r][r] ** 2
                H[blk(0), blk(r)] += Bk[r][0] * Bk[r][r]
                for s_ in range(r + 1, R):
                    H[blk(0), blk(r)] += Bk[s_][0] * Bk[s_][r]
                H[blk(r), blk(r)] += Bk[r][0] * Bk[r][0].T
                H[blk(r), blk(r)] += Bk[0][0] * Bk[r][r]
                for r2 in range(r + 1, R):
                    H[blk(r), blk(r2)] += Bk[r][0] * Bk[r2][0].T
                    H[blk(r), blk(r2)] += Bk[0][0] * Bk[r2][r].T
        H[:, U:] *= 2.0
        self.hess_mat = np.triu(H) + np.triu(H, 1).T  # Symmetric(H, :U)
        self.hess_updated = True

    def dder3(self, direction):  # :298-396
        U, R = self.U, self.R
        d = self._views(np.asarray(direction, dtype=np.float64))
        out = np.zeros(self.dim)
        for k in range(len(self.Ps)):
            A = self.L11iP[k]
            LL = (A * d[0][None, :]) @ A.T
            LU = (np.triu(LL) + np.triu(LL, 1).T) @ A  # Hermitian(LL): its upper triangle
            out[:U] += np.sum(LU * LU, axis=0)
        out[:U] *= 2 - R
        for k in range(len(self.Ps)):
            A, Bm, E = self.L11iP[k], self.matLiP[k], self.edge[k]
            L = A.shape[0]
            scaled_diag = A * d[0][None, :]
            scaled_pt = Bm * d[0][None, :]
            scaled_row, scaled_col = [], []
            for r in range(1, R):
                scaled_pt = scaled_pt + E[r - 1] * d[r][None, :]
                scaled_row.append(Bm * d[r][None, :] + E[r - 1] * d[0][None, :])
                scaled_col.append(A * d[r][None, :])
            col = np.zeros((R * L, L))
            row = []
            col[:L] = scaled_pt @ Bm.T
            diag = np.zeros((L, L))
            for r in range(1, R):
                col[:L] += scaled_row[r - 1] @ E[r - 1].T
                row.append(scaled_row[r - 1] @ A.T)
                col[r * L : (r + 1) * L] = scaled_col[r - 1] @ Bm.T + scaled_diag @ E[r - 1].T
                diag = scaled_diag @ A.T
            ctgs = np.hstack([Bm] + [E[r - 1] for r in range(1, R)])  # L x (U R)
            half = col @ ctgs  # (R L) x (U R)
            for r in range(1, R):
                half[:L, r * U : (r + 1) * U] += row[r - 1] @ A
                half[r * L : (r + 1) * L, r * U : (r + 1) * U] += diag @ A
            out[:U] += np.sum(half[:, :U] ** 2, axis=0)
            for r in range(1, R):
                hb = half[:, r * U : (r + 1) * U]
                out[r * U : (r + 1) * U] += 2.0 * np.sum(hb * half[:, :U], axis=0)
                out[:U] += np.sum(hb * hb, axis=0)
        return out


class WSOSInterpEpiNormOne(_GenericHessCone):
    """src/Cones/wsosinterpepinormone.jl: one Schur complement mat_r = Lam(p_1) - Lam(p_r) Lam(p_1)^-1 Lam(p_r) per
    polynomial r and P_k.  The reference's hess_prod! / inv_hess_prod! exploit the arrow shape of the Hessian
    (:332-427); here they are the generic products with the same stored Hessian."""

    def setup_extra_data(self):  # :101-142
        self.R, self.U, self.Ps = self.spec.R, self.spec.U, self.spec.Ps

    def _views(self, vec):
        U = self.U
        return [vec[r * U : (r + 1) * U] for r in range(self.R)]

    def update_feas(self):  # :161-218
        pv = self._views(self.point)
        self.L11, self.Lm, self.LiL = [], [], []
        self.is_feas_flag = True
        for P in self.Ps:
            Lam11 = P.T @ (pv[0][:, None] * P)
            c, info = lapack.dpotrf(Lam11, lower=1, clean=1)
            if info != 0:
                self.is_feas_flag = False
                break
            LiL_k, Lm_k = [], []
            for r in range(1, self.R):
                M = sla.solve_triangular(c, P.T @ (pv[r][:, None] * P), lower=True)
                cm, info = lapack.dpotrf(Lam11 - M.T @ M, lower=1, clean=1)
                if info != 0:
                    self.is_feas_flag = False
                    break
                LiL_k.append(M)
                Lm_k.append(cm)
            if not self.is_feas_flag:
                break
            self.L11.append(c)
            self.LiL.append(LiL_k)
            self.Lm.append(Lm_k)
        self.feas_updated = True
        return self.is_feas_flag

    def is_dual_feas(self):  # :220-224
        return bool(np.all(self.dual_point[: self.U] > EPS))

    def update_grad(self):  # :226-265
        U, R = self.U, self.R
        g = np.zeros(self.dim)
        self.L11iP, self.S11, self.S12 = [], [], []
        for k, P in enumerate(self.Ps):
            A = sla.solve_triangular(self.L11[k], P.T, lower=True)
            S12 = [sla.solve_triangular(self.Lm[k][r], -(self.LiL[k][r].T @ A), lower=True) for r in range(R - 1)]
            S11 = [sla.solve_triangular(self.Lm[k][r], P.T, lower=True) for r in range(R - 1)]
            self.L11iP.append(A)
            self.S11.append(S11)
            self.S12.append(S12)
            g[:U] -= R * np.sum(A * A, axis=0)
            for r in range(R - 1):
                g[:U] -= 2.0 * np.sum(S12[r] * S12[r], axis=0)
                g[(r + 1) * U : (r + 2) * U] -= 2.0 * np.sum(S12[r] * S11[r], axis=0)
        self.grad_vec = g
        self.grad_updated = True
        return g

    def update_hess(self):  # :267-330
        self.grad()
        U, R = self.U, self.R
        diag = [np.zeros((U, U)) for _ in range(R)]
        edge = [np.zeros((U, U)) for _ in range(R - 1)]
        for k in range(len(self.Ps)):
            A = self.L11iP[k]
            P11 = A.T @ A
            for r in range(R - 1):
                P2 = self.S12[k][r].T @ self.S11[k][r]
                P1 = self.S12[k][r].T @ self.S12[k][r] + P11
                uu = np.triu(2.0 * (P1 * P1 + P2 * P2))
                diag[0] += uu
                diag[r + 1] += uu
                edge[r] += np.triu(4.0 * P1 * P2)
            diag[0] -= np.triu(P11 * P11) * (R - 2)
        H = np.zeros((self.dim, self.dim))
        H[:U, :U] = diag[0]
        for r in range(R - 1):
            sl = slice((r + 1) * U, (r + 2) * U)
            H[:U, sl] = edge[r] + np.triu(edge[r], 1).T  # copytri!(edge, 'U')
            H[sl, sl] = diag[r + 1]
        self.hess_mat = np.triu(H) + np.triu(H, 1).T
        self.hess_updated = True

    def dder3(self, direction):  # :429-521
        U, R = self.U, self.R
        d = self._views(np.asarray(direction, dtype=np.float64))
        out = np.zeros(self.dim)
        for k in range(len(self.Ps)):
            A = self.L11iP[k]
            LL = (A * d[0][None, :]) @ A.T
            LU = (np.triu(LL) + np.triu(LL, 1).T) @ A
            out[:U] += np.sum(LU * LU, axis=0)
        out[:U] *= 2 - R
        for k in range(len(self.Ps)):
            A, S11, S12 = self.L11iP[k], self.S11[k], self.S12[k]
            L = A.shape[0]
            dir22 = A * d[0][None, :]
            dir22_A = dir22 @ A.T
            LU_last = dir22_A @ A
            for s_ in range(R - 1):
                D = d[s_ + 1][None, :]
                d11 = S11[s_] * d[0][None, :] + S12[s_] * D
                d12 = S11[s_] * D + S12[s_] * d[0][None, :]
                d21 = A * D
                half = np.zeros((2 * L, 2 * U))
                LL = d11 @ S12[s_].T
                half[:L, :U] = LL @ S12[s_]
                half[:L, U:] = LL @ S11[s_]
                LL = d12 @ S12[s_].T
                half[:L, :U] += LL @ S11[s_]
                half[:L, U:] += LL @ S12[s_]
                LL = d21 @ S12[s_].T
                half[L:, :U] = LL @ S12[s_]
                half[L:, U:] = LL @ S11[s_]
                LL = dir22 @ S12[s_].T
                half[L:, :U] += LL @ S11[s_]
                half[L:, U:] += LL @ S12[s_]
                half[:L, :U] += (d11 @ A.T) @ A
                half[:L, U:] += (d12 @ A.T) @ A
                half[L:, :U] += (d21 @ A.T) @ A
                half[L:, U:] += LU_last
                out[:U] += np.sum(half[:, :U] ** 2, axis=0) + np.sum(half[:, U:] ** 2, axis=0)
                out[(s_ + 1) * U : (s_ + 2) * U] += 2.0 * np.sum(half[:, U:] * half[:, :U], axis=0)
        return out


_CLASSES = {
    cs.NONNEGATIVE: Nonnegative,
    cs.EPINORMEUCL: EpiNormEucl,
    cs.EPIPERSQUARE: EpiPerSquare,
    cs.POSSEMIDEFTRI: PosSemidefTri,
    cs.WSOSINTERPNONNEGATIVE: WSOSInterpNonnegative,
    cs.WSOSINTERPPOSSEMIDEFTRI: WSOSInterpPosSemidefTri,
    cs.EPINORMINF: EpiNormInf,
    cs.DOUBLYNONNEGATIVETRI: DoublyNonnegativeTri,
    cs.POSSEMIDEFTRISPARSE: PosSemidefTriSparse,
    cs.WSOSINTERPEPINORMEUCL: WSOSInterpEpiNormEucl,
    cs.WSOSINTERPEPINORMONE: WSOSInterpEpiNormOne,
}


def make_cone(spec: cs.ConeSpec) -> Cone:
    return _CLASSES[spec.type_code](spec)


# ---------------------------------------------------------------------------
class CpuConeSet(ConeSet):
    """Per-cone loop over CPU cone objects, in the reference's order."""

    def __init__(self, specs: List[cs.ConeSpec]):
        super().__init__(specs)
        self.cones = [make_cone(s) for s in self.specs]
        self._setup_done = False

    def _sl(self, k):
        return slice(int(self.offs[k]), int(self.offs[k + 1]))

    def setup(self):
        for c in self.cones:
            c.setup_data()
        self._setup_done = True

    def load_point(self, primal, scal=1.0):
        for k, c in enumerate(self.cones):
            c.load_point(primal[self._sl(k)], None if scal == 1.0 else scal)

    def load_dual_point(self, dual):
        for k, c in enumerate(self.cones):
            c.load_dual_point(dual[self._sl(k)])

    def reset_data(self):
        for c in self.cones:
            c.reset_data()

    def is_feas(self):
        return all(c.is_feas() for c in self.cones)

    def is_dual_feas(self):
        return all(c.is_dual_feas() for c in self.cones)

    def grad(self):
        return np.concatenate([c.grad() for c in self.cones]) if self.K else np.zeros(0)

    def point(self):
        return np.concatenate([c.point for c in self.cones]) if self.K else np.zeros(0)

    def check_neighborhood(self, irtmu, use_max_prox, agg0, proxsqr_bound):  # search.jl:141-158
        agg = agg0
        for c in self.cones:
            in_prox = False
            if c.is_feas() and c.is_dual_feas() and c.check_numerics():
                proxsqr = c.get_proxsqr(irtmu, use_max_prox)
                agg = max(agg, proxsqr) if use_max_prox else agg + proxsqr
                in_prox = agg < proxsqr_bound
            if not in_prox:
                return False, agg
        return True, agg

    def _map(self, arr, name):
        out = np.empty_like(arr)
        for k, c in enumerate(self.cones):
            out[self._sl(k)] = getattr(c, name)(arr[self._sl(k)])
        return out

    def hess_prod(self, arr):
        return self._map(arr, "hess_prod")

    def inv_hess_prod(self, arr):
        return self._map(arr, "inv_hess_prod")

    def hess_prod_slow(self, arr):
        return self._map(arr, "hess_prod_slow")

    def block_hess_prod(self, arr):  # qrchol.jl:94-105
        out = np.empty_like(arr)
        for k, c in enumerate(self.cones):
            fn = c.inv_hess_prod if c.use_dual_barrier else c.hess_prod
            out[self._sl(k)] = fn(arr[self._sl(k)])
        return out

    def dder3(self, direction):
        return self._map(direction, "dder3")

    def use_dder3(self):
        return np.array([c.use_dder3() for c in self.cones], dtype=bool)
