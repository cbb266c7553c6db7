"""Cone-set interface the host driver talks to.

The reference loops over ``model.cones`` and calls the per-cone oracle set of
src/Cones/Cones.jl:34-304.  Here the driver talks to ONE object that owns all K
cones of the model and takes / returns full q-vectors (or q x m arrays) laid out
by ``cone_offs``.  Two backends implement it: ``oracle.cones.CpuConeSet`` (the
CPU checker, a per-cone loop like the reference) and ``gpu.GpuConeBatch`` (the
product: batched CUDA kernels behind the C-ABI).
"""
from __future__ import annotations

from typing import List, Tuple

import numpy as np

from ..cones_spec import ConeSpec, build_cone_idxs


class ConeSet:
    def __init__(self, specs: List[ConeSpec]):
        self.specs = list(specs)
        self.K = len(self.specs)
        self.q = int(sum(s.dim for s in self.specs))
        self.offs = build_cone_idxs(self.q, self.specs)
        self.starts = self.offs[:-1].copy()
        self.dims = np.diff(self.offs)
        self.nus = np.array([s.nu for s in self.specs], dtype=np.float64)
        self.nu = float(self.nus.sum()) if self.K else 0.0
        self.dual_mask = np.array([s.use_dual_barrier for s in self.specs], dtype=bool)
        # row-level mask: True where the cone's "primal" slice is z (dual barrier), point.jl:52-59
        self.dual_rows = np.repeat(self.dual_mask, self.dims) if self.K else np.zeros(0, bool)
        self.any_dual = bool(self.dual_mask.any())
        # runs of consecutive cones of equal dimension: (first cone, number of cones, dim, first row)
        self._runs = []
        k = 0
        while k < self.K:
            k2 = k
            while k2 + 1 < self.K and self.dims[k2 + 1] == self.dims[k]:
                k2 += 1
            self._runs.append((k, k2 - k + 1, int(self.dims[k]), int(self.offs[k])))
            k = k2 + 1

    # ---- view helpers (point.jl:50-59) ---------------------------------
    def primal_of(self, z: np.ndarray, s: np.ndarray) -> np.ndarray:
        return np.where(self.dual_rows, z, s) if self.any_dual else s

    def dual_of(self, z: np.ndarray, s: np.ndarray) -> np.ndarray:
        return np.where(self.dual_rows, s, z) if self.any_dual else z

    def cone_dots(self, a: np.ndarray, b: np.ndarray) -> np.ndarray:
        """K per-cone inner products <a_k, b_k> (fixed left-to-right order)."""
        if self.K == 0:
            return np.zeros(0)
        if 4 * len(self._runs) <= self.K:
            # many small cones of equal size (cfg4: 10^5 second-order cones): one reshaped row sum per
            # run instead of a reduceat over 10^5 tiny segments (several times faster)
            out = np.empty(self.K)
            for k0, cnt, dim, r0 in self._runs:
                sl = slice(r0, r0 + cnt * dim)
                prod = (a[sl] * b[sl]).reshape(cnt, dim)
                if dim <= 16:
                    acc = prod[:, 0].copy()
                    for j in range(1, dim):
                        acc += prod[:, j]
                    out[k0 : k0 + cnt] = acc
                else:
                    out[k0 : k0 + cnt] = prod.sum(axis=1)
            return out
        return np.add.reduceat(a * b, self.starts)

    def expand(self, per_cone: np.ndarray) -> np.ndarray:
        return np.repeat(per_cone, self.dims)

    def initial_point(self) -> np.ndarray:
        """set_initial_point! per cone, concatenated."""
        out = np.zeros(self.q)
        cache = {}  # the initial point depends on the cone's type and shape only (10^5 equal cones: one evaluation)
        for k, spec in enumerate(self.specs):
            key = (spec.type_code, spec.dim, getattr(spec, "R", 0), getattr(spec, "U", 0), id(spec) if hasattr(spec, "rows0") else 0)
            v = cache.get(key)
            if v is None:
                v = cache[key] = spec.initial_point()
            out[self.offs[k] : self.offs[k + 1]] = v
        return out

    # ---- oracle interface (Cones.jl) ------------------------------------
    def setup(self) -> None:  # setup_data!
        raise NotImplementedError

    def load_point(self, primal: np.ndarray, scal: float = 1.0) -> None:
        raise NotImplementedError

    def load_dual_point(self, dual: np.ndarray) -> None:
        raise NotImplementedError

    def reset_data(self) -> None:
        raise NotImplementedError

    def is_feas(self) -> bool:
        raise NotImplementedError

    def is_dual_feas(self) -> bool:
        raise NotImplementedError

    def grad(self) -> np.ndarray:
        raise NotImplementedError

    def point(self) -> np.ndarray:
        """Currently loaded (scaled) primal point, cone.point per cone."""
        raise NotImplementedError

    def check_neighborhood(
        self, irtmu: float, use_max_prox: bool, agg0: float, proxsqr_bound: float
    ) -> Tuple[bool, float]:
        """Cone loop of check_cone_points (search.jl:141-158) on the loaded
        candidate: feas, dual feas, check_numerics, proxsqr aggregation."""
        raise NotImplementedError

    def hess_prod(self, arr: np.ndarray) -> np.ndarray:
        raise NotImplementedError

    def inv_hess_prod(self, arr: np.ndarray) -> np.ndarray:
        raise NotImplementedError

    def hess_prod_slow(self, arr: np.ndarray) -> np.ndarray:
        raise NotImplementedError

    def block_hess_prod(self, arr: np.ndarray) -> np.ndarray:
        """qrchol.jl:94-105: inv_hess_prod for dual-barrier cones, else hess_prod."""
        raise NotImplementedError

    def dder3(self, direction: np.ndarray) -> np.ndarray:
        raise NotImplementedError

    def use_dder3(self) -> np.ndarray:
        return np.ones(self.K, dtype=bool)
