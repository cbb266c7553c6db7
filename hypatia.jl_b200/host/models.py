"""Model and Point containers (host side).

Mirrors src/Models/Models.jl:21-86 (Model) and src/Solvers/point.jl:12-62
(Point layout ``vec = [x(n); y(p); z(q); tau; s(q); kap]``).
"""
from __future__ import annotations

from typing import List

import numpy as np

from ..cones_spec import ConeSpec, build_cone_idxs


def _dense(M, rows: int, cols: int) -> np.ndarray:
    """densify! (Models.jl:78-86): accept ndarray, scipy sparse, or a scalar
    standing for a UniformScaling (lambda * I)."""
    if np.isscalar(M):
        assert rows == cols, "UniformScaling needs a square block"
        return np.asfortranarray(float(M) * np.eye(rows))
    if hasattr(M, "toarray"):
        M = M.toarray()
    M = np.asarray(M, dtype=np.float64)
    if M.ndim == 1:
        M = M.reshape(rows, cols)
    assert M.shape == (rows, cols), f"shape {M.shape} != {(rows, cols)}"
    return np.asfortranarray(M)


class Model:
    """Primal general conic form  min c'x : b - Ax = 0, h - Gx in K  (Models.jl:21-61)."""

    def __init__(self, c, A, b, G, h, cones: List[ConeSpec], obj_offset: float = 0.0):
        self.c = np.array(c, dtype=np.float64).ravel()
        self.b = np.array(b, dtype=np.float64).ravel()
        self.h = np.array(h, dtype=np.float64).ravel()
        self.n = self.c.size
        self.p = self.b.size
        self.q = self.h.size
        self.obj_offset = float(obj_offset)
        self.A = _dense(A, self.p, self.n)
        self.G = _dense(G, self.q, self.n)
        self.cones = list(cones)
        self.cone_offs = build_cone_idxs(self.q, self.cones)
        self.nu = float(sum(cone.nu for cone in self.cones)) if self.cones else 0.0

    def copy_shallow(self) -> "Model":
        m = Model.__new__(Model)
        m.__dict__.update(self.__dict__)
        return m


class Point:
    """point.jl:31-62.  All fields are views into ``vec``."""

    def __init__(self, n: int, p: int, q: int):
        self.n, self.p, self.q = n, p, q
        self.tau_idx = n + p + q
        self.vec = np.zeros(self.tau_idx + q + 2)
        v = self.vec
        self.x = v[:n]
        self.y = v[n : n + p]
        self.z = v[n + p : n + p + q]
        self.s = v[self.tau_idx + 1 : self.tau_idx + 1 + q]
        self.ztsk = v[n + p :]

    @property
    def tau(self) -> float:
        return float(self.vec[self.tau_idx])

    @tau.setter
    def tau(self, val: float):
        self.vec[self.tau_idx] = val

    @property
    def kap(self) -> float:
        return float(self.vec[-1])

    @kap.setter
    def kap(self, val: float):
        self.vec[-1] = val


class SubPoint:
    """3x3-subsystem point [x; y; z] (systemsolvers/common.jl:191-209)."""

    def __init__(self, n: int, p: int, q: int):
        self.vec = np.zeros(n + p + q)
        self.x = self.vec[:n]
        self.y = self.vec[n : n + p]
        self.z = self.vec[n + p :]
