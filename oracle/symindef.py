"""ORACLE — TEST INFRASTRUCTURE ONLY.

CPU restatement of the reference's dense symmetric-indefinite system solver,
src/Solvers/systemsolvers/symindef.jl:
  setup_rhs3        :38-62
  load              :231-247  (lower triangle of [0 . .; A 0 .; G 0 -H~])
  update_lhs        :249-270  (z-z blocks: -inv_hess for primal-barrier cones, -hess for dual-barrier ones;
                               symm_fact_copy!, src/linearalgebra/dense.jl:205-219)
  solve_subsystem3  :272-280
The factorization is LAPACK's dsytrf_rook on the LOWER triangle, like `bunchkaufman!(Symmetric(lhs, :L), true)`.
The explicit Hessian blocks (`Cones.hess` / `Cones.inv_hess`) are the oracle cones' products with the identity.
solve_subsystem4 / solve_system are the shared code of systemsolvers/common.jl (host/syscommon.py).
"""
from __future__ import annotations

import numpy as np

from hypatia_b200.host.syscommon import QRCholSystemSolver as _SystemSolverBase

from .dense import increase_diag
from .lapack_rook import dsytrf_rook, dsytrs_rook


class _SymFact:
    def __init__(self, ldl, ipiv, success, shifted):
        self.ldl, self.ipiv, self.success, self.shifted = ldl, ipiv, success, shifted

    @property
    def kind_code(self) -> int:
        return -1 if not self.success else (2 if self.shifted else 1)

    def solve(self, b: np.ndarray) -> np.ndarray:
        return dsytrs_rook(self.ldl, self.ipiv, b, uplo=b"L")


def symm_fact_copy(mat: np.ndarray) -> _SymFact:  # dense.jl:205-219 (lower triangle referenced)
    ldl, ipiv, info = dsytrf_rook(mat, uplo=b"L")
    if info == 0:
        return _SymFact(ldl, ipiv, True, False)
    ldl, ipiv, info = dsytrf_rook(increase_diag(np.array(mat, order="F")), uplo=b"L")
    return _SymFact(ldl, ipiv, info == 0, True)


class SymIndefDenseSystemSolver(_SystemSolverBase):
    def load(self, solver) -> None:  # symindef.jl:231-247
        model = solver.model
        n, p, q = model.n, model.p, model.q
        lhs = np.zeros((n + p + q, n + p + q), order="F")
        if p:
            lhs[n : n + p, :n] = model.A
        lhs[n + p :, :n] = model.G
        self.lhs_sub = lhs
        self.fact = None
        self.fact_kinds = []
        self.setup_point_sub(solver)

    def update_lhs(self, solver) -> None:  # symindef.jl:249-270
        model, cs = solver.model, solver.cones
        z0 = model.n + model.p
        lhs = self.lhs_sub
        for k, c in enumerate(cs.cones):
            lo, hi = z0 + int(cs.offs[k]), z0 + int(cs.offs[k + 1])
            eye = np.eye(hi - lo)
            H_k = c.hess_prod(eye) if c.use_dual_barrier else c.inv_hess_prod(eye)
            lhs[lo:hi, lo:hi] = -0.5 * (H_k + H_k.T)
        self.fact = symm_fact_copy(lhs)
        self.fact_kinds.append(self.fact.kind_code)
        if not self.fact.success:
            print("symmetric linear system factorization failed")
        self.solve_subsystem3(solver, self.sol_const, self.rhs_const)

    def setup_rhs3(self, solver, rhs, sol, rhs_sub) -> None:  # symindef.jl:38-62
        cs = solver.cones
        for k, c in enumerate(cs.cones):
            sl = slice(int(cs.offs[k]), int(cs.offs[k + 1]))
            if c.use_dual_barrier:
                rhs_sub.z[sl] = -rhs.z[sl] - rhs.s[sl]
            else:
                rhs_sub.z[sl] = -c.inv_hess_prod(rhs.s[sl]) - rhs.z[sl]

    def solve_subsystem3(self, solver, sol, rhs) -> None:  # symindef.jl:272-280
        sol.vec[:] = self.fact.solve(rhs.vec)
