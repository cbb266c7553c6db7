"""ORACLE — TEST INFRASTRUCTURE ONLY.

CPU restatement of the reference's dense QR-Cholesky system solver,
src/Solvers/systemsolvers/qrchol.jl:
  setup_rhs3        :23-44
  solve_subsystem3  :46-92   (p == 0 after reduction: r = x + G'z; x = chol \\ r; z = H~(Gx) - z)
  block_hess_prod!  :94-105
  load              :145-187
  update_lhs        :189-207
  update_lhs_fact   :209-265 (sqrt-cone rows packed from row 0 -> syrk; others gemm-accumulated; posdef_fact_copy!)
Models always reach this solver reduced (process.jl:301-338), so p == 0 and
Ap_Q == I; the p > 0 branches of the reference are not restated.
"""
from __future__ import annotations

import time

import numpy as np

from hypatia_b200.host.syscommon import QRCholSystemSolver

from .dense import outer_prod_upper, posdef_fact_copy


class QRCholDenseSystemSolver(QRCholSystemSolver):
    def load(self, solver) -> None:
        model = solver.model
        assert model.p == 0, "oracle QRChol expects a reduced model (p == 0)"
        n, q = model.n, model.q
        self.GQ2 = model.G  # Ap_Q == I
        self.HGQ2 = np.zeros((q, n), order="F")
        self.lhs = np.zeros((n, n), order="F")
        self.fact = None
        self.time_assemble = 0.0
        self.fact_kinds = []  # per update_lhs: 0 Cholesky, 1 Bunch-Kaufman, 2 shifted + BK, -1 failed (dense.jl:228-249)
        self.setup_point_sub(solver)

    def update_lhs(self, solver) -> None:
        model, cones = solver.model, solver.cones
        if model.n:
            self.update_lhs_fact(solver)
        self.rhs_const.z[:] = cones.block_hess_prod(model.h)
        self.solve_subsystem3(solver, self.sol_const, self.rhs_const)

    def update_lhs_fact(self, solver) -> None:
        model = solver.model
        cs = solver.cones  # CpuConeSet
        nmp = model.n
        t0 = time.time()
        use_sqrt = [c.use_sqrt_hess_oracles(nmp) for c in cs.cones]
        self.use_sqrt_hess_cones = use_sqrt
        GQ2, HGQ2 = self.GQ2, self.HGQ2
        if any(use_sqrt):
            idx = 0
            for k, c in enumerate(cs.cones):
                if not use_sqrt[k]:
                    continue
                arr_k = GQ2[cs.offs[k] : cs.offs[k + 1], :]
                q_k = arr_k.shape[0]
                if c.use_dual_barrier:
                    HGQ2[idx : idx + q_k, :] = c.inv_sqrt_hess_prod(arr_k)
                else:
                    HGQ2[idx : idx + q_k, :] = c.sqrt_hess_prod(arr_k)
                idx += q_k
            lhs = outer_prod_upper(HGQ2[:idx, :])
        else:
            lhs = np.zeros((nmp, nmp), order="F")
        for k, c in enumerate(cs.cones):
            if use_sqrt[k]:
                continue
            arr_k = GQ2[cs.offs[k] : cs.offs[k + 1], :]
            prod_k = c.inv_hess_prod(arr_k) if c.use_dual_barrier else c.hess_prod(arr_k)
            lhs += arr_k.T @ prod_k
        self.lhs = lhs
        t1 = time.time()
        self.fact = posdef_fact_copy(lhs)
        t2 = time.time()
        self.fact_kinds.append(self.fact.kind_code)
        self.time_assemble += t1 - t0
        solver.time_upfact += t2 - t1
        if not self.fact.success:
            print("positive definite linear system factorization failed")

    def setup_rhs3(self, solver, rhs, sol, rhs_sub) -> None:
        cones = solver.cones
        if cones.any_dual:
            # dual-barrier cones: H^-1(-zrhs - srhs); primal: H zrhs - srhs
            out = np.empty(cones.q)
            for k, c in enumerate(cones.cones):
                sl = slice(int(cones.offs[k]), int(cones.offs[k + 1]))
                if c.use_dual_barrier:
                    out[sl] = c.inv_hess_prod(-rhs.z[sl] - rhs.s[sl])
                else:
                    out[sl] = -c.hess_prod(rhs.z[sl]) - rhs.s[sl]
            # note axpby!(-1, rhs_s, -1, prod): -(H zrhs) - srhs (qrchol.jl:40)
            rhs_sub.z[:] = out
        else:
            rhs_sub.z[:] = -cones.hess_prod(rhs.z) - rhs.s

    def solve_subsystem3(self, solver, sol, rhs) -> None:
        model, cones = solver.model, solver.cones
        sol.vec[:] = rhs.vec
        x, z = sol.x, sol.z
        if model.n:
            r = x + self.GQ2.T @ z
            x[:] = self.fact.solve(r)
            Gx = self.GQ2 @ x
        else:
            Gx = np.zeros(model.q)
        z[:] = cones.block_hess_prod(Gx) - z
