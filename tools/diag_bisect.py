"""Bisect GPU-vs-CPU deviation: variants of the GPU system solver where single steps are
replaced by host numpy (diagnostic only; never part of the product path)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import scipy.linalg as sla

import hypatia_b200  # noqa: F401
from hypatia_b200 import instances
from hypatia_b200.gpu import B200QRCholSystemSolver
from hypatia_b200.host.solver import Solver
from hypatia_b200.lib import Context

mode = sys.argv[2]


class Variant(B200QRCholSystemSolver):
    def load(self, solver):
        super().load(solver)
        self.Ghost = np.array(solver.model.G)

    def solve_subsystem3(self, solver, sol, rhs):
        if mode == "gpu":
            return super().solve_subsystem3(solver, sol, rhs)
        cones = solver.cones
        S = self.schur_matrix()
        S = np.triu(S) + np.triu(S, 1).T
        if mode == "hostsolve":  # everything on host except cones and S
            r = rhs.x + self.Ghost.T @ rhs.z
            x = sla.cho_solve(sla.cho_factor(S), r)
            Gx = self.Ghost @ x
        elif mode == "devsolve":  # device potrs, host gemv
            r = rhs.x + self.Ghost.T @ rhs.z
            x = np.empty_like(r)
            from hypatia_b200.lib import dptr, f64
            rr = f64(r)
            self.ctx.check(self.lib.hyb_schur_solve(self.ctx.ptr, dptr(rr), dptr(x), 1), "solve")
            xh = sla.cho_solve(sla.cho_factor(S), r)
            err = np.abs(x - xh).max() / max(np.abs(xh).max(), 1e-300)
            if solver.num_iters >= 5:
                print("SOLVE MISMATCH iter", solver.num_iters, "err", err, "\nS=", S, "\nr=", r, "\nx_dev=", x, "\nx_host=", xh, "kind", self.last_fact_kind)
            Gx = self.Ghost @ x
        elif mode == "devgemv":  # host solve, device gemv
            r = rhs.x + self.Gt_mul(rhs.z)
            x = sla.cho_solve(sla.cho_factor(S), r)
            Gx = self.G_mul(x)
        z = cones.block_hess_prod(Gx) - rhs.z
        sol.x[:] = x
        sol.z[:] = z


name = sys.argv[1]
fn = lambda: instances.KNOWN_ANSWER[name]()[0]
ctx = Context(0)
ss = Variant(ctx)
s = Solver(ss.cones_factory, ss, trace=True)
s.load(fn())
s.solve()
print(mode, "status", s.status, "iters", s.num_iters, "obj", s.primal_obj)
