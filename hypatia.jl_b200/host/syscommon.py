"""Direction solving shared by QRChol system solvers (host side).

Follows src/Solvers/systemsolvers/common.jl:22-83 (get_directions with
iterative refinement), :86-125 (apply_lhs), :133-158 (solve_system), :161-189
(solve_subsystem4), :191-224 (sub points, dot_obj) and the abstract
``QRCholSystemSolver`` (qrchol.jl:21).  Backends subclass ``QRCholSystemSolver``
and provide load / update_lhs / setup_rhs3 / solve_subsystem3 / free_memory:
``oracle.qrchol.QRCholDenseSystemSolver`` (CPU checker) and
``gpu.B200QRCholSystemSolver`` (product).
"""
from __future__ import annotations

import math

import numpy as np

from .models import Point, SubPoint


import os as _os

_DEBUG = bool(_os.environ.get("HYB_DEBUG"))


def _norm_inf(v: np.ndarray) -> float:
    return float(np.max(np.abs(v))) if v.size else 0.0


class QRCholSystemSolver:
    device_G = False  # True when the backend serves G / G' products from the device

    # ---- to be provided by the backend ------------------------------------
    def load(self, solver) -> None:
        raise NotImplementedError

    def update_lhs(self, solver) -> None:
        raise NotImplementedError

    def setup_rhs3(self, solver, rhs: Point, sol: Point, rhs_sub: SubPoint) -> None:
        raise NotImplementedError

    def solve_subsystem3(self, solver, sol: SubPoint, rhs: SubPoint) -> None:
        raise NotImplementedError

    def free_memory(self) -> None:
        pass

    # ---- common.jl:191-220 -------------------------------------------------
    def setup_point_sub(self, solver) -> None:
        m = solver.model
        self.sol_sub = SubPoint(m.n, m.p, m.q)
        self.rhs_sub = SubPoint(m.n, m.p, m.q)
        self.rhs_const = SubPoint(m.n, m.p, m.q)
        self.sol_const = SubPoint(m.n, m.p, m.q)
        self.set_point_sub_rhs(solver)

    def set_point_sub_rhs(self, solver) -> None:
        m = solver.model
        self.rhs_const.x[:] = -m.c
        self.rhs_const.y[:] = m.b
        self.rhs_const.z[:] = m.h

    # ---- common.jl:161-189 -------------------------------------------------
    def solve_subsystem4(self, solver, sol: Point, rhs: Point) -> None:
        model = solver.model
        rhs_sub, sol_sub = self.rhs_sub, self.sol_sub
        rhs_sub.x[:] = rhs.x
        rhs_sub.y[:] = -rhs.y
        self.setup_rhs3(solver, rhs, sol, rhs_sub)
        self.solve_subsystem3(solver, sol_sub, rhs_sub)
        sol_const = self.sol_const
        tau_num = rhs.tau + rhs.kap + dot_obj(model, sol_sub)
        taubar = solver.point.tau
        tau_denom = solver.mu / taubar / taubar - dot_obj(model, sol_const)
        sol_tau = tau_num / tau_denom
        dim3 = sol_sub.vec.size
        sol.vec[:dim3] = sol_sub.vec + sol_tau * sol_const.vec
        sol.tau = sol_tau

    # ---- common.jl:133-158 -------------------------------------------------
    def solve_system(self, solver, sol: Point, rhs: Point) -> None:
        model = solver.model
        self.solve_subsystem4(solver, sol, rhs)
        tau = sol.tau
        sol.s[:] = model.h * tau - rhs.z - solver.G_mul(sol.x)
        taubar = solver.point.tau
        sol.kap = -solver.mu / taubar / taubar * tau + rhs.kap


def dot_obj(model, point) -> float:  # common.jl:222-224
    return float(model.c @ point.x) + float(model.b @ point.y) + float(model.h @ point.z)


def apply_lhs(stepper, solver) -> Point:  # common.jl:86-125
    model = solver.model
    dir_, res = stepper.dir, stepper.temp
    cones = solver.cones
    tau_dir, kap_dir = dir_.tau, dir_.kap
    res.x[:] = solver.Gt_mul(dir_.z) + tau_dir * model.c
    res.z[:] = model.h * tau_dir - dir_.s - solver.G_mul(dir_.x)
    res.tau = -float(model.c @ dir_.x) - float(model.h @ dir_.z) - kap_dir
    if model.p:
        res.x[:] += model.A.T @ dir_.y
        res.y[:] = tau_dir * model.b - model.A @ dir_.x
        res.tau = res.tau - float(model.b @ dir_.y)
    prim_dir = cones.primal_of(dir_.z, dir_.s)
    dual_dir = cones.dual_of(dir_.z, dir_.s)
    res.s[:] = cones.hess_prod_slow(prim_dir) + dual_dir
    tau = solver.point.tau
    res.kap = solver.mu / tau * tau_dir / tau + kap_dir
    return res


def get_directions(stepper, solver, min_impr_tol: float = 0.5) -> Point:  # common.jl:22-83
    rhs, dir_, dir_temp, res = stepper.rhs, stepper.dir, stepper.dir_temp, stepper.temp
    ss = solver.syssolver
    res_norm_cutoff = solver.res_norm_cutoff
    max_ref_steps = solver.max_ref_steps

    ss.solve_system(solver, dir_, rhs)
    if max_ref_steps == 0:
        return dir_

    dir_temp[:] = dir_.vec
    apply_lhs(stepper, solver)
    res.vec -= rhs.vec
    res_norm = _norm_inf(res.vec)

    if res_norm > res_norm_cutoff:
        is_prev_slow = False
        prev_res_norm = res_norm
        for _ in range(max_ref_steps):
            stepper.last_num_ref += 1
            ss.solve_system(solver, dir_, res)
            dir_.vec[:] = dir_temp - dir_.vec  # axpby!(true, dir_temp, -1, dir.vec)
            apply_lhs(stepper, solver)
            res.vec -= rhs.vec
            res_norm_new = _norm_inf(res.vec)
            if _DEBUG:
                print(f"    refine: {res_norm:.3e} -> {res_norm_new:.3e}")
            if res_norm_new >= res_norm:
                dir_.vec[:] = dir_temp
                break
            dir_temp[:] = dir_.vec
            res_norm = res_norm_new
            if res_norm < res_norm_cutoff:
                break
            is_curr_slow = res_norm > min_impr_tol * prev_res_norm
            if is_prev_slow and is_curr_slow:
                break
            prev_res_norm = res_norm
            is_prev_slow = is_curr_slow

    if _DEBUG:
        print(f"  get_directions: res_norm {res_norm:.3e} cutoff {res_norm_cutoff:.3e} refs {stepper.last_num_ref} |dir| {_norm_inf(dir_.vec):.3e}")
    assert not math.isnan(res_norm), "NaN residual in direction solve"
    solver.worst_dir_res = max(solver.worst_dir_res, res_norm)
    return dir_
