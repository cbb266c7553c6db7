"""Combined predict-and-center stepper, RHS builders and step-distance search
(host side; callers of the hot path — SURVEY §8a-D).

Follows src/Solvers/steppers/combined.jl:13-196, steppers/common.jl:14-119 and
src/Solvers/search.jl:12-162.  Per-cone scalar work (s'z gates, dder3 identity
gate) is vectorised over the cone offsets with fixed-order segment sums.
"""
from __future__ import annotations

import math
import time

import numpy as np

from .models import Point
from .syscommon import get_directions

EPS = float(np.finfo(np.float64).eps)

# search.jl:48-69
DEFAULT_ALPHA_SCHED = (
    0.9999, 0.999, 0.99, 0.97, 0.95, 0.9, 0.85, 0.8, 0.7, 0.6, 0.5, 0.3, 0.1, 0.05, 0.01, 0.005, 0.001, 0.0005,
)


class StepSearcher:
    """search.jl:12-46"""

    def __init__(self, nu: float, min_prox=0.01, prox_bound=0.99, use_max_prox=True, alpha_sched=DEFAULT_ALPHA_SCHED):
        self.min_prox = min_prox
        self.prox_bound = prox_bound
        self.use_max_prox = use_max_prox
        self.alpha_sched = tuple(alpha_sched)
        self.nup1 = nu + 1.0
        self.prev_sched = 0
        self.prox = 0.0


class CombinedStepper:
    def __init__(self, shift_sched: int = 0, **searcher_options):
        self.shift_sched = shift_sched
        self.searcher_options = searcher_options
        self.last_num_cand = 0
        self.last_num_ref = 0

    def load(self, solver) -> None:  # combined.jl:41-58
        m = solver.model
        self.prev_alpha = 1.0
        self.rhs = Point(m.n, m.p, m.q)
        self.rhs2 = Point(m.n, m.p, m.q)  # second right-hand side of a batched pair
        self.dir = Point(m.n, m.p, m.q)
        self.temp = Point(m.n, m.p, m.q)
        self.dir_cent = Point(m.n, m.p, m.q)
        self.dir_centadj = Point(m.n, m.p, m.q)
        self.dir_pred = Point(m.n, m.p, m.q)
        self.dir_predadj = Point(m.n, m.p, m.q)
        self.dir_temp = np.zeros_like(self.rhs.vec)
        self.searcher = StepSearcher(m.nu, **self.searcher_options)
        self.unadj_only = self.cent_only = False

    def expect_improvement(self) -> bool:
        return True

    # ---- one iteration (combined.jl:60-125) ------------------------------
    def step(self, solver) -> bool:
        point = solver.point
        rhs, dir_ = self.rhs, self.dir
        self.last_num_cand = 0
        self.last_num_ref = 0

        t0 = time.time()
        solver.syssolver.update_lhs(solver)
        solver.time_upsys += time.time() - t0

        dev = bool(getattr(solver, "device_stepper", False))
        if dev:
            # Device-resident stepper (hyb_step_*): right-hand sides are built on the device from the device
            # iterate, the directions stay there; the host sees their (tau, kap) entries only.
            ss = solver.syssolver
            t0 = time.time()
            ss.step_rhs(0, 0, solver)  # cent
            ss.step_rhs(2, 1, solver)  # pred
            solver.time_uprhs += time.time() - t0
            t0 = time.time()
            (self.dir_cent.tau, self.dir_cent.kap), (self.dir_pred.tau, self.dir_pred.kap) = ss.step_directions(self, solver, 2)
            ss.dirs_keep(0, 0)
            ss.dirs_keep(1, 2)
            solver.time_getdir += time.time() - t0
            t0 = time.time()
            ss.step_rhs(1, 0, solver, self.dir_cent.tau)  # centadj
            ss.step_rhs(3, 1, solver, self.dir_pred.tau)  # predadj
            solver.time_uprhs += time.time() - t0
            t0 = time.time()
            (self.dir_centadj.tau, self.dir_centadj.kap), (self.dir_predadj.tau, self.dir_predadj.kap) = ss.step_directions(
                self, solver, 2)
            ss.dirs_keep(0, 1)
            ss.dirs_keep(1, 3)
            solver.time_getdir += time.time() - t0
            builders = ()
        elif getattr(solver.syssolver, "batched_dirs", False):
            # Device-resident direction solves: the two unadjusted directions are independent of
            # each other, and each adjustment needs only its own unadjusted direction
            # (combined.jl:74-95), so the four solves run as two batched pairs.
            ss = solver.syssolver
            # direction slots on the device for the line search: 0 cent, 1 centadj, 2 pred, 3 predadj
            for (b1, d1), (b2, d2), needs_dir, slots in (
                ((update_rhs_cent, self.dir_cent), (update_rhs_pred, self.dir_pred), False, (0, 2)),
                ((update_rhs_centadj, self.dir_centadj), (update_rhs_predadj, self.dir_predadj), True, (1, 3)),
            ):
                t0 = time.time()
                if needs_dir:
                    b1(solver, self.rhs, self.dir_cent)
                    b2(solver, self.rhs2, self.dir_pred)
                else:
                    b1(solver, self.rhs)
                    b2(solver, self.rhs2)
                solver.time_uprhs += time.time() - t0
                t0 = time.time()
                ss.get_directions_batch(self, solver, (self.rhs, self.rhs2), (d1, d2), slots)
                solver.time_getdir += time.time() - t0
            builders = ()
        else:
            builders = (
                (update_rhs_cent, self.dir_cent, False),
                (update_rhs_centadj, self.dir_centadj, True),
                (update_rhs_pred, self.dir_pred, False),
                (update_rhs_predadj, self.dir_predadj, True),
            )
        for builder, store, needs_dir in builders:
            t0 = time.time()
            if needs_dir:
                builder(solver, rhs, dir_)
            else:
                builder(solver, rhs)
            solver.time_uprhs += time.time() - t0
            t0 = time.time()
            get_directions(self, solver)
            solver.time_getdir += time.time() - t0
            store.vec[:] = dir_.vec

        self.unadj_only = self.cent_only = False
        t0 = time.time()
        self._device_search = bool(getattr(solver.syssolver, "device_search", False)) and solver.cones.K > 0
        self._device_stepper = dev
        if self._device_search and not dev:
            solver.syssolver.search_load_point(point)
        alpha = self.search_alpha(point, solver)
        if alpha == 0.0:
            self.unadj_only = True  # combined without adjustment
            alpha = self.search_alpha(point, solver)
            if alpha == 0.0:
                self.cent_only = True  # centering with adjustment
                self.unadj_only = False
                alpha = self.search_alpha(point, solver)
                if alpha == 0.0:
                    self.unadj_only = True  # centering without adjustment
                    alpha = self.search_alpha(point, solver)
                    if alpha == 0.0:
                        print("warning: cannot step in centering direction")
                        solver.status = "NumericalFailure"
                        self.prev_alpha = alpha
                        solver.time_search += time.time() - t0
                        return False
        solver.time_search += time.time() - t0
        if dev:
            cf = self._coefs(alpha)
            dtau, dkap = self._combine_taukap(cf)
            solver._zs_dev = solver.syssolver.step_point_update(cf)
            point.tau += dtau
            point.kap += dkap
        else:
            self.update_stepper_points(alpha, point, False)
        self.prev_alpha = alpha
        return True

    # ---- combined.jl:129-177 ----------------------------------------------
    def update_stepper_points(self, alpha: float, point: Point, ztsk_only: bool) -> None:
        if ztsk_only:
            cand = self.temp.ztsk
            cand[:] = point.ztsk
            sel = lambda pt: pt.ztsk  # noqa: E731
        else:
            cand = point.vec
            sel = lambda pt: pt.vec  # noqa: E731
        dir_cent, dir_pred = sel(self.dir_cent), sel(self.dir_pred)
        if self.unadj_only:
            if self.cent_only:
                cand += alpha * dir_cent
            else:
                cand += alpha * dir_pred + (1 - alpha) * dir_cent
        else:
            dir_centadj = sel(self.dir_centadj)
            alpha_sqr = alpha * alpha
            if self.cent_only:
                cand += alpha * dir_cent + alpha_sqr * dir_centadj
            else:
                dir_predadj = sel(self.dir_predadj)
                am1 = 1 - alpha
                cand += alpha * dir_pred + alpha_sqr * dir_predadj + am1 * dir_cent + (am1 * am1) * dir_centadj

    def start_sched(self) -> int:  # combined.jl:179-182 (1-based)
        if self.shift_sched <= 0:
            return 1
        return max(1, self.searcher.prev_sched - self.shift_sched)

    # ---- search.jl:72-95 ---------------------------------------------------
    def search_alpha(self, point: Point, solver) -> float:
        searcher = self.searcher
        sched = self.start_sched()
        if getattr(self, "_device_search", False) and getattr(solver.syssolver, "spec_search", False):
            return self._search_alpha_speculative(point, solver, sched)
        while sched <= len(searcher.alpha_sched):
            alpha = searcher.alpha_sched[sched - 1]
            self.last_num_cand += 1
            if getattr(self, "_device_search", False):
                accepted = self.check_cone_points_device(solver, point, alpha)
            else:
                self.update_stepper_points(alpha, point, True)
                accepted = self.check_cone_points(solver)
            if accepted:
                searcher.prev_sched = sched
                return alpha
            sched += 1
        searcher.prev_sched = sched
        if getattr(self, "_device_search", False) and not getattr(self, "_device_stepper", False):
            solver.cones.reset_data()  # the device holds a rejected candidate: reload on next use
        return 0.0

    def _search_alpha_speculative(self, point: Point, solver, sched: int) -> float:
        """search.jl:72-98 with several schedule entries per device call (hyb_search_alpha): the entries are
        evaluated in schedule order with every accept / reject decision taken on the device; the first accepted
        one wins, exactly as in the one-at-a-time loop.  The first window reaches one entry past the one accepted
        by the previous search (step lengths change slowly between iterations), later windows take four."""
        searcher, ss, cones = self.searcher, solver.syssolver, solver.cones
        n = len(searcher.alpha_sched)
        min_prox, bound = searcher.min_prox, searcher.prox_bound**2
        hi = min(n, max(sched + 1, getattr(searcher, "prev_sched", sched) + 1))
        while sched <= n:
            cands = []
            for k in range(sched, hi + 1):
                alpha = searcher.alpha_sched[k - 1]
                cf = self._coefs(alpha)
                dtau, dkap = self._combine_taukap(cf)
                tau, kap = point.tau + dtau, point.kap + dkap
                taukap = tau * kap
                if min(tau, kap, taukap) < EPS:  # search.jl:108-111, decided without the device
                    continue
                cands.append((k, alpha, cf, taukap))
            if cands:
                idx, mu, agg = ss.search_alpha([c[2] for c in cands], [c[3] for c in cands], searcher.nup1, min_prox, bound,
                                               searcher.use_max_prox)
                if idx >= 0:
                    k, alpha = cands[idx][0], cands[idx][1]
                    self.last_num_cand += k - sched + 1
                    searcher.prox = math.sqrt(agg)
                    searcher.prev_sched = k
                    if getattr(self, "_device_stepper", False):
                        cones.note_device_resident(1.0 / math.sqrt(mu))
                    else:  # the device holds the candidate; refresh the host mirror of the cone batch once
                        self.update_stepper_points(alpha, point, True)
                        cand = self.temp
                        cones.note_device_point(cones.primal_of(cand.z, cand.s), cones.dual_of(cand.z, cand.s),
                                                1.0 / math.sqrt(mu))
                    return alpha
            self.last_num_cand += hi - sched + 1
            sched = hi + 1
            hi = min(n, sched + 3)
        searcher.prev_sched = sched
        if not getattr(self, "_device_stepper", False):
            cones.reset_data()  # the device holds a rejected candidate: reload on next use
        return 0.0

    def _coefs(self, alpha: float):
        """Coefficients of (cent, centadj, pred, predadj) in update_stepper_points (combined.jl:146-173)."""
        if self.unadj_only:
            return (alpha, 0.0, 0.0, 0.0) if self.cent_only else (1 - alpha, 0.0, alpha, 0.0)
        if self.cent_only:
            return (alpha, alpha * alpha, 0.0, 0.0)
        am1 = 1 - alpha
        return (am1, am1 * am1, alpha, alpha * alpha)

    def _combine_taukap(self, cf):
        """(d tau, d kap) of update_stepper_points for coefficients cf, same left-to-right sum (combined.jl:146-173)."""
        dirs = (self.dir_cent, self.dir_centadj, self.dir_pred, self.dir_predadj)
        dtau = dkap = None
        for k in (2, 3, 0, 1):
            if cf[k] != 0.0:
                dtau = cf[k] * dirs[k].tau if dtau is None else dtau + cf[k] * dirs[k].tau
                dkap = cf[k] * dirs[k].kap if dkap is None else dkap + cf[k] * dirs[k].kap
        return (dtau or 0.0), (dkap or 0.0)

    # ---- search.jl:100-162 with the O(q) work on the device (hyb_search_gates / hyb_search_check) ----
    def check_cone_points_device(self, solver, point: Point, alpha: float) -> bool:
        searcher, ss, cones = self.searcher, solver.syssolver, solver.cones
        min_prox = searcher.min_prox
        proxsqr_bound = searcher.prox_bound**2
        cf = self._coefs(alpha)
        dirs = (self.dir_cent, self.dir_centadj, self.dir_pred, self.dir_predadj)
        # tau, kap of the candidate: same left-to-right sum as update_stepper_points
        dtau = dkap = None
        for k in (2, 3, 0, 1):
            if cf[k] != 0.0:
                dtau = cf[k] * dirs[k].tau if dtau is None else dtau + cf[k] * dirs[k].tau
                dkap = cf[k] * dirs[k].kap if dkap is None else dkap + cf[k] * dirs[k].kap
        tau, kap = point.tau + (dtau or 0.0), point.kap + (dkap or 0.0)
        taukap = tau * kap
        if min(tau, kap, taukap) < EPS:
            return False
        min_szk, sum_szk = ss.search_gates(cf)
        if min_szk < EPS:
            return False
        mu = (sum_szk + taukap) / searcher.nup1
        if mu < EPS:
            return False
        taukap_rel = taukap / mu
        if taukap_rel < min_prox:
            return False
        taukap_proxsqr = (taukap_rel - 1.0) ** 2
        if taukap_proxsqr > proxsqr_bound:
            return False
        ok, agg = ss.search_check(mu, min_prox, proxsqr_bound, searcher.use_max_prox, taukap_proxsqr)
        if not ok:
            return False
        searcher.prox = math.sqrt(agg)
        if getattr(self, "_device_stepper", False):
            cones.note_device_resident(1.0 / math.sqrt(mu))
            return True
        # accepted: the device holds the candidate; refresh the host mirror of the cone batch once
        self.update_stepper_points(alpha, point, True)
        cand = self.temp
        cones.note_device_point(cones.primal_of(cand.z, cand.s), cones.dual_of(cand.z, cand.s), 1.0 / math.sqrt(mu))
        return True

    # ---- search.jl:100-162 -------------------------------------------------
    def check_cone_points(self, solver) -> bool:
        searcher = self.searcher
        cand = self.temp
        cones = solver.cones
        min_prox = searcher.min_prox
        proxsqr_bound = searcher.prox_bound**2

        tau, kap = cand.tau, cand.kap
        taukap = tau * kap
        if min(tau, kap, taukap) < EPS:
            return False
        primal = cones.primal_of(cand.z, cand.s)
        dual = cones.dual_of(cand.z, cand.s)
        szk = cones.cone_dots(primal, dual)
        if szk.size and float(szk.min()) < EPS:
            return False
        mu = (float(szk.sum()) + taukap) / searcher.nup1
        if mu < EPS:
            return False
        taukap_rel = taukap / mu
        if taukap_rel < min_prox:
            return False
        taukap_proxsqr = (taukap_rel - 1.0) ** 2
        if taukap_proxsqr > proxsqr_bound:
            return False
        if szk.size:
            sz_rel = szk / (mu * cones.nus)
            if np.any(sz_rel < min_prox) or np.any(cones.nus * (sz_rel - 1.0) ** 2 > proxsqr_bound):
                return False

        irtmu = 1.0 / math.sqrt(mu)
        cones.load_point(primal, irtmu)
        cones.load_dual_point(dual)
        cones.reset_data()
        ok, agg = cones.check_neighborhood(irtmu, searcher.use_max_prox, taukap_proxsqr, proxsqr_bound)
        if not ok:
            return False
        searcher.prox = math.sqrt(agg)
        return True


# ---------------------------------------------------------------------------
# RHS builders (steppers/common.jl)
# ---------------------------------------------------------------------------
def update_rhs_pred(solver, rhs: Point) -> Point:  # common.jl:14-28
    rhs.x[:] = solver.x_residual
    rhs.y[:] = solver.y_residual
    rhs.z[:] = solver.z_residual
    rhs.tau = solver.tau_residual
    cones, point = solver.cones, solver.point
    rhs.s[:] = -cones.dual_of(point.z, point.s)
    rhs.kap = -point.kap
    return rhs


def update_rhs_cent(solver, rhs: Point) -> Point:  # common.jl:66-83
    rhs.x[:] = 0
    rhs.y[:] = 0
    rhs.z[:] = 0
    rhs.tau = 0.0
    cones, point = solver.cones, solver.point
    rtmu = math.sqrt(solver.mu)
    rhs.s[:] = -cones.dual_of(point.z, point.s) - rtmu * cones.grad()
    rhs.kap = -point.kap + solver.mu / point.tau
    return rhs


def _adj_common(solver, rhs: Point, dir_: Point, pred: bool) -> Point:
    """update_rhs_predadj (common.jl:30-63) / update_rhs_centadj (:86-119)."""
    rhs.vec[:] = 0
    cones = solver.cones
    rteps = math.sqrt(EPS)
    irtrtmu = 1.0 / math.sqrt(math.sqrt(solver.mu))
    prim_dir = cones.primal_of(dir_.z, dir_.s)
    prim_scal = irtrtmu * prim_dir
    if pred:
        H_prim = cones.hess_prod_slow(prim_dir)
    else:
        H_prim = cones.hess_prod_slow(prim_scal)
    dder3 = cones.dder3(prim_scal)
    dot1 = cones.cone_dots(dder3, cones.point())
    dot2 = cones.cone_dots(prim_scal, H_prim)
    if pred:
        dot2 = irtrtmu * dot2
    viol = np.abs(dot1 - dot2) / (rteps + np.abs(dot2))
    keep = cones.expand((viol < 1e-4) & cones.use_dder3())
    if pred:
        rhs.s[:] = np.where(keep, H_prim + dder3, 0.0)
    else:
        rhs.s[:] = np.where(keep, dder3, 0.0)
    taubar = solver.point.tau
    tau_dir_tau = dir_.tau / taubar
    if pred:
        rhs.kap = tau_dir_tau * solver.mu / taubar * (1 + tau_dir_tau)
    else:
        rhs.kap = tau_dir_tau * solver.mu / taubar * tau_dir_tau
    return rhs


def update_rhs_predadj(solver, rhs: Point, dir_: Point) -> Point:
    return _adj_common(solver, rhs, dir_, True)


def update_rhs_centadj(solver, rhs: Point, dir_: Point) -> Point:
    return _adj_common(solver, rhs, dir_, False)
