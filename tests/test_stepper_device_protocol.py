"""CPU test of the host side of the device-resident stepper protocol (batched direction pairs with
kept direction slots, hyb_search_gates / hyb_search_check): a numpy stand-in implements the calls the
GPU system solver makes, with the semantics include/hypatia_b200.h documents, on top of the CPU
oracle.  Driving the stepper through that protocol must reproduce the classic path (one direction at
a time, check_cone_points on the host) step for step."""
import math

import numpy as np
import pytest

import hypatia_b200  # noqa: F401
from hypatia_b200 import instances
from hypatia_b200.host.solver import Solver
from hypatia_b200.host.syscommon import get_directions
from oracle.cones import CpuConeSet
from oracle.qrchol import QRCholDenseSystemSolver


class ProtocolConeSet(CpuConeSet):
    def note_device_point(self, primal, dual, scal):
        # the stand-in's search_check has loaded exactly this point already
        assert np.allclose(self.point(), np.asarray(primal) * scal, rtol=1e-14, atol=0)


class ProtocolSystemSolver(QRCholDenseSystemSolver):
    """Numpy stand-in for B200QRCholSystemSolver's batched / device-search entry points."""

    batched_dirs = True
    device_search = True

    def __init__(self):
        super().__init__()
        self.kept = {}
        self.calls = dict(batch=0, gates=0, check=0)

    def get_directions_batch(self, stepper, solver, rhs_points, dir_points, slots=None):
        self.calls["batch"] += 1
        for j, (rhs, out) in enumerate(zip(rhs_points, dir_points)):
            stepper.rhs.vec[:] = rhs.vec  # rhs_points[0] may be stepper.rhs itself
            get_directions(stepper, solver)
            out.vec[:] = stepper.dir.vec
            if slots is not None:  # hyb_dirs_keep
                self.kept[slots[j]] = (out.z.copy(), out.s.copy())

    def search_load_point(self, point):
        self.pz, self.ps = point.z.copy(), point.s.copy()
        self._solver_cones = None

    def search_gates(self, coefs):
        self.calls["gates"] += 1
        z, s = self.pz.copy(), self.ps.copy()
        dz = ds = None
        for k in (2, 3, 0, 1):  # pred, predadj, cent, centadj: the order of update_stepper_points
            if coefs[k] != 0.0:
                kz, ks = self.kept[k]
                dz = coefs[k] * kz if dz is None else dz + coefs[k] * kz
                ds = coefs[k] * ks if ds is None else ds + coefs[k] * ks
        if dz is not None:
            z += dz
            s += ds
        cones = self.cones_ref
        self.cand_primal, self.cand_dual = cones.primal_of(z, s), cones.dual_of(z, s)
        self.szk = cones.cone_dots(self.cand_primal, self.cand_dual)
        return float(self.szk.min()), float(self.szk.sum())

    def search_check(self, mu, min_prox, proxsqr_bound, use_max_prox, agg0):
        self.calls["check"] += 1
        cones = self.cones_ref
        sz_rel = self.szk / (mu * cones.nus)
        if np.any(sz_rel < min_prox) or np.any(cones.nus * (sz_rel - 1.0) ** 2 > proxsqr_bound):
            return False, agg0
        irtmu = 1.0 / math.sqrt(mu)
        cones.load_point(self.cand_primal, irtmu)
        cones.load_dual_point(self.cand_dual)
        cones.reset_data()
        return cones.check_neighborhood(irtmu, use_max_prox, agg0, proxsqr_bound)

    def load(self, solver):
        super().load(solver)
        self.cones_ref = solver.cones


CASES = {
    "block_multicone": lambda: instances.block_multicone(N=48, seed=2, psd_side=4),
    "linearopt": lambda: instances.linearopt(m=20, n=60, seed=3),
    "epipersquare4": lambda: instances.KNOWN_ANSWER["epipersquare4"]()[0],
    "wsos_dual_barrier": lambda: instances.KNOWN_ANSWER["wsosinterpnonnegative1"]()[0],
    "primalinfeas2": lambda: instances.KNOWN_ANSWER["primalinfeas2"]()[0],
}


@pytest.mark.parametrize("name", list(CASES))
def test_protocol_path_reproduces_classic_path(name):
    classic = Solver(CpuConeSet, QRCholDenseSystemSolver(), trace=True)
    classic.load(CASES[name]())
    classic.solve()
    ss = ProtocolSystemSolver()
    proto = Solver(ProtocolConeSet, ss, trace=True)
    proto.load(CASES[name]())
    proto.solve()
    assert ss.calls["batch"] == 2 * proto.num_iters and ss.calls["gates"] > 0 and ss.calls["check"] > 0
    assert proto.status == classic.status
    assert proto.num_iters == classic.num_iters
    for key in ("alpha", "n_cand", "n_ref"):
        assert [r[key] for r in proto.trace] == [r[key] for r in classic.trace], key
    assert np.allclose([r["mu"] for r in proto.trace], [r["mu"] for r in classic.trace], rtol=1e-10)
    assert np.allclose(proto.get_x(), classic.get_x(), rtol=1e-9, atol=1e-12)


class SpeculativeSystemSolver(ProtocolSystemSolver):
    """... plus hyb_search_alpha: several schedule entries per call, the first accepted one wins; every decision the
    device takes (scalar gates of search.jl:108-130, neighbourhood check) restated here with the same arithmetic."""

    spec_search = True

    def __init__(self):
        super().__init__()
        self.calls["alpha"] = 0
        self.windows = []

    def search_alpha(self, coefs, taukaps, nup1, min_prox, proxsqr_bound, use_max_prox):
        self.calls["alpha"] += 1
        self.windows.append(len(taukaps))
        eps = float(np.finfo(np.float64).eps)
        for i, (cf, taukap) in enumerate(zip(coefs, taukaps)):
            min_szk, sum_szk = self.search_gates(cf)
            if min_szk < eps:
                continue
            mu = (sum_szk + taukap) / nup1
            if mu < eps:
                continue
            rel = taukap / mu
            tp = (rel - 1.0) ** 2
            if rel < min_prox or tp > proxsqr_bound:
                continue
            ok, agg = self.search_check(mu, min_prox, proxsqr_bound, use_max_prox, tp)
            if ok:
                return i, mu, agg
        return -1, float("nan"), float("nan")


@pytest.mark.parametrize("name", list(CASES))
def test_speculative_search_windows_reproduce_the_sequential_search(name):
    """host/stepper.py:_search_alpha_speculative (window sizes, candidate counting, prev_sched, the fall-back chain of
    combined.jl:93-118 when a whole schedule is rejected) against the one-candidate-per-call loop of search.jl:72-98."""
    seq_ss = ProtocolSystemSolver()
    seq = Solver(ProtocolConeSet, seq_ss, trace=True)
    seq.load(CASES[name]())
    seq.solve()
    ss = SpeculativeSystemSolver()
    spec = Solver(ProtocolConeSet, ss, trace=True)
    spec.load(CASES[name]())
    spec.solve()
    assert ss.calls["alpha"] > 0 and max(ss.windows) >= 2
    assert ss.calls["alpha"] < seq_ss.calls["gates"]  # fewer host round trips than candidates
    assert spec.status == seq.status and spec.num_iters == seq.num_iters
    for key in ("alpha", "n_cand", "n_ref", "mu", "prox"):
        assert [r[key] for r in spec.trace] == [r[key] for r in seq.trace], key
    assert np.array_equal(spec.get_x(), seq.get_x())
