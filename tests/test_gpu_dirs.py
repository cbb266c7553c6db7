"""Device-resident, batched get_directions (hyb_get_directions, csrc/dirs.cu) against the
reference's one-direction-at-a-time pattern (systemsolvers/common.jl:22-189) driven from the
host through the per-call C-ABI, and against the CPU oracle."""
import ctypes as C

import numpy as np
import pytest

import hypatia_b200  # noqa: F401
from hypatia_b200 import instances
from hypatia_b200.gpu import make_gpu_solver
from hypatia_b200.host.models import Point
from hypatia_b200.host.syscommon import apply_lhs, get_directions
from hypatia_b200.host import stepper as st

pytestmark = pytest.mark.gpu


def _host_residual(stepper, s, dir_vec, rhs_vec):
    """inf-norm of LHS * dir - rhs evaluated by the host code (apply_lhs, common.jl:86-125)."""
    stepper.dir.vec[:] = dir_vec
    res = apply_lhs(stepper, s)
    return float(np.max(np.abs(res.vec - rhs_vec)))


def _same_solution(stepper, s, dev_vec, host_vec, rhs_vec):
    """Two solves of the same (possibly ill-conditioned) system agree when they have the same
    residual quality; their difference is then bounded by cond * eps, checked loosely."""
    scale = np.max(np.abs(host_vec)) + 1e-300
    assert np.max(np.abs(dev_vec - host_vec)) <= 1e-6 * scale
    r_dev = _host_residual(stepper, s, dev_vec, rhs_vec)
    r_host = _host_residual(stepper, s, host_vec, rhs_vec)
    floor = 1e-13 * max(1.0, np.max(np.abs(rhs_vec)), scale)
    assert r_dev <= 10.0 * r_host + floor, (r_dev, r_host)


def _capture_state(model_fn, n_iters, **kw):
    """Run n_iters iterations on the GPU path and return the solver frozen at the next
    iteration's state (point accepted, cones loaded, residuals computed)."""
    import os

    # these tests drive the host-vector protocol (right-hand sides built by host/stepper.py): the host must
    # hold the iterate and the cone batch's mirror, which the device-resident stepper does not keep
    old = os.environ.get("HYB_HOST_STEPPER")
    os.environ["HYB_HOST_STEPPER"] = "1"
    try:
        s = make_gpu_solver(0, iter_limit=n_iters, **kw)
    finally:
        if old is None:
            del os.environ["HYB_HOST_STEPPER"]
        else:
            os.environ["HYB_HOST_STEPPER"] = old
    s.load(model_fn())
    s.solve()
    return s


# The two-solver protocol comparisons below run under the tolerances of the reference's own benchmark
# protocol (benchmarks/natvsext/run.jl:44-60): at the default eps^0.5 tolerances the last iterations of
# block_multicone sit on the rounding floor, where the iteration count depends on the last bits of every
# kernel, and "the same solve on two paths" stops being a well-posed comparison.
KW = dict(tol_rel_opt=1e-7, tol_feas=1e-7, tol_abs_opt=1e-10, tol_infeas=1e-10)

MODELS = {
    "block_multicone": lambda: instances.block_multicone(N=96, seed=3, psd_side=8),
    "linearopt": lambda: instances.linearopt(m=30, n=80, seed=1),
    "polymin_wsos": lambda: instances.polymin(n=1, halfdeg=4, seed=2),
    # primal-form WSOS instance: the cone uses the dual barrier (z/s roles swapped, inverse products)
    "wsos_dual_barrier": lambda: instances.KNOWN_ANSWER["wsosinterpnonnegative1"]()[0],
    "epipersquare4": lambda: instances.KNOWN_ANSWER["epipersquare4"]()[0],
}


@pytest.mark.parametrize("name", list(MODELS))
@pytest.mark.parametrize("n_iters", [1, 6])
def test_batched_directions_match_host_pattern(name, n_iters):
    s = _capture_state(MODELS[name], n_iters)
    ss, stepper, m = s.syssolver, s.stepper, s.model
    assert ss.batched_dirs
    # a fresh factorization at the frozen point, like the top of CombinedStepper.step
    ss.update_lhs(s)
    rhs1, rhs2 = Point(m.n, m.p, m.q), Point(m.n, m.p, m.q)
    d1, d2 = Point(m.n, m.p, m.q), Point(m.n, m.p, m.q)
    st.update_rhs_cent(s, rhs1)
    st.update_rhs_pred(s, rhs2)
    s.max_ref_steps = 5
    for cutoff in (s.res_norm_cutoff, 0.0):  # 0.0 forces the refinement rounds
        s.res_norm_cutoff = cutoff
        stepper.last_num_ref = 0
        ss._dev_const_stale = True
        ss.get_directions_batch(stepper, s, (rhs1, rhs2), (d1, d2))
        nref_dev = stepper.last_num_ref
        # host pattern: one direction at a time through solve_system / apply_lhs
        ss._host_const_stale = True
        got = []
        nref_host = 0
        for rhs in (rhs1, rhs2):
            stepper.rhs.vec[:] = rhs.vec
            stepper.last_num_ref = 0
            get_directions(stepper, s)
            nref_host += stepper.last_num_ref
            got.append(stepper.dir.vec.copy())
        for dev, host, rhs in zip((d1, d2), got, (rhs1, rhs2)):
            _same_solution(stepper, s, dev.vec.copy(), host, rhs.vec)
        if cutoff > 0.0:
            assert nref_dev == nref_host
        else:
            assert nref_dev >= 2  # both columns went through at least one refinement round
    # adjustment pair
    st.update_rhs_centadj(s, rhs1, d1)
    st.update_rhs_predadj(s, rhs2, d2)
    a1, a2 = Point(m.n, m.p, m.q), Point(m.n, m.p, m.q)
    ss.get_directions_batch(stepper, s, (rhs1, rhs2), (a1, a2))
    for rhs, dev in ((rhs1, a1), (rhs2, a2)):
        stepper.rhs.vec[:] = rhs.vec
        get_directions(stepper, s)
        _same_solution(stepper, s, dev.vec.copy(), stepper.dir.vec.copy(), rhs.vec)


@pytest.mark.parametrize("name", ["block_multicone", "linearopt"])
def test_batched_and_unbatched_solves_agree(name):
    a = make_gpu_solver(0, trace=True, batched_dirs=True)
    b = make_gpu_solver(0, trace=True, batched_dirs=False)
    for s in (a, b):
        s.load(MODELS[name]())
        s.solve()
    assert a.status == b.status == "Optimal"
    assert abs(a.num_iters - b.num_iters) <= 1
    assert abs(a.primal_obj - b.primal_obj) <= 1e-7 * (1 + abs(b.primal_obj))  # north_star objective tolerance
    n = len(b.trace) // 2
    assert np.allclose([r["mu"] for r in a.trace[:n]], [r["mu"] for r in b.trace[:n]], rtol=1e-8)


@pytest.mark.parametrize("N,R", [(100, 2), (300, 3), (1000, 4), (129, 1)])
def test_multi_rhs_wavefront_solves(N, R):
    """potrs with R right-hand sides in one pair of wavefront kernels == R single solves."""
    from hypatia_b200.lib import Context, dptr, f64

    rng = np.random.default_rng(N + R)
    A = rng.standard_normal((N, N))
    A = A @ A.T + N * np.eye(N)
    B = rng.standard_normal((N, R))
    ctx = Context(0)
    _, X, info = ctx.posdef_solve(A, B)
    assert info == 0
    ref = np.linalg.solve(A, B)
    assert np.max(np.abs(X - ref)) <= 1e-10 * np.max(np.abs(ref))


@pytest.mark.parametrize("name", list(MODELS))
def test_device_line_search_matches_host_loop(name, monkeypatch):
    """check_cone_points on the device (hyb_search_gates / hyb_search_check) takes the same
    accept / reject decisions as the host loop of search.jl:72-162."""
    dev = make_gpu_solver(0, trace=True, **KW)
    assert dev.syssolver.device_search
    monkeypatch.setenv("HYB_HOST_SEARCH", "1")
    host = make_gpu_solver(0, trace=True, **KW)
    assert not host.syssolver.device_search
    for s in (dev, host):
        s.load(MODELS[name]())
        s.solve()
    assert dev.status == host.status
    assert abs(dev.num_iters - host.num_iters) <= 1
    if host.status == "Optimal":
        assert abs(dev.primal_obj - host.primal_obj) <= 1e-7 * (1 + abs(host.primal_obj))
    # same step lengths and candidate counts while the iterates are away from the rounding floor
    n = max(3, min(len(dev.trace), len(host.trace)) // 2)
    assert [r["alpha"] for r in dev.trace[:n]] == [r["alpha"] for r in host.trace[:n]]
    assert [r["n_cand"] for r in dev.trace[:n]] == [r["n_cand"] for r in host.trace[:n]]
    assert np.allclose([r["mu"] for r in dev.trace[:n]], [r["mu"] for r in host.trace[:n]], rtol=1e-9)
    assert np.allclose([r["prox"] for r in dev.trace[:n]], [r["prox"] for r in host.trace[:n]], rtol=1e-6, atol=1e-12)


@pytest.mark.parametrize("name", list(MODELS))
def test_device_resident_stepper_matches_host_vector_protocol(name, monkeypatch):
    """hyb_step_* (iterate, right-hand-side builders, point update on the device) against the same solve with
    the host building the right-hand sides and holding the iterate (steppers/common.jl:14-119, combined.jl:129-177)."""
    dev = make_gpu_solver(0, trace=True, **KW)
    assert dev.syssolver.device_stepper
    monkeypatch.setenv("HYB_HOST_STEPPER", "1")
    host = make_gpu_solver(0, trace=True, **KW)
    assert not host.syssolver.device_stepper and host.syssolver.device_search
    for s in (dev, host):
        s.load(MODELS[name]())
        s.solve()
    assert dev.status == host.status
    assert abs(dev.num_iters - host.num_iters) <= 1
    if host.status == "Optimal":
        assert abs(dev.primal_obj - host.primal_obj) <= 1e-7 * (1 + abs(host.primal_obj))
        assert np.abs(dev.get_x() - host.get_x()).max() <= 1e-6 * (1 + np.abs(host.get_x()).max())
    n = max(3, min(len(dev.trace), len(host.trace)) // 2)
    assert [r["alpha"] for r in dev.trace[:n]] == [r["alpha"] for r in host.trace[:n]]
    for key in ("mu", "x_norm_res", "z_norm_res", "gap", "tau", "kap"):
        a = np.array([r[key] for r in dev.trace[:n]])
        b = np.array([r[key] for r in host.trace[:n]])
        assert np.all(np.abs(a - b) <= 1e-8 * np.maximum(np.abs(b), 1e-5)), key


def _at_point(solver, model_fn, point_vec):
    """Bring a freshly loaded solver to the state the top of CombinedStepper.step sees at the given iterate:
    point set, mu, cones loaded at point / sqrt(mu) (search.jl:144-145), convergence parameters, cutoff."""
    import math

    solver.load(model_fn())
    solver._setup_solver()
    solver._setup_loaded()
    solver._initialize_point()
    solver._setup_stepping()
    solver.point.vec[:] = point_vec
    solver.calc_mu()
    cones, pt = solver.cones, solver.point
    cones.load_point(cones.primal_of(pt.z, pt.s), 1.0 / math.sqrt(solver.mu))
    cones.load_dual_point(cones.dual_of(pt.z, pt.s))
    cones.reset_data()
    assert cones.is_feas()
    solver._calc_convergence_params()
    solver.res_norm_cutoff = 1e-4 * max(solver.x_norm_res, solver.y_norm_res, solver.z_norm_res, solver.tau_feas)
    solver.syssolver.update_lhs(solver)


@pytest.mark.parametrize("name", ["block_multicone", "linearopt", "wsos_dual_barrier", "epipersquare4"])
@pytest.mark.parametrize("n_iters", [2, 5])
def test_directions_against_the_oracle_at_the_same_iterate(name, n_iters, monkeypatch):
    """The four directions of one iteration (cent, pred, centadj, predadj; systemsolvers/common.jl:22-189,
    steppers/common.jl:14-119) computed by the CUDA path and by the CPU oracle from the SAME iterate (the oracle's
    after n_iters iterations): same right-hand sides, same directions up to the conditioning of the system, same
    residual quality."""
    from hypatia_b200.host.solver import Solver
    from oracle.cones import CpuConeSet
    from oracle.qrchol import QRCholDenseSystemSolver

    ref = Solver(CpuConeSet, QRCholDenseSystemSolver(), iter_limit=n_iters)
    ref.load(MODELS[name]())
    ref.solve()
    point_vec = ref.point.vec.copy()
    cpu = Solver(CpuConeSet, QRCholDenseSystemSolver())
    _at_point(cpu, MODELS[name], point_vec)
    monkeypatch.setenv("HYB_HOST_STEPPER", "1")  # right-hand sides from the host builders on both sides
    gpu = make_gpu_solver(0)
    _at_point(gpu, MODELS[name], point_vec)
    assert abs(gpu.mu - cpu.mu) <= 1e-14 * cpu.mu
    m = cpu.model
    dirs = {}
    for s in (cpu, gpu):
        stp = s.stepper
        out = []
        for builder, base in ((st.update_rhs_cent, None), (st.update_rhs_pred, None), (st.update_rhs_centadj, 0),
                              (st.update_rhs_predadj, 1)):
            if base is None:
                builder(s, stp.rhs)
            else:
                builder(s, stp.rhs, out[base][1])
            rhs = stp.rhs.vec.copy()
            get_directions(stp, s)
            d = Point(m.n, m.p, m.q)
            d.vec[:] = stp.dir.vec
            out.append((rhs, d))
        dirs[id(s)] = out
    for i, ((rhs_c, d_c), (rhs_g, d_g)) in enumerate(zip(dirs[id(cpu)], dirs[id(gpu)])):
        scale_r = np.max(np.abs(rhs_c)) + 1e-300
        # oracle products agree: grad alone in the cent / pred right-hand sides; the adjustment right-hand sides
        # are dder3 of the backend's OWN cent / pred direction, so they inherit those directions' agreement
        # (1e-6 of their scale, below) through a product that is linear in it
        assert np.max(np.abs(rhs_g - rhs_c)) <= (1e-9 if i < 2 else 1e-7) * scale_r
        scale = np.max(np.abs(d_c.vec)) + 1e-300
        assert np.max(np.abs(d_g.vec - d_c.vec)) <= 1e-6 * scale
        # both satisfy their 6 x 6 system equally well.  Each residual is evaluated with the backend's own
        # oracles: the two backends' Hessians differ by ~1e-13 relative (the cancellation in dist = (u^2 - |w|^2)/2
        # and in X^-1 near the boundary amplifies the summation-order difference; measured, tools/diag_dir_acc.py),
        # so a direction is only self-consistent to that level with the OTHER backend's operator
        r_g = _host_residual(gpu.stepper, gpu, d_g.vec, rhs_g)
        r_c = _host_residual(cpu.stepper, cpu, d_c.vec, rhs_c)
        floor = 1e-13 * max(1.0, scale_r, scale)
        assert r_g <= 10.0 * r_c + floor, (r_g, r_c)
        r_x = _host_residual(cpu.stepper, cpu, d_g.vec, rhs_c)  # cross check, at the operators' own agreement
        assert r_x <= 10.0 * r_c + 1e-10 * max(1.0, scale_r, scale), (r_x, r_c)


@pytest.mark.parametrize("name", ["block_multicone", "linearopt", "epipersquare4"])
@pytest.mark.parametrize("host_stepper", [False, True])
def test_speculative_line_search_is_the_sequential_one(name, host_stepper, monkeypatch):
    """hyb_search_alpha (several schedule entries per call, decisions on the device, search.jl:72-98) takes the
    same decisions with the same numbers as one candidate per call through hyb_search_gates / hyb_search_check:
    identical step lengths, candidate counts and mu over the whole solve."""
    if host_stepper:
        monkeypatch.setenv("HYB_HOST_STEPPER", "1")
    spec = make_gpu_solver(0, trace=True, **KW)
    spec.load(MODELS[name]())
    spec.solve()
    assert spec.syssolver.spec_search
    monkeypatch.setenv("HYB_SPEC_SEARCH", "0")
    seq = make_gpu_solver(0, trace=True, **KW)
    seq.load(MODELS[name]())
    seq.solve()
    assert not seq.syssolver.spec_search
    assert spec.status == seq.status and spec.num_iters == seq.num_iters
    assert [r["alpha"] for r in spec.trace] == [r["alpha"] for r in seq.trace]
    assert [r["n_cand"] for r in spec.trace] == [r["n_cand"] for r in seq.trace]
    assert [r["mu"] for r in spec.trace] == [r["mu"] for r in seq.trace]
    assert [r["prox"] for r in spec.trace] == [r["prox"] for r in seq.trace]
    assert spec.primal_obj == seq.primal_obj
