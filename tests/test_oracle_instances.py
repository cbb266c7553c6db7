"""Pins the CPU oracle (oracle/ + host driver) against the reference's RNG-free
known-answer instances (test/nativeinstances.jl; tolerance eps^(1/4), :37) and
the certificate checks of build_solve_check (:40-106)."""
import numpy as np
import pytest

import hypatia_b200  # noqa: F401
from hypatia_b200 import instances
from hypatia_b200.host.models import Model
from hypatia_b200.host.solver import Solver
from hypatia_b200 import cones_spec as cs
from oracle.cones import CpuConeSet
from oracle.qrchol import QRCholDenseSystemSolver

TOL = float(np.finfo(np.float64).eps) ** 0.25


def cpu_solver(**kw):
    return Solver(CpuConeSet, QRCholDenseSystemSolver(), **kw)


def check_certificates(solver, model, tol=TOL):
    """build_solve_check, test/nativeinstances.jl:40-106."""
    c, A, b, G, h = model.c, model.A, model.b, model.G, model.h
    x, y, z, s = solver.get_x(), solver.get_y(), solver.get_z(), solver.get_s()
    rt_tol = np.sqrt(tol)
    close = lambda a, bb, t=tol: np.allclose(a, bb, atol=t, rtol=t)  # noqa: E731
    if solver.status == "Optimal":
        assert close(solver.primal_obj, solver.dual_obj)
        assert close(c @ x + model.obj_offset, solver.primal_obj)
        assert close(-(b @ y) - h @ z + model.obj_offset, solver.dual_obj)
        assert close(A @ x, b)
        assert close(G @ x + s, h)
        assert close(G.T @ z + A.T @ y, -c)
        assert close(s @ z, 0.0, rt_tol)
    elif solver.status == "PrimalInfeasible":
        assert close(-(b @ y) - h @ z, solver.dual_obj)
        assert close(G.T @ z, -A.T @ y, rt_tol)
    elif solver.status == "DualInfeasible":
        assert close(c @ x, solver.primal_obj)
        assert close(G @ x, -s, rt_tol)
        assert close(A @ x, np.zeros(y.size), rt_tol)


def check_expected(solver, exp, tol=TOL):
    tol = tol * exp.get("tol_factor", 1.0)  # a few reference instances test at 10 * test_tol
    assert solver.status == exp["status"]
    close = lambda a, bb: np.allclose(a, bb, atol=tol, rtol=tol)  # noqa: E731
    if "primal_obj" in exp:
        assert close(solver.primal_obj, exp["primal_obj"])
    for key in ("x", "y", "s", "z"):
        if key in exp:
            assert close(getattr(solver, "get_" + key)(), exp[key]), key
    for idx, val in exp.get("x_at", {}).items():
        assert close(solver.get_x()[idx], val)
    if "x_norm" in exp:
        assert close(np.linalg.norm(solver.get_x()), exp["x_norm"])
    if "x_abs_sum" in exp:
        assert close(np.sum(np.abs(solver.get_x())), exp["x_abs_sum"])


@pytest.mark.parametrize("name", sorted(instances.KNOWN_ANSWER))
def test_known_answer(name):
    model, exp = instances.KNOWN_ANSWER[name]()
    solver = cpu_solver()
    solver.load(model)
    solver.solve()
    check_expected(solver, exp)
    check_certificates(solver, model)


def test_modify1():
    """test/nativeinstances.jl:2866-2949 — re-solve after modify_* reuses preprocessing."""
    model = Model([1, 0], [[1, 1]], [-2], -np.eye(2), [0, 0], [cs.Nonnegative(2)])
    solver = cpu_solver()
    solver.load(model)
    solver.solve()
    assert solver.status == "PrimalInfeasible"

    solver.modify_b([-1.0])
    assert solver.status == "Modified"
    solver.solve()
    assert solver.status == "PrimalInfeasible"
    assert solver.time_rescale == 0 and solver.time_loadsys == 0

    solver.modify_c([2.0, 1.0])
    solver.modify_b([2.0])
    solver.modify_obj_offset(1.0)
    solver.solve()
    assert solver.status == "Optimal"
    assert np.isclose(solver.primal_obj, 3, atol=TOL, rtol=TOL)
    assert np.allclose(solver.get_x(), [0, 2], atol=TOL, rtol=TOL)
    check_certificates(solver, solver.orig_model)
    assert solver.time_rescale == 0 and solver.time_loadsys == 0

    solver.modify_c([3.0, 1.0])
    solver.modify_h([1.0, -2.0])
    solver.modify_obj_offset(-1.0)
    solver.solve()
    assert solver.status == "Optimal"
    assert np.isclose(solver.primal_obj, -1, atol=TOL, rtol=TOL)
    assert np.allclose(solver.get_x(), [-1, 3], atol=TOL, rtol=TOL)
    check_certificates(solver, solver.orig_model)


@pytest.mark.parametrize("seed", [1, 2])
def test_possemideftri3_4_eigmax(seed):
    """RNG-dependent but property-checked (test/nativeinstances.jl:359-396)."""
    rng = np.random.default_rng(seed)
    B = rng.random((3, 3))
    mat = np.triu(B) + np.triu(B, 1).T
    eig_max = np.linalg.eigvalsh(mat).max()
    dim = 6
    c = -cs.smat_to_svec(mat)
    A = cs.smat_to_svec(np.eye(3)).reshape(1, dim)
    model = Model(c, A, [1.0], -np.eye(dim), np.zeros(dim), [cs.PosSemidefTri(dim)])
    solver = cpu_solver()
    solver.load(model)
    solver.solve()
    assert solver.status == "Optimal"
    assert np.isclose(solver.primal_obj, -eig_max, atol=TOL, rtol=TOL)


def test_consistent_dependent_equalities():
    """consistent1-like (test/nativeinstances.jl:130-151): dependent but consistent
    primal and dual equalities are removed by the pivoted-QR preprocessing."""
    rng = np.random.default_rng(1)
    n, p, q = 30, 15, 30
    c = np.zeros(n)
    A = rng.integers(-9, 10, (p, n)).astype(float)
    G = 10 * np.eye(q, n)
    r1, r2 = rng.random(2)
    A[10:15, :] = r1 * A[0:5, :] - r2 * A[5:10, :]
    b = A.sum(axis=1)
    r1, r2 = rng.random(2)
    A[:, 10:15] = r1 * A[:, 0:5] - r2 * A[:, 5:10]
    G[:, 10:15] = r1 * G[:, 0:5] - r2 * G[:, 5:10]
    c[10:15] = r1 * c[0:5] - r2 * c[5:10]
    model = Model(c, A, b, G, np.zeros(q), [cs.Nonnegative(q)])
    solver = cpu_solver()
    solver.load(model)
    solver.solve()
    assert solver.status == "Optimal"


def test_inconsistent_primal():
    """inconsistent1 (test/nativeinstances.jl:153-170)."""
    rng = np.random.default_rng(1)
    n, p, q = 30, 15, 30
    c = rng.integers(0, 10, n).astype(float)
    A = rng.integers(-9, 10, (p, n)).astype(float)
    b = rng.random(p)
    r1, r2 = rng.random(2)
    A[10:15, :] = r1 * A[0:5, :] - r2 * A[5:10, :]
    b[10:15] = 2 * (r1 * b[0:5] - r2 * b[5:10])
    model = Model(c, A, b, -np.eye(q, n), np.zeros(q), [cs.Nonnegative(q)])
    solver = cpu_solver()
    solver.load(model)
    solver.solve()
    assert solver.status == "PrimalInconsistent"
