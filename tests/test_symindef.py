"""SymIndefDenseSystemSolver (src/Solvers/systemsolvers/symindef.jl:216-280).

CPU: the oracle restatement (oracle/symindef.py) against the reference's known answers -- the reference runs
this solver over its `inst_minimal` and `inst_cones_few` lists (test/runnativetests.jl:56-68,88-107) -- and against
the QRChol oracle on the same models.
GPU: the device path (hyb_symindef_*: assembly, rook-pivoted Bunch-Kaufman, solves) against the oracle: the
assembled system matrix entry by entry, one solve, and whole solves."""
import numpy as np
import pytest

import hypatia_b200  # noqa: F401
from hypatia_b200 import instances
from hypatia_b200.host.solver import Solver

NAMES = ["nonnegative1", "nonnegative4", "epinormeucl1", "epinormeucl2", "epipersquare1", "epipersquare3", "epinorminf1",
         "epinorminf3_dual", "possemideftri1", "possemideftri2", "wsosinterpnonnegative1", "wsosinterpnonnegative2",
         "wsosinterppossemideftri1"]
NAMES = [n for n in NAMES if n in instances.KNOWN_ANSWER]
TOL = float(np.finfo(np.float64).eps) ** 0.25  # test/nativeinstances.jl tolerance for Float64


def _solve_cpu(name, ss):
    from oracle.cones import CpuConeSet

    model, want = instances.KNOWN_ANSWER[name]()[:2]
    s = Solver(CpuConeSet, ss)
    s.load(model)
    s.solve()
    return s, want


@pytest.mark.parametrize("name", NAMES)
def test_oracle_symindef_reproduces_known_answers(name):
    from oracle.qrchol import QRCholDenseSystemSolver
    from oracle.symindef import SymIndefDenseSystemSolver

    s, want = _solve_cpu(name, SymIndefDenseSystemSolver())
    assert s.status == want.get("status", "Optimal")
    if "primal_obj" in want:
        assert abs(s.primal_obj - want["primal_obj"]) <= TOL * (1 + abs(want["primal_obj"]))
    assert all(k == 1 for k in s.syssolver.fact_kinds)  # plain Bunch-Kaufman, no diagonal shift
    r, _ = _solve_cpu(name, QRCholDenseSystemSolver())
    assert r.status == s.status and abs(r.num_iters - s.num_iters) <= 1
    assert abs(r.primal_obj - s.primal_obj) <= 1e-7 * (1 + abs(r.primal_obj))


@pytest.mark.gpu
@pytest.mark.parametrize("name", NAMES)
def test_gpu_symindef_matches_oracle(name):
    from hypatia_b200.gpu import make_gpu_solver
    from oracle.symindef import SymIndefDenseSystemSolver

    o, want = _solve_cpu(name, SymIndefDenseSystemSolver())
    g = make_gpu_solver(0, syssolver="symindef", trace=True)
    g.load(instances.KNOWN_ANSWER[name]()[0])
    g.solve()
    print(f"{name}: gpu {g.status}/{g.num_iters} obj {g.primal_obj:.12g}; oracle {o.status}/{o.num_iters} obj {o.primal_obj:.12g}")
    assert g.status == o.status
    assert abs(g.num_iters - o.num_iters) <= 1
    assert abs(g.primal_obj - o.primal_obj) <= 1e-7 * (1 + abs(o.primal_obj))
    assert g.syssolver.last_fact_kind == 1


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["block_multicone", "epinorminf3_dual", "wsosinterpnonnegative1"])
def test_gpu_symindef_system_and_solve_at_the_same_iterate(name):
    """The assembled system matrix and one solve of the CUDA path against the oracle from the SAME iterate."""
    import math

    from hypatia_b200.gpu import make_gpu_solver
    from oracle.cones import CpuConeSet
    from oracle.symindef import SymIndefDenseSystemSolver

    def build():
        if name == "block_multicone":
            return instances.block_multicone(N=64, seed=5, psd_side=6)
        return instances.KNOWN_ANSWER[name]()[0]

    ref = Solver(CpuConeSet, SymIndefDenseSystemSolver(), iter_limit=3)
    ref.load(build())
    ref.solve()
    point_vec = ref.point.vec.copy()
    solvers = [Solver(CpuConeSet, SymIndefDenseSystemSolver()), make_gpu_solver(0, syssolver="symindef")]
    sols = []
    for s in solvers:
        s.load(build())
        s._setup_solver()
        s._setup_loaded()
        s._initialize_point()
        s._setup_stepping()
        s.point.vec[:] = point_vec
        s.calc_mu()
        cones, pt = s.cones, s.point
        cones.load_point(cones.primal_of(pt.z, pt.s), 1.0 / math.sqrt(s.mu))
        cones.load_dual_point(cones.dual_of(pt.z, pt.s))
        cones.reset_data()
        assert cones.is_feas()
        s.syssolver.update_lhs(s)
        sols.append(s.syssolver.sol_const.vec.copy())
    cpu, gpu = solvers
    K_cpu = np.tril(cpu.syssolver.lhs_sub)
    K_cpu = K_cpu + np.tril(K_cpu, -1).T
    K_gpu = np.triu(gpu.syssolver.system_matrix())
    K_gpu = K_gpu + np.triu(K_gpu, 1).T
    scale = np.abs(K_cpu).max()
    assert np.abs(K_gpu - K_cpu).max() <= 1e-10 * scale
    # the solve, relative to the conditioning the system has (residual of each solution in the oracle's matrix)
    rhs = cpu.syssolver.rhs_const.vec
    for x in sols:
        assert np.abs(K_cpu @ x - rhs).max() <= 1e-9 * (np.abs(K_cpu) @ np.abs(x) + np.abs(rhs)).max()
    assert np.abs(sols[0] - sols[1]).max() <= 1e-6 * (1 + np.abs(sols[0]).max())
