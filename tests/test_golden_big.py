"""Parity at the FULL sizes the measurements are quoted on (BASELINE.json configs 1-5).

tests/golden/trajectories_big.json holds the pinned CPU oracle's whole trajectory on those models
(status, iteration count, objective, and per iteration: mu, x/z residual norms, gap, tau, kap, step
length), written by tests/golden/make_golden_big.py in the build container (the oracle needs minutes
per instance there; the GPU box only reads the committed numbers).

CPU (`-m "not gpu"`): the fixture is well-formed and the oracle still reproduces the smallest entry.
GPU: the CUDA path through the C-ABI solves every model and is compared with the committed numbers at
north_star's tolerances: same status, iterations +-1, objective 1e-7 relative, per-iteration residuals /
mu / gap / tau / kap 1e-8 relative on the first half of the iterations (asserted) -- the first deviating
iteration of the whole run is printed, not hidden (SURVEY §7 hard part 7).
"""
import importlib.util
import json
import os

import numpy as np
import pytest

import hypatia_b200  # noqa: F401
from hypatia_b200.host.solver import Solver

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
BIG = json.load(open(os.path.join(HERE, "trajectories_big.json")))
KEYS = ("x_norm_res", "z_norm_res", "mu", "gap", "tau", "kap")


def _mg():
    spec = importlib.util.spec_from_file_location("make_golden_big", os.path.join(HERE, "make_golden_big.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    return mg


def deviation(trace, gold, upto=None):
    """(worst relative deviation, first (iteration, key, deviation) above 1e-8) over the common
    iterations.  Residual norms are compared relative to max(|value|, 1e-5): a residual is a sum of terms
    of size O(1)-O(100), so evaluating it in a different summation order moves it by ~1e-13 whatever its
    own size (cfg1's z residual IS that noise, 5e-14, from iteration 0 on -- the fixture written on one
    machine and the same oracle run on another differ by 1.6e-14 there), and 1e-8 relative is only
    defined above 1e-13 / 1e-8 = 1e-5."""
    n = min(len(trace), len(gold["mu"]))
    if upto is not None:
        n = min(n, upto)
    worst, first = 0.0, None
    for i in range(n):
        for k in KEYS:
            floor = 1e-5 if k.endswith("_res") else 1e-12
            b = gold[k][i]
            d = abs(trace[i][k] - b) / max(abs(b), floor)
            worst = max(worst, d)
            if d > 1e-8 and first is None:
                first = (i, k, d)
    return worst, first


def stall_onset(alphas):
    """Index of the first iteration whose step length collapsed below 0.1 (None if it never does)."""
    return next((i for i, a in enumerate(alphas) if i > 0 and a < 0.1), None)


def check_against_gold(solver, gold, name):
    worst, first = deviation(solver.trace, gold)
    print(f"{name}: status {solver.status}/{gold['status']}, iterations {solver.num_iters}/{gold['num_iters']}, "
          f"objective {solver.primal_obj:.12g}/{gold['primal_obj']:.12g}, worst trajectory deviation {worst:.2e}, "
          f"first > 1e-8: {first}")
    assert solver.status == gold["status"]
    if gold.get("exit") == "SlowProgress":
        # At the DEFAULT tolerances (eps^0.5) cfg4 / cfg5 stop a hair short of the feasibility tolerance: the
        # oracle itself stalls (x_feas 2.5e-8 against 1.5e-8, step lengths collapse to 0.005 ... 0.0005, mu
        # frozen to three digits) and leaves through SlowProgress -> NearOptimal.  The iterations spent INSIDE
        # the stall make no progress, and how many of them pass before three consecutive "slow" ones are
        # counted is decided by rounding (the same oracle gives 20 or 22 on two machines' BLAS).  What is
        # held to +-1 here is the iteration at which the stall begins; the `_benchtol` entries (the
        # reference's own benchmark tolerances, where these models end Optimal) carry the plain +-1 check.
        assert solver.exit_reason == "SlowProgress"
        got, want = stall_onset([r["alpha"] for r in solver.trace]), stall_onset(gold["alpha"])
        print(f"{name}: stall begins at iteration {got} (oracle {want})")
        assert got is not None and abs(got - want) <= 1
        assert abs(solver.num_iters - gold["num_iters"]) <= 4
    else:
        assert abs(solver.num_iters - gold["num_iters"]) <= 1
    assert abs(solver.primal_obj - gold["primal_obj"]) <= 1e-7 * (1 + abs(gold["primal_obj"]))
    half = max(3, gold["num_iters"] // 2)
    _, early = deviation(solver.trace, gold, upto=half)
    assert early is None, early


def test_fixture_is_complete():
    mg = _mg()
    assert set(BIG) <= set(mg.BIG)
    for name, g in BIG.items():
        n = g["num_iters"] + 1  # one trace row per convergence check
        for k in KEYS + ("alpha",):
            assert len(g[k]) == n, (name, k)
        assert g["status"] in ("Optimal", "NearOptimal")


def test_oracle_reproduces_smallest_entry():
    from oracle.cones import CpuConeSet
    from oracle.qrchol import QRCholDenseSystemSolver

    mg = _mg()
    name = "cfg1_linearopt_500x1000"
    s = Solver(CpuConeSet, QRCholDenseSystemSolver(), trace=True, **mg.solver_kwargs(name))
    s.load(mg.build(name))
    s.solve()
    check_against_gold(s, BIG[name], name)


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(BIG))
def test_gpu_path_against_full_size_golden(name):
    from hypatia_b200.gpu import make_gpu_solver

    mg = _mg()
    s = make_gpu_solver(0, trace=True, **mg.solver_kwargs(name))
    s.load(mg.build(name))
    s.solve()
    print(f"{name}: exit {getattr(s, 'exit_reason', None)}, oracle exit {BIG[name].get('exit')}, "
          f"fact kind of the last update {s.syssolver.last_fact_kind}")
    check_against_gold(s, BIG[name], name)
