"""hyb_geqp3 (csrc/qr.cu), the column-pivoted QR of the preprocessing (process.jl:109,233), against
LAPACK dgeqp3 on the same inputs, and whole solves with every preprocessing QR forced onto the GPU."""
import numpy as np
import pytest
import scipy.linalg as sla

import hypatia_b200  # noqa: F401
from hypatia_b200 import instances
from hypatia_b200.gpu import make_gpu_solver
from hypatia_b200.host.solver import Solver, _PivotedQR
from oracle.cones import CpuConeSet
from oracle.qrchol import QRCholDenseSystemSolver

from gpu_util import context

pytestmark = pytest.mark.gpu
EPS = np.finfo(np.float64).eps


def _matrix(kind, m, n, rng):
    if kind == "random":
        return rng.standard_normal((m, n))
    if kind == "graded":  # column norms over 12 orders of magnitude
        return rng.standard_normal((m, n)) * np.logspace(0, -12, n)[None, :]
    if kind == "rank_deficient":
        r = max(1, min(m, n) // 2)
        return rng.standard_normal((m, r)) @ rng.standard_normal((r, n))
    if kind == "zero_and_duplicate_columns":
        M = rng.standard_normal((m, n))
        M[:, n // 3] = 0.0
        M[:, n // 2] = M[:, 0]
        return M
    raise ValueError(kind)


@pytest.mark.parametrize("kind", ["random", "graded", "rank_deficient", "zero_and_duplicate_columns"])
@pytest.mark.parametrize("m,n", [(1, 1), (7, 3), (3, 7), (64, 64), (300, 130), (130, 300), (1000, 257)])
def test_geqp3_matches_lapack(kind, m, n):
    rng = np.random.default_rng(m * 31 + n)
    M = _matrix(kind, m, n, rng)
    gpu = _PivotedQR(M, context().geqp3)
    cpu = _PivotedQR(M)
    k = min(m, n)
    scale = np.abs(M).max() + 1e-300
    # Q R = M P : rebuild from the factored form with the host's dormqr
    R = np.triu(gpu.qr[:k, :])
    QR = gpu.apply_q(np.vstack([R, np.zeros((m - k, n))]), trans=False)
    assert np.abs(QR - M[:, gpu.piv]).max() <= 200 * max(m, n) * EPS * scale
    assert sorted(gpu.piv.tolist()) == list(range(n))
    # the pivoted diagonal (what get_rank_est reads, process.jl:403-412) agrees with LAPACK's
    dg, dc = np.abs(np.diagonal(gpu.qr))[:k], np.abs(np.diagonal(cpu.qr))[:k]
    assert np.abs(dg - dc).max() <= 1e-8 * scale * np.sqrt(m)
    tol = 1000 * EPS
    assert gpu.rank_est(tol * scale * np.sqrt(m)) == cpu.rank_est(tol * scale * np.sqrt(m))
    if kind in ("random", "graded"):  # no ties: the same pivot order
        assert np.array_equal(gpu.piv, cpu.piv)


def _oracle(model, **kw):
    s = Solver(CpuConeSet, QRCholDenseSystemSolver(), **kw)
    s.load(model)
    s.solve()
    return s


@pytest.mark.parametrize(
    "build",
    [
        lambda: instances.linearopt(m=40, n=120, seed=2),  # p > 0: both QRs (A' and [A; G] after the reduce)
        lambda: instances.block_multicone(N=64, seed=4, psd_side=6),
        lambda: instances.KNOWN_ANSWER["dimension1"]()[0],
        lambda: instances.KNOWN_ANSWER["primalinfeas1"]()[0],
        lambda: instances.KNOWN_ANSWER["dualinfeas2"]()[0],
        lambda: instances.KNOWN_ANSWER["epipersquare4"]()[0],
    ],
    ids=["linearopt", "block_multicone", "dimension1", "primalinfeas1", "dualinfeas2", "epipersquare4"],
)
def test_solves_with_gpu_preprocessing_qr(build, monkeypatch):
    monkeypatch.setenv("HYB_GPU_QR_MIN", "0")
    # the reference's benchmark tolerances (benchmarks/natvsext/run.jl:44-60): this test is about the
    # preprocessing, and at the default eps^0.5 the block model reaches x_feas = 2.0e-8 against 1.5e-8 at
    # iteration 10, after which the count of remaining iterations is decided by rounding (12 +- 2)
    kw = dict(tol_rel_opt=1e-7, tol_feas=1e-7, tol_abs_opt=1e-10, tol_infeas=1e-10)
    gpu = make_gpu_solver(0, **kw)
    assert gpu.qr_backend_min_cols == 0
    gpu.load(build())
    gpu.solve()
    ref = _oracle(build(), **kw)
    assert gpu.status == ref.status
    assert abs(gpu.num_iters - ref.num_iters) <= 1
    if ref.status == "Optimal":
        assert abs(gpu.primal_obj - ref.primal_obj) <= 1e-7 * (1 + abs(ref.primal_obj))
        assert np.abs(gpu.get_x() - ref.get_x()).max() <= 1e-5 * (1 + np.abs(ref.get_x()).max())


@pytest.mark.parametrize("m,n,k", [(1, 1, 1), (5, 7, 3), (300, 130, 130), (257, 300, 64), (1000, 64, 64), (70, 2500, 700)])
def test_ormqr_right_matches_lapack(m, n, k):
    """hyb_ormqr_right (C <- C Q, the G * Ap_Q of the reduce, process.jl:310-338) against LAPACK dormqr('R', 'N') on
    the reflectors of a QR of a random n x k matrix; n = 2500 walks the reflector through shared memory in chunks."""
    from scipy.linalg import lapack

    rng = np.random.default_rng(m + 7 * n + 13 * k)
    (qr, tau), _, _ = sla.qr(rng.standard_normal((n, k)), mode="raw", pivoting=True)
    refl = np.asfortranarray(qr[:, :k])
    if k > 2:
        tau = tau.copy()
        tau[k // 2] = 0.0  # an identity reflector (what dlarfg returns for a zero column)
    C0 = np.asfortranarray(rng.standard_normal((m, n)))
    _, lw, _ = lapack.dormqr("R", "N", refl, tau, C0.copy(order="F"), -1)
    want, _, info = lapack.dormqr("R", "N", refl, tau, C0.copy(order="F"), int(lw[0]))
    assert info == 0
    got = context().ormqr_right(refl, tau, C0)
    assert np.abs(got - want).max() <= 50 * k * EPS * max(1.0, np.abs(want).max())
    # Q is orthogonal: row norms are preserved
    assert np.allclose(np.linalg.norm(got, axis=1), np.linalg.norm(C0, axis=1), rtol=1e-12)
