"""Parity of the batched CUDA cone oracles against the CPU oracle cones on the same
seeded points, through the C-ABI (GpuConeBatch).  Mirrors the identities the
reference tests (test/cone.jl:31-129) plus direct value comparison."""
import numpy as np
import pytest

import hypatia_b200  # noqa: F401
from hypatia_b200 import cones_spec as cs
from hypatia_b200.gpu import GpuConeBatch
from oracle.cones import CpuConeSet

from gpu_util import context, rel_err

pytestmark = pytest.mark.gpu
TOL = 1e3 * np.finfo(np.float64).eps


def interior_point(specs, rng, noise=0.1, scale=1.0):
    out = []
    for s in specs:
        p = s.initial_point()
        if s.type_code == cs.POSSEMIDEFTRISPARSE:
            p = p * 2.0 + noise * (2 * rng.random(s.dim) - 1) / max(1.0, np.sqrt(s.side))
        elif s.type_code == cs.POSSEMIDEFTRI:
            B = rng.standard_normal((s.side, s.side)) * noise
            p = p + cs.smat_to_svec(B @ B.T)
        else:
            p = p + noise * (2 * rng.random(s.dim) - 1) * (0.5 if s.type_code != cs.NONNEGATIVE else 1.0) / max(1.0, np.sqrt(s.dim) / 2)
        out.append(p * scale)
    return np.concatenate(out)


def _wsos(n, halfdeg, use_dual=False):
    from hypatia_b200 import polyutils as pu

    U, _, Ps = pu.interpolate(pu.BoxDomain(-np.ones(n), np.ones(n)), halfdeg, sample=False)
    return cs.WSOSInterpNonnegative(U, Ps, use_dual=use_dual)


def _wsospsd(n, halfdeg, R):
    from hypatia_b200 import polyutils as pu

    U, _, Ps = pu.interpolate(pu.BoxDomain(-np.ones(n), np.ones(n)), halfdeg, sample=False)
    return cs.wsos_psd(R, U, Ps)


def _wsoseuc(n, halfdeg, R, use_dual=False):
    from hypatia_b200 import polyutils as pu

    U, _, Ps = pu.interpolate(pu.BoxDomain(-np.ones(n), np.ones(n)), halfdeg, sample=False)
    return cs.wsos_epinormeucl(R, U, Ps, use_dual=use_dual)


def _wsosone(n, halfdeg, R, use_dual=False):
    from hypatia_b200 import polyutils as pu

    U, _, Ps = pu.interpolate(pu.BoxDomain(-np.ones(n), np.ones(n)), halfdeg, sample=False)
    return cs.wsos_epinormone(R, U, Ps, use_dual=use_dual)


def _spsd(side, seed, use_dual=False):
    rng = np.random.default_rng(seed)
    mask = np.tril(rng.random((side, side)) < 1.0 / np.sqrt(side), -1) | np.eye(side, dtype=bool)
    cols, rows = np.nonzero(mask.T)
    return cs.psd_sparse(side, rows + 1, cols + 1, use_dual=use_dual)


SPEC_SETS = {
    "wsospsd": [_wsospsd(1, 1, 1), _wsospsd(1, 2, 2), cs.EpiNormEucl(4), _wsospsd(2, 2, 3), _wsospsd(1, 1, 4)],
    "wsos": [_wsos(1, 3), cs.Nonnegative(3), _wsos(2, 2, use_dual=True), _wsos(2, 4)],
    "nonneg": [cs.Nonnegative(1), cs.Nonnegative(6), cs.Nonnegative(300)],
    "soc_small": [cs.EpiNormEucl(2)] + [cs.EpiNormEucl(3)] * 5 + [cs.EpiNormEucl(8)] * 40,
    "rsoc_small": [cs.EpiPerSquare(3)] * 3 + [cs.EpiPerSquare(8)] * 33 + [cs.EpiPerSquare(4)],
    # many small cones and >= 32 columns: the coalesced tile kernel (ragged last tile, odd offsets, d up to 32)
    "soc_wide": [cs.Nonnegative(3)] + [cs.EpiNormEucl(8)] * 300 + [cs.EpiPerSquare(5)] * 70 + [cs.EpiNormEucl(32)] * 40
    + [cs.EpiPerSquare(3)] * 33,
    "soc_large": [cs.EpiNormEucl(33), cs.EpiNormEucl(100), cs.EpiPerSquare(70), cs.EpiPerSquare(257)],
    "psd": [cs.PosSemidefTri(cs.svec_length(s)) for s in (1, 2, 3, 5, 16, 16, 16, 32)],
    # even offsets and dims: products with >= 64 columns take the symmetric-Kronecker DMMA path
    "psd_wide": [cs.Nonnegative(4)] + [cs.PosSemidefTri(cs.svec_length(8))] * 3 + [cs.PosSemidefTri(136)] * 5
    + [cs.EpiNormEucl(6)] + [cs.PosSemidefTri(cs.svec_length(32))],
    "psd_large": [cs.PosSemidefTri(cs.svec_length(s)) for s in (33, 64, 100, 150)] + [cs.EpiNormEucl(5)],
    # EpiNormInf (no sqrt oracles; use_dual = epigraph of the 1-norm), thread-per-cone and warp-per-cone sizes,
    # odd row offsets
    "norminf_wide": [cs.Nonnegative(1)] + [cs.EpiNormInf(4)] * 300 + [cs.EpiNormInf(7, use_dual=True)] * 50,
    "norminf_small": [cs.EpiNormInf(2)] + [cs.EpiNormInf(4)] * 40 + [cs.EpiNormInf(7, use_dual=True)] * 5 + [cs.EpiNormInf(3)],
    "norminf_large": [cs.EpiNormInf(40), cs.EpiNormInf(129, use_dual=True), cs.EpiNormInf(33)],
    "norminf_mixed": [cs.Nonnegative(3)] + [cs.EpiNormInf(4)] * 5 + [cs.EpiNormEucl(5), cs.EpiNormInf(6, use_dual=True),
                                                                     cs.PosSemidefTri(6), cs.EpiNormInf(9)],
    # DoublyNonnegativeTri: stored Hessian + its Cholesky (the generic oracles), odd offsets, next to other families
    "dnn": [cs.DoublyNonnegativeTri(cs.svec_length(s)) for s in (1, 2, 3)] + [cs.Nonnegative(2)]
    + [cs.DoublyNonnegativeTri(cs.svec_length(s)) for s in (6, 12)] + [cs.EpiNormEucl(5), cs.DoublyNonnegativeTri(cs.svec_length(20))],
    # PosSemidefTriSparse, dense implementation: seeded patterns (test/cone.jl:296-300), primal and dual barrier
    "psd_sparse": [_spsd(1, 0), _spsd(2, 1), cs.Nonnegative(1), _spsd(10, 2), _spsd(25, 3, use_dual=True), cs.EpiNormEucl(4),
                   _spsd(40, 4)],
    # WSOSInterpEpiNormEucl (test/cone.jl:848-853 shapes), dual barrier by default, one primal-barrier instance
    "wsos_epinormeucl": [_wsoseuc(1, 1, 2), cs.Nonnegative(3), _wsoseuc(1, 2, 4), _wsoseuc(2, 2, 3), _wsoseuc(3, 1, 2),
                         _wsoseuc(1, 3, 3, use_dual=True)],
    "wsos_epinormone": [_wsosone(1, 1, 2), cs.Nonnegative(3), _wsosone(1, 2, 4), _wsosone(2, 2, 3), _wsosone(3, 1, 2),
                        _wsosone(1, 3, 3, use_dual=True)],
    "mixed": [cs.Nonnegative(10)] + [cs.EpiNormEucl(8)] * 4 + [cs.EpiPerSquare(8)] * 4
    + [cs.PosSemidefTri(136)] * 3 + [cs.Nonnegative(3), cs.EpiNormEucl(5)],
}


def make_pair(specs):
    cpu = CpuConeSet(specs)
    gpu = GpuConeBatch(specs, context())
    cpu.setup()
    gpu.setup()
    return cpu, gpu


@pytest.mark.parametrize("name", sorted(SPEC_SETS))
def test_oracles_match_cpu(name):
    specs = SPEC_SETS[name]
    rng = np.random.default_rng(sorted(SPEC_SETS).index(name) + 100)
    cpu, gpu = make_pair(specs)
    tol_scale = 1.0  # the same tolerance for every family (measured WSOS errors: 2e-16 ... 1.6e-15, tools/diag_wsos_tol.py)
    q = cpu.q
    primal = interior_point(specs, rng, scale=0.7)
    dual = interior_point(specs, rng, scale=1.3)
    irtmu = 1.1
    for c in (cpu, gpu):
        c.load_point(primal, irtmu)
        c.load_dual_point(dual)
        c.reset_data()
    assert cpu.is_feas() and gpu.is_feas()
    assert cpu.is_dual_feas() and gpu.is_dual_feas()
    assert rel_err(gpu.grad(), cpu.grad()) < TOL * tol_scale
    assert rel_err(gpu.point(), cpu.point()) == 0.0
    ok_c, agg_c = cpu.check_neighborhood(irtmu, True, 0.01, 1e6)
    ok_g, agg_g = gpu.check_neighborhood(irtmu, True, 0.01, 1e6)
    assert ok_c == ok_g
    assert abs(agg_c - agg_g) <= 1e-9 * tol_scale * max(1.0, abs(agg_c))
    ok_c, agg_c = cpu.check_neighborhood(irtmu, False, 0.01, 1e9)
    ok_g, agg_g = gpu.check_neighborhood(irtmu, False, 0.01, 1e9)
    assert ok_c == ok_g and abs(agg_c - agg_g) <= 1e-9 * tol_scale * max(1.0, abs(agg_c))

    for m in (1, 7) + ((80, 200) if name in ("psd_wide", "soc_wide", "norminf_wide") else ()):
        arr = rng.standard_normal(q) if m == 1 else np.asfortranarray(rng.standard_normal((q, m)))
        for fn in ("hess_prod", "inv_hess_prod", "block_hess_prod", "hess_prod_slow"):
            assert rel_err(getattr(gpu, fn)(arr), getattr(cpu, fn)(arr)) < 1e4 * TOL * tol_scale, (fn, m)
        if any(sp.type_code == cs.EPINORMINF for sp in specs):
            # EpiNormInf has no sqrt oracles (epinorminf.jl:73): the batch refuses instead of inventing one
            from hypatia_b200.lib import HybError

            with pytest.raises(HybError):
                gpu.sqrt_hess_prod(arr)
            continue
        # sqrt factors against the per-cone CPU oracles
        want_s = np.empty_like(arr)
        want_i = np.empty_like(arr)
        for k, c in enumerate(cpu.cones):
            sl = slice(int(cpu.offs[k]), int(cpu.offs[k + 1]))
            assert c.use_sqrt_hess_oracles(c.dim + 1)
            want_s[sl] = c.sqrt_hess_prod(arr[sl])
            want_i[sl] = c.inv_sqrt_hess_prod(arr[sl])
        assert rel_err(gpu.sqrt_hess_prod(arr), want_s) < 1e4 * TOL * tol_scale
        assert rel_err(gpu.inv_sqrt_hess_prod(arr), want_i) < 1e4 * TOL * tol_scale
    d = 0.1 * rng.standard_normal(q)
    assert rel_err(gpu.dder3(d), cpu.dder3(d)) < 1e4 * TOL * tol_scale


def test_oracle_identities_on_gpu():
    """test/cone.jl:84-126 identities evaluated with the GPU oracles only."""
    specs = SPEC_SETS["mixed"]
    rng = np.random.default_rng(5)
    _, gpu = make_pair(specs)
    q = gpu.q
    point = interior_point(specs, rng)
    dualp = interior_point(specs, rng)
    gpu.load_point(point)
    gpu.load_dual_point(dualp)
    gpu.reset_data()
    g = gpu.grad()
    nu_k = gpu.cone_dots(point, g)
    assert np.allclose(nu_k, -gpu.nus, atol=TOL, rtol=TOL)
    assert rel_err(gpu.hess_prod(point), -g) < 1e3 * TOL
    assert rel_err(gpu.inv_hess_prod(g), -point) < 1e3 * TOL
    E = np.asfortranarray(np.eye(q))
    H = gpu.hess_prod(E)
    Hi = gpu.inv_hess_prod(E)
    assert rel_err(H @ Hi, np.eye(q)) < 1e4 * TOL
    W = gpu.sqrt_hess_prod(E)
    assert rel_err(W.T @ W, H) < 1e4 * TOL
    Wi = gpu.inv_sqrt_hess_prod(E)
    assert rel_err(Wi.T @ Wi, Hi) < 1e4 * TOL
    assert rel_err(-gpu.dder3(point), g) < 1e3 * TOL
    d = 0.1 * rng.standard_normal(q)
    d3 = gpu.dder3(d)
    assert np.allclose(gpu.cone_dots(d3, point), gpu.cone_dots(d, H @ d), rtol=1e-9, atol=1e-9)


@pytest.mark.parametrize("which", ["primal", "dual"])
def test_infeasible_points_are_rejected(which):
    specs = SPEC_SETS["mixed"]
    rng = np.random.default_rng(11)
    cpu, gpu = make_pair(specs)
    for k in range(len(specs)):
        primal = interior_point(specs, rng)
        dual = interior_point(specs, rng)
        bad = primal if which == "primal" else dual
        lo, hi = int(cpu.offs[k]), int(cpu.offs[k + 1])
        bad[lo:hi] = -bad[lo:hi]  # push cone k's point out of the cone
        for c in (cpu, gpu):
            c.load_point(primal, 1.0)
            c.load_dual_point(dual)
            c.reset_data()
        assert cpu.is_feas() == gpu.is_feas()
        assert cpu.is_dual_feas() == gpu.is_dual_feas()
        ok_c, _ = cpu.check_neighborhood(1.0, True, 0.0, 1e9)
        ok_g, _ = gpu.check_neighborhood(1.0, True, 0.0, 1e9)
        assert ok_c == ok_g and not ok_g


def test_many_small_cones_batch():
    """cfg4-shaped batch: thousands of EpiNormEucl(4) cones, HBM-bound kernels."""
    K = 5000
    specs = [cs.Nonnegative(64)] + [cs.EpiNormEucl(4)] * K
    rng = np.random.default_rng(2)
    cpu, gpu = make_pair(specs)
    primal = interior_point(specs, rng)
    dual = interior_point(specs, rng)
    for c in (cpu, gpu):
        c.load_point(primal, 0.9)
        c.load_dual_point(dual)
        c.reset_data()
    assert cpu.is_feas() and gpu.is_feas()
    assert rel_err(gpu.grad(), cpu.grad()) < TOL
    arr = np.asfortranarray(rng.standard_normal((cpu.q, 3)))
    assert rel_err(gpu.hess_prod(arr), cpu.hess_prod(arr)) < 1e3 * TOL
    ok_c, agg_c = cpu.check_neighborhood(0.9, True, 0.0, 1e9)
    ok_g, agg_g = gpu.check_neighborhood(0.9, True, 0.0, 1e9)
    assert ok_c == ok_g and abs(agg_c - agg_g) <= 1e-10 * max(1, abs(agg_c))


def test_sparse_psd_factor_based_hess_prod_slow(monkeypatch):
    """hess_prod_slow! of PosSemidefTriSparse through the Cholesky factor of the pattern matrix
    (possemideftrisparse/denseimpl.jl:142-163), forced on at a benign point, against the stored-Hessian product."""
    specs = SPEC_SETS["psd_sparse"]
    rng = np.random.default_rng(11)
    cpu, gpu = make_pair(specs)
    primal = interior_point(specs, rng, scale=0.9)
    dual = interior_point(specs, rng, scale=1.1)
    for c in (cpu, gpu):
        c.load_point(primal, 1.0)
        c.load_dual_point(dual)
        c.reset_data()
    assert cpu.is_feas() and gpu.is_feas()
    arr = np.asfortranarray(rng.standard_normal((cpu.q, 5)))
    want = cpu.hess_prod(arr)
    monkeypatch.setenv("HYB_FORCE_HESS_PROD_SLOW", "1")
    got = gpu.hess_prod_slow(arr)
    assert rel_err(got, want) < 1e4 * TOL
    got1 = gpu.hess_prod_slow(arr[:, 0].copy())
    assert rel_err(got1, want[:, 0]) < 1e4 * TOL
