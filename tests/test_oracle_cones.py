"""Oracle identities of the reference's cone tests (test/cone.jl:31-129) applied to
the CPU oracle cones at the reference's sizes (test/runconetests.jl).  Tolerance
1e3*eps as in the reference (:35) unless noted."""
import math
import numpy as np
import pytest

import hypatia_b200  # noqa: F401
from hypatia_b200 import cones_spec as cs
from hypatia_b200 import polyutils as pu
from oracle.cones import make_cone

EPS = float(np.finfo(np.float64).eps)


def perturb_scale(point, rng, noise, scale):
    point = point + 2 * noise * rng.random(point.size) - noise
    return point * scale


def run_oracle_identities(spec, tol=1e3 * EPS, noise=0.1, scale=0.1, init_tol=None):
    rng = np.random.default_rng(1)
    init_tol = tol if init_tol is None else init_tol
    cone = make_cone(spec)
    dim = spec.dim
    cone.setup_data()
    cone.reset_data()
    close = lambda a, b, t=tol: np.allclose(a, b, atol=t, rtol=t)  # noqa: E731

    point = spec.initial_point()
    cone.load_point(point)
    assert cone.is_feas()
    dual_point = -cone.grad()
    cone.load_dual_point(dual_point)
    assert cone.is_dual_feas()
    assert cone.get_proxsqr(1.0, True) <= 1
    assert cone.get_proxsqr(1.0, False) <= dim
    if np.isfinite(init_tol):
        assert close(point, dual_point, init_tol)  # centrality of the initial point

    hess = cone.hess()
    assert close(hess @ point, dual_point)
    assert close(cone.hess_prod(point), dual_point)
    inv_hess = cone.inv_hess()
    assert close(inv_hess @ dual_point, point)
    assert close(cone.inv_hess_prod(dual_point), point)
    assert close(hess @ inv_hess, np.eye(dim))

    point = perturb_scale(point, rng, noise, scale)
    dual_point = perturb_scale(dual_point, rng, noise, 1 / scale)
    cone.reset_data()
    cone.load_point(point)
    assert cone.is_feas()
    cone.load_dual_point(dual_point)
    assert cone.is_dual_feas()

    nu = spec.nu
    grad = cone.grad()
    assert close(point @ grad, -nu)
    hess = np.array(cone.hess())
    inv_hess = np.array(cone.inv_hess())
    assert close(hess @ inv_hess, np.eye(dim))
    assert close(hess @ point, -grad)
    assert close(cone.hess_prod(point), -grad)
    assert close(cone.inv_hess_prod(grad), -point)
    assert close(cone.hess_prod(inv_hess), np.eye(dim))
    assert close(cone.inv_hess_prod(hess), np.eye(dim))

    psi = dual_point + grad
    proxsqr = psi @ cone.inv_hess_prod(psi)
    assert close(cone.get_proxsqr(1.0, False), proxsqr)

    if hasattr(cone, "use_hess_prod_slow"):
        cone.update_use_hess_prod_slow()
        assert cone.use_hess_prod_slow_updated and not cone.use_hess_prod_slow
        cone.use_hess_prod_slow = True
        assert close(cone.hess_prod_slow(inv_hess), np.eye(dim))

    if cone.use_sqrt_hess_oracles(dim + 1):
        prod_mat2 = cone.sqrt_hess_prod(inv_hess).T.copy()
        assert close(cone.sqrt_hess_prod(prod_mat2), np.eye(dim))
        prod_mat2 = cone.inv_sqrt_hess_prod(np.eye(dim))
        assert close(prod_mat2.T @ prod_mat2, inv_hess)

    assert close(-cone.dder3(point), grad)
    direction = perturb_scale(np.zeros(dim), rng, noise, 1.0)
    dder3 = cone.dder3(direction)
    assert close(dder3 @ point, direction @ hess @ direction)


@pytest.mark.parametrize("dim", [1, 2, 6])
def test_nonnegative(dim):  # test/cone.jl:343-347
    run_oracle_identities(cs.Nonnegative(dim))


@pytest.mark.parametrize("side", [1, 2, 3, 5])
def test_possemideftri(side):  # :357-361
    run_oracle_identities(cs.PosSemidefTri(cs.svec_length(side)))


@pytest.mark.parametrize("dw", [1, 2, 5])
def test_epinormeucl(dw):  # :485-489
    run_oracle_identities(cs.EpiNormEucl(1 + dw))


@pytest.mark.parametrize("dw", [1, 2, 5])
def test_epinorminf(dw):  # :465-470
    run_oracle_identities(cs.EpiNormInf(1 + dw))


def test_epinorminf_barrier_derivatives():
    """test/cone.jl:472-480 gives the barrier -sum log(u^2 - w_i^2) + (d - 1) log u: the oracle's grad,
    Hessian product and dder3 are checked against finite differences of it."""
    dw = 3
    spec = cs.EpiNormInf(1 + dw)
    cone = make_cone(spec)
    cone.setup_data()
    rng = np.random.default_rng(4)
    s = np.concatenate([[2.0], rng.uniform(-1, 1, dw)])

    def barrier(v):
        return -np.sum(np.log(v[0] ** 2 - v[1:] ** 2)) + (dw - 1) * np.log(v[0])

    def grad_at(v):
        c = make_cone(spec)
        c.setup_data()
        c.load_point(v)
        assert c.is_feas()
        return c.grad().copy()

    cone.load_point(s)
    assert cone.is_feas()
    g = cone.grad()
    h = 1e-6
    fd_g = np.array([(barrier(s + h * e) - barrier(s - h * e)) / (2 * h) for e in np.eye(1 + dw)])
    assert np.allclose(g, fd_g, rtol=1e-7, atol=1e-7)
    d = rng.standard_normal(1 + dw)
    fd_H = (grad_at(s + h * d) - grad_at(s - h * d)) / (2 * h)
    assert np.allclose(cone.hess_prod(d), fd_H, rtol=1e-6, atol=1e-6)
    # dder3(d) = -1/2 * third directional derivative: d/dt [H(s + t d) d] = -2 dder3(d)
    def Hd_at(v):
        c = make_cone(spec)
        c.setup_data()
        c.load_point(v)
        assert c.is_feas()
        return c.hess_prod(d)

    h3 = 1e-5
    fd_T = (Hd_at(s + h3 * d) - Hd_at(s - h3 * d)) / (2 * h3)
    assert np.allclose(cone.dder3(d), -0.5 * fd_T, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("dw", [1, 2, 5])
def test_epipersquare(dw):  # :502-506
    run_oracle_identities(cs.EpiPerSquare(2 + dw))


@pytest.mark.parametrize("n,halfdeg", [(1, 1), (1, 3), (2, 1), (2, 2), (3, 1)])
def test_wsosinterpnonnegative(n, halfdeg):  # :795-800
    U, _, Ps = pu.interpolate(pu.BoxDomain(-np.ones(n), np.ones(n)), halfdeg, sample=False)  # rand_interp, :323-328
    run_oracle_identities(cs.WSOSInterpNonnegative(U, Ps), init_tol=np.inf)


@pytest.mark.parametrize("n,halfdeg,R", [(1, 1, 1), (1, 1, 4), (2, 2, 1), (3, 1, 2)])
def test_wsosinterppossemideftri(n, halfdeg, R):  # :813-818
    U, _, Ps = pu.interpolate(pu.BoxDomain(-np.ones(n), np.ones(n)), halfdeg, sample=False)
    run_oracle_identities(cs.wsos_psd(R, U, Ps), init_tol=np.inf)


def test_index_maps():
    """arrayutilities.jl index maps, bit-exact."""
    for side in range(1, 40):
        assert cs.svec_side(cs.svec_length(side)) == side
    assert [cs.svec_idx(r, c) for r, c in [(1, 1), (2, 1), (1, 2), (2, 2), (3, 1), (3, 3)]] == [1, 2, 2, 3, 4, 6]
    assert list(cs.block_idxs(5, 3)) == [10, 11, 12, 13, 14]
    M = np.arange(16.0).reshape(4, 4)
    M = M + M.T
    v = cs.smat_to_svec(M)
    assert v.size == 10 and v[0] == M[0, 0] and v[1] == M[0, 1] * np.sqrt(2) and v[2] == M[1, 1]
    assert np.allclose(cs.svec_to_smat(v), M)
    assert np.array_equal(cs.PosSemidefTri(10).initial_point(), cs.smat_to_svec(np.eye(4)))
    w = np.ones(12)
    cs.scale_svec_incr(w, 2.0, 4)
    assert np.array_equal(w, [1, 1, 1, 1, 2, 2, 2, 2, 1, 1, 1, 1])


@pytest.mark.parametrize("side", [1, 2, 5])
def test_doublynonnegativetri(side):  # test/cone.jl:373-381 (init_tol = sqrt(eps))
    run_oracle_identities(cs.DoublyNonnegativeTri(cs.svec_length(side)), init_tol=math.sqrt(EPS))


@pytest.mark.parametrize("side", [10, 20])
def test_doublynonnegativetri_initial_point(side):  # test/cone.jl:378-380 (init_only)
    spec = cs.DoublyNonnegativeTri(cs.svec_length(side))
    cone = make_cone(spec)
    cone.setup_data()
    point = spec.initial_point()
    cone.load_point(point)
    assert cone.is_feas()
    assert np.allclose(point, -cone.grad(), atol=math.sqrt(EPS), rtol=math.sqrt(EPS))


def test_doublynonnegativetri_barrier_derivatives():
    """test/cone.jl:383-391: -logdet(W) - sum log(off-diagonal s); grad by finite differences of it."""
    side = 3
    spec = cs.DoublyNonnegativeTri(cs.svec_length(side))
    od = spec.offdiag_idxs()
    rng = np.random.default_rng(7)
    B = rng.uniform(0.1, 1.0, (side, side))
    W = B @ B.T + side * np.eye(side)
    s = cs.smat_to_svec(W)

    def barrier(v):
        return -np.linalg.slogdet(cs.svec_to_smat(v))[1] - np.sum(np.log(v[od]))

    cone = make_cone(spec)
    cone.setup_data()
    cone.load_point(s)
    assert cone.is_feas()
    h = 1e-6
    fd = np.array([(barrier(s + h * e) - barrier(s - h * e)) / (2 * h) for e in np.eye(spec.dim)])
    assert np.allclose(cone.grad(), fd, rtol=1e-7, atol=1e-7)


def rand_sppsd_pattern(side, rng):
    """test/cone.jl:296-300: lower triangle of a random pattern of density 1/sqrt(side) plus the diagonal,
    column-major order (findnz); 1-based."""
    mask = np.tril(rng.random((side, side)) < 1.0 / math.sqrt(side), -1) | np.eye(side, dtype=bool)
    cols, rows = np.nonzero(mask.T)
    return (rows + 1).tolist(), (cols + 1).tolist()


@pytest.mark.parametrize("side", [1, 2, 10, 25, 40])
def test_possemideftrisparse(side):  # test/cone.jl:396-400
    rows, cols = rand_sppsd_pattern(side, np.random.default_rng(side))
    run_oracle_identities(cs.psd_sparse(side, rows, cols))


def test_possemideftrisparse_barrier_derivatives():
    """test/cone.jl:402-415: -logdet of the pattern matrix; gradient by finite differences of it."""
    side = 8
    rng = np.random.default_rng(2)
    rows, cols = rand_sppsd_pattern(side, rng)
    spec = cs.psd_sparse(side, rows, cols)
    r0, c0 = spec.rows0, spec.cols0
    scale = np.where(r0 != c0, 1 / math.sqrt(2.0), 1.0)

    def mat(v):
        M = np.zeros((side, side))
        M[r0, c0] = v * scale
        M[c0, r0] = v * scale
        return M

    s = spec.initial_point() * 3.0 + 0.2 * rng.uniform(-1, 1, spec.dim)
    cone = make_cone(spec)
    cone.setup_data()
    cone.load_point(s)
    assert cone.is_feas()
    h = 1e-6
    fd = np.array([(-np.linalg.slogdet(mat(s + h * e))[1] + np.linalg.slogdet(mat(s - h * e))[1]) / (2 * h) for e in np.eye(spec.dim)])
    assert np.allclose(cone.grad(), fd, rtol=1e-6, atol=1e-7)


def _interp(n, halfdeg):
    from hypatia_b200 import polyutils as pu

    U, _, Ps = pu.interpolate(pu.BoxDomain(-np.ones(n), np.ones(n)), halfdeg, sample=False)
    return U, Ps


@pytest.mark.parametrize("n,halfdeg,ds", [(1, 1, 1), (1, 2, 3), (2, 2, 2), (3, 1, 1)])
def test_wsosinterpepinormeucl(n, halfdeg, ds):  # test/cone.jl:848-853 (init_tol = Inf)
    U, Ps = _interp(n, halfdeg)
    run_oracle_identities(cs.wsos_epinormeucl(1 + ds, U, Ps), init_tol=np.inf, tol=1e5 * EPS)


def test_wsosinterpepinormeucl_barrier_derivatives():
    """test/cone.jl:855-873: the barrier as a function of the point; gradient and Hessian-vector product of the
    oracle against finite differences of it."""
    U, Ps = _interp(1, 1)
    ds = 2
    spec = cs.wsos_epinormeucl(1 + ds, U, Ps)

    def barrier(s):
        tot = 0.0
        for P in Ps:
            Lam = P.T @ (s[:U, None] * P)
            L1 = np.linalg.cholesky(Lam)
            PL1 = np.linalg.solve(L1, P.T)
            for i in range(ds):
                LLi = PL1 @ (s[(1 + i) * U : (2 + i) * U, None] * P)
                Lam = Lam - LLi.T @ LLi
            tot += -2.0 * np.sum(np.log(np.diag(L1))) - np.linalg.slogdet(Lam)[1]
        return tot

    rng = np.random.default_rng(3)
    s = spec.initial_point() + 0.1 * rng.uniform(-1, 1, spec.dim)
    cone = make_cone(spec)
    cone.setup_data()
    cone.load_point(s)
    assert cone.is_feas()
    h = 1e-5
    eye = np.eye(spec.dim)
    fd = np.array([(barrier(s + h * e) - barrier(s - h * e)) / (2 * h) for e in eye])
    assert np.allclose(cone.grad(), fd, rtol=1e-6, atol=1e-6)
    d = rng.standard_normal(spec.dim)

    def grad_at(v):
        c = make_cone(spec)
        c.setup_data()
        c.load_point(v)
        assert c.is_feas()
        return c.grad().copy()

    fd_h = (grad_at(s + h * d) - grad_at(s - h * d)) / (2 * h)
    assert np.allclose(cone.hess_prod(d), fd_h, rtol=1e-5, atol=1e-5)
    fd_3 = (grad_at(s + h * d) - 2 * grad_at(s) + grad_at(s - h * d)) / (h * h)  # = grad'''[d, d]
    assert np.allclose(-0.5 * fd_3, cone.dder3(d), rtol=1e-3, atol=1e-3)


@pytest.mark.parametrize("n,halfdeg,ds", [(1, 1, 1), (1, 2, 3), (2, 2, 2), (3, 1, 1)])
def test_wsosinterpepinormone(n, halfdeg, ds):  # test/cone.jl:880-885 (init_tol = Inf)
    U, Ps = _interp(n, halfdeg)
    run_oracle_identities(cs.wsos_epinormone(1 + ds, U, Ps), init_tol=np.inf, tol=1e5 * EPS)


def test_wsosinterpepinormone_barrier_derivatives():
    """test/cone.jl:887-903: the barrier; gradient, Hessian product and dder3 against finite differences."""
    U, Ps = _interp(1, 1)
    ds = 2
    spec = cs.wsos_epinormone(1 + ds, U, Ps)

    def barrier(s):
        tot = 0.0
        for P in Ps:
            Lam = P.T @ (s[:U, None] * P)
            L1 = np.linalg.cholesky(Lam)
            PL1 = np.linalg.solve(L1, P.T)
            tot += -2.0 * np.sum(np.log(np.diag(L1)))
            for i in range(ds):
                LLi = PL1 @ (s[(1 + i) * U : (2 + i) * U, None] * P)
                tot -= np.linalg.slogdet(Lam - LLi.T @ LLi)[1]
        return tot

    rng = np.random.default_rng(5)
    s = spec.initial_point() + 0.1 * rng.uniform(-1, 1, spec.dim)

    def cone_at(v):
        c = make_cone(spec)
        c.setup_data()
        c.load_point(v)
        assert c.is_feas()
        return c

    cone = cone_at(s)
    h = 1e-5
    fd = np.array([(barrier(s + h * e) - barrier(s - h * e)) / (2 * h) for e in np.eye(spec.dim)])
    assert np.allclose(cone.grad(), fd, rtol=1e-6, atol=1e-6)
    d = rng.standard_normal(spec.dim)
    fd_h = (cone_at(s + h * d).grad() - cone_at(s - h * d).grad()) / (2 * h)
    assert np.allclose(cone.hess_prod(d), fd_h, rtol=1e-5, atol=1e-5)
    fd_3 = (cone_at(s + h * d).grad() - 2 * cone.grad() + cone_at(s - h * d).grad()) / (h * h)
    assert np.allclose(-0.5 * fd_3, cone.dder3(d), rtol=1e-3, atol=1e-3)
