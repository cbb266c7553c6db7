"""Symmetric-indefinite fallback of the Schur factorization on the GPU (csrc/sytrf.cu) against LAPACK's
dsytrf_rook / dsytrs_rook (oracle/lapack_rook.py), i.e. the reference's symm_fact! = bunchkaufman!(A, true)
(src/linearalgebra/dense.jl:199-203), and the whole posdef_fact_copy! chain (dense.jl:228-249):
Cholesky -> Bunch-Kaufman of the same matrix -> diagonal shift + Bunch-Kaufman."""
import os
import subprocess
import sys

import numpy as np
import pytest

import hypatia_b200  # noqa: F401
from hypatia_b200 import instances
from hypatia_b200.gpu import make_gpu_solver
from oracle.dense import posdef_fact_copy
from oracle.lapack_rook import dsytrf_rook, dsytrs_rook

from gpu_util import context, rel_err

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _indefinite(n, seed, sparse=False):
    rng = np.random.default_rng(seed)
    M = rng.standard_normal((n, n))
    M = M + M.T
    if sparse:  # many exact zeros: 2 x 2 pivots and long rook searches
        M[np.abs(M) < 1.2] = 0.0
    return M


@pytest.mark.parametrize("n,seed,sparse", [(1, 0, False), (2, 1, False), (3, 2, False), (7, 3, False), (50, 4, False),
                                           (50, 5, True), (130, 6, False), (300, 7, True), (515, 8, False)])
def test_rook_factorization_matches_lapack(n, seed, sparse):
    A = _indefinite(n, seed, sparse)
    rng = np.random.default_rng(100 + seed)
    B = rng.standard_normal((n, 3))
    ldu_g, ipiv_g, X_g, info_g = context().symm_solve(np.triu(A), B)
    ldu_c, ipiv_c, info_c = dsytrf_rook(np.triu(A))
    assert info_g == info_c == 0
    assert np.array_equal(ipiv_g, ipiv_c), (ipiv_g, ipiv_c)  # same pivot choices, 1 x 1 and 2 x 2 alike
    assert rel_err(np.triu(ldu_g), np.triu(ldu_c)) < 1e-10
    X_c = dsytrs_rook(ldu_c, ipiv_c, B)
    assert rel_err(X_g, X_c) < 1e-9
    assert rel_err(A @ X_g, B) < 1e-9 * max(1.0, np.linalg.cond(A))


def test_rook_reports_an_exactly_singular_matrix_like_lapack():
    A = _indefinite(12, 9)
    A[:, 4] = 0.0
    A[4, :] = 0.0
    _, ipiv_g, _, info_g = context().symm_solve(np.triu(A))
    _, ipiv_c, info_c = dsytrf_rook(np.triu(A))
    assert info_g == info_c != 0
    assert np.array_equal(ipiv_g, ipiv_c)


def _loaded_solver():
    s = make_gpu_solver(0)
    s.load(instances.block_multicone(N=96, seed=0, soc_dim=8, psd_side=6))
    s._setup_solver()
    s._setup_loaded()
    s._initialize_point()
    s._setup_stepping()
    s._calc_convergence_params()
    return s


def _factor(ss):
    import ctypes as C

    tf, kind = C.c_double(0), C.c_int(0)
    rc = ss.ctx.check(ss.lib.hyb_schur_factor(ss.ctx.ptr, C.byref(tf), C.byref(kind)), "hyb_schur_factor")
    return rc, kind.value


def _schur_solve(ss, rhs):
    from hypatia_b200.lib import dptr

    x = np.empty_like(rhs)
    ss.ctx.check(ss.lib.hyb_schur_solve(ss.ctx.ptr, dptr(np.asfortranarray(rhs)), dptr(x), 1), "hyb_schur_solve")
    return x


def test_posdef_fact_copy_chain_on_the_schur_matrix():
    """fact_kind 0 / 1 / 2 on a definite, an indefinite and a singular Schur matrix, solves against the oracle."""
    from hypatia_b200.lib import dptr

    s = _loaded_solver()
    ss = s.syssolver
    ss.update_lhs(s)
    assert ss.last_fact_kind == 0
    N = ss.N
    S = np.triu(ss.schur_matrix())
    S = S + np.triu(S, 1).T
    rng = np.random.default_rng(5)
    rhs = rng.standard_normal(N)
    # (1) indefinite: flip the sign of a few eigen-directions
    w, V = np.linalg.eigh(S)
    w2 = w.copy()
    w2[::7] *= -1.0
    S_ind = (V * w2) @ V.T
    S_ind = 0.5 * (S_ind + S_ind.T)
    for mat, expect_kind in ((S, 0), (S_ind, 1), (np.zeros((N, N)), 2)):
        up = np.asfortranarray(np.triu(mat))
        ss.ctx.check(ss.lib.hyb_schur_set(ss.ctx.ptr, dptr(up)), "hyb_schur_set")
        rc, kind = _factor(ss)
        fact = posdef_fact_copy(np.array(up, order="F"))
        assert kind == expect_kind == fact.kind_code, (kind, expect_kind, fact.kind_code)
        x_gpu = _schur_solve(ss, rhs)
        x_cpu = fact.solve(rhs)
        assert rel_err(x_gpu, x_cpu) < 1e-8, (expect_kind, rel_err(x_gpu, x_cpu))


def test_whole_solve_through_the_bunch_kaufman_path():
    """HYB_FORCE_BK makes every update_lhs take the fallback (Bunch-Kaufman of a positive definite matrix is
    a valid factorization of it): the solve must reach the same answer with fact_kind 1 throughout."""
    code = (
        "import sys; sys.path.insert(0, %r)\n"
        "import hypatia_b200\n"
        "from hypatia_b200 import instances\n"
        "from hypatia_b200.gpu import make_gpu_solver\n"
        "s = make_gpu_solver(0); s.load(instances.block_multicone(N=64, seed=0, psd_side=8)); s.solve()\n"
        "print('RESULT', s.status, s.num_iters, repr(s.primal_obj), s.syssolver.last_fact_kind)\n" % ROOT
    )
    outs = []
    for force in (False, True):
        env = dict(os.environ)
        env.pop("HYB_FORCE_BK", None)
        if force:
            env["HYB_FORCE_BK"] = "1"
        r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=600)
        line = [ln for ln in r.stdout.splitlines() if ln.startswith("RESULT")]
        assert line, r.stdout + r.stderr
        outs.append(line[0].split())
    (_, st0, it0, obj0, k0), (_, st1, it1, obj1, k1) = outs
    assert (k0, k1) == ("0", "1")
    assert st0 == st1 == "Optimal"
    assert abs(int(it0) - int(it1)) <= 1
    assert abs(float(obj0) - float(obj1)) <= 1e-7 * (1 + abs(float(obj0)))
