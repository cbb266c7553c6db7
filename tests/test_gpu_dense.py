"""Parity of the dense engine (DMMA gemm/syrk, blocked Cholesky, wavefront solves)
against numpy/scipy on the same inputs, through the C-ABI.  FP64; tolerances are
a small multiple of K*eps relative to the result scale (summation-order only)."""
import numpy as np
import pytest
import scipy.linalg as sla

from gpu_util import context, rel_err

pytestmark = pytest.mark.gpu
EPS = np.finfo(np.float64).eps


# the last two shapes have 150 tiles and a long K: on a 148-SM part the 2 tiles of the last wave are
# split along K (gemm_dmma.cu tail splitting) and summed by the fixed-order reduce kernel
@pytest.mark.parametrize("M,N,K", [(1, 1, 1), (3, 5, 7), (128, 128, 16), (129, 127, 33), (200, 300, 1000), (64, 513, 257), (500, 500, 1001),
                                   (1280, 1920, 2100), (1275, 1900, 4096)])
def test_gemm_tn(M, N, K):
    rng = np.random.default_rng(M * 7 + N * 3 + K)
    A = rng.standard_normal((K, M))
    B = rng.standard_normal((K, N))
    C0 = rng.standard_normal((M, N))
    got = context().gemm_tn(A, B, C0, alpha=0.75, beta=-0.5)
    want = 0.75 * A.T @ B - 0.5 * C0
    assert rel_err(got, want) < 50 * K * EPS


# N = 2176 / 2170: 153 upper tiles = one wave + 5 split tiles
@pytest.mark.parametrize("K,N", [(1, 1), (5, 3), (1000, 500), (777, 257), (4096, 384), (130, 1025), (2048, 2176), (3001, 2170)])
def test_outer_prod_upper(K, N):
    rng = np.random.default_rng(K + N)
    A = rng.standard_normal((K, N))
    C0 = rng.standard_normal((N, N))
    got = context().outer_prod(A, C0, alpha=1.0, beta=0.0)
    want = A.T @ A
    iu = np.triu_indices(N)
    assert rel_err(got[iu], want[iu]) < 50 * K * EPS
    # strictly lower triangle untouched (dsyrk 'U' semantics)
    il = np.tril_indices(N, -1)
    assert np.array_equal(got[il], C0[il])
    got2 = context().outer_prod(A, C0, alpha=-1.0, beta=1.0)
    assert rel_err(got2[iu], (C0 - want)[iu]) < 50 * K * EPS


@pytest.mark.parametrize("N", [1, 2, 31, 128, 129, 300, 1000, 2100])
def test_posdef_solve(N):
    rng = np.random.default_rng(N)
    B = rng.standard_normal((N + 5, N))
    A = B.T @ B + N * 1e-3 * np.eye(N)
    rhs = rng.standard_normal((N, 3))
    U, X, info = context().posdef_solve(np.triu(A), rhs)
    assert info == 0
    Uref = sla.cholesky(A, lower=False)
    assert rel_err(U, Uref) < 1e-10
    resid = np.linalg.norm(A @ X - rhs) / (np.linalg.norm(A) * np.linalg.norm(X))
    assert resid < 100 * N * EPS
    Xref = sla.cho_solve((Uref, False), rhs)
    assert rel_err(X, Xref) < 1e-8 * np.linalg.cond(A) ** 0.5 + 1e-9


def test_posdef_detects_indefinite():
    N = 300
    rng = np.random.default_rng(0)
    B = rng.standard_normal((N, N))
    A = B + B.T  # indefinite
    _, _, info = context().posdef_solve(np.triu(A), np.ones(N))
    _, info_ref = sla.lapack.dpotrf(A, lower=0)
    assert info == info_ref and info > 0
    # failure beyond the first panel
    A2 = np.eye(N) * 4.0
    A2[200, 200] = -1.0
    _, _, info2 = context().posdef_solve(A2, np.ones(N))
    assert info2 == 201


@pytest.mark.parametrize("N,nrhs", [(5, 4), (128, 65), (129, 64), (300, 200), (700, 130)])
def test_posdef_solve_many_rhs(N, nrhs):
    """Blocked multi-RHS triangular solves (forward with U', backward with U)."""
    rng = np.random.default_rng(N + nrhs)
    B = rng.standard_normal((N + 5, N))
    A = B.T @ B + N * 1e-3 * np.eye(N)
    rhs = rng.standard_normal((N, nrhs))
    _, X, info = context().posdef_solve(np.triu(A), rhs)
    assert info == 0
    resid = np.linalg.norm(A @ X - rhs) / (np.linalg.norm(A) * np.linalg.norm(X))
    assert resid < 100 * N * EPS
    Xref = sla.cho_solve(sla.cho_factor(A), rhs)
    assert rel_err(X, Xref) < 1e-8 * np.linalg.cond(A) ** 0.5 + 1e-9
