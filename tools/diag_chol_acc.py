"""Componentwise backward error (Oettli-Prager) of factor + solve on graded SPD matrices: this library vs LAPACK."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import scipy.linalg as sla
import hypatia_b200
from hypatia_b200 import lib as _lib
if os.environ.get('HYB_LIB'):
    for _k in ('hyb_symm_solve', 'hyb_schur_set'):
        _lib.SIGNATURES.pop(_k, None)
from hypatia_b200.lib import Context
ctx = Context(0)
rng = np.random.default_rng(0)
def omega(A, x, b):
    return np.max(np.abs(b - A @ x) / (np.abs(A) @ np.abs(x) + np.abs(b)))
for n in (60, 64, 200, 500, 1500):
    for grade in (0, 4, 8):
        G = rng.standard_normal((2 * n, n))
        d = 10.0 ** (grade * (rng.random(2 * n) - 0.5))
        A = (G.T * d) @ G
        A = 0.5 * (A + A.T)
        b = rng.standard_normal(n)
        U, x, info = ctx.posdef_solve(A, b)
        c, low = sla.cho_factor(A)
        xl = sla.cho_solve((c, low), b)
        print(f"n={n:5d} grade 1e{grade}: cond {np.linalg.cond(A):.1e} info {info} omega gpu {omega(A, x, b):.2e} lapack {omega(A, xl, b):.2e}  |x-xl|/|xl| {np.linalg.norm(x-xl)/np.linalg.norm(xl):.2e}")
