"""Verbose GPU solve of one small instance (diagnostics)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import hypatia_b200
from hypatia_b200 import instances
from hypatia_b200 import lib as _lib
if os.environ.get('HYB_LIB'):
    for _k in ('hyb_symm_solve', 'hyb_schur_set'):
        _lib.SIGNATURES.pop(_k, None)
from hypatia_b200.gpu import make_gpu_solver
from hypatia_b200.lib import Context
s = make_gpu_solver(0, verbose=True)
s.load(instances.block_multicone(N=64, seed=4, psd_side=6))
s.iter_callback = lambda sv: print('   fact_kind', sv.syssolver.last_fact_kind, 'nref', sv.stepper.last_num_ref, 'ncand', sv.stepper.last_num_cand)
s.solve()
print(s.status, s.num_iters, s.primal_obj)
# backward error of the factorization
ctx = Context(0)
rng = np.random.default_rng(0)
for n in (64, 200, 1000, 3000):
    B = rng.standard_normal((n, n)); A = B @ B.T + 1e-3 * n * np.eye(n)
    U, X, info = ctx.posdef_solve(A, rng.standard_normal(n))
    print(n, info, np.abs(U.T @ U - A).max() / np.abs(A).max())
