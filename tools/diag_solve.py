import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, scipy.linalg as sla
import hypatia_b200
from hypatia_b200.lib import Context
ctx = Context(0)
rng = np.random.default_rng(0)
def check(A, b, tag):
    U, x, info = ctx.posdef_solve(np.triu(A), b)
    xr = sla.cho_solve(sla.cho_factor(A), b)
    Ur = sla.cholesky(A)
    be = np.linalg.norm(A @ x - b) / (np.linalg.norm(A) * np.linalg.norm(x) + np.linalg.norm(b))
    ber = np.linalg.norm(A @ xr - b) / (np.linalg.norm(A) * np.linalg.norm(xr) + np.linalg.norm(b))
    print(f"{tag}: info {info} cond {np.linalg.cond(A):.2e} U relerr {np.abs(U-Ur).max()/np.abs(Ur).max():.2e} x relerr {np.abs(x-xr).max()/np.abs(xr).max():.2e} backward err gpu {be:.2e} lapack {ber:.2e}")
a, c = 1.26992725e9, 2.54293585e9
b_ = 1.796e9
check(np.array([[a, b_], [b_, c]]), np.array([1.0, -2.0]), "2x2 iter6-like")
check(np.array([[a, b_], [b_, c]]), np.array([3.3e8, -7.1e9]), "2x2 big rhs")
for N, cond in [(2, 1e6), (3, 1e8), (50, 1e8), (128, 1e10), (300, 1e10), (1000, 1e12)]:
    Q, _ = np.linalg.qr(rng.standard_normal((N, N)))
    d = np.logspace(0, np.log10(cond), N)
    A = (Q * d) @ Q.T
    A = (A + A.T) / 2
    check(A, rng.standard_normal(N), f"N={N} cond~{cond:.0e}")
