"""One Schur-shaped syrk launch (cfg5 N=8192: K=q=16368) for ncu captures."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hypatia_b200
from hypatia_b200.lib import Context
ctx = Context(0)
N, K = 8192, 16368
ms = ctx.bench_gemm_tn(N, N, K, True, reps=1)
print(f"syrk N={N} K={K}: {ms:.3f} ms  {K*N*(N+1)/ms/1e9:.2f} TFLOP/s")
