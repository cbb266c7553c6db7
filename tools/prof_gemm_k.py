"""Rate of the update form C (+)= A'B of the DMMA engine against the contraction length K (the Cholesky's
trailing updates: K = 128 near updates, K = 512 far updates)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hypatia_b200
from hypatia_b200.lib import Context
ctx = Context(0)
for beta in ("0", "1"):
    os.environ["HYB_BENCH_GEMM_BETA"] = beta
    for (M, N, upper) in ((1024, 18944, False), (5120, 5120, True)):
        for K in (128, 256, 512, 1024, 4096):
            ms = ctx.bench_gemm_tn(M, N, K, upper, reps=5)
            fl = K * N * (N + 1) if upper else 2 * M * N * K
            tiles = (N // 128) * (N // 128 + 1) // 2 if upper else (M // 128) * (N // 128)
            print(f"beta={beta} M={M} N={N} upper={upper} K={K}: {ms*1e3:.0f} us {fl/ms/1e9:.1f} TFLOP/s  tiles {tiles} ({tiles/148:.2f} waves) -> {ms*1e3/ (-(-tiles//148)):.1f} us per wave")
