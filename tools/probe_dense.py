"""GPU-box probe: FP64 GEMM measuring stick (cuBLAS through torch) beside this library's
DMMA gemm/syrk, Cholesky and triangular solves.  Writes gpurun_out/probe_dense.json."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import hypatia_b200  # noqa: F401
from hypatia_b200.lib import Context

out = {}
dev = torch.device("cuda:0")
print(torch.cuda.get_device_name(0))
for n in (4096, 8192):
    a = torch.randn(n, n, dtype=torch.float64, device=dev)
    b = torch.randn(n, n, dtype=torch.float64, device=dev)
    for _ in range(2):
        torch.matmul(a.T, b)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    e0.record()
    for _ in range(reps):
        torch.matmul(a.T, b)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    out[f"cublas_dgemm_tn_{n}_tflops"] = 2 * n**3 / ms / 1e9
    print(f"cuBLAS dgemm TN {n}: {ms:.2f} ms  {2*n**3/ms/1e9:.2f} TFLOP/s")
    if n == 8192:
        A = torch.randn(n, n, dtype=torch.float64, device=dev)
        S = A @ A.T + n * torch.eye(n, dtype=torch.float64, device=dev)
        torch.linalg.cholesky(S)
        torch.cuda.synchronize()
        e0.record()
        torch.linalg.cholesky(S)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        out["cusolver_potrf_8192_ms"] = ms
        print(f"cuSOLVER potrf 8192: {ms:.2f} ms  {n**3/3/ms/1e9:.2f} TFLOP/s")
    del a, b
torch.cuda.empty_cache()

ctx = Context(0)
for (M, N, K, upper) in [(4096, 4096, 4096, False), (8192, 8192, 8192, False), (8192, 8192, 16384, True), (2048, 2048, 128, True), (8192, 8192, 128, True)]:
    ms = ctx.bench_gemm_tn(M, N, K, upper, reps=3)
    flops = (K * N * (N + 1) if upper else 2 * M * N * K)
    out[f"hyb_gemm_{M}_{N}_{K}_{'U' if upper else 'F'}"] = dict(ms=ms, tflops=flops / ms / 1e9)
    print(f"hyb gemm_tn M={M} N={N} K={K} upper={upper}: {ms:.3f} ms  {flops/ms/1e9:.2f} TFLOP/s")
for N in (2048, 8192):
    mf, ms = ctx.bench_potrf(N, reps=2)
    out[f"hyb_potrf_{N}"] = dict(ms_factor=mf, ms_solve=ms, tflops=N**3 / 3 / mf / 1e9, solve_gbs=8 * N * N / ms / 1e6)
    print(f"hyb potrf N={N}: {mf:.2f} ms ({N**3/3/mf/1e9:.2f} TFLOP/s); potrs {ms:.3f} ms ({8*N*N/ms/1e6:.0f} GB/s)")
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/probe_dense.json", "w"), indent=1)
