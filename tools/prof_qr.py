"""Time the GPU column-pivoted QR (hyb_geqp3) against LAPACK dgeqp3 on the host."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, scipy.linalg as sla
import hypatia_b200  # noqa
from hypatia_b200.lib import Context
ctx = Context(0)
rng = np.random.default_rng(0)
for m, n, cpu in ((4092, 2048, True), (8184, 4096, True), (16368, 8192, False)):
    M = np.asfortranarray(rng.standard_normal((m, n)) / np.sqrt(n))
    ctx.geqp3(M[:64, :32])
    ctx.stats(reset=True)
    t = time.time(); qr, tau, piv = ctx.geqp3(M); tg = time.time() - t
    dev_ms = ctx.stats(reset=True)[0]
    line = f"geqp3 {m}x{n}: GPU {tg:.3f} s wall ({dev_ms:.0f} ms on the device)"
    if cpu:
        t = time.time(); (q2, t2), _, p2 = sla.qr(M, mode="raw", pivoting=True); tc = time.time() - t
        line += f"; LAPACK {tc:.2f} s ({os.cpu_count()} cores); same pivots: {np.array_equal(piv, p2)}; max |diag| diff {np.abs(np.abs(np.diagonal(qr)) - np.abs(np.diagonal(q2))).max():.2e}"
    print(line, flush=True)
