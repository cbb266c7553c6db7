"""Allocation-free gemv entry point (hyb_gemv_n_noalloc), collected last on purpose.

Some hosts cannot let this library allocate: their own allocator already owns the GPU
(a Julia session with CUDA.jl's pool, a PyTorch caching allocator) or the call sits in a
section where cudaMalloc's device-wide synchronisation is not acceptable.  For them the
library offers y = G x on caller-owned device pointers with its scratch in a fixed device
array.  This test drives that entry point through the C-ABI, twice in a row (the scratch is
shared and keeps the previous call's contents), and compares each result with the regular
entry point hyb_G_mul on the same data, at the tolerance the other gemv tests use.
"""
import ctypes as C

import numpy as np
import pytest

import hypatia_b200  # noqa: F401
from hypatia_b200 import cones_spec as cs
from hypatia_b200.lib import ConeDesc, Context, dptr, f64

pytestmark = pytest.mark.gpu


def reference_product(ctx, G, x):
    """y = G x through the existing entry point (model load + hyb_G_mul)."""
    q, N = G.shape
    desc = (ConeDesc * 1)()
    desc[0].type = cs.NONNEGATIVE
    desc[0].offset = 0
    desc[0].dim = q
    Gf = f64(G)
    ctx.check(ctx.lib.hyb_model_load(ctx.ptr, q, N, dptr(Gf), 1, desc), "hyb_model_load")
    out = np.empty(q)
    xx = f64(x)
    ctx.check(ctx.lib.hyb_G_mul(ctx.ptr, b"N", dptr(xx), dptr(out)), "hyb_G_mul")
    return out


def test_noalloc_gemv_matches_regular_entry_point_twice():
    import torch

    ctx = Context(0)
    rng = np.random.default_rng(7)
    shapes = [(1000, 333), (4097, 129)]  # second call reuses the scratch the first one dirtied
    for q, N in shapes:
        G = np.asfortranarray(rng.standard_normal((q, N)))
        x = rng.standard_normal(N)
        want = reference_product(ctx, G, x)
        # caller-owned device buffers (column-major G with ld = q)
        dG = torch.from_numpy(np.ascontiguousarray(G.T)).cuda()  # memory layout == column-major q x N
        dx = torch.from_numpy(x).cuda()
        dy = torch.full((q,), float("nan"), dtype=torch.float64, device="cuda")
        torch.cuda.synchronize()
        rc = ctx.lib.hyb_gemv_n_noalloc(dG.data_ptr(), q, N, q, dx.data_ptr(), dy.data_ptr(), None)
        assert rc == 0, ctx.lib.hyb_last_error(ctx.ptr)
        torch.cuda.synchronize()
        got = dy.cpu().numpy()
        assert np.max(np.abs(got - want)) <= 1e-13 * max(1.0, np.max(np.abs(want)))
        assert np.max(np.abs(got - G @ x)) <= 1e-12 * np.max(np.abs(want))


def test_noalloc_gemv_rejects_oversized_q():
    ctx = Context(0)
    rc = ctx.lib.hyb_gemv_n_noalloc(None, 65537, 4, 65537, None, None, None)
    assert rc < 0
    assert b"65536" in ctx.lib.hyb_last_error(ctx.ptr)
