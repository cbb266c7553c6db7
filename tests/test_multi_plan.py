"""Partition of the cones over several GPUs (csrc/multi.cu: plan_shards, through hyb_multi_plan) -- host logic, no
GPU: contiguous row ranges that cover [0, q), cones kept whole except Nonnegative ones (cut exactly at the even
target row so that LPs shard too), balanced to within the largest atom."""
import ctypes as C

import numpy as np
import pytest

import hypatia_b200  # noqa: F401
from hypatia_b200 import cones_spec as cs
from hypatia_b200.lib import ConeDesc, load


def _descs(specs):
    arr = (ConeDesc * max(len(specs), 1))()
    off = 0
    for k, s in enumerate(specs):
        arr[k].type = s.type_code
        arr[k].use_dual_barrier = int(s.use_dual_barrier)
        arr[k].offset = off
        arr[k].dim = s.dim
        off += s.dim
    return arr, off


def _plan(specs, ndev):
    lib = load()
    descs, q = _descs(specs)
    lo, hi, cnt = (C.c_int64 * ndev)(), (C.c_int64 * ndev)(), (C.c_int64 * ndev)()
    assert lib.hyb_multi_plan(q, len(specs), descs, ndev, lo, hi, cnt) == 0
    return q, list(lo), list(hi), list(cnt)


@pytest.mark.parametrize("ndev", [1, 2, 3, 4, 8])
def test_block_model_partition(ndev):
    specs = ([cs.Nonnegative(4096)] + [cs.EpiNormEucl(8)] * 512 + [cs.EpiPerSquare(8)] * 512 + [cs.PosSemidefTri(136)] * 30)
    q, lo, hi, cnt = _plan(specs, ndev)
    assert lo[0] == 0 and hi[-1] == q
    assert all(hi[d] == lo[d + 1] for d in range(ndev - 1))
    rows = np.array(hi) - np.array(lo)
    assert rows.min() > 0
    assert rows.max() - rows.min() <= 2 * 136  # balanced to within the largest atom (a PSD cone of 136 rows)
    # a boundary is either inside the orthant (then even) or at a cone boundary
    bounds = set(np.cumsum([s.dim for s in specs]).tolist())
    for b in hi[:-1]:
        assert (b < 4096 and b % 2 == 0) or b in bounds
    assert sum(cnt) >= len(specs)  # split orthants count once per piece


def test_lp_is_cut_evenly():
    q, lo, hi, cnt = _plan([cs.Nonnegative(1000)], 4)
    assert [h - l for l, h in zip(lo, hi)] == [250, 250, 250, 250] and cnt == [1, 1, 1, 1]


def test_one_giant_cone_stays_whole():
    q, lo, hi, cnt = _plan([cs.PosSemidefTri(cs.svec_length(60)), cs.Nonnegative(10)], 3)
    assert hi[-1] == q and sum(cnt) >= 2
    # the PSD cone is an atom: it lives on exactly one device
    owners = [d for d in range(3) if lo[d] <= 0 < hi[d]]
    assert len(owners) == 1 and hi[owners[0]] >= cs.svec_length(60)


def test_more_devices_than_rows():
    q, lo, hi, cnt = _plan([cs.EpiNormEucl(3)], 4)
    assert hi[-1] == q == 3 and sum(h - l for l, h in zip(lo, hi)) == 3
