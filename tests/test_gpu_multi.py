"""Several GPUs from ONE host process through the C-ABI alone (hyb_multi_*, csrc/multi.cu): no torch, no
torch.distributed -- what a Julia host would do with ccalls.  With one visible GPU the devices are [0, 0]
(two shards on the same GPU: the same code path except that the peer stores stay on the device)."""
import numpy as np
import pytest

import hypatia_b200  # noqa: F401
from hypatia_b200 import cones_spec as cs
from hypatia_b200 import instances
from hypatia_b200.gpu import make_gpu_solver
from hypatia_b200.host.solver import Solver
from hypatia_b200.lib import load
from oracle.cones import CpuConeSet
from oracle.qrchol import QRCholDenseSystemSolver

from gpu_util import rel_err

pytestmark = pytest.mark.gpu
KW = dict(tol_rel_opt=1e-7, tol_feas=1e-7, tol_abs_opt=1e-10, tol_infeas=1e-10)  # benchmarks/natvsext/run.jl:44-60


def _devices(n):
    lib = load()
    # device count through the library's own context creation: device k exists iff hyb_ctx_create(k) works
    have = 0
    for k in range(8):
        p = lib.hyb_ctx_create(k)
        if not p:
            break
        lib.hyb_ctx_destroy(p)
        have += 1
    return [k % have for k in range(n)]


# (N, ndev): N = 160 is two tile columns (with three devices one owner holds nothing); 520 and 700 spread five and six
# tile columns, the last one ragged, over three and four owners with two compact column blocks each
@pytest.mark.parametrize("N,ndev", [(160, 2), (160, 3), (520, 3), (700, 4)])
def test_partial_schur_matrices_sum_to_the_single_gpu_matrix(N, ndev):
    def model():
        return instances.block_multicone(N=N, seed=5, psd_side=8)

    def prepared(**kw):
        s = make_gpu_solver(0, **kw)
        s.load(model())
        s._setup_solver()
        s._setup_loaded()
        s._initialize_point()
        s._setup_stepping()
        s._calc_convergence_params()
        s.syssolver.update_lhs(s)
        return s

    one = prepared()
    many = prepared(devices=_devices(ndev))
    ranges = many.syssolver.row_ranges
    assert len(ranges) == ndev and ranges[0][0] == 0 and ranges[-1][1] == many.model.q
    assert all(a[1] == b[0] for a, b in zip(ranges, ranges[1:])) and all(hi > lo for lo, hi in ranges)
    S1 = np.triu(one.syssolver.schur_matrix())
    Sm = np.triu(many.syssolver.schur_matrix())
    assert rel_err(Sm, S1) < 1e-13  # same products, summed in a different order
    assert many.syssolver.last_fact_kind == 0


@pytest.mark.parametrize(
    "which",
    [("block_multicone", dict(N=160, seed=5, psd_side=8)), ("linearopt", dict(m=60, n=200, seed=3)),
     ("portfolio_norminf", dict(K=120, d=4, n=24, seed=2))],
    ids=["block_multicone", "linearopt_split_nonnegative", "norminf_gemm_branch"],
)
def test_single_process_multi_gpu_solve_matches_oracle(which):
    build = lambda: getattr(instances, which[0])(**which[1])  # noqa: E731
    gpu = make_gpu_solver(0, devices=_devices(2), trace=True, **KW)
    gpu.load(build())
    gpu.solve()
    ref = Solver(CpuConeSet, QRCholDenseSystemSolver(), trace=True, **KW)
    ref.load(build())
    ref.solve()
    assert gpu.status == ref.status == "Optimal"
    assert abs(gpu.num_iters - ref.num_iters) <= 1
    assert abs(gpu.primal_obj - ref.primal_obj) <= 1e-7 * (1 + abs(ref.primal_obj))
    n = max(3, min(len(gpu.trace), len(ref.trace)) // 2)
    assert np.allclose([r["mu"] for r in gpu.trace[:n]], [r["mu"] for r in ref.trace[:n]], rtol=1e-8)
    assert all(hi > lo for lo, hi in gpu.syssolver.row_ranges)


def test_one_device_list_is_the_single_gpu_path():
    gpu = make_gpu_solver(0, devices=[0], mode="single_process", **KW)
    gpu.load(instances.block_multicone(N=64, seed=0, psd_side=8))
    gpu.solve()
    assert gpu.status == "Optimal"
    assert gpu.syssolver.row_ranges == [(0, gpu.model.q)]
