"""world_size-2 CPU (gloo) test of the multi-GPU layer's host logic: cone partition,
row slicing, all-gather / all-reduce stitching and the reduce-to-root of the partial
Schur matrices, with a numpy local backend standing in for the rank's GPU.  The sharded
solve must reproduce the single-process oracle solve."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, which, out):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch
    import torch.distributed as dist

    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import hypatia_b200  # noqa: F401
        from hypatia_b200 import instances
        from hypatia_b200.host.solver import Solver
        from hypatia_b200.sharded import Comm, ShardedQRCholSystemSolver
        from oracle.cones import CpuConeSet
        from oracle.dense import posdef_fact_copy

        class NumpyLocalBackend:
            """Rank-local shard on the CPU (test double for GpuLocalBackend)."""

            def local_cones_factory(self, specs):
                return CpuConeSet(specs)

            def load(self, G_rows, local_cones):
                self.G = np.array(G_rows)
                self.cones = local_cones
                self.N = G_rows.shape[1]

            def assemble(self):
                acc = np.zeros((self.N, self.N))
                for k, c in enumerate(self.cones.cones):
                    Gk = self.G[self.cones.offs[k] : self.cones.offs[k + 1], :]
                    W = c.inv_sqrt_hess_prod(Gk) if c.use_dual_barrier else c.sqrt_hess_prod(Gk)
                    acc += W.T @ W
                self.S = torch.from_numpy(acc)
                return 0.0

            def reduce_to_root(self, comm):
                comm.reduce_tensor_to_root(self.S)

            def factor(self):
                self.fact = posdef_fact_copy(np.triu(self.S.numpy()))
                return 0.0, 0, not self.fact.success

            def solve(self, rhs):
                return self.fact.solve(rhs)

            def G_mul(self, x):
                return self.G @ x

            def Gt_mul(self, z_local):
                return self.G.T @ z_local

        comm = Comm(dist)
        model = getattr(instances, which[0])(**which[1])
        ss = ShardedQRCholSystemSolver(NumpyLocalBackend(), comm)
        solver = Solver(ss.cones_factory, ss, trace=True)
        solver.setup_comm = comm
        solver.load(model)
        solver.solve()
        if rank == 0:
            out.put((solver.status, solver.num_iters, solver.primal_obj, solver.get_x(), [r["mu"] for r in solver.trace],
                     solver.cones.row_counts))
    finally:
        dist.destroy_process_group()


def _run_sharded(which, world=2):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, which, out)) for r in range(world)]
    for p in procs:
        p.start()
    res = out.get(timeout=300)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    return res


def _run_single(which):
    import hypatia_b200  # noqa: F401
    from hypatia_b200 import instances
    from hypatia_b200.host.solver import Solver
    from oracle.cones import CpuConeSet
    from oracle.qrchol import QRCholDenseSystemSolver

    solver = Solver(CpuConeSet, QRCholDenseSystemSolver(), trace=True)
    solver.load(getattr(instances, which[0])(**which[1]))
    solver.solve()
    return solver


@pytest.mark.parametrize(
    "which",
    [("block_multicone", dict(N=48, seed=4, psd_side=4)), ("linearopt", dict(m=20, n=60, seed=2))],
    ids=["block_multicone", "linearopt_split_nonnegative"],
)
def test_sharded_solve_matches_single_process(which):
    status, iters, obj, x, mus, row_counts = _run_sharded(which, world=2)
    ref = _run_single(which)
    assert min(row_counts) > 0 and sum(row_counts) == ref.model.q  # both ranks own rows
    assert status == ref.status == "Optimal"
    assert abs(iters - ref.num_iters) <= 1
    assert abs(obj - ref.primal_obj) <= 1e-7 * (1 + abs(ref.primal_obj))
    assert np.allclose(x, ref.get_x(), rtol=1e-5, atol=1e-6)
    n = min(len(mus), len(ref.trace)) // 2
    assert np.allclose(mus[:n], [r["mu"] for r in ref.trace[:n]], rtol=1e-8)


def test_partition_is_contiguous_and_balanced():
    import hypatia_b200  # noqa: F401
    from hypatia_b200 import cones_spec as cs
    from hypatia_b200.sharded import partition_cones, shard_atoms

    specs = [cs.Nonnegative(1000)] + [cs.EpiNormEucl(8)] * 100 + [cs.PosSemidefTri(136)] * 5
    for world in (1, 2, 4, 8):
        atoms = shard_atoms(specs, world)
        assert sum(a.dim for a in atoms) == sum(s.dim for s in specs)
        parts = partition_cones(atoms, world)
        assert parts[0].start == 0 and parts[-1].stop == len(atoms)
        assert all(parts[i].stop == parts[i + 1].start for i in range(world - 1))
        rows = [sum(atoms[k].dim for k in p) for p in parts]
        assert max(rows) - min(rows) <= 136 + 1000 // world
