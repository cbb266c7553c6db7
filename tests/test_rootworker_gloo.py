"""world_size-2 CPU (gloo) test of the root/worker multi-GPU protocol (rootworker.py): the
partition, the row hand-out, the per-iteration point broadcast, the slot hand-back and the
stop / abort messages, with numpy shards and a file-backed memory map standing in for the GPU
shards and the CUDA-IPC peer slots.  The root's solve must reproduce the single-process solve."""
import os
import socket
import sys
import tempfile

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _build(which):
    import hypatia_b200  # noqa: F401
    from hypatia_b200 import instances
    from hypatia_b200.host.models import Model

    if which[0] == "dual_inconsistent":
        # c outside the range of G' with a rank-deficient G: rejected during preprocessing
        from hypatia_b200.cones_spec import Nonnegative

        G = -np.array([[1.0, 1.0], [1.0, 1.0], [2.0, 2.0]])
        return Model(np.array([1.0, -1.0]), np.zeros((0, 2)), np.zeros(0), G, np.zeros(3), [Nonnegative(3)])
    if which[0] == "known":
        return instances.KNOWN_ANSWER[which[1]]()[0]
    return getattr(instances, which[0])(**which[1])


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
        from hypatia_b200.host.solver import Solver
        from hypatia_b200.rootworker import RootAssembly, WorkerQRCholSystemSolver
        from hypatia_b200.sharded import Comm
        from oracle.cones import CpuConeSet
        from oracle.dense import posdef_fact_copy
        from oracle.qrchol import QRCholDenseSystemSolver

        class RecordingConeSet(CpuConeSet):
            def load_point(self, primal, scal=1.0):
                self.last_primal, self.last_scal = np.array(primal), float(scal)
                super().load_point(primal, scal)

            def load_dual_point(self, dual):
                self.last_dual = np.array(dual)
                super().load_dual_point(dual)

        class NumpyShard:
            """A rank's shard on the CPU; its slot is a slice of a memory-mapped file."""

            def __init__(self, atoms):
                self.cones = CpuConeSet(atoms)
                self.cones.setup()

            def load(self, G_rows):
                self.G = np.array(G_rows)
                self.N = G_rows.shape[1]

            def attach(self, payload, rank, world):
                mm = np.memmap(payload.decode(), dtype=np.float64, mode="r+", shape=(world, self.N, self.N))
                self.mm, self.slot = mm, mm[rank]
                return True

            def assemble_into_slot(self, primal, dual, scal):
                cs_ = self.cones
                cs_.load_point(primal, scal)
                cs_.load_dual_point(dual)
                cs_.reset_data()
                assert cs_.is_feas()
                acc = np.zeros((self.N, self.N))
                for k, c in enumerate(cs_.cones):
                    Gk = self.G[cs_.offs[k] : cs_.offs[k + 1], :]
                    if c.use_sqrt_hess_oracles(self.N):
                        W = c.inv_sqrt_hess_prod(Gk) if c.use_dual_barrier else c.sqrt_hess_prod(Gk)
                        acc += W.T @ W
                    else:
                        acc += Gk.T @ (c.inv_hess_prod(Gk) if c.use_dual_barrier else c.hess_prod(Gk))
                self.slot[:] = acc
                self.mm.flush()
                return 0.0

        class CpuRoot(RootAssembly, QRCholDenseSystemSolver):
            def __init__(self, comm, tmpdir):
                QRCholDenseSystemSolver.__init__(self)
                self._rw_init(comm)
                self.path = os.path.join(tmpdir, "slots.bin")

            def cones_factory(self, specs):
                return RecordingConeSet(specs)

            def _make_shard(self, atoms):
                return NumpyShard(atoms)

            def _publish_slots(self):
                n = self._n
                self.mm = np.memmap(self.path, dtype=np.float64, mode="w+", shape=(self.comm.world, n, n))
                self.mm.flush()
                return self.path.encode()

            def _attach_own_shard(self):
                self.shard.attach(self.path.encode(), 0, self.comm.world)

            def _current_point(self, cones):
                return cones.last_scal, cones.last_primal, cones.last_dual

            def load(self, solver):
                QRCholDenseSystemSolver.load(self, solver)
                self._rw_setup(solver)

            def update_lhs_fact(self, solver):
                self._rw_collect(solver)
                mm = np.memmap(self.path, dtype=np.float64, mode="r", shape=self.mm.shape)
                lhs = np.zeros(mm.shape[1:], order="F")
                for r in range(mm.shape[0]):  # rank order, like hyb_peer_root_sum
                    lhs += mm[r]
                if os.environ.get("HYB_DEBUG_RW"):
                    QRCholDenseSystemSolver.update_lhs_fact(self, solver)
                    print("lhs diff", np.abs(np.triu(lhs) - np.triu(self.lhs)).max(), np.abs(self.lhs).max(), flush=True)
                self.lhs = lhs
                self.fact = posdef_fact_copy(np.triu(lhs))
                self.num_collects = getattr(self, "num_collects", 0) + 1

        comm = Comm(dist)
        model = _build(which)
        if rank == 0:
            with tempfile.TemporaryDirectory() as tmp:
                ss = CpuRoot(comm, tmp)
                solver = Solver(ss.cones_factory, ss, trace=True)
                solver.load(model)
                solver.solve()
                out.put(("root", solver.status, solver.num_iters, getattr(solver, "primal_obj", None),
                         [r["mu"] for r in solver.trace], getattr(ss, "row_counts", None), getattr(ss, "num_collects", 0)))
        else:
            ss = WorkerQRCholSystemSolver(comm, shard_factory=NumpyShard)
            solver = Solver(ss.cones_factory, ss)
            solver.load(model)
            solver.solve()
            out.put(("worker", solver.status, ss.num_updates))
    finally:
        dist.destroy_process_group()


def _run(which, world=2):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, which, out)) for r in range(world)]
    for p in procs:
        p.start()
    res = [out.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    root = [r for r in res if r[0] == "root"][0]
    workers = [r for r in res if r[0] == "worker"]
    return root, workers


@pytest.mark.parametrize(
    "which",
    [("block_multicone", dict(N=48, seed=2, psd_side=4)), ("linearopt", dict(m=20, n=60, seed=3))],
    ids=["block_multicone", "linearopt_split_nonnegative"],
)
def test_rootworker_solve_matches_single_process(which):
    import hypatia_b200  # noqa: F401
    from hypatia_b200.host.solver import Solver
    from oracle.cones import CpuConeSet
    from oracle.qrchol import QRCholDenseSystemSolver

    root, workers = _run(which)
    _, status, iters, obj, mus, row_counts, collects = root
    ref = Solver(CpuConeSet, QRCholDenseSystemSolver(), trace=True)
    ref.load(_build(which))
    ref.solve()
    assert status == ref.status == "Optimal"
    assert iters == ref.num_iters
    assert abs(obj - ref.primal_obj) <= 1e-9 * (1 + abs(ref.primal_obj))
    assert np.allclose(mus, [r["mu"] for r in ref.trace], rtol=1e-6)
    assert min(row_counts) > 0
    # every Schur update of the root was served by the worker, which then got the stop message
    assert workers == [("worker", "Worker", collects)] and collects >= iters


def test_rootworker_with_an_empty_shard():
    """Fewer cone atoms than ranks: a worker owns no rows, serves every update with a zero slot."""
    import hypatia_b200  # noqa: F401
    from hypatia_b200.host.solver import Solver
    from oracle.cones import CpuConeSet
    from oracle.qrchol import QRCholDenseSystemSolver

    which = ("known", "epinormeucl1")
    root, workers = _run(which, world=3)
    _, status, iters, obj, mus, row_counts, collects = root
    ref = Solver(CpuConeSet, QRCholDenseSystemSolver(), trace=True)
    ref.load(_build(which))
    ref.solve()
    assert min(row_counts) == 0 and sum(row_counts) == ref.model.q
    assert status == ref.status == "Optimal" and iters == ref.num_iters
    assert abs(obj - ref.primal_obj) <= 1e-9 * (1 + abs(ref.primal_obj))
    assert sorted(workers) == [("worker", "Worker", collects)] * 2


def test_rootworker_abort_releases_workers():
    root, workers = _run(("dual_inconsistent", {}))
    assert root[1] == "DualInconsistent"
    assert workers == [("worker", "Worker", 0)]
