"""Multi-GPU, root/worker form (the north-star shape of SURVEY §8e).

Rank 0 is the only rank that runs the interior-point driver: it holds the full model on
its GPU and serves every oracle, solve and line-search check exactly like the single-GPU
path.  What is distributed is the expensive, embarrassingly parallel part of
`update_lhs_fact` (qrchol.jl:209-265):

    lhs = sum_k (W_k GQ2_k)' (W_k GQ2_k)        summed over the cones k

Every rank (the root included, through a second "shard" context) owns a contiguous range
of cones and the matching rows of G*Q2.  Per iteration:

  1. root broadcasts the accepted point (one 2q+2 vector)           -- 1 collective
  2. every rank loads its slice, runs the W*GQ2 product kernels and the DMMA syrk whose
     epilogue stores each finished tile straight into the rank's slot in the ROOT's memory
     (CUDA IPC peer pointer, NVLink)                                 -- no collective
  3. barrier; root adds the slots in rank order (deterministic) and factorizes.

Workers sit in `serve()` between iterations.  Nothing else is exchanged: the solves, the
cone oracles on vectors and the neighbourhood checks all stay on the factorizing GPU.
"""
from __future__ import annotations

import ctypes as C
import time

import numpy as np

from .gpu import B200QRCholSystemSolver, GpuConeBatch
from .lib import Context, dptr, f64
from .sharded import Comm, partition_cones, shard_atoms

CMD_STOP, CMD_ASSEMBLE = 0.0, 1.0


class ShardEngine:
    """One rank's shard: its rows of G*Q2 and its cone atoms in a context of their own."""

    def __init__(self, device: int, atoms):
        self.ctx = Context(device)
        self.lib = self.ctx.lib
        # a shard only assembles into the root's memory: no Schur matrix / Cholesky workspace of its own
        self.ctx.check(self.lib.hyb_ctx_set_assemble_only(self.ctx.ptr, 1), "hyb_ctx_set_assemble_only")
        self.cones = GpuConeBatch(atoms, self.ctx)
        self.q = self.cones.q

    def load(self, G_rows: np.ndarray) -> None:
        G = f64(G_rows)
        self.N = G_rows.shape[1]
        self.ctx.check(
            self.lib.hyb_model_load(self.ctx.ptr, self.q, self.N, dptr(G), self.cones.K, self.cones.descs), "hyb_model_load"
        )
        self.cones.model_loaded = True
        self.cones._pending_primal = None
        self.cones._pending_dual = np.zeros(self.q)
        self.cones._pending_scal = 1.0

    def attach(self, payload: bytes, rank: int, world: int) -> bool:
        """Map the root's slots (CUDA IPC handle) and point this shard at slot `rank`."""
        self._err = ""
        rc = self.lib.hyb_peer_attach(self.ctx.ptr, payload, rank, world)
        if rc != 0:
            self._err = self.lib.hyb_last_error(self.ctx.ptr).decode()
        return rc == 0

    def assemble_into_slot(self, primal: np.ndarray, dual: np.ndarray, scal: float) -> float:
        self.cones.load_point(primal, scal)
        self.cones.load_dual_point(dual)
        self.cones._sync()
        ta = C.c_double(0)
        self.ctx.check(self.lib.hyb_schur_assemble_peer(self.ctx.ptr, C.byref(ta)), "hyb_schur_assemble_peer")
        return ta.value


def _partition(specs, world):
    atoms = shard_atoms(specs, world)
    parts = partition_cones(atoms, world)
    rows = [int(sum(atoms[k].dim for k in part)) for part in parts]
    lo = np.concatenate([[0], np.cumsum(rows)]).astype(int)
    return atoms, parts, rows, lo


class RootAssembly:
    """Root side of the protocol, independent of where a shard computes (mixed into a QRChol
    system solver).  The concrete class supplies `_make_shard(atoms)`, `_publish_slots()` ->
    bytes the workers attach with, `_attach_own_shard()` and `_current_point(cones)`."""

    is_worker = False

    def _rw_init(self, comm: Comm) -> None:
        self.comm = comm
        self.time_reduce_ms = 0.0
        self.time_bcast_ms = 0.0
        self._msg_len = None

    def _rw_setup(self, solver) -> None:
        model, comm = solver.model, self.comm
        self._n = model.n
        atoms, parts, rows, lo = _partition(model.cones, comm.world)
        self.row_counts, self.row_lo = rows, lo
        self.shard = self._make_shard([atoms[k] for k in parts[0]])
        self.shard.load(np.asfortranarray(model.G[lo[0] : lo[1], :]))
        comm.bcast_object(("load", model.n, self._publish_slots()))
        # the workers never preprocess: they get their rows of the reduced G*Q2 from the root
        torch = comm.torch
        for r in range(1, comm.world):
            if rows[r] == 0:
                continue
            blk = np.ascontiguousarray(model.G[lo[r] : lo[r + 1], :].T)  # (n, rows_r) row-major == column-major rows
            comm.dist.send(torch.from_numpy(blk).to(comm.device), dst=r)
        self._attach_own_shard()
        ok = comm.allreduce_min(np.array([1.0]))[0] > 0.5
        if not ok:
            raise RuntimeError("a worker rank could not map the root's peer slots (CUDA IPC / P2P unavailable)")
        self._q = model.q
        self._msg_len = 2 * model.q + 2

    def _broadcast_point(self, cmd: float, point) -> None:
        torch = self.comm.torch
        q = self._q
        msg = np.zeros(self._msg_len)
        msg[0] = cmd
        if cmd == CMD_ASSEMBLE:
            msg[1], msg[2 : 2 + q], msg[2 + q :] = point
        t = torch.from_numpy(msg).to(self.comm.device)
        self.comm.dist.broadcast(t, src=0)

    def _rw_collect(self, solver) -> float:
        """Broadcast the point, assemble the root's own share, wait for every slot.  Returns the
        device time (ms) of the root's own assembly."""
        scal, primal, dual = point = self._current_point(solver.cones)
        t0 = time.perf_counter()
        self._broadcast_point(CMD_ASSEMBLE, point)
        self.time_bcast_ms += 1e3 * (time.perf_counter() - t0)
        lo = self.row_lo
        ta = self.shard.assemble_into_slot(primal[lo[0] : lo[1]], dual[lo[0] : lo[1]], scal)
        t0 = time.perf_counter()
        self.comm.barrier()
        self.time_reduce_ms += 1e3 * (time.perf_counter() - t0)
        return ta

    def free_memory(self) -> None:  # Solvers.jl:390 -- also releases the workers
        if self._msg_len:
            self._broadcast_point(CMD_STOP, None)
            self._msg_len = None

    def abort_workers(self) -> None:
        """The root found the model inconsistent during preprocessing: no system is ever loaded."""
        self.comm.bcast_object(("abort", 0, b""))


class RootQRCholSystemSolver(RootAssembly, B200QRCholSystemSolver):
    """Rank 0: the single-GPU system solver plus the distributed Schur assembly."""

    def __init__(self, ctx: Context, comm: Comm, device: int, batched_dirs: bool = True):
        B200QRCholSystemSolver.__init__(self, ctx, batched_dirs=batched_dirs)
        self.device_stepper = False  # the accepted point is broadcast to the workers from host memory
        self._rw_init(comm)
        self.device = device

    def _make_shard(self, atoms):
        return ShardEngine(self.device, atoms)

    def _publish_slots(self) -> bytes:
        # peer slots live in the root context; workers map them through CUDA IPC
        handle = C.create_string_buffer(64)
        self.ctx.check(self.lib.hyb_peer_root_create(self.ctx.ptr, self.comm.world, handle), "hyb_peer_root_create")
        return bytes(handle.raw)

    def _attach_own_shard(self) -> None:
        ptr, r_, c_, ld_ = C.c_void_p(0), C.c_int64(0), C.c_int64(0), C.c_int64(0)
        self.ctx.check(self.lib.hyb_device_buffer(self.ctx.ptr, 1, C.byref(ptr), C.byref(r_), C.byref(c_), C.byref(ld_)), "hyb_device_buffer")
        self.shard.ctx.check(self.lib.hyb_peer_set_slot(self.shard.ctx.ptr, ptr), "hyb_peer_set_slot")

    def _current_point(self, cones):
        cones._sync()
        return cones._pending_scal, cones._pending_primal, cones._pending_dual

    def load(self, solver) -> None:
        super().load(solver)  # full model on the root context (solves, oracles, factorization)
        self._rw_setup(solver)

    def update_lhs(self, solver) -> None:
        model, cones = solver.model, solver.cones
        self.time_assemble_ms += self._rw_collect(solver)
        t0 = time.perf_counter()
        ts = C.c_double(0)
        self.ctx.check(self.lib.hyb_peer_root_sum(self.ctx.ptr, C.byref(ts)), "hyb_peer_root_sum")
        self.time_reduce_ms += 1e3 * (time.perf_counter() - t0)
        tf, kind = C.c_double(0), C.c_int(0)
        rc = self.ctx.check(self.lib.hyb_schur_factor(self.ctx.ptr, C.byref(tf), C.byref(kind)), "hyb_schur_factor")
        self.time_fact_ms += tf.value
        self._after_factor(solver, tf.value, kind.value, rc)


class WorkerQRCholSystemSolver:
    """Ranks > 0: assemble the partial Schur matrix of their cone range on request."""

    is_worker = True
    device_G = False

    def __init__(self, comm: Comm, device: int = 0, shard_factory=None):
        self.comm = comm
        self.device = device
        self.shard_factory = shard_factory or (lambda atoms: ShardEngine(device, atoms))
        self.time_assemble_ms = 0.0
        self.num_updates = 0

    def cones_factory(self, specs):  # the worker's host driver never evaluates an oracle
        return None

    def serve(self, solver) -> None:
        """Everything a worker does: wait for the root's reduced model, then assemble on request."""
        comm, torch = self.comm, self.comm.torch
        specs = solver.orig_model.cones
        what, n, raw = comm.bcast_object(None)
        if what == "abort":
            return
        atoms, parts, rows, lo = _partition(specs, comm.world)
        r = comm.rank
        q = self.q = int(lo[-1])
        self.lo, self.hi = int(lo[r]), int(lo[r + 1])
        self.shard = self.shard_factory([atoms[k] for k in parts[r]])
        if rows[r] > 0:
            t = torch.empty((n, rows[r]), dtype=torch.float64, device=comm.device)
            comm.dist.recv(t, src=0)
            G_rows = np.asfortranarray(t.cpu().numpy().T)
        else:
            G_rows = np.zeros((0, n), order="F")
        self.shard.load(G_rows)
        mine = self.shard.attach(raw, r, comm.world)
        ok = comm.allreduce_min(np.array([1.0 if mine else 0.0]))[0] > 0.5
        if not ok:
            raise RuntimeError("peer slot mapping failed: " + getattr(self.shard, "_err", ""))
        while True:
            t = torch.empty(2 * q + 2, dtype=torch.float64, device=comm.device)
            comm.dist.broadcast(t, src=0)
            msg = t.cpu().numpy()
            if msg[0] == CMD_STOP:
                break
            primal, dual = msg[2 : 2 + q], msg[2 + q :]
            self.time_assemble_ms += self.shard.assemble_into_slot(primal[self.lo : self.hi], dual[self.lo : self.hi], float(msg[1]))
            self.num_updates += 1
            comm.barrier()

    def free_memory(self) -> None:
        pass


def make_rootworker_solver(device: int, dist, batched_dirs: bool = True, **solver_kwargs):
    import torch

    from .host.solver import Solver

    backend = dist.get_backend()
    comm = Comm(dist, torch.device("cuda", device) if backend == "nccl" else torch.device("cpu"))
    if comm.rank == 0:
        ctx = Context(device)
        ss = RootQRCholSystemSolver(ctx, comm, device, batched_dirs=batched_dirs)
        solver = Solver(ss.cones_factory, ss, **solver_kwargs)
        from .gpu import attach_gpu_qr

        attach_gpu_qr(solver, ctx)
    else:
        ss = WorkerQRCholSystemSolver(comm, device)
        solver = Solver(ss.cones_factory, ss, **solver_kwargs)
    return solver
