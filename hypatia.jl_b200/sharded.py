"""Multi-GPU layer: cones partitioned by block across ranks.

    lhs = sum_k GQ2_k' H~_k GQ2_k        (qrchol.jl:229-254)

is a sum over independent cones, so rank g owns a contiguous range of cones, the
matching rows of G*Q2 and the matching slices of the point; it builds its partial
Schur matrix, one reduce (sum) brings the partials onto the factorizing rank (root),
which factorizes and solves; N-vectors are all-reduced / broadcast and q-vectors
all-gathered so every rank's (replicated, deterministic) host driver sees full vectors.

One process per GPU, ``torch.distributed`` for the plumbing (NCCL on GPUs, gloo in the
CPU tests).  The layer is generic over a *local backend* (the rank's shard):
``GpuLocalBackend`` below is the product; ``tests/`` plugs a numpy backend in to check
the sharding / stitching / collective logic with world_size 2 on CPU.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Sequence

import numpy as np

from . import cones_spec as cs
from .host.coneset import ConeSet
from .host.syscommon import QRCholSystemSolver


# ---------------------------------------------------------------------------------
def shard_atoms(specs: Sequence[cs.ConeSpec], world: int) -> List[cs.ConeSpec]:
    """The units that can be assigned to a rank.  Every cone is one atom, except that a
    Nonnegative cone (whose oracles are coordinate-wise: nonnegative.jl:55-141) is cut into
    up to `world` row ranges so that LPs shard too.  The host-level gates that look at a
    cone as a whole (s'z per cone, search.jl:112-130) still use the original cones."""
    atoms: List[cs.ConeSpec] = []
    for s in specs:
        if s.type_code == cs.NONNEGATIVE and world > 1 and s.dim >= 2 * world:
            base, rem = divmod(s.dim, world)
            for r in range(world):
                atoms.append(cs.Nonnegative(base + (1 if r < rem else 0)))
        else:
            atoms.append(s)
    return atoms


def partition_cones(specs: Sequence[cs.ConeSpec], world: int) -> List[range]:
    """Contiguous ranges of atoms balanced by rows (the syrk cost of a cone is q_k * N^2):
    rank r ends at the first atom boundary at or after r/world of the rows."""
    dims = np.array([s.dim for s in specs], dtype=np.int64)
    K = len(dims)
    world = max(1, world)
    cum = np.concatenate([[0], np.cumsum(dims)])
    total = int(cum[-1])
    bounds = [0]
    for r in range(1, world):
        target = total * r / world
        k = int(np.searchsorted(cum, target, side="left"))  # first boundary with cum >= target
        k = min(max(k, bounds[-1]), K)
        bounds.append(k)
    bounds.append(K)
    return [range(bounds[i], bounds[i + 1]) for i in range(world)]


class Comm:
    """Thin wrapper over torch.distributed for float64 vectors held in numpy."""

    def __init__(self, dist, device=None):
        import torch

        self.torch = torch
        self.dist = dist
        self.rank = dist.get_rank()
        self.world = dist.get_world_size()
        self.device = device if device is not None else torch.device("cpu")

    def _t(self, a: np.ndarray):
        return self.torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).to(self.device)

    def allreduce_sum(self, a: np.ndarray) -> np.ndarray:
        t = self._t(a)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return t.cpu().numpy()

    def allreduce_max(self, a: np.ndarray) -> np.ndarray:
        t = self._t(a)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return t.cpu().numpy()

    def allreduce_min(self, a: np.ndarray) -> np.ndarray:
        t = self._t(a)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MIN)
        return t.cpu().numpy()

    def bcast(self, a: np.ndarray, src: int = 0) -> np.ndarray:
        t = self._t(a)
        self.dist.broadcast(t, src=src)
        return t.cpu().numpy()

    def allgather_rows(self, local: np.ndarray, row_counts: Sequence[int]) -> np.ndarray:
        """Stack the ranks' row blocks (1-D or 2-D, row counts differ) in rank order."""
        local = np.asarray(local, dtype=np.float64)
        one_d = local.ndim == 1
        loc2 = local.reshape(local.shape[0], -1)
        m = loc2.shape[1]
        rmax = int(max(row_counts)) if len(row_counts) else 0
        pad = np.zeros((rmax, m))
        pad[: loc2.shape[0]] = loc2
        t = self._t(pad)
        outs = [self.torch.empty_like(t) for _ in range(self.world)]
        self.dist.all_gather(outs, t)
        full = np.concatenate([o.cpu().numpy()[: row_counts[r]] for r, o in enumerate(outs)], axis=0)
        return full[:, 0] if one_d else np.asfortranarray(full)

    def reduce_tensor_to_root(self, t):
        """In-place sum of a (device) tensor onto rank 0: NCCL reduce over NVLink on GPUs.
        With the gloo backend (CPU tests, or two test ranks sharing one GPU) a CUDA tensor is
        staged through host memory."""
        if t.is_cuda and self.dist.get_backend() == "gloo":
            h = t.cpu()
            self.dist.reduce(h, dst=0, op=self.dist.ReduceOp.SUM)
            if self.rank == 0:
                t.copy_(h)
            return
        self.dist.reduce(t, dst=0, op=self.dist.ReduceOp.SUM)

    def bcast_object(self, obj, src: int = 0):
        box = [obj]
        self.dist.broadcast_object_list(box, src=src)
        return box[0]

    def barrier(self):
        self.dist.barrier()


# ---------------------------------------------------------------------------------
class ShardedConeSet(ConeSet):
    """Full-model cone set whose oracles run on the owning ranks."""

    def __init__(self, specs, comm: Comm, local_factory):
        super().__init__(specs)
        self.comm = comm
        self.atoms = shard_atoms(self.specs, comm.world)
        self.parts = partition_cones(self.atoms, comm.world)
        self.mine = self.parts[comm.rank]
        self.row_counts = [int(sum(self.atoms[k].dim for k in part)) for part in self.parts]
        self.row_lo = int(sum(self.row_counts[: comm.rank]))
        self.row_hi = self.row_lo + self.row_counts[comm.rank]
        self.local = local_factory([self.atoms[k] for k in self.mine])
        self._point = np.zeros(self.q)

    def _loc(self, v):
        return v[self.row_lo : self.row_hi]

    def _gather(self, local):
        return self.comm.allgather_rows(local, self.row_counts)

    def setup(self):
        self.local.setup()

    def load_point(self, primal, scal=1.0):
        self._point = primal * scal if scal != 1.0 else np.array(primal, dtype=np.float64)
        self.local.load_point(self._loc(primal), scal)

    def load_dual_point(self, dual):
        self.local.load_dual_point(self._loc(dual))

    def reset_data(self):
        self.local.reset_data()

    def point(self):
        return self._point

    def is_feas(self):
        ok = 1.0 if self.local.is_feas() else 0.0
        return bool(self.comm.allreduce_min(np.array([ok]))[0] > 0.5)

    def is_dual_feas(self):
        ok = 1.0 if self.local.is_dual_feas() else 0.0
        return bool(self.comm.allreduce_min(np.array([ok]))[0] > 0.5)

    def grad(self):
        # a rank whose shard is infeasible still has to take part in the collective
        g = self.local.grad() if self.local.q else np.zeros(0)
        return self._gather(g)

    def check_neighborhood(self, irtmu, use_max_prox, agg0, proxsqr_bound):
        # search.jl:141-158 with the aggregation split over ranks: max is order independent;
        # the sum variant adds the ranks' partial sums in rank order
        base = agg0 if use_max_prox else 0.0
        ok, agg = self.local.check_neighborhood(irtmu, use_max_prox, base, np.inf)
        all_ok = self.comm.allreduce_min(np.array([1.0 if ok else 0.0]))[0] > 0.5
        if use_max_prox:
            total = float(self.comm.allreduce_max(np.array([agg if np.isfinite(agg) else np.finfo(float).max]))[0])
            total = max(total, agg0)
        else:
            parts = self._gather_scalars(agg)
            total = agg0 + float(sum(parts))
        return bool(all_ok and total < proxsqr_bound), total

    def _gather_scalars(self, val):
        return list(self.comm.allgather_rows(np.array([val]), [1] * self.comm.world))

    def _map(self, name, arr):
        loc = np.asfortranarray(self._loc(arr)) if arr.ndim == 2 else self._loc(arr)
        return self._gather(getattr(self.local, name)(loc))

    def hess_prod(self, arr):
        return self._map("hess_prod", arr)

    def inv_hess_prod(self, arr):
        return self._map("inv_hess_prod", arr)

    def hess_prod_slow(self, arr):
        return self._map("hess_prod_slow", arr)

    def block_hess_prod(self, arr):
        return self._map("block_hess_prod", arr)

    def dder3(self, direction):
        return self._map("dder3", direction)

    def use_dder3(self):
        return np.ones(self.K, dtype=bool)


# ---------------------------------------------------------------------------------
class ShardedQRCholSystemSolver(QRCholSystemSolver):
    """QRChol system solver over a sharded model (qrchol.jl:111-265 split over ranks).

    ``backend`` is the rank-local engine with:
      load(G_rows), assemble() -> ms, reduce_to_root(comm), factor() -> (ms, kind, failed),
      solve(rhs) -> x  (root only), G_mul(x) -> local rows, Gt_mul(z_local) -> partial N-vector
    """

    device_G = True

    def __init__(self, backend, comm: Comm):
        self.backend = backend
        self.comm = comm
        self.time_assemble_ms = 0.0
        self.time_reduce_ms = 0.0
        self.time_fact_ms = 0.0
        self.num_updates = 0
        self.last_fact_kind = 0

    def cones_factory(self, specs):
        return ShardedConeSet(specs, self.comm, self.backend.local_cones_factory)

    def load(self, solver) -> None:
        model = solver.model
        assert model.p == 0, "sharded QRChol expects the reduced model (p == 0)"
        cones: ShardedConeSet = solver.cones
        self.q, self.N = model.q, model.n
        self.row_lo, self.row_hi = cones.row_lo, cones.row_hi
        self.row_counts = cones.row_counts
        G_rows = np.asfortranarray(model.G[self.row_lo : self.row_hi, :])
        self.backend.load(G_rows, cones.local)
        self.setup_point_sub(solver)

    def update_lhs(self, solver) -> None:
        import time

        model, cones = solver.model, solver.cones
        self.time_assemble_ms += self.backend.assemble()
        t0 = time.perf_counter()
        self.backend.reduce_to_root(self.comm)
        self.time_reduce_ms += 1e3 * (time.perf_counter() - t0)
        failed = 0.0
        if self.comm.rank == 0:
            ms, kind, bad = self.backend.factor()
            self.time_fact_ms += ms
            self.last_fact_kind = kind
            solver.time_upfact += ms * 1e-3
            failed = 1.0 if bad else 0.0
        failed = self.comm.bcast(np.array([failed]))[0]
        if failed and self.comm.rank == 0:
            print("positive definite linear system factorization failed")
        self.num_updates += 1
        self.rhs_const.z[:] = cones.block_hess_prod(model.h)
        self.solve_subsystem3(solver, self.sol_const, self.rhs_const)

    def setup_rhs3(self, solver, rhs, sol, rhs_sub) -> None:
        cones = solver.cones
        if cones.any_dual:
            dr = cones.dual_rows
            arr = np.where(dr, -rhs.z - rhs.s, rhs.z)
            prod = cones.block_hess_prod(arr)
            rhs_sub.z[:] = np.where(dr, prod, -prod - rhs.s)
        else:
            rhs_sub.z[:] = -cones.hess_prod(rhs.z) - rhs.s

    def solve_subsystem3(self, solver, sol, rhs) -> None:
        cones: ShardedConeSet = solver.cones
        # r = x + G'z : partial products summed over ranks
        r = rhs.x + self.Gt_mul(rhs.z)
        # x = fact \ r on the factorizing rank, then broadcast
        x = self.backend.solve(r) if self.comm.rank == 0 else np.zeros(self.N)
        x = self.comm.bcast(x, src=0) if self.N else x
        # z = H~(Gx) - z, shard-local, gathered
        Gx_loc = self.backend.G_mul(x)
        z_loc = cones.local.block_hess_prod(Gx_loc) - rhs.z[self.row_lo : self.row_hi]
        sol.x[:] = x
        sol.z[:] = self.comm.allgather_rows(z_loc, self.row_counts)

    def G_mul(self, x):
        return self.comm.allgather_rows(self.backend.G_mul(x), self.row_counts)

    def Gt_mul(self, z):
        part = self.backend.Gt_mul(z[self.row_lo : self.row_hi])
        return self.comm.allreduce_sum(part) if self.N else part


# ---------------------------------------------------------------------------------
class GpuLocalBackend:
    """The rank's shard on its B200, through the C-ABI."""

    def __init__(self, ctx, comm=None, use_peer=True):
        from .gpu import GpuConeBatch

        self.ctx = ctx
        self.lib = ctx.lib
        self._GpuConeBatch = GpuConeBatch
        self.S_tensor = None
        self.comm = comm
        self.want_peer = use_peer
        self.peer = False
        self.time_peer_sum_ms = 0.0

    def local_cones_factory(self, specs):
        return self._GpuConeBatch(specs, self.ctx)

    def load(self, G_rows: np.ndarray, local_cones) -> None:
        import torch

        from .lib import dptr, f64

        self.cones = local_cones
        self.q, self.N = G_rows.shape
        G = f64(G_rows)
        self.ctx.check(
            self.lib.hyb_model_load(self.ctx.ptr, self.q, self.N, dptr(G), local_cones.K, local_cones.descs), "hyb_model_load"
        )
        local_cones.mark_model_loaded()
        # torch view of the device-resident partial Schur matrix for the NCCL reduce
        ptr, rows, cols, ld = C.c_void_p(0), C.c_int64(0), C.c_int64(0), C.c_int64(0)
        self.ctx.check(self.lib.hyb_device_buffer(self.ctx.ptr, 0, C.byref(ptr), C.byref(rows), C.byref(cols), C.byref(ld)), "hyb_device_buffer")
        n_elem = int(ld.value) * max(int(cols.value), 1)

        class _Iface:
            __cuda_array_interface__ = {"shape": (n_elem,), "typestr": "<f8", "data": (int(ptr.value), False), "version": 2}

        self._iface = _Iface()
        self.S_tensor = torch.as_tensor(self._iface, device=torch.device("cuda", torch.cuda.current_device())) if self.N else None
        self._setup_peer()

    def _setup_peer(self) -> None:
        """Peer-memory reduce: every rank's syrk epilogue stores into its slot in the root's
        memory (CUDA IPC + NVLink P2P).  Needs one GPU per rank on one node and the NCCL
        backend's device topology; otherwise the NCCL / gloo reduce stays in use."""
        self.peer = False
        comm = self.comm
        if not (self.want_peer and comm is not None and comm.world > 1 and self.N and comm.dist.get_backend() == "nccl"):
            return
        import ctypes as C

        handle = C.create_string_buffer(64)
        ok = 1
        if comm.rank == 0:
            rc = self.lib.hyb_peer_root_create(self.ctx.ptr, comm.world, handle)
            ok = 1 if rc == 0 else 0
        payload = comm.bcast_object((ok, bytes(handle.raw)) if comm.rank == 0 else None)
        ok, raw = payload
        if ok and comm.rank != 0:
            rc = self.lib.hyb_peer_attach(self.ctx.ptr, raw, comm.rank, comm.world)
            ok = 1 if rc == 0 else 0
        all_ok = comm.allreduce_min(np.array([float(ok)]))[0] > 0.5
        self.peer = bool(all_ok)
        if not self.peer and comm.rank == 0:
            print("hypatia_b200: peer-memory Schur reduce unavailable (%s); using the NCCL reduce"
                  % self.lib.hyb_last_error(self.ctx.ptr).decode())

    def assemble(self) -> float:
        self.cones._sync()
        ta = C.c_double(0)
        if self.peer:
            self.ctx.check(self.lib.hyb_schur_assemble_peer(self.ctx.ptr, C.byref(ta)), "hyb_schur_assemble_peer")
        else:
            self.ctx.check(self.lib.hyb_schur_assemble(self.ctx.ptr, C.byref(ta)), "hyb_schur_assemble")
        return ta.value

    def reduce_to_root(self, comm: Comm) -> None:
        if self.peer:
            # every rank's tiles are already in the root's memory (the assemble call returned after
            # its kernels completed); one host barrier, then the root adds the slots in rank order
            comm.barrier()
            if comm.rank == 0:
                t = C.c_double(0)
                self.ctx.check(self.lib.hyb_peer_root_sum(self.ctx.ptr, C.byref(t)), "hyb_peer_root_sum")
                self.time_peer_sum_ms += t.value
            return
        if self.S_tensor is not None and comm.world > 1:
            comm.reduce_tensor_to_root(self.S_tensor)
            comm.torch.cuda.synchronize()

    def factor(self):
        tf, kind = C.c_double(0), C.c_int(0)
        rc = self.ctx.check(self.lib.hyb_schur_factor(self.ctx.ptr, C.byref(tf), C.byref(kind)), "hyb_schur_factor")
        return tf.value, kind.value, rc > 0

    def solve(self, rhs: np.ndarray) -> np.ndarray:
        from .lib import dptr, f64

        r = f64(rhs)
        x = np.empty(self.N)
        self.ctx.check(self.lib.hyb_schur_solve(self.ctx.ptr, dptr(r), dptr(x), 1), "hyb_schur_solve")
        return x

    def G_mul(self, x: np.ndarray) -> np.ndarray:
        from .lib import dptr, f64

        out = np.empty(self.q)
        xx = f64(x)
        self.ctx.check(self.lib.hyb_G_mul(self.ctx.ptr, b"N", dptr(xx), dptr(out)), "hyb_G_mul")
        return out

    def Gt_mul(self, z_local: np.ndarray) -> np.ndarray:
        from .lib import dptr, f64

        out = np.zeros(self.N)
        zz = f64(z_local)
        self.ctx.check(self.lib.hyb_G_mul(self.ctx.ptr, b"T", dptr(zz), dptr(out)), "hyb_G_mul")
        return out
