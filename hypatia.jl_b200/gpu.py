"""Product path: GPU-resident cone batch + B200QRCholSystemSolver, both thin
wrappers over the C-ABI (include/hypatia_b200.h).  Mirrors, on the host side,
the Julia types INTEGRATION.md describes:

  B200QRCholSystemSolver{T} <: Solvers.QRCholSystemSolver{T}   (qrchol.jl:21,111-265)
  B200ConeBatch{T}          <: Cones.Cone{T}                   (Cones.jl:34-304)

No CPU arithmetic for the hot path happens here and there is no fallback: every
oracle / assembly / factorization / solve is a C-ABI call into CUDA kernels.
"""
from __future__ import annotations

import ctypes as C
import os
import time
from typing import List

import numpy as np

from . import cones_spec as cs
from .host.coneset import ConeSet
from .host.syscommon import QRCholSystemSolver
from .lib import ConeDesc, Context, HybError, c_double_p, dptr, f64


class GpuConeBatch(ConeSet):
    """All cones of a model as one device-resident batch."""

    def __init__(self, specs: List[cs.ConeSpec], ctx: Context):
        super().__init__(specs)
        self.ctx = ctx
        self.lib = ctx.lib
        self._grad = None
        self._point = np.zeros(self.q)
        self._keep = []  # keeps Ps arrays alive while descs reference them
        self.descs = self._make_descs()
        self.model_loaded = False

    def _make_descs(self):
        arr = (ConeDesc * max(self.K, 1))()
        for k, s in enumerate(self.specs):
            d = arr[k]
            d.type = s.type_code
            d.use_dual_barrier = int(s.use_dual_barrier)
            d.offset = int(self.offs[k])
            d.dim = int(s.dim)
            d.side_or_R = int(getattr(s, "side", 0) or getattr(s, "R", 0) or 0)
            d.U = int(getattr(s, "U", 0) or (s.dim if s.type_code == cs.WSOSINTERPNONNEGATIVE else 0))
            Ps = getattr(s, "Ps", None)
            if Ps:
                mats = [f64(P) for P in Ps]
                ptrs = (c_double_p * len(mats))(*[dptr(P) for P in mats])
                Ls = (C.c_int64 * len(mats))(*[P.shape[1] for P in mats])
                self._keep.append((mats, ptrs, Ls))
                d.nPs = len(mats)
                d.Ps = C.cast(ptrs, C.POINTER(c_double_p))
                d.Ls = C.cast(Ls, C.POINTER(C.c_int64))
            else:
                d.nPs = 0
            if s.type_code == cs.POSSEMIDEFTRISPARSE:  # the pattern travels in Ls: rows then columns, 0-based
                pat = np.ascontiguousarray(np.concatenate([s.rows0, s.cols0]), dtype=np.int64)
                Ls = (C.c_int64 * pat.size)(*pat.tolist())
                self._keep.append(Ls)
                d.Ls = C.cast(Ls, C.POINTER(C.c_int64))
        return arr

    # The device model (G and the cones) is loaded by the system solver's load();
    # until then the batch loads a cones-only model (N = 0) so the initial-point
    # oracles of Solvers.jl:749-767 can run.
    def setup(self):
        if not self.model_loaded:
            dummy = np.zeros(1)
            self.ctx.check(
                self.lib.hyb_model_load(self.ctx.ptr, self.q, 0, dptr(dummy), self.K, self.descs), "hyb_model_load"
            )
            self.model_loaded = True
        self._pending_primal = None
        self._pending_dual = np.zeros(self.q)
        self._pending_scal = 1.0

    def mark_model_loaded(self):
        self.model_loaded = True
        self._reload()

    def _reload(self):
        if self._pending_primal is not None:
            self._push()

    def _push(self):
        p = f64(self._pending_primal)
        d = f64(self._pending_dual)
        self.ctx.check(self.lib.hyb_cones_load_point(self.ctx.ptr, dptr(p), dptr(d), float(self._pending_scal)), "hyb_cones_load_point")
        self._grad = None
        self._dirty = False

    def load_point(self, primal, scal=1.0):
        self._pending_primal = np.array(primal, dtype=np.float64)
        self._pending_scal = float(scal)
        self._point = self._pending_primal * scal if scal != 1.0 else self._pending_primal.copy()
        self._dirty = True

    def load_dual_point(self, dual):
        self._pending_dual = np.array(dual, dtype=np.float64)
        self._dirty = True

    def reset_data(self):
        self._dirty = True

    def note_device_resident(self, scal):
        """Device-resident stepper: the device holds the accepted candidate and the host keeps no copy."""
        self._pending_primal = None
        self._pending_scal = float(scal)
        self._point = None
        self._grad = None
        self._dirty = False

    def note_device_point(self, primal, dual, scal):
        """The device already holds this point (loaded by hyb_search_check from a candidate it formed
        itself): bring the host mirror up to date without pushing anything."""
        self._pending_primal = np.array(primal, dtype=np.float64)
        self._pending_dual = np.array(dual, dtype=np.float64)
        self._pending_scal = float(scal)
        self._point = self._pending_primal * scal if scal != 1.0 else self._pending_primal.copy()
        self._grad = None
        self._dirty = False

    def _sync(self):
        if getattr(self, "_dirty", False):
            self._push()

    def point(self):
        return self._point

    def is_feas(self):
        self._sync()
        a, b = C.c_int(0), C.c_int(0)
        self.ctx.check(self.lib.hyb_cones_is_feas(self.ctx.ptr, C.byref(a), C.byref(b)), "hyb_cones_is_feas")
        return bool(a.value)

    def is_dual_feas(self):
        self._sync()
        a, b = C.c_int(0), C.c_int(0)
        self.ctx.check(self.lib.hyb_cones_is_feas(self.ctx.ptr, C.byref(a), C.byref(b)), "hyb_cones_is_feas")
        return bool(b.value)

    def grad(self):
        self._sync()
        if self._grad is None:
            g = np.empty(self.q)
            self.ctx.check(self.lib.hyb_cones_grad(self.ctx.ptr, dptr(g)), "hyb_cones_grad")
            self._grad = g
        return self._grad

    def check_neighborhood(self, irtmu, use_max_prox, agg0, proxsqr_bound):
        self._sync()
        ok, agg = C.c_int(0), C.c_double(0.0)
        self.ctx.check(
            self.lib.hyb_cones_check(self.ctx.ptr, float(irtmu), int(use_max_prox), float(agg0), float(proxsqr_bound), C.byref(ok), C.byref(agg)),
            "hyb_cones_check",
        )
        return bool(ok.value), float(agg.value)

    def _prod(self, which, arr):
        self._sync()
        a = f64(arr)
        m = 1 if a.ndim == 1 else a.shape[1]
        out = np.empty_like(a, order="F")
        self.ctx.check(self.lib.hyb_cones_hess_prod(self.ctx.ptr, which, dptr(a), dptr(out), m), "hyb_cones_hess_prod")
        return out

    def hess_prod(self, arr):
        return self._prod(0, arr)

    def inv_hess_prod(self, arr):
        return self._prod(1, arr)

    def sqrt_hess_prod(self, arr):
        return self._prod(2, arr)

    def inv_sqrt_hess_prod(self, arr):
        return self._prod(3, arr)

    def block_hess_prod(self, arr):
        return self._prod(4, arr)

    def hess_prod_slow(self, arr):
        return self._prod(6, arr)

    def dder3(self, direction):
        self._sync()
        d = f64(direction)
        out = np.empty(self.q)
        self.ctx.check(self.lib.hyb_cones_dder3(self.ctx.ptr, dptr(d), dptr(out)), "hyb_cones_dder3")
        return out


class B200QRCholSystemSolver(QRCholSystemSolver):
    """Drop-in for QRCholDenseSystemSolver (qrchol.jl:111-265) on one B200."""

    device_G = True

    def __init__(self, ctx: Context, batched_dirs: bool = True):
        self.ctx = ctx
        self.lib = ctx.lib
        # device-resident get_directions in batched pairs (hyb_get_directions); False keeps the
        # reference's one-direction-at-a-time call pattern through the generic host code
        self.batched_dirs = batched_dirs
        # line-search candidates formed and gated on the device (hyb_search_*); needs the directions
        # the batched solves leave there.  HYB_HOST_SEARCH=1 keeps the host loop of search.jl
        import os

        self.device_search = bool(batched_dirs) and not os.environ.get("HYB_HOST_SEARCH")
        # the whole stepper inner loop on the device (hyb_step_*): the iterate never leaves the GPU between
        # initialize_point and postprocess.  HYB_HOST_STEPPER=1 keeps the host-vector protocol above.
        self.device_stepper = self.device_search and not os.environ.get("HYB_HOST_STEPPER")
        self._dev_const_stale = True
        self._host_const_stale = True
        self.time_assemble_ms = 0.0
        self.time_fact_ms = 0.0
        self.num_updates = 0
        self.last_fact_kind = 0

    def cones_factory(self, specs):
        return GpuConeBatch(specs, self.ctx)

    # ---- load (qrchol.jl:145-187): G*Q2 goes to the device once ------------------
    def load(self, solver) -> None:
        model = solver.model
        assert model.p == 0, "B200QRCholSystemSolver expects the reduced model (p == 0, Ap_Q == I)"
        cones = solver.cones
        assert isinstance(cones, GpuConeBatch), "B200QRCholSystemSolver needs the GPU cone batch"
        G = f64(model.G)
        self.q, self.N = model.q, model.n
        self.ctx.check(
            self.lib.hyb_model_load(self.ctx.ptr, model.q, model.n, dptr(G), cones.K, cones.descs), "hyb_model_load"
        )
        cones.mark_model_loaded()
        # several schedule entries per call with the decisions on the device (hyb_search_alpha); HYB_SPEC_SEARCH=0
        # keeps one candidate per call
        self.spec_search = (self.device_search and not os.environ.get("HYB_SPEC_SEARCH") == "0"
                            and bool(self.lib.hyb_search_alpha_supported(self.ctx.ptr)))
        self.setup_point_sub(solver)

    def set_point_sub_rhs(self, solver) -> None:  # common.jl:211-220 (also on modify)
        super().set_point_sub_rhs(solver)
        m = solver.model
        cv, hv = f64(m.c), f64(m.h)
        self.ctx.check(self.lib.hyb_dirs_setup(self.ctx.ptr, dptr(cv), dptr(hv)), "hyb_dirs_setup")
        nus = f64(solver.cones.nus)
        self.ctx.check(self.lib.hyb_search_setup(self.ctx.ptr, dptr(nus) if nus.size else None), "hyb_search_setup")
        self._dev_const_stale = self._host_const_stale = True

    # ---- update_lhs (qrchol.jl:189-207) ------------------------------------------
    def update_lhs(self, solver) -> None:
        cones = solver.cones
        cones._sync()
        ta, tf, kind = C.c_double(0), C.c_double(0), C.c_int(0)
        rc = self.ctx.check(
            self.lib.hyb_schur_update(self.ctx.ptr, C.byref(ta), C.byref(tf), C.byref(kind)), "hyb_schur_update"
        )
        self.time_assemble_ms += ta.value
        self.time_fact_ms += tf.value
        self._after_factor(solver, tf.value, kind.value, rc)

    def _after_factor(self, solver, tf_ms, kind, rc) -> None:
        self.num_updates += 1
        self.last_fact_kind = kind
        solver.time_upfact += tf_ms * 1e-3
        if rc > 0:
            print("positive definite linear system factorization failed")
        # the constant system (qrchol.jl:199-206) is solved with the first batch of directions
        # (hyb_get_directions, refresh_const) or, for the host call pattern, right here
        self._dev_const_stale = self._host_const_stale = True
        if not self.batched_dirs:
            self._ensure_host_const(solver)

    def _ensure_host_const(self, solver) -> None:
        if self._host_const_stale:
            self.rhs_const.z[:] = solver.cones.block_hess_prod(solver.model.h)
            self.solve_subsystem3(solver, self.sol_const, self.rhs_const)
            self._host_const_stale = False

    def solve_subsystem4(self, solver, sol, rhs) -> None:
        self._ensure_host_const(solver)
        super().solve_subsystem4(solver, sol, rhs)

    # ---- get_directions (common.jl:22-83), device-resident and batched --------------
    # ---- check_cone_points (search.jl:100-162) on the device ---------------------------
    def search_load_point(self, point) -> None:
        z, s_ = f64(point.z), f64(point.s)
        self.ctx.check(self.lib.hyb_search_load_point(self.ctx.ptr, dptr(z), dptr(s_)), "hyb_search_load_point")

    def search_gates(self, coefs):
        cf = (C.c_double * 4)(*coefs)
        mn, sm = C.c_double(0), C.c_double(0)
        self.ctx.check(self.lib.hyb_search_gates(self.ctx.ptr, cf, C.byref(mn), C.byref(sm)), "hyb_search_gates")
        return mn.value, sm.value

    def search_check(self, mu, min_prox, proxsqr_bound, use_max_prox, agg0):
        ok, agg = C.c_int(0), C.c_double(0)
        self.ctx.check(
            self.lib.hyb_search_check(self.ctx.ptr, float(mu), float(min_prox), float(proxsqr_bound), int(use_max_prox),
                                      float(agg0), C.byref(ok), C.byref(agg)),
            "hyb_search_check",
        )
        return bool(ok.value), float(agg.value)

    def search_alpha(self, coefs, taukaps, nup1, min_prox, proxsqr_bound, use_max_prox):
        """First accepted of several schedule entries, decided on the device (hyb_search_alpha)."""
        cf = f64(np.asarray(coefs, dtype=np.float64).reshape(-1))
        tk = f64(np.asarray(taukaps, dtype=np.float64))
        acc, mu, agg = C.c_int(-1), C.c_double(0), C.c_double(0)
        self.ctx.check(
            self.lib.hyb_search_alpha(self.ctx.ptr, int(tk.size), dptr(cf), dptr(tk), float(nup1), float(min_prox),
                                      float(proxsqr_bound), int(use_max_prox), C.byref(acc), C.byref(mu), C.byref(agg)),
            "hyb_search_alpha",
        )
        return int(acc.value), float(mu.value), float(agg.value)

    # ---- device-resident stepper (hyb_step_*): steppers/common.jl, combined.jl, Solvers.jl:631-693 ------
    def step_point_set(self, point) -> None:
        x, z, s_ = f64(point.x), f64(point.z), f64(point.s)
        self.ctx.check(self.lib.hyb_step_point_set(self.ctx.ptr, dptr(x), dptr(z), dptr(s_)), "hyb_step_point_set")

    def step_point_get(self, point) -> None:
        x, z, s_ = np.empty(self.N), np.empty(self.q), np.empty(self.q)
        self.ctx.check(self.lib.hyb_step_point_get(self.ctx.ptr, dptr(x), dptr(z), dptr(s_)), "hyb_step_point_get")
        point.x[:] = x
        point.z[:] = z
        point.s[:] = s_

    def step_convergence(self, tau: float):
        out = (C.c_double * 7)()
        self.ctx.check(self.lib.hyb_step_convergence(self.ctx.ptr, float(tau), out), "hyb_step_convergence")
        return [float(v) for v in out]

    def step_rhs(self, kind: int, col: int, solver, dir_tau: float = 0.0) -> None:
        pt = solver.point
        self.ctx.check(
            self.lib.hyb_step_rhs(self.ctx.ptr, int(kind), int(col), float(solver.mu), float(pt.tau), float(pt.kap),
                                  float(solver.tau_residual), float(dir_tau)),
            "hyb_step_rhs",
        )

    def step_directions(self, stepper, solver, m: int):
        """get_directions on the m right-hand sides built by step_rhs; returns [(tau, kap)] of the directions."""
        import math

        solver.cones._sync()
        nref = (C.c_int * m)()
        rn = (C.c_double * m)()
        tk = (C.c_double * (2 * m))()
        self.ctx.check(
            self.lib.hyb_step_directions(
                self.ctx.ptr, m, int(self._dev_const_stale), float(solver.point.tau), float(solver.mu),
                float(solver.res_norm_cutoff), int(solver.max_ref_steps), 0.5, nref, rn, tk,
            ),
            "hyb_step_directions",
        )
        self._dev_const_stale = False
        self.last_res_norms = [float(v) for v in rn]
        self.last_num_refs = [int(v) for v in nref]
        for j in range(m):
            stepper.last_num_ref += int(nref[j])
            assert not math.isnan(rn[j]), "NaN residual in direction solve"
            solver.worst_dir_res = max(solver.worst_dir_res, float(rn[j]))
        return [(float(tk[2 * j]), float(tk[2 * j + 1])) for j in range(m)]

    def dirs_keep(self, col: int, slot: int) -> None:
        self.ctx.check(self.lib.hyb_dirs_keep(self.ctx.ptr, int(col), int(slot)), "hyb_dirs_keep")

    def step_point_update(self, coefs) -> float:
        cf = (C.c_double * 4)(*coefs)
        zs = C.c_double(0)
        self.ctx.check(self.lib.hyb_step_point_update(self.ctx.ptr, cf, C.byref(zs)), "hyb_step_point_update")
        return zs.value

    def get_directions_batch(self, stepper, solver, rhs_points, dir_points, slots=None) -> None:
        import math

        solver.cones._sync()
        m = len(rhs_points)
        L = rhs_points[0].vec.size
        rhs = np.empty((L, m), order="F")
        for j, r in enumerate(rhs_points):
            rhs[:, j] = r.vec
        out = np.empty((L, m), order="F")
        nref = (C.c_int * m)()
        rn = (C.c_double * m)()
        sc = np.empty(self.N + self.q) if self._dev_const_stale else None
        self.ctx.check(
            self.lib.hyb_get_directions(
                self.ctx.ptr, m, dptr(rhs), dptr(out), L, int(self._dev_const_stale), float(solver.point.tau), float(solver.mu),
                float(solver.res_norm_cutoff), int(solver.max_ref_steps), 0.5, nref, rn, dptr(sc) if sc is not None else None,
            ),
            "hyb_get_directions",
        )
        if sc is not None:
            self.sol_const.x[:] = sc[: self.N]
            self.sol_const.z[:] = sc[self.N :]
            self._dev_const_stale = self._host_const_stale = False
        if slots is not None and self.device_search:
            for j, slot in enumerate(slots):
                self.ctx.check(self.lib.hyb_dirs_keep(self.ctx.ptr, j, int(slot)), "hyb_dirs_keep")
        self.last_res_norms = [float(v) for v in rn]
        self.last_num_refs = [int(v) for v in nref]
        for j, d in enumerate(dir_points):
            d.vec[:] = out[:, j]
            stepper.last_num_ref += int(nref[j])
            assert not math.isnan(rn[j]), "NaN residual in direction solve"
            solver.worst_dir_res = max(solver.worst_dir_res, float(rn[j]))

    # ---- setup_rhs3 (qrchol.jl:23-44) --------------------------------------------
    def setup_rhs3(self, solver, rhs, sol, rhs_sub) -> None:
        cones = solver.cones
        if cones.any_dual:
            dr = cones.dual_rows
            arr = np.where(dr, -rhs.z - rhs.s, rhs.z)
            prod = cones.block_hess_prod(arr)
            rhs_sub.z[:] = np.where(dr, prod, -prod - rhs.s)
        else:
            rhs_sub.z[:] = -cones.hess_prod(rhs.z) - rhs.s

    # ---- solve_subsystem3 (qrchol.jl:46-92), p == 0 ------------------------------
    def solve_subsystem3(self, solver, sol, rhs) -> None:
        solver.cones._sync()
        rx, rz = f64(rhs.x), f64(rhs.z)
        sx = np.empty(self.N)
        sz = np.empty(self.q)
        self.ctx.check(
            self.lib.hyb_solve_subsystem3(self.ctx.ptr, dptr(rx), dptr(rz), dptr(sx), dptr(sz)), "hyb_solve_subsystem3"
        )
        sol.x[:] = sx
        sol.z[:] = sz

    # ---- G products (K7 call sites) ------------------------------------------------
    def G_mul(self, x):
        out = np.empty(self.q)
        xx = f64(x)
        self.ctx.check(self.lib.hyb_G_mul(self.ctx.ptr, b"N", dptr(xx), dptr(out)), "hyb_G_mul")
        return out

    def Gt_mul(self, z):
        out = np.empty(self.N)
        zz = f64(z)
        self.ctx.check(self.lib.hyb_G_mul(self.ctx.ptr, b"T", dptr(zz), dptr(out)), "hyb_G_mul")
        return out

    def schur_matrix(self) -> np.ndarray:
        out = np.zeros((self.N, self.N), order="F")
        self.ctx.check(self.lib.hyb_schur_get(self.ctx.ptr, dptr(out)), "hyb_schur_get")
        return out

    def free_memory(self) -> None:  # Solvers.jl:390 — device memory is owned by the context
        pass


class B200SymIndefDenseSystemSolver(B200QRCholSystemSolver):
    """Drop-in for SymIndefDenseSystemSolver (symindef.jl:216-280) on one B200: cone oracles, the assembly of
    the (n + q) x (n + q) system, its rook-pivoted Bunch-Kaufman factorization and the solves run on the device
    (hyb_symindef_*); the stepper follows the reference's one-direction-at-a-time call pattern
    (systemsolvers/common.jl) on host vectors."""

    def __init__(self, ctx: Context):
        super().__init__(ctx, batched_dirs=False)
        self.device_search = self.device_stepper = False

    def load(self, solver) -> None:  # symindef.jl:231-247
        super().load(solver)
        self.ctx.check(self.lib.hyb_symindef_load(self.ctx.ptr), "hyb_symindef_load")

    def update_lhs(self, solver) -> None:  # symindef.jl:249-270
        solver.cones._sync()
        tf, kind = C.c_double(0), C.c_int(0)
        rc = self.ctx.check(self.lib.hyb_symindef_update(self.ctx.ptr, C.byref(tf), C.byref(kind)), "hyb_symindef_update")
        self.time_fact_ms += tf.value
        self.num_updates += 1
        self.last_fact_kind = kind.value
        solver.time_upfact += tf.value * 1e-3
        if rc > 0:
            print("symmetric linear system factorization failed")
        self.solve_subsystem3(solver, self.sol_const, self.rhs_const)

    def solve_subsystem4(self, solver, sol, rhs) -> None:  # common.jl:161-189, no QRChol constant-system laziness
        QRCholSystemSolver.solve_subsystem4(self, solver, sol, rhs)

    def setup_rhs3(self, solver, rhs, sol, rhs_sub) -> None:  # symindef.jl:38-62
        cones = solver.cones
        prod = cones.inv_hess_prod(rhs.s)
        if cones.any_dual:
            rhs_sub.z[:] = np.where(cones.dual_rows, -rhs.z - rhs.s, -prod - rhs.z)
        else:
            rhs_sub.z[:] = -prod - rhs.z

    def solve_subsystem3(self, solver, sol, rhs) -> None:  # symindef.jl:272-280
        solver.cones._sync()
        r = f64(rhs.vec)
        out = np.empty_like(r)
        self.ctx.check(self.lib.hyb_symindef_solve(self.ctx.ptr, dptr(r), dptr(out), 1), "hyb_symindef_solve")
        sol.vec[:] = out

    def system_matrix(self) -> np.ndarray:
        d = self.N + self.q
        out = np.zeros((d, d), order="F")
        self.ctx.check(self.lib.hyb_symindef_get(self.ctx.ptr, dptr(out)), "hyb_symindef_get")
        return out


class B200MultiQRCholSystemSolver(B200QRCholSystemSolver):
    """`B200QRCholSystemSolver{Float64}(devices = 0:7)` of INTEGRATION.md: one host process, several GPUs, through
    hyb_multi_* alone (no torch.distributed).  Only `load` and `update_lhs` differ from the single-GPU solver:
    the cones are partitioned by block over the devices for the Schur assembly; everything else runs in the
    root device's context."""

    def __init__(self, mctx, batched_dirs: bool = True):
        super().__init__(mctx.root, batched_dirs=batched_dirs)
        self.mctx = mctx
        self.time_reduce_ms = 0.0

    def load(self, solver) -> None:
        model = solver.model
        assert model.p == 0, "B200QRCholSystemSolver expects the reduced model (p == 0, Ap_Q == I)"
        cones = solver.cones
        G = f64(model.G)
        self.q, self.N = model.q, model.n
        self.ctx.check(
            self.lib.hyb_multi_model_load(self.mctx.ptr, model.q, model.n, dptr(G), cones.K, cones.descs), "hyb_multi_model_load"
        )
        cones.mark_model_loaded()
        self.row_ranges = self.mctx.shard_rows()
        self.setup_point_sub(solver)

    def update_lhs(self, solver) -> None:
        solver.cones._sync()
        ta, tr, tf, kind = C.c_double(0), C.c_double(0), C.c_double(0), C.c_int(0)
        rc = self.ctx.check(
            self.lib.hyb_multi_schur_update(self.mctx.ptr, C.byref(ta), C.byref(tr), C.byref(tf), C.byref(kind)),
            "hyb_multi_schur_update",
        )
        self.time_assemble_ms += ta.value
        self.time_reduce_ms += tr.value
        self.time_fact_ms += tf.value
        self.last_assemble_ms = ta.value
        self._after_factor(solver, tf.value, kind.value, rc)


def make_gpu_solver(device: int = 0, dist=None, use_peer: bool = True, mode: str = "rootworker", batched_dirs: bool = True,
                    devices=None, syssolver: str = "qrchol", **solver_kwargs):
    """Solver wired to the GPU backends (what `Hypatia.Optimizer(syssolver =
    B200QRCholSystemSolver{Float64}(), use_dense_model = true)` selects in Julia).

    With several ranks, `mode="rootworker"` (default) runs the algorithm on rank 0 and uses the
    other GPUs as Schur-assembly workers over peer memory (rootworker.py); `mode="replicated"`
    runs the same host driver on every rank with sharded cone oracles (sharded.py).  ``devices=[0, 1, ...]``
    instead drives several GPUs from THIS process through hyb_multi_* (no torch.distributed at all)."""
    from .host.solver import Solver

    if syssolver == "symindef":  # syssolver = SymIndefDenseSystemSolver{Float64}() in Julia; one GPU
        ctx = Context(device)
        ss = B200SymIndefDenseSystemSolver(ctx)
        solver = Solver(ss.cones_factory, ss, **solver_kwargs)
        attach_gpu_qr(solver, ctx)
        return solver
    assert syssolver == "qrchol", syssolver
    if dist is not None and dist.get_world_size() > 1 and mode == "rootworker":
        from .rootworker import make_rootworker_solver

        return make_rootworker_solver(device, dist, batched_dirs=batched_dirs, **solver_kwargs)
    if devices is not None and len(devices) > 1 or (devices is not None and mode == "single_process"):
        from .lib import MultiContext

        mctx = MultiContext(devices)
        ss = B200MultiQRCholSystemSolver(mctx, batched_dirs=batched_dirs)
        solver = Solver(ss.cones_factory, ss, **solver_kwargs)
        attach_gpu_qr(solver, mctx.root)
        return solver
    ctx = Context(device)
    if dist is not None and dist.get_world_size() > 1:
        import torch

        from .sharded import Comm, GpuLocalBackend, ShardedQRCholSystemSolver

        comm = Comm(dist, torch.device("cuda", device))
        ss = ShardedQRCholSystemSolver(GpuLocalBackend(ctx, comm, use_peer=use_peer), comm)
        ss.ctx = ctx
        solver = Solver(ss.cones_factory, ss, **solver_kwargs)
        solver.setup_comm = comm  # rank 0 does the preprocessing QR and broadcasts its outcome
        return solver
    ss = B200QRCholSystemSolver(ctx, batched_dirs=batched_dirs)
    solver = Solver(ss.cones_factory, ss, **solver_kwargs)
    attach_gpu_qr(solver, ctx)
    return solver


def attach_gpu_qr(solver, ctx: Context) -> None:
    """Preprocessing QRs (process.jl:109,233) of matrices with >= HYB_GPU_QR_MIN (default 256) rows and
    columns run on the GPU (hyb_geqp3); smaller ones stay with LAPACK, where they take microseconds."""
    import os

    solver.qr_backend = ctx.geqp3
    solver.ormqr_backend = ctx.ormqr_right  # G * Ap_Q of the reduce (process.jl:310-338), same size threshold ...
    solver.ormqr_min_size = int(os.environ.get("HYB_GPU_ORMQR_MIN_SIZE", str(1 << 22)))  # ... or G of 4M entries and more
    solver.qr_backend_min_cols = int(os.environ.get("HYB_GPU_QR_MIN", "256"))
