"""Host-side interior-point driver (stand-in for the Julia host).

In production the caller of the C-ABI is Hypatia's own unmodified Julia driver;
Julia is absent from this image, so this file restates that driver in Python so
the hot path can be exercised end to end.  It is generic over a ``ConeSet`` and
a ``SystemSolver`` backend and contains no cone or linear-system arithmetic of
its own beyond O(n+q) vector work.

Follows, function by function:
  src/Solvers/Solvers.jl:188-284 (options), :352-393 (solve), :395-452 (setup),
  :455-497 (initial point), :499-564, :566-623 (step_and_check), :625-693
  (mu, convergence params), :696-747 (check_converged), :749-767
  src/Solvers/process.jl:15-66 (rescale), :68-200 (dual eq), :202-396 (primal
  eq / reduce), :403-412 (rank), :415-505 (postprocess)
"""
from __future__ import annotations

import math
import time
from typing import Callable, Optional

import numpy as np
import scipy.linalg as sla
from scipy.linalg import lapack

from .models import Model, Point

EPS = float(np.finfo(np.float64).eps)

# Solvers.jl:40-60
STATUSES = (
    "NotLoaded Loaded Modified SolveCalled PrimalInconsistent DualInconsistent Optimal "
    "PrimalInfeasible DualInfeasible IllPosed NearOptimal NearPrimalInfeasible "
    "NearDualInfeasible NearIllPosed SlowProgress IterationLimit TimeLimit NumericalFailure "
    "UnknownStatus"
).split()
INFEAS_STATUSES = ("PrimalInfeasible", "DualInfeasible", "NearPrimalInfeasible", "NearDualInfeasible")
CHECK_NEAR_STATUSES = ("SlowProgress", "IterationLimit", "TimeLimit", "NumericalFailure")


def _norm_inf(v: np.ndarray) -> float:
    return float(np.max(np.abs(v))) if v.size else 0.0


class _PivotedQR:
    """Column-pivoted Householder QR (LAPACK dgeqp3), the ``qr!(M, ColumnNorm())``
    of process.jl:109,233.  Q is kept in factored form and applied with dormqr."""

    def __init__(self, M: np.ndarray, backend=None):
        """``backend(M) -> (qr, tau, piv)`` replaces LAPACK's dgeqp3 with the same output
        conventions (the GPU path passes ``Context.geqp3``); Q is still applied on the host."""
        M = np.array(M, dtype=np.float64, order="F")
        self.m, self.ncols = M.shape
        if M.size == 0:
            self.qr = M
            self.tau = np.zeros(0)
            self.piv = np.arange(self.ncols)
            return
        if backend is not None:
            try:
                qr, tau, piv = backend(M, overwrite_a=True)  # M is this object's private copy
            except TypeError:
                qr, tau, piv = backend(M)
        else:
            (qr, tau), _, piv = sla.qr(M, mode="raw", pivoting=True, overwrite_a=True)
        self.qr, self.tau, self.piv = np.asfortranarray(qr), tau, piv
        self.refl = np.asfortranarray(self.qr[:, : tau.size])  # Householder vectors for dormqr

    def rank_est(self, tol: float) -> int:
        """get_rank_est (process.jl:403-412): count |R_ii| > tol."""
        d = np.abs(np.diagonal(self.qr))
        return int(np.count_nonzero(d > tol))

    def R(self, rank: int) -> np.ndarray:
        return np.triu(self.qr[:rank, :rank])

    def apply_q(self, c: np.ndarray, trans: bool) -> np.ndarray:
        """c <- Q c or Q' c with the full m x m Q."""
        c2 = np.array(c, dtype=np.float64, order="F")
        one_d = c2.ndim == 1
        if one_d:
            c2 = c2.reshape(-1, 1, order="F")
        if self.tau.size == 0:
            return c2.ravel() if one_d else c2
        t = "T" if trans else "N"
        _, lw, info = lapack.dormqr("L", t, self.refl, self.tau, c2, -1)
        out, _, info = lapack.dormqr("L", t, self.refl, self.tau, c2, int(lw[0]), overwrite_c=1)
        assert info == 0
        return out.ravel() if one_d else out

    def right_apply_q(self, M: np.ndarray) -> np.ndarray:
        """M <- M Q.  ``ormqr_backend(refl, tau, M) -> M Q`` (set by the GPU path next to the QR backend) replaces
        LAPACK's dormqr for the O(q n k) product G * Ap_Q of the reduce (process.jl:310-338)."""
        M2 = np.array(M, dtype=np.float64, order="F")
        if self.tau.size == 0:
            return M2
        be = getattr(self, "ormqr_backend", None)
        if be is not None:
            try:
                return be(self.refl, self.tau, M2, overwrite=True)  # M2 is a private copy
            except TypeError:
                return be(self.refl, self.tau, M2)
        _, lw, info = lapack.dormqr("R", "N", self.refl, self.tau, M2, -1)
        out, _, info = lapack.dormqr("R", "N", self.refl, self.tau, M2, int(lw[0]), overwrite_c=1)
        assert info == 0
        return out


class Solver:
    """Solvers.jl:79-284.  ``cones_factory(specs) -> ConeSet`` chooses the oracle
    backend; ``syssolver`` the linear-system backend."""

    def __init__(
        self,
        cones_factory: Callable,
        syssolver,
        stepper=None,
        verbose: bool = False,
        iter_limit: int = 1000,
        time_limit: float = math.inf,
        tol_rel_opt: Optional[float] = None,
        tol_abs_opt: Optional[float] = None,
        tol_feas: Optional[float] = None,
        tol_infeas: Optional[float] = None,
        tol_tau_infeas: Optional[float] = None,
        tol_illposed: Optional[float] = None,
        default_tol_power: Optional[float] = None,
        default_tol_relax: Optional[float] = None,
        tol_slow: float = 1e-3,
        near_factor: float = 1000.0,
        preprocess: bool = True,
        reduce: bool = True,
        rescale: bool = True,
        init_tol_qr: float = 1000 * EPS,
        tol_inconsistent: Optional[float] = None,
        trace: bool = False,
    ):
        from .stepper import CombinedStepper

        assert preprocess and reduce, "QRChol system solvers need preprocess+reduce (Solvers.jl:212-217)"
        assert near_factor >= 1
        power = 0.5 if default_tol_power is None else float(default_tol_power)
        loose = EPS**power
        tight = EPS ** (1.5 * power)
        if default_tol_relax is not None:
            loose *= default_tol_relax
            tight *= default_tol_relax
        self.tol_rel_opt = loose if tol_rel_opt is None else tol_rel_opt
        self.tol_abs_opt = tight if tol_abs_opt is None else tol_abs_opt
        self.tol_feas = loose if tol_feas is None else tol_feas
        self.tol_infeas = tight if tol_infeas is None else tol_infeas
        self.tol_tau_infeas = 1e-2 if tol_tau_infeas is None else tol_tau_infeas
        self.tol_illposed = tight if tol_illposed is None else tol_illposed
        self.tol_slow = tol_slow
        self.near_factor = near_factor
        self.verbose = verbose
        self.iter_limit = iter_limit
        self.time_limit = time_limit
        self.preprocess, self.reduce, self.rescale = preprocess, reduce, rescale
        self.init_tol_qr = init_tol_qr
        self.tol_inconsistent = init_tol_qr if tol_inconsistent is None else tol_inconsistent
        self.cones_factory = cones_factory
        self.syssolver = syssolver
        self.stepper = stepper if stepper is not None else CombinedStepper()
        self.status = "NotLoaded"
        self.trace_enabled = trace
        self.trace = []
        self.iter_callback = None  # called once per completed iteration (bench hooks)

    # ------------------------------------------------------------------ load / modify
    def load(self, model: Model) -> None:
        self.orig_model = model
        self.status = "Loaded"

    def modify_obj_offset(self, offset_new: float) -> None:
        self.orig_model.obj_offset = float(offset_new)
        self.status = "Modified"

    def modify_c(self, c_new, idxs=None) -> None:
        if idxs is None:
            assert len(c_new) == self.orig_model.n
            self.orig_model.c = np.array(c_new, dtype=np.float64)
        else:
            self.orig_model.c[np.asarray(idxs)] = c_new
        self.status = "Modified"

    def modify_b(self, b_new, idxs=None) -> None:
        if idxs is None:
            assert len(b_new) == self.orig_model.p
            self.orig_model.b = np.array(b_new, dtype=np.float64)
        else:
            self.orig_model.b[np.asarray(idxs)] = b_new
        self.status = "Modified"

    def modify_h(self, h_new, idxs=None) -> None:
        if idxs is None:
            assert len(h_new) == self.orig_model.q
            self.orig_model.h = np.array(h_new, dtype=np.float64)
        else:
            self.orig_model.h[np.asarray(idxs)] = h_new
        self.status = "Modified"

    # ------------------------------------------------------------------ G products (K7 sites)
    def G_mul(self, x: np.ndarray) -> np.ndarray:
        ss = self.syssolver
        if getattr(ss, "device_G", False):
            return ss.G_mul(x)
        return self.model.G @ x

    def Gt_mul(self, z: np.ndarray) -> np.ndarray:
        ss = self.syssolver
        if getattr(ss, "device_G", False):
            return ss.Gt_mul(z)
        return self.model.G.T @ z

    # ------------------------------------------------------------------ solve
    def solve(self) -> None:
        init_status = self.status
        if init_status not in ("Loaded", "Modified"):
            raise RuntimeError(f"solve called when solver status is {init_status}")
        self._setup_solver()
        if getattr(self.syssolver, "is_worker", False):
            # multi-GPU root/worker form (rootworker.py): this rank only assembles partial Schur
            # matrices for the rank that runs the algorithm
            self.syssolver.serve(self)
            self.status = "Worker"
            self.solve_time = time.time() - self.solve_time
            return
        if init_status == "Modified":
            self._setup_modified()
        else:
            self._setup_loaded()

        if self.status not in ("PrimalInconsistent", "DualInconsistent"):
            self._initialize_point()
            self._setup_stepping()
            while True:
                if self._step_and_check():
                    break
            if self.status in CHECK_NEAR_STATUSES:
                self.exit_reason = self.status  # which exit led to the near-convergence recheck (Solvers.jl:375-377)
                self._check_converged(True)
            self._postprocess()
        elif hasattr(self.syssolver, "abort_workers"):
            self.syssolver.abort_workers()
        self.solve_time = time.time() - self.solve_time
        if self.verbose:
            print(f"\nstatus is {self.status} after {self.num_iters} iterations and {self.solve_time:.3f} seconds\n")
        self.syssolver.free_memory()

    def _setup_solver(self) -> None:  # Solvers.jl:529-564
        self.status = "SolveCalled"
        self.num_iters = 0
        self.solve_time = time.time()
        for name in ("rescale", "initx", "inity", "loadsys", "upsys", "upfact", "uprhs", "getdir", "search"):
            setattr(self, "time_" + name, 0.0)
        self.res_norm_cutoff = 0.0
        self.max_ref_steps = 5
        nan = math.nan
        self.x_norm_res_t = self.y_norm_res_t = self.z_norm_res_t = nan
        self.x_norm_res = self.y_norm_res = self.z_norm_res = nan
        self.primal_obj_t = self.dual_obj_t = self.primal_obj = self.dual_obj = nan
        self.gap = self.x_feas = self.y_feas = self.z_feas = self.tau_feas = nan
        self.trace = []
        self.exit_reason = None

    def _setup_loaded(self) -> None:  # Solvers.jl:395-429
        orig = self.orig_model
        self.result = Point(orig.n, orig.p, orig.q)
        model = self.model = Model(orig.c, orig.A, orig.b, orig.G, orig.h, orig.cones, orig.obj_offset)
        t0 = time.time()
        self.used_rescaling = self._rescale_data()
        self.time_rescale = time.time() - t0

        self.cones = self.cones_factory(model.cones)
        self._setup_point()
        if self.status != "SolveCalled":
            return
        self.stepper.load(self)
        t0 = time.time()
        self.syssolver.load(self)
        self.time_loadsys = time.time() - t0
        self.x_residual = np.zeros(model.n)
        self.y_residual = np.zeros(model.p)
        self.z_residual = np.zeros(model.q)
        self.tau_residual = 0.0

    def _setup_modified(self) -> None:  # Solvers.jl:431-452
        orig, model = self.orig_model, self.model
        model.obj_offset = orig.obj_offset
        model.c = orig.c.copy()
        model.b = orig.b.copy()
        model.h = orig.h.copy()
        if self.used_rescaling:
            model.c = model.c / self.c_scale
            model.b = model.b / self.b_scale
            model.h = model.h / self.h_scale
        self._update_initial_point()
        self.syssolver.set_point_sub_rhs(self)

    # ------------------------------------------------------------------ preprocessing
    def _qr_backend_for(self, shape):
        """The pivoted-QR implementation for a matrix of this shape: ``self.qr_backend`` (set by
        the GPU path) when the matrix has at least ``qr_backend_min_cols`` columns, else LAPACK."""
        be = getattr(self, "qr_backend", None)
        if be is not None and min(shape) >= getattr(self, "qr_backend_min_cols", 0):
            return be
        return None

    def _rescale_data(self) -> bool:  # process.jl:15-66
        if not self.rescale:
            return False
        model = self.model
        c, A, b, G, h = model.c, model.A, model.b, model.G, model.h
        minval = math.sqrt(EPS)

        # one |.| pass per matrix; row and column maxima from the same temporary (G can be gigabytes)
        def absmax(M):
            if M.size == 0:
                return np.full(M.shape[1], minval), np.full(M.shape[0], minval)
            aM = np.abs(M)
            return np.maximum(aM.max(axis=0), minval), np.maximum(aM.max(axis=1), minval)

        Acol, Arow = absmax(A)
        Gcol, Grow = absmax(G)
        self.c_scale = c_scale = np.sqrt(np.maximum(np.abs(c), np.maximum(Acol, Gcol)))
        self.b_scale = b_scale = np.sqrt(np.maximum(np.abs(b), Arow))
        h_scale = self.h_scale = np.ones(model.q)
        offs = np.asarray(model.cone_offs, dtype=np.int64)
        if len(model.cones):
            # per-row cones (Nonnegative) scale row by row, every other cone by the maximum over its rows
            # (process.jl:36-52); segment maxima instead of a Python loop over 10^5 cones
            habs = np.abs(h)
            per_row = np.repeat(np.array([bool(cn.per_row_h_scale) for cn in model.cones]), np.diff(offs))
            hmax = np.maximum(np.maximum.reduceat(habs, offs[:-1]), minval)
            gmax = np.maximum.reduceat(Grow, offs[:-1])
            cone_scale = np.sqrt(np.maximum(hmax, gmax))
            h_scale[:] = np.where(per_row, np.sqrt(np.maximum(habs, Grow)), np.repeat(cone_scale, np.diff(offs)))
        model.c = c / c_scale
        # (M ./ c_scale') ./ row_scale with one result array (same two roundings per entry as the two-temporary form)
        def scaled(M, row_scale):
            out = np.empty(M.shape, order="F")
            np.divide(M, c_scale[None, :], out=out)
            np.divide(out, row_scale[:, None], out=out)
            return out

        model.A = scaled(A, b_scale)
        model.G = scaled(G, h_scale)
        model.b = b / b_scale
        model.h = h / h_scale
        return True

    def _initialize_cone_point(self):  # Solvers.jl:749-767
        cones = self.cones
        cones.setup()
        primal = cones.initial_point()
        cones.load_point(primal)
        assert cones.is_feas(), "initial point is not feasible"
        dual = -cones.grad()
        cones.load_dual_point(dual)
        assert cones.is_dual_feas(), "initial dual point is not feasible"
        init_z = np.where(cones.dual_rows, primal, dual)
        init_s = np.where(cones.dual_rows, dual, primal)
        return init_z, init_s

    def _setup_point(self) -> None:  # Solvers.jl:455-482 (reduce branch)
        init_z, init_s = self._initialize_cone_point()
        t0 = time.time()
        self._handle_primal_eq()
        self.time_inity = time.time() - t0
        if self.status == "PrimalInconsistent":
            return
        init_y = self._update_primal_eq(init_z)
        t0 = time.time()
        init_x = self._dual_eq_and_init_x(init_s)
        self.time_initx = time.time() - t0
        if self.status == "DualInconsistent":
            return
        model = self.model
        point = self.point = Point(model.n, model.p, model.q)
        point.x[:] = init_x
        point.y[:] = init_y
        point.z[:] = init_z
        point.s[:] = init_s

    def _dual_eq_and_init_x(self, init_s: np.ndarray) -> np.ndarray:
        """handle_dual_eq + update_dual_eq (process.jl:68-200).  With several ranks running the
        same (replicated, deterministic) host driver, only rank 0 does the pivoted QR of [A; G] and
        broadcasts its outcome (status, kept columns, initial x): identical results everywhere
        without every rank repeating the same LAPACK factorization on a share of the cores."""
        comm = getattr(self, "setup_comm", None)
        if comm is None or comm.world == 1:
            self._handle_dual_eq()
            if self.status == "DualInconsistent":
                return np.zeros(0)
            return self._update_dual_eq(init_s)
        model = self.model
        payload = None
        if comm.rank == 0:
            self._handle_dual_eq()
            if self.status == "DualInconsistent":
                payload = ("DualInconsistent", None, None)
            else:
                init_x = self._update_dual_eq(init_s)
                payload = (self.status, np.asarray(self.x_keep_idxs), init_x)
        status, keep, init_x = comm.bcast_object(payload)
        if comm.rank != 0:
            self.AG_fact = None
            if status == "DualInconsistent":
                self.status = status
                return np.zeros(0)
            self.x_keep_idxs = keep
            if keep.size != model.n:
                model.A = np.asfortranarray(model.A[:, keep])
                model.G = np.asfortranarray(model.G[:, keep])
                model.n = keep.size
            model.c = model.c[keep]
        return init_x

    def _update_initial_point(self) -> None:  # Solvers.jl:484-497
        init_z, init_s = self._initialize_cone_point()
        point = self.point
        point.y[:] = self._update_primal_eq(init_z)
        point.x[:] = self._update_dual_eq(init_s)
        point.z[:] = init_z
        point.s[:] = init_s

    def _handle_primal_eq(self) -> None:  # process.jl:202-339
        model = self.model
        p = model.p
        self.reduced = False
        if p == 0:
            self.y_keep_idxs = np.zeros(0, dtype=np.int64)
            return
        fact = self.Ap_fact = _PivotedQR(model.A.T, self._qr_backend_for(model.A.T.shape))
        # G * Ap_Q on the same backend as the QR, and in any case when G is large (the product is O(q n rank): one
        # pass over a multi-gigabyte G per reflector on the host)
        if self._qr_backend_for(model.A.T.shape) is not None or model.G.size >= getattr(self, "ormqr_min_size", 1 << 62):
            fact.ormqr_backend = getattr(self, "ormqr_backend", None)
        rank = self.Ap_rank = fact.rank_est(self.init_tol_qr)
        Ap_R = fact.R(rank)
        y_keep = fact.piv[:rank].copy()
        if rank < p:
            x_sub = np.zeros(model.n)
            x_sub[:rank] = sla.solve_triangular(Ap_R, model.b[y_keep], trans="T", lower=False)
            x_sub = fact.apply_q(x_sub, trans=False)
            residual = _norm_inf(model.A @ x_sub - model.b)
            if residual > self.tol_inconsistent:
                if self.verbose:
                    print(f"some primal equality constraints are inconsistent (residual {residual})")
                self.status = "PrimalInconsistent"
                return
        # reduce (process.jl:301-338)
        model.n -= rank
        GQ = fact.right_apply_q(model.G)  # [GQ1 GQ2] = G0 * Q
        self.reduce_GQ1 = np.asfortranarray(GQ[:, :rank])
        model.G = np.asfortranarray(GQ[:, rank:])
        model.p = 0
        model.A = np.zeros((0, model.n), order="F")
        self.reduce_Ap_R = Ap_R
        self.reduce_y_keep_idxs = y_keep
        self.y_keep_idxs = np.arange(p)
        self.reduced = True

    def _update_primal_eq(self, init_z: np.ndarray) -> np.ndarray:  # process.jl:343-396
        if self.orig_model.p == 0:
            return np.zeros(0)
        model = self.model
        rank = self.Ap_rank
        cQ = self.Ap_fact.apply_q(model.c, trans=True)  # c' Q
        self.reduce_cQ1 = cQ[:rank].copy()
        model.c = cQ[rank:].copy()
        Rpib0 = sla.solve_triangular(self.reduce_Ap_R, model.b[self.reduce_y_keep_idxs], trans="T", lower=False)
        self.reduce_Rpib0 = Rpib0
        model.obj_offset += float(self.reduce_cQ1 @ Rpib0)
        model.b = np.zeros(0)
        model.h = model.h - self.reduce_GQ1 @ Rpib0
        return np.zeros(0)

    def _handle_dual_eq(self) -> None:  # process.jl:68-165
        model = self.model
        n = model.n
        self.x_keep_idxs = np.arange(n)
        self.AG_fact = None
        if n == 0:
            return
        AG = model.G if model.p == 0 else np.vstack([model.A, model.G])
        preset = getattr(self, "preset_AG_fact", None)  # bench only: reuse an identical factorization
        if preset is not None and (preset.m, preset.ncols) == AG.shape:
            fact = self.AG_fact = preset
        else:
            fact = self.AG_fact = _PivotedQR(AG, self._qr_backend_for(AG.shape))
        rank = self.AG_rank = fact.rank_est(self.init_tol_qr)
        if rank == n:
            return
        AG_R = self.AG_R = fact.R(rank)
        x_keep = fact.piv[:rank].copy()
        p = model.p
        yz_sub = np.zeros(p + model.q)
        yz_sub[:rank] = sla.solve_triangular(AG_R, model.c[x_keep], trans="T", lower=False)
        yz_sub = fact.apply_q(yz_sub, trans=False)
        residual = model.G.T @ yz_sub[p:] - model.c
        if p:
            residual += model.A.T @ yz_sub[:p]
        res_norm = _norm_inf(residual)
        if res_norm > self.tol_inconsistent:
            if self.verbose:
                print(f"some dual equality constraints are inconsistent (residual norm {res_norm})")
            self.status = "DualInconsistent"
            return
        model.A = np.asfortranarray(model.A[:, x_keep])
        model.G = np.asfortranarray(model.G[:, x_keep])
        model.n = rank
        self.x_keep_idxs = x_keep

    def _update_dual_eq(self, init_s: np.ndarray) -> np.ndarray:  # process.jl:169-200
        model = self.model
        model.c = model.c[self.x_keep_idxs]
        if model.n == 0:
            return np.zeros(0)
        rhs_x = np.concatenate([model.b, model.h - init_s])
        fact = self.AG_fact
        rhs_x = fact.apply_q(rhs_x, trans=True)
        if self.AG_rank == fact.ncols:
            # AG_fact \ rhs_x for a full-column-rank pivoted QR
            sol = sla.solve_triangular(fact.R(self.AG_rank), rhs_x[: self.AG_rank], lower=False)
            init_x = np.empty(fact.ncols)
            init_x[fact.piv] = sol
            return init_x
        return sla.solve_triangular(self.AG_R, rhs_x[: model.n], lower=False)

    # ------------------------------------------------------------------ iteration
    def _initialize_point(self) -> None:  # Solvers.jl:499-516
        point = self.point
        point.tau = 1.0
        point.kap = 1.0
        self.calc_mu()
        if math.isnan(self.mu) or abs(1.0 - self.mu) > math.sqrt(EPS):
            print(f"warning: initial mu is {self.mu} but should be 1")
        cones = self.cones
        cones.load_point(cones.primal_of(point.z, point.s))
        cones.load_dual_point(cones.dual_of(point.z, point.s))
        # device-resident stepper (gpu.py, hyb_step_*): from here to postprocess the iterate's x, z, s live on
        # the device; the host keeps tau, kap, mu
        self.device_stepper = bool(getattr(self.syssolver, "device_stepper", False)) and cones.K > 0
        if self.device_stepper:
            cones._sync()
            self.syssolver.step_point_set(point)
            self._zs_dev = float(point.z @ point.s)

    def _setup_stepping(self) -> None:  # Solvers.jl:518-527
        model = self.model
        self.x_conv_tol = 1.0 / (1.0 + _norm_inf(model.c))
        self.y_conv_tol = 1.0 / (1.0 + _norm_inf(model.b))
        self.z_conv_tol = 1.0 / (1.0 + _norm_inf(model.h))
        self.prev_is_slow = False
        self.prev2_is_slow = False
        self.worst_dir_res = 0.0

    def calc_mu(self) -> float:  # Solvers.jl:625-629
        point = self.point
        zs = self._zs_dev if getattr(self, "device_stepper", False) else float(point.z @ point.s)
        self.mu = (zs + point.tau * point.kap) / (self.model.nu + 1.0)
        return self.mu

    def _step_and_check(self) -> bool:  # Solvers.jl:566-623
        improv = self._calc_convergence_params()
        if self.trace_enabled:
            self.trace.append(self.trace_row())
        if self.verbose:
            self._print_iteration()
        if self._check_converged():
            return True
        if self.num_iters == self.iter_limit:
            self.status = "IterationLimit"
            return True
        if time.time() - self.solve_time >= self.time_limit:
            self.status = "TimeLimit"
            return True
        if self.stepper.expect_improvement():
            if improv < self.tol_slow:
                if self.prev_is_slow and self.prev2_is_slow:
                    self.exit_reason = "SlowProgress"
                    self.status = "SlowProgress"
                    return True
                self.prev2_is_slow = self.prev_is_slow
                self.prev_is_slow = True
            else:
                self.prev2_is_slow = self.prev_is_slow
                self.prev_is_slow = False
        self.res_norm_cutoff = 1e-4 * max(self.x_norm_res, self.y_norm_res, self.z_norm_res, self.tau_feas)
        self.worst_dir_res = 0.0
        if not self.stepper.step(self):
            return True
        self.calc_mu()
        if min(self.point.tau, self.point.kap, self.mu) <= 0:
            print("warning: tau, kappa, or mu is nonpositive; terminating")
            self.status = "NumericalFailure"
            return True
        self.num_iters += 1
        if self.iter_callback is not None:
            self.iter_callback(self)
        return False

    def _calc_convergence_params(self) -> float:  # Solvers.jl:631-693
        model, point = self.model, self.point
        tau = point.tau
        if getattr(self, "device_stepper", False):
            return self._calc_convergence_params_device()
        xr = self.Gt_mul(point.z)
        if model.p:
            xr = xr + model.A.T @ point.y
        self.x_norm_res_t = _norm_inf(xr)
        xr = xr + model.c * tau
        self.x_norm_res = _norm_inf(xr) / tau
        self.x_residual = -xr
        x_feas = self.x_norm_res * self.x_conv_tol

        yr = model.A @ point.x if model.p else np.zeros(0)
        self.y_norm_res_t = _norm_inf(yr)
        yr = yr - model.b * tau
        self.y_norm_res = _norm_inf(yr) / tau
        self.y_residual = yr
        y_feas = self.y_norm_res * self.y_conv_tol

        zr = self.G_mul(point.x) + point.s
        self.z_norm_res_t = _norm_inf(zr)
        zr = zr - model.h * tau
        self.z_norm_res = _norm_inf(zr) / tau
        self.z_residual = zr
        z_feas = self.z_norm_res * self.z_conv_tol

        self.primal_obj_t = float(model.c @ point.x)
        self.dual_obj_t = -float(model.b @ point.y) - float(model.h @ point.z)
        self.tau_residual = self.primal_obj_t - self.dual_obj_t + point.kap
        tau_feas = abs(self.tau_residual)

        improv = 0.0
        for curr, prev in ((x_feas, self.x_feas), (y_feas, self.y_feas), (z_feas, self.z_feas), (tau_feas, self.tau_feas)):
            if math.isnan(prev) or math.isnan(curr):
                continue
            improv = max(improv, (prev - curr) / (abs(prev) + EPS))
        self.x_feas, self.y_feas, self.z_feas, self.tau_feas = x_feas, y_feas, z_feas, tau_feas
        self.primal_obj = self.primal_obj_t / tau + model.obj_offset
        self.dual_obj = self.dual_obj_t / tau + model.obj_offset
        self.gap = float(point.z @ point.s)
        return improv

    def _calc_convergence_params_device(self) -> float:
        """Solvers.jl:631-693 with the O(qN) and O(q) work on the device (hyb_step_convergence); the residual
        vectors stay there as the prediction right-hand side.  Reduced model: p = 0."""
        model, point = self.model, self.point
        tau = point.tau
        o = self.syssolver.step_convergence(tau)
        self.x_norm_res_t = o[0]
        self.x_norm_res = o[1] / tau
        x_feas = self.x_norm_res * self.x_conv_tol
        self.y_norm_res_t = self.y_norm_res = 0.0
        y_feas = 0.0
        self.z_norm_res_t = o[2]
        self.z_norm_res = o[3] / tau
        z_feas = self.z_norm_res * self.z_conv_tol
        self.primal_obj_t = o[4]
        self.dual_obj_t = -o[5]
        self.tau_residual = self.primal_obj_t - self.dual_obj_t + point.kap
        tau_feas = abs(self.tau_residual)
        improv = 0.0
        for curr, prev in ((x_feas, self.x_feas), (y_feas, self.y_feas), (z_feas, self.z_feas), (tau_feas, self.tau_feas)):
            if math.isnan(prev) or math.isnan(curr):
                continue
            improv = max(improv, (prev - curr) / (abs(prev) + EPS))
        self.x_feas, self.y_feas, self.z_feas, self.tau_feas = x_feas, y_feas, z_feas, tau_feas
        self.primal_obj = self.primal_obj_t / tau + model.obj_offset
        self.dual_obj = self.dual_obj_t / tau + model.obj_offset
        self.gap = o[6]
        self._zs_dev = o[6]
        return improv

    def _check_converged(self, check_near: bool = False) -> bool:  # Solvers.jl:696-747
        nf = self.near_factor if check_near else 1.0
        tau = self.point.tau
        pt, dt = self.primal_obj_t, self.dual_obj_t
        worst_feas = max(self.x_feas, self.y_feas, self.z_feas)
        if worst_feas <= nf * self.tol_feas:
            is_abs_opt = self.gap <= nf * self.tol_abs_opt
            worst_gap = min(self.gap / tau, abs(pt - dt))
            max_tau_obj = max(tau, min(abs(pt), abs(dt)))
            is_rel_opt = worst_gap <= nf * self.tol_rel_opt * max_tau_obj
            if is_abs_opt or is_rel_opt:
                self.status = "NearOptimal" if check_near else "Optimal"
                return True
        if tau <= nf * self.tol_tau_infeas:
            tol = nf * self.tol_infeas
            if dt > EPS and self.x_norm_res_t <= tol * dt:
                self.status = "NearPrimalInfeasible" if check_near else "PrimalInfeasible"
                self.primal_obj, self.dual_obj = pt, dt
                return True
            yz_res = max(self.y_norm_res_t, self.z_norm_res_t)
            if pt < -EPS and yz_res <= tol * -pt:
                self.status = "NearDualInfeasible" if check_near else "DualInfeasible"
                self.primal_obj, self.dual_obj = pt, dt
                return True
        if max(tau, self.point.kap) <= nf * self.tol_illposed:
            self.status = "NearIllPosed" if check_near else "IllPosed"
            return True
        return False

    # ------------------------------------------------------------------ postprocess
    def _postprocess(self) -> None:  # process.jl:415-505
        point, result = self.point, self.result
        if getattr(self, "device_stepper", False):
            self.syssolver.step_point_get(point)  # the iterate comes back from the device once
            self.device_stepper = False
        tau = point.tau
        if tau <= 0 or np.any(np.isnan(point.vec)):
            result.vec[:] = math.nan
            return
        result.s[:] = point.s
        result.z[:] = point.z
        x = point.x.copy()
        y = point.y.copy()
        if self.status not in INFEAS_STATUSES:
            result.s[:] /= tau
            result.z[:] /= tau
            x /= tau
            y /= tau
        orig = self.orig_model
        if orig.n:
            if orig.p:
                self._unreduce_x(x)
            else:
                result.x[self.x_keep_idxs] = x
        if orig.p:
            self._unreduce_y()
        if self.used_rescaling:
            result.s[:] *= self.h_scale
            result.z[:] /= self.h_scale
            result.y[:] /= self.b_scale
            result.x[:] /= self.c_scale

    def _unreduce_x(self, x: np.ndarray) -> None:  # process.jl:471-489
        Rpib0 = np.zeros_like(self.reduce_Rpib0) if self.status in INFEAS_STATUSES else self.reduce_Rpib0
        xa = np.zeros(self.orig_model.n - Rpib0.size)
        xa[self.x_keep_idxs] = x
        xb = np.concatenate([Rpib0, xa])
        self.result.x[:] = self.Ap_fact.apply_q(xb, trans=False)

    def _unreduce_y(self) -> None:  # process.jl:491-505
        ya = self.reduce_GQ1.T @ self.result.z
        if self.status not in INFEAS_STATUSES:
            ya = ya + self.reduce_cQ1
        keep = self.reduce_y_keep_idxs
        if keep.size == 0:
            return
        ya_sub = sla.solve_triangular(self.reduce_Ap_R, ya[: keep.size], lower=False)
        self.result.y[keep] = -ya_sub

    # ------------------------------------------------------------------ getters / logging
    def get_x(self):
        return self.result.x.copy()

    def get_y(self):
        return self.result.y.copy()

    def get_z(self):
        return self.result.z.copy()

    def get_s(self):
        return self.result.s.copy()

    def trace_row(self) -> dict:
        """Per-iteration tuple the reference prints (Solvers.jl:819-841)."""
        st = self.stepper
        return dict(
            iter=self.num_iters,
            primal_obj=self.primal_obj,
            dual_obj=self.dual_obj,
            gap=self.gap,
            x_feas=self.x_feas,
            y_feas=self.y_feas,
            z_feas=self.z_feas,
            x_norm_res=self.x_norm_res,
            y_norm_res=self.y_norm_res,
            z_norm_res=self.z_norm_res,
            tau=self.point.tau,
            kap=self.point.kap,
            mu=self.mu,
            worst_dir_res=self.worst_dir_res,
            prox=getattr(getattr(st, "searcher", None), "prox", 0.0),
            alpha=getattr(st, "prev_alpha", 1.0),
            n_cand=getattr(st, "last_num_cand", 0),
            n_ref=getattr(st, "last_num_ref", 0),
        )

    def _print_iteration(self) -> None:
        r = self.trace_row()
        if self.num_iters == 0:
            print(f"{'iter':>5} {'p_obj':>12} {'d_obj':>12} |{'abs_gap':>9} {'x_feas':>9} {'z_feas':>9} |{'tau':>9} {'kap':>9} {'mu':>9} |{'dir_res':>8} {'prox':>8} {'alpha':>9}")
        print(
            f"{r['iter']:5d} {r['primal_obj']:12.4e} {r['dual_obj']:12.4e} |{r['gap']:9.2e} {r['x_feas']:9.2e} {r['z_feas']:9.2e} "
            f"|{r['tau']:9.2e} {r['kap']:9.2e} {r['mu']:9.2e} |{r['worst_dir_res']:8.1e} {r['prox']:8.1e} {r['alpha']:9.2e}"
        )
