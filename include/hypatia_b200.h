/* hypatia_b200 — C ABI of the B200-native interior-point hot path.
 *
 * Plain C: pointers and sizes only.  Every entry point is what a Hypatia.jl
 * `ccall` would bind for this path; the reference interface each one replaces is
 * cited as file:line into jump-dev/Hypatia.jl v0.9.2.  All host arrays are
 * column-major Float64 and owned by the caller; the library copies what it needs
 * to the device and keeps no host pointer after a call returns.  All calls are
 * synchronous at return.  One context drives one GPU (one process per GPU).
 *
 * Return convention: 0 = ok; > 0 = numerical condition reported through the out
 * parameters (mirrors the reference, where numerical failure is not an
 * exception, qrchol.jl:260-262); < 0 = CUDA / usage error, message available from
 * hyb_last_error().
 */
#ifndef HYPATIA_B200_H
#define HYPATIA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct hyb_ctx hyb_ctx;

/* cone type codes (src/Cones/*.jl) */
enum hyb_cone_type {
    HYB_CONE_NONNEGATIVE = 0,             /* nonnegative.jl */
    HYB_CONE_EPINORMEUCL = 1,             /* epinormeucl.jl */
    HYB_CONE_EPIPERSQUARE = 2,            /* epipersquare.jl */
    HYB_CONE_POSSEMIDEFTRI = 3,           /* possemideftri.jl (real svec) */
    HYB_CONE_WSOSINTERPNONNEGATIVE = 4,   /* wsosinterpnonnegative.jl */
    HYB_CONE_WSOSINTERPPOSSEMIDEFTRI = 5, /* wsosinterppossemideftri.jl */
    HYB_CONE_EPINORMINF = 6,              /* epinorminf.jl (real; use_dual_barrier: epigraph of the 1-norm) */
    HYB_CONE_DOUBLYNONNEGATIVETRI = 7,    /* doublynonnegativetri.jl (real svec; side_or_R = side) */
    HYB_CONE_POSSEMIDEFTRISPARSE = 8,     /* possemideftrisparse/denseimpl.jl (real): side_or_R = side, dim = pattern size,
                                             Ls = dim 0-based rows then dim 0-based columns of the pattern (row >= col) */
    HYB_CONE_WSOSINTERPEPINORMEUCL = 9,   /* wsosinterpepinormeucl.jl: side_or_R = R, U, Ps / Ls as for the other WSOS cones */
    HYB_CONE_WSOSINTERPEPINORMONE = 10    /* wsosinterpepinormone.jl: the same fields */
};

/* One entry per cone of model.cones, in order; offsets follow build_cone_idxs
 * (src/Models/Models.jl:63-73), 0-based. */
typedef struct hyb_cone_desc {
    int32_t type;             /* hyb_cone_type */
    int32_t use_dual_barrier; /* Cones.use_dual_barrier(cone) */
    int64_t offset;           /* first row of the cone in s / z */
    int64_t dim;              /* Cones.dimension(cone) */
    int64_t side_or_R;        /* PSD side, or R for WSOS-PSD; 0 otherwise */
    int64_t U;                /* WSOS: number of interpolation points; 0 otherwise */
    int32_t nPs;              /* WSOS: number of P matrices */
    int32_t _pad;
    const double* const* Ps;  /* WSOS: nPs column-major U x L[k] matrices */
    const int64_t* Ls;        /* WSOS: nPs column counts; PosSemidefTriSparse: the pattern (2 dim entries) */
} hyb_cone_desc;

/* ---- context --------------------------------------------------------------- */
/* Creates a context on CUDA device `device`.  NULL on failure (see hyb_last_error(NULL)). */
hyb_ctx* hyb_ctx_create(int device);
/* Solvers.free_memory(syssolver), Solvers.jl:390,796 */
void hyb_ctx_destroy(hyb_ctx* ctx);
const char* hyb_last_error(hyb_ctx* ctx);
/* number of kernels this library has launched so far in this process */
int64_t hyb_launch_count(void);

/* ---- load: Solvers.load(syssolver, solver), qrchol.jl:145-187 ----------------
 * GQ2 = (G*Ap_Q)[:, p+1:n] is the q x N matrix of the reduced model (ld = q). */
int hyb_model_load(hyb_ctx* ctx, int64_t q, int64_t N, const double* GQ2, int64_t ncones, const hyb_cone_desc* descs);

/* ---- cone batch: Cones oracle interface (src/Cones/Cones.jl:34-304) ----------- */
/* load_point(cone, primal, scal) + load_dual_point + reset_data for every cone
 * (Cones.jl:164-190; search.jl:144-146).  primal/dual are full q-vectors in
 * primal-view / dual-view order (point.jl:52-59). */
int hyb_cones_load_point(hyb_ctx* ctx, const double* primal, const double* dual, double scal);
/* is_feas / is_dual_feas over all cones (Cones.jl:63,70) */
int hyb_cones_is_feas(hyb_ctx* ctx, int* is_feas, int* is_dual_feas);
/* cone loop of check_cone_points (search.jl:141-158): feasibility, dual feasibility,
 * check_numerics (Cones.jl:267-284) and get_proxsqr (Cones.jl:288-304) aggregated with
 * max (use_max_prox) or sum starting from agg0.  *ok = 1 iff every cone passes and the
 * aggregate stays below proxsqr_bound. */
int hyb_cones_check(hyb_ctx* ctx, double irtmu, int use_max_prox, double agg0, double proxsqr_bound, int* ok,
                    double* agg);
/* grad(cone) for every cone (Cones.jl:78) */
int hyb_cones_grad(hyb_ctx* ctx, double* grad);
/* hess_prod! / inv_hess_prod! / hess_prod_slow! / block_hess_prod! on a q x m array
 * (Cones.jl:108-125,229-231; qrchol.jl:94-105).  which: 0 hess, 1 inv_hess,
 * 2 sqrt_hess, 3 inv_sqrt_hess, 4 block (inv_hess for dual-barrier cones else hess),
 * 6 hess_prod_slow. */
int hyb_cones_hess_prod(hyb_ctx* ctx, int which, const double* arr, double* prod, int64_t m);
/* dder3(cone, dir) for every cone (Cones.jl:139) */
int hyb_cones_dder3(hyb_ctx* ctx, const double* dir, double* out);

/* ---- Schur complement: update_lhs_fact, qrchol.jl:209-265 -----------------------
 * Assembles lhs = sum_k (W_k GQ2_k)'(W_k GQ2_k) with the currently loaded cone state
 * and factorizes it (posdef_fact_copy!, dense.jl:228-249).  fact_kind: 0 Cholesky,
 * 2 Cholesky after the diagonal shift of increase_diag! (dense.jl:100-107), -1 failed.
 * Times are CUDA-event milliseconds (time_upsys / time_upfact, Solvers.jl:112-113). */
int hyb_schur_update(hyb_ctx* ctx, double* t_assemble_ms, double* t_fact_ms, int* fact_kind);
/* The two halves of hyb_schur_update, for the multi-GPU path: every rank assembles
 * the partial sum over ITS cones (qrchol.jl:228-254 restricted to a cone range), the partials are
 * reduced onto the factorizing rank (NCCL reduce on the buffer hyb_device_buffer(0) exposes), and
 * only that rank factorizes (qrchol.jl:257-258). */
int hyb_schur_assemble(hyb_ctx* ctx, double* t_assemble_ms);
int hyb_schur_factor(hyb_ctx* ctx, double* t_fact_ms, int* fact_kind);
/* device address of an internal matrix for collectives: which = 0 -> assembled Schur matrix
 * (rows x cols, leading dimension ld, column-major FP64, upper triangle valid); which = 1 -> base of
 * the peer slots created by hyb_peer_root_create. */
int hyb_device_buffer(hyb_ctx* ctx, int which, void** ptr, int64_t* rows, int64_t* cols, int64_t* ld);
/* Fused assembly + NVLink transfer for the multi-GPU reduce (no counterpart in the reference,
 * which is single-process): the factorizing rank owns one N x N slot per rank
 * (hyb_peer_root_create exports a 64-byte CUDA IPC handle); the other ranks map it
 * (hyb_peer_attach) and hyb_schur_assemble_peer makes the syrk epilogue store each finished
 * tile directly into the rank's slot in the ROOT's memory (peer stores over NVLink, overlapped
 * with the remaining tiles' math).  After a host barrier the root adds the slots in rank order
 * (hyb_peer_root_sum) - bitwise deterministic, unlike atomics. */
int hyb_peer_root_create(hyb_ctx* ctx, int world, unsigned char* handle64_out);
int hyb_peer_attach(hyb_ctx* ctx, const unsigned char* handle64, int rank, int world);
int hyb_peer_set_slot(hyb_ctx* ctx, void* slot_device_ptr); /* same-process attach (root's own shard context) */
int hyb_schur_assemble_peer(hyb_ctx* ctx, double* t_assemble_ms);
int hyb_peer_root_sum(hyb_ctx* ctx, double* t_ms);
/* ldiv!(x, fact, rhs), qrchol.jl:75; rhs and x are N x nrhs */
int hyb_schur_solve(hyb_ctx* ctx, const double* rhs, double* x, int64_t nrhs);
/* copy of the assembled N x N matrix (upper triangle valid), for tests / fallbacks */
int hyb_schur_get(hyb_ctx* ctx, double* lhs);
/* mul!(out, G, v) ('N', out has q entries) or mul!(out, G', v) ('T', N entries):
 * qrchol.jl:59,80; systemsolvers/common.jl:95,98,151; Solvers.jl:638,656 */
int hyb_G_mul(hyb_ctx* ctx, char trans, const double* v, double* out);
/* solve_subsystem3 for the reduced model (p = 0), qrchol.jl:46-92:
 *   r = rhs_x + G' rhs_z ;  sol_x = fact \ r ;  sol_z = Htilde (G sol_x) - rhs_z */
int hyb_solve_subsystem3(hyb_ctx* ctx, const double* rhs_x, const double* rhs_z, double* sol_x, double* sol_z);

/* ---- device-resident direction solves: src/Solvers/systemsolvers/common.jl -------- */
/* set_point_sub_rhs (common.jl:211-220): the model vectors c (N) and h (q) the direction solves
 * need (rhs_const = (-c, h), dot_obj, apply_lhs).  Call after hyb_model_load and again whenever
 * c or h change (modify, Solvers.jl:431-452). */
int hyb_dirs_setup(hyb_ctx* ctx, const double* c, const double* h);
/* get_directions (common.jl:22-83) for `nrhs` right-hand sides at once: solve_system
 * (common.jl:133-158: setup_rhs3 qrchol.jl:23-44, solve_subsystem3 qrchol.jl:46-92,
 * solve_subsystem4 common.jl:161-189), the residual apply_lhs(dir) - rhs (common.jl:86-125) and
 * the refinement rounds (accepted while the inf-norm residual decreases; stops below
 * res_norm_cutoff, after two slow rounds (> min_impr_tol x previous) or max_ref_steps), all on the
 * device.  rhs / dir: nrhs columns of [x(N); z(q); tau; s(q); kap] (point.jl:31-62, p = 0) with
 * leading dimension ldv.  The columns must be independent of each other (the stepper pairs
 * (cent, pred) and (centadj, predadj), combined.jl:74-95).  refresh_const != 0 also solves the
 * constant system of update_lhs (qrchol.jl:199-206: rhs_const.z = Htilde h, sol_const =
 * solve_subsystem3(rhs_const)) in the same batch; it must be set on the first call after every
 * hyb_schur_factor.  nrhs + (refresh_const ? 1 : 0) <= 4.  tau_bar = point.tau, mu = solver.mu.
 * Out: num_ref[j] refinement solves used, res_norm[j] final residual inf-norm (NaN if the
 * residual is NaN; the caller asserts like common.jl:81), sol_const [x(N); z(q)] if not NULL. */
int hyb_get_directions(hyb_ctx* ctx, int nrhs, const double* rhs, double* dir, int64_t ldv, int refresh_const,
                       double tau_bar, double mu, double res_norm_cutoff, int max_ref_steps, double min_impr_tol,
                       int* num_ref, double* res_norm, double* sol_const);

/* ---- device-resident line search: src/Solvers/search.jl, steppers/combined.jl ------------ */
/* Barrier parameters nu_k of the K cones (Cones.get_nu, used by the gates of search.jl:124-130).
 * Call once after hyb_dirs_setup. */
int hyb_search_setup(hyb_ctx* ctx, const double* nus);
/* Keep column `col` of the last hyb_get_directions result on the device as direction `slot`
 * (0 cent, 1 centadj, 2 pred, 3 predadj: the four directions of combined.jl:74-95). */
int hyb_dirs_keep(hyb_ctx* ctx, int col, int slot);
/* point.z, point.s of the current iterate (q each). */
int hyb_search_load_point(hyb_ctx* ctx, const double* z, const double* s);
/* First half of check_cone_points (search.jl:100-122) for the candidate
 *   cand = point + coef[0] cent + coef[1] centadj + coef[2] pred + coef[3] predadj
 * (update_stepper_points, combined.jl:129-177; tau and kap stay with the caller): forms the
 * candidate's primal / dual cone slices on the device and returns min_k <s_k, z_k> and sum_k <s_k, z_k>. */
int hyb_search_gates(hyb_ctx* ctx, const double* coef, double* min_szk, double* sum_szk);
/* Second half (search.jl:124-158) for the same candidate, given its mu: the per-cone gates
 * sz_rel >= min_prox and nu (sz_rel - 1)^2 <= proxsqr_bound, then load_point(1/sqrt(mu)) /
 * load_dual_point / reset_data and the batched oracle checks of hyb_cones_check with the
 * aggregate started at agg0.  *ok = 1 iff the candidate is accepted; the cone batch is then
 * loaded at the candidate like after hyb_cones_load_point. */
int hyb_search_check(hyb_ctx* ctx, double mu, double min_prox, double proxsqr_bound, int use_max_prox, double agg0,
                     int* ok, double* agg);

/* Several entries of the step-length schedule in one call (search_alpha, search.jl:72-98): candidates
 * i = 0 .. ncand-1 (coefs[4 i .. 4 i + 3] as for hyb_search_gates, taukap[i] = tau_i * kap_i of the candidate) are
 * enqueued back to back -- candidate point, <s_k, z_k>, the scalar and per-cone gates, the oracle update, the
 * neighbourhood check -- with every decision taken on the device: once a candidate is accepted the kernels
 * of the later ones return at once, so the outcome (*accepted = index of the first accepted candidate or -1, its
 * mu and proximity aggregate, the state of the cone batch) is that of evaluating them one after the other with
 * hyb_search_gates / hyb_search_check, at one host synchronisation instead of up to three per candidate.
 * For models all of whose cones are served by the closed-form kernel families
 * (hyb_search_alpha_supported() == 1: no PosSemidefTri of side > 32, no WSOS cone). */
int hyb_search_alpha_supported(hyb_ctx* ctx);
int hyb_search_alpha(hyb_ctx* ctx, int ncand, const double* coefs, const double* taukap, double nup1, double min_prox,
                     double proxsqr_bound, int use_max_prox, int* accepted, double* mu, double* agg);

/* ---- device-resident stepper: src/Solvers/Solvers.jl, steppers/common.jl, steppers/combined.jl ----
 * The iterate's x (N), z (q), s (q) stay on the device for the whole solve; tau, kap, mu stay with the caller.
 * With these and the line search above an iteration moves only scalars across PCIe.  Needs hyb_dirs_setup. */
/* the iterate after initialize_point (Solvers.jl:499-516) / for postprocess (process.jl:415-469; NULL = skip) */
int hyb_step_point_set(hyb_ctx* ctx, const double* x, const double* z, const double* s);
int hyb_step_point_get(hyb_ctx* ctx, double* x, double* z, double* s);
/* calc_convergence_params (Solvers.jl:631-693, reduced model p = 0) at the device iterate and the caller's tau.
 * Keeps x_residual = -(G'z + c tau) and z_residual = G x + s - h tau on the device (update_rhs_pred uses them).
 * out7 = { |G'z|_inf, |G'z + c tau|_inf, |G x + s|_inf, |G x + s - h tau|_inf, c'x, h'z, z's }. */
int hyb_step_convergence(hyb_ctx* ctx, double tau, double* out7);
/* Right-hand side builders of steppers/common.jl into column `col` (0..3) of the direction engine:
 * kind 0 update_rhs_cent (:66-83), 1 update_rhs_centadj (:86-119), 2 update_rhs_pred (:14-28),
 * 3 update_rhs_predadj (:30-63).  mu, tau, kap: solver.mu, point.tau, point.kap; tau_residual (pred only):
 * solver.tau_residual; dir_tau (adjustments only): the tau entry of the kept unadjusted direction (slot 0 for
 * centadj, slot 2 for predadj, see hyb_dirs_keep), whose z / s parts are read on the device.  The adjustments
 * apply the reference's per-cone identity gate |<dder3, s> - <d, H d>| / (sqrt(eps) + |.|) < 1e-4. */
int hyb_step_rhs(hyb_ctx* ctx, int kind, int col, double mu, double tau, double kap, double tau_residual, double dir_tau);
/* hyb_get_directions on the right-hand sides built by hyb_step_rhs in columns 0..nrhs-1; the directions stay on
 * the device (keep them with hyb_dirs_keep); dir_taukap receives their (tau, kap) entries, 2 per column. */
int hyb_step_directions(hyb_ctx* ctx, int nrhs, int refresh_const, double tau_bar, double mu, double res_norm_cutoff,
                        int max_ref_steps, double min_impr_tol, int* num_ref, double* res_norm, double* dir_taukap);
/* update_stepper_points(alpha, point, false) (combined.jl:129-177) on x, z, s:
 * point += coef[0] cent + coef[1] centadj + coef[2] pred + coef[3] predadj (the kept directions);
 * *zs = <z, s> of the new iterate (calc_mu, Solvers.jl:625-629). */
int hyb_step_point_update(hyb_ctx* ctx, const double* coef, double* zs);

/* ---- preprocessing: src/Solvers/process.jl ------------------------------------------ */
/* qr!(M, ColumnNorm()) = LAPACK.geqp3!, process.jl:109 ([A; G], handle_dual_eq) and :233 (A',
 * handle_primal_eq): column-pivoted Householder QR of the m x n host matrix A (lda), in place, with
 * LAPACK's output conventions -- R on and above the diagonal, Householder vectors below it (unit
 * first entry implicit), tau (min(m,n)), jpvt (n, 0-based: column j of A*P is column jpvt[j] of A)
 * -- so dormqr / triangular solves on the host consume it unchanged.  Needs no loaded model. */
int hyb_geqp3(hyb_ctx* ctx, int64_t m, int64_t n, double* A, int64_t lda, int* jpvt, double* tau);
/* C (m x n, host, in / out) <- C Q with the Q of a Householder QR in factored form: LAPACK dormqr('R', 'N', ...).
 * V (n x k, leading dimension ldv) holds the reflectors below its diagonal as dgeqrf / dgeqp3 / hyb_geqp3 leave them,
 * tau their k scalars.  Replaces `rmul!(G, Ap_fact.Q)` of the reduce, src/Solvers/process.jl:310-338 (the
 * reference's own comment there: "very inefficient method used for sparse Ap_fact"). */
int hyb_ormqr_right(hyb_ctx* ctx, int64_t m, int64_t n, int64_t k, const double* V, int64_t ldv, const double* tau, double* C,
                    int64_t ldc);

/* ---- dense helpers: src/linearalgebra/dense.jl ---------------------------------- */
/* outer_prod!(A, B, alpha, beta) = BLAS.syrk!('U','T'), dense.jl:72-79:
 * C(N x N, upper) = alpha A'A + beta C with A K x N (lda), C (ldc) */
int hyb_outer_prod(hyb_ctx* ctx, int64_t K, int64_t N, double alpha, const double* A, int64_t lda, double beta,
                   double* C, int64_t ldc);
/* mul!(C, A', B, alpha, beta), qrchol.jl:253: C(M x N) = alpha A'B + beta C */
int hyb_gemm_tn_host(hyb_ctx* ctx, int64_t M, int64_t N, int64_t K, double alpha, const double* A, int64_t lda,
                     const double* B, int64_t ldb, double beta, double* C, int64_t ldc);
/* posdef_fact_copy!(mat2, mat, false) + ldiv!: factor the N x N matrix A (upper triangle
 * referenced, overwritten by U) and solve for nrhs right-hand sides in B (N x nrhs, ldb).
 * *info = LAPACK dpotrf info (0 success, j = leading minor j not positive definite). */
int hyb_posdef_solve(hyb_ctx* ctx, int64_t N, double* A, int64_t lda, double* B, int64_t ldb, int64_t nrhs, int* info);
/* symm_fact!(A) = bunchkaufman!(A, true) + ldiv!, dense.jl:199-203 (LAPACK dsytrf_rook / dsytrs_rook): factor
 * the symmetric N x N matrix A (upper triangle referenced) as U D U' with rook pivoting, in LAPACK's storage
 * (A overwritten by the multipliers and D, ipiv[N] 1-based with negative entries for 2 x 2 pivots), and solve
 * for nrhs <= 4 right-hand sides in B.  *info = LAPACK info (0, or the 1-based column of an exactly zero pivot). */
int hyb_symm_solve(hyb_ctx* ctx, int64_t N, double* A, int64_t lda, int* ipiv, double* B, int64_t ldb, int64_t nrhs,
                   int* info);
/* Replace the assembled Schur matrix by a host matrix (N x N, upper triangle referenced) before
 * hyb_schur_factor: the `lhs` field of QRCholDenseSystemSolver (qrchol.jl:111-143) is plain data a caller may
 * overwrite, and the tests use it to drive posdef_fact_copy!'s fallback chain with an indefinite matrix. */
int hyb_schur_set(hyb_ctx* ctx, const double* lhs);

/* ---- dense symmetric-indefinite system solver ------------------------------------------
 * Drop-in for SymIndefDenseSystemSolver (src/Solvers/systemsolvers/symindef.jl:216-280) on the model loaded by
 * hyb_model_load (reduced: p = 0).  The (n + q) x (n + q) system  [0 G'; G -H~]  (symindef.jl:8-37; H~_k =
 * inv_hess for primal-barrier cones, hess for dual-barrier ones, :255-259) lives on the device.
 *   hyb_symindef_load    load (:231-247): storage for the system and its factorization.
 *   hyb_symindef_update  update_lhs (:249-265) for the point last loaded into the cones: assemble, then
 *                        symm_fact_copy! (dense.jl:205-219): rook-pivoted Bunch-Kaufman, on failure the diagonal
 *                        shift and a second attempt.  *fact_kind = 1, 2 (shifted) or -1 (failed; returns 1).
 *   hyb_symindef_solve   solve_subsystem3 (:272-280): sol = fact \ rhs for nrhs <= 4 host vectors of length n + q
 *                        laid out [x; z] (column-major, leading dimension n + q).
 *   hyb_symindef_get     the assembled system (upper triangle valid, (n + q)^2 doubles) for the parity tests. */
int hyb_symindef_load(hyb_ctx* ctx);
int hyb_symindef_update(hyb_ctx* ctx, double* t_ms, int* fact_kind);
int hyb_symindef_solve(hyb_ctx* ctx, const double* rhs, double* sol, int64_t nrhs);
int hyb_symindef_get(hyb_ctx* ctx, double* lhs);

/* ---- several GPUs from one host process ----------------------------------------------
 * a Julia host uses a whole node with ccalls alone.  The cones are partitioned by block over the
 * devices; each device assembles the partial Schur matrix of its cones (update_lhs_fact, qrchol.jl:209-254, is
 * a sum over cones) and its syrk epilogue stores every tile into the memory of the tile's OWNER device over NVLink
 * (peer access; tile column j belongs to device j mod ndev); every owner adds the ndev partial copies of its columns
 * in device order and stores the sums into the FIRST device's Schur matrix, which factorizes (posdef_fact_copy!,
 * qrchol.jl:257-258) and serves everything else.  devices[0] is the factorizing GPU; every pair of devices needs a
 * peer-access path.  hyb_multi_root() is its ordinary context: load
 * points, run oracles, directions and the line search through it exactly as on one GPU, and call
 * hyb_multi_schur_update where a single GPU calls hyb_schur_update. */
typedef struct hyb_multi hyb_multi;
hyb_multi* hyb_multi_create(int ndev, const int* devices);
void hyb_multi_destroy(hyb_multi* m);
hyb_ctx* hyb_multi_root(hyb_multi* m);
/* Solvers.load(syssolver, solver), qrchol.jl:145-187: the full model on the root plus every device's row range
 * (Nonnegative cones are cut at the range boundaries so that LPs shard too) */
int hyb_multi_model_load(hyb_multi* m, int64_t q, int64_t N, const double* GQ2, int64_t ncones, const hyb_cone_desc* descs);
/* the partition hyb_multi_model_load would make, computed on the host alone (no GPU is touched): row ranges and
 * the number of (possibly split) cones per device; arrays of ndev entries, NULL = not wanted */
int hyb_multi_plan(int64_t q, int64_t ncones, const hyb_cone_desc* descs, int ndev, int64_t* row_lo, int64_t* row_hi,
                   int64_t* cones_per_dev);
/* rows [row_lo[d], row_hi[d]) of s / z owned by device d (arrays of ndev entries) */
int hyb_multi_shard_rows(hyb_multi* m, int64_t* row_lo, int64_t* row_hi);
/* update_lhs_fact (qrchol.jl:209-265) for the point loaded in the root context: t_assemble_ms = slowest device's
 * assembly, t_reduce_ms = the fixed-order sum of the partials, then as hyb_schur_factor */
int hyb_multi_schur_update(hyb_multi* m, double* t_assemble_ms, double* t_reduce_ms, double* t_fact_ms, int* fact_kind);
/* A context that will only ever assemble partial Schur matrices into another context's memory
 * (hyb_schur_assemble_peer) can skip its own Schur matrix and Cholesky workspace: set before hyb_model_load. */
int hyb_ctx_set_assemble_only(hyb_ctx* ctx, int flag);

/* ---- allocation-free product (no counterpart in the reference) -----------------------
 * y = G x on DEVICE pointers owned by the caller (column-major G, q x N, leading dimension
 * ld >= q; all on the current device), launched on `stream` (a cudaStream_t, NULL for the
 * default stream) and not synchronised.  For hosts that must not see a cudaMalloc from this
 * library: another runtime's pool owns the GPU memory, or the call sits in a latency-critical
 * section.  Its split-sum scratch is a fixed 4 MiB device array inside the library, so
 * q is limited to 65536 rows (larger q returns -4), every GPU that loads the library keeps
 * that array until exit, and calls that overlap on one GPU race on it.  The FIRST call on a
 * device is a warm-up: it makes the CUDA runtime load this library's code on that device (which
 * is when the 4 MiB array itself comes into being -- an implicit device allocation that may
 * synchronise) and looks up the array's address.  Make that call at initialisation, outside any
 * latency-critical section and before any stream capture; every later call allocates nothing. */
int hyb_gemv_n_noalloc(const double* G_dev, int64_t q, int64_t N, int64_t ld, const double* x_dev, double* y_dev,
                       void* stream);

/* ---- measurement helpers (device-resident synthetic operands, CUDA-event timing on the
 * library's own stream).  They time the same kernels the entry points above launch:
 * the dsyrk / dgemm replacement (dense.jl:72-79, qrchol.jl:242,253) and the dpotrf /
 * dpotrs replacement (dense.jl:226, qrchol.jl:75). */
/* Accumulated since the last reset: device milliseconds spent between the uploads and the
 * downloads of every compute call (CUDA events on the library's stream), and the bytes
 * the calls moved host->device / device->host.  Feeds solver.time_* (Solvers.jl:107-116). */
int hyb_stats(hyb_ctx* ctx, int reset, double* device_ms, int64_t* h2d_bytes, int64_t* d2h_bytes);
/* CUDA-event split of the last hyb_schur_update: W*GQ2 products, G'HG syrk, factorization */
int hyb_last_schur_timings(hyb_ctx* ctx, double* t_apply_w_ms, double* t_syrk_ms, double* t_factor_ms);
int hyb_bench_gemm_tn(hyb_ctx* ctx, int64_t M, int64_t N, int64_t K, int upper, int reps, double* ms_per_launch);
int hyb_bench_potrf(hyb_ctx* ctx, int64_t N, int reps, double* ms_factor, double* ms_solve);

#ifdef __cplusplus
}
#endif
#endif /* HYPATIA_B200_H */
