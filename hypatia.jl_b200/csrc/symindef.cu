// Dense symmetric-indefinite system solver on the device: the drop-in for the reference's
// SymIndefDenseSystemSolver (src/Solvers/systemsolvers/symindef.jl:216-277).
//
// The 3 x 3 system in (x, y, z) after eliminating s, kap and tau (symindef.jl:8-37), with the model reduced
// (p = 0, as every model reaches this library):
//
//        [ 0    G'  ] [x]   [ rhs_x ]
//        [ G   -H~  ] [z] = [ rhs_z ]       H~_k = (mu H_k)^-1 for primal-barrier cones, mu H_k for dual-barrier ones
//
// * load (symindef.jl:231-247): the constant blocks.  Here the matrix is rebuilt on the device at every update
//   (the factorization works in place), G' by one transpose of the resident G*Q2.
// * update_lhs (symindef.jl:249-270): the cone blocks -H~_k are the columns of the block products applied to
//   the identity (selector 8 of hybi_cones_prod: hess for dual-barrier cones, inv_hess for the others), written
//   straight into the z-z block and negated; then symm_fact_copy! (dense.jl:205-219): rook-pivoted Bunch-Kaufman,
//   and on failure the diagonal shift and a second attempt.
// * solve_subsystem3 (symindef.jl:272-280): dsytrs_rook-equivalent solves with the factor.
// The upper triangle is used (the reference stores the lower one): the same symmetric matrix, the same
// factorization up to the order in which LAPACK visits it.
#include <cmath>

#include "ctx.cuh"

struct HybSymIndef {
    HybChol fact;          // (n + q) x (n + q): the system matrix, then its factorization
    double* eye = nullptr; // q x q identity (ld = ldq)
    double* rhs = nullptr; // (n + q) x HYB_MAX_RHS staging of the right-hand sides (ld = ldr)
    int64_t n = 0, q = 0, ldq = 0, ldr = 0;
    bool ok = false;
};

namespace {

__global__ void eye_kernel(int64_t q, int64_t ld, double* __restrict__ E) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t j = blockIdx.y;
    if (i < q) E[i + j * ld] = (i == j) ? 1.0 : 0.0;
}

// z-z block <- -(z-z block), upper triangle incl. diagonal; the rest of it is never read
__global__ void negate_upper_kernel(int64_t q, int64_t ld, double* __restrict__ B) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t j = blockIdx.y;
    if (i < q && i <= j) B[i + j * ld] = -B[i + j * ld];
}

// increase_diag! (dense.jl:100-107) on the whole diagonal of the system matrix
__global__ void si_increase_diag_kernel(int64_t n, double* A, int64_t ld) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double diag_pert = 1.0 + 1e-5, diag_min = 1000.0 * 2.220446049250313e-16;
    A[i + i * ld] = diag_pert * fmax(A[i + i * ld], diag_min);
}

int assemble(hyb_ctx* c, HybSymIndef& si) {
    const int64_t n = si.n, q = si.q, dim = n + q, ld = si.fact.ld;
    HYB_CUDA(cudaMemsetAsync(si.fact.F, 0, sizeof(double) * ld * dim, c->st));
    if (n > 0 && q > 0)  // x-z block: G' (n x q) at rows 0.., columns n..
        HYB_CHECK(hyb_transpose(c->st, c->GQ2, c->ldq, q, n, si.fact.F + n * ld, ld, 0));
    if (q > 0) {
        double* Z = si.fact.F + n + n * ld;
        HYB_CHECK(hybi_cones_prod(c->st, c->cones, 8, si.eye, si.ldq, Z, ld, q));
        dim3 grid((unsigned)((q + 255) / 256), (unsigned)q);
        negate_upper_kernel<<<grid, 256, 0, c->st>>>(q, ld, Z);
        HYB_CUDA(cudaGetLastError());
    }
    return 0;
}

}  // namespace

void hyb_symindef_free(hyb_ctx* c) {
    if (!c->symindef) return;
    hyb_chol_free(c->symindef->fact);
    cudaFree(c->symindef->eye);
    cudaFree(c->symindef->rhs);
    delete c->symindef;
    c->symindef = nullptr;
}

extern "C" {

int hyb_symindef_load(hyb_ctx* c) {
    if (!c || !c->loaded) return -1;
    if (c->assemble_only) {
        hyb_set_error("hyb_symindef_load: this context was loaded assemble-only");
        return -3;
    }
    HYB_CUDA(cudaSetDevice(c->device));
    hyb_symindef_free(c);
    if (c->q > 65535 || c->N + c->q > 65535) {  // the dense (n + q)^2 system: a solver for moderate sizes (grid.y limits too)
        hyb_set_error("hyb_symindef_load: n + q = %lld is beyond what the dense symmetric-indefinite solver is meant for",
                      (long long)(c->N + c->q));
        return -3;
    }
    HybSymIndef* si = new HybSymIndef();
    c->symindef = si;
    si->n = c->N;
    si->q = c->q;
    si->ldq = hyb_round_up(c->q > 0 ? c->q : 1, 16);
    HYB_CHECK(hyb_chol_alloc(si->fact, si->n + si->q));
    HYB_CUDA(cudaMalloc(&si->eye, sizeof(double) * si->ldq * (si->q > 0 ? si->q : 1)));
    si->ldr = hyb_round_up(si->n + si->q > 0 ? si->n + si->q : 1, 16);
    HYB_CUDA(cudaMalloc(&si->rhs, sizeof(double) * si->ldr * HYB_MAX_RHS));
    if (si->q > 0) {
        dim3 grid((unsigned)((si->q + 255) / 256), (unsigned)si->q);
        eye_kernel<<<grid, 256, 0, c->st>>>(si->q, si->ldq, si->eye);
        HYB_CUDA(cudaGetLastError());
    }
    return 0;
}

int hyb_symindef_update(hyb_ctx* c, double* t_ms, int* fact_kind) {
    if (!c || !c->loaded || !c->symindef) {
        hyb_set_error("hyb_symindef_update: call hyb_symindef_load first");
        return -1;
    }
    HYB_CUDA(cudaSetDevice(c->device));
    HybSymIndef& si = *c->symindef;
    if (c->updated_irtmu != c->updated_irtmu) HYB_CHECK(hyb_ensure_updated(c, c->loaded_scal));
    if (t_ms) *t_ms = 0;
    si.ok = false;
    const int64_t dim = si.n + si.q;
    if (dim == 0) {
        si.ok = true;
        if (fact_kind) *fact_kind = 1;
        return 0;
    }
    HYB_CUDA(cudaEventRecord(c->ev1, c->st));
    // symm_fact_copy! (dense.jl:205-219)
    int info = 0, kind = 1;
    HYB_CHECK(assemble(c, si));
    HYB_CHECK(hyb_sytrf_rook(c->st, si.fact, &info));
    if (info != 0) {
        HYB_CHECK(assemble(c, si));
        si_increase_diag_kernel<<<(unsigned)((dim + 255) / 256), 256, 0, c->st>>>(dim, si.fact.F, si.fact.ld);
        HYB_CUDA(cudaGetLastError());
        HYB_CHECK(hyb_sytrf_rook(c->st, si.fact, &info));
        kind = (info == 0) ? 2 : -1;
    }
    HYB_CUDA(cudaEventRecord(c->ev2, c->st));
    HYB_CUDA(cudaEventSynchronize(c->ev2));
    float tf = 0;
    HYB_CUDA(cudaEventElapsedTime(&tf, c->ev1, c->ev2));
    c->dev_ms += tf;
    if (t_ms) *t_ms = tf;
    if (fact_kind) *fact_kind = kind;
    si.ok = (kind >= 0);
    return kind < 0 ? 1 : 0;
}

int hyb_symindef_solve(hyb_ctx* c, const double* rhs, double* sol, int64_t nrhs) {
    if (!c || !c->symindef) return -1;
    HybSymIndef& si = *c->symindef;
    if (!si.ok) {
        hyb_set_error("hyb_symindef_solve: no valid factorization (hyb_symindef_update failed or was not called)");
        return -3;
    }
    if (nrhs < 0 || nrhs > HYB_MAX_RHS) {
        hyb_set_error("hyb_symindef_solve: %lld right-hand sides (max %d)", (long long)nrhs, HYB_MAX_RHS);
        return -3;
    }
    HYB_CUDA(cudaSetDevice(c->device));
    const int64_t dim = si.n + si.q;
    if (dim == 0 || nrhs == 0) return 0;
    const int64_t ldb = si.ldr;
    double* dB = si.rhs;  // persistent staging: no allocation on the solve path
    cudaError_t e = cudaMemcpy2DAsync(dB, sizeof(double) * ldb, rhs, sizeof(double) * dim, sizeof(double) * dim, nrhs,
                                      cudaMemcpyHostToDevice, c->st);
    c->h2d_bytes += (int64_t)sizeof(double) * dim * nrhs;
    int rc = 0;
    if (e == cudaSuccess) {
        KBEGIN(c);
        rc = hyb_sytrs_rook(c->st, si.fact, dB, ldb, (int)nrhs);
        KEND(c);
    }
    if (e == cudaSuccess && rc == 0) {
        e = cudaMemcpy2DAsync(sol, sizeof(double) * dim, dB, sizeof(double) * ldb, sizeof(double) * dim, nrhs,
                              cudaMemcpyDeviceToHost, c->st);
        c->d2h_bytes += (int64_t)sizeof(double) * dim * nrhs;
        if (e == cudaSuccess) e = cudaStreamSynchronize(c->st);
        if (e == cudaSuccess) ksync(c);
    }
    if (e != cudaSuccess) {
        hyb_set_error("hyb_symindef_solve: %s", cudaGetErrorString(e));
        return -1;
    }
    return rc;
}

// the assembled system matrix (upper triangle valid), (n + q) x (n + q) column-major, for tests: rebuilt from
// the current cone state, not the factorized copy
int hyb_symindef_get(hyb_ctx* c, double* lhs) {
    if (!c || !c->symindef) return -1;
    HYB_CUDA(cudaSetDevice(c->device));
    HybSymIndef& si = *c->symindef;
    const int64_t dim = si.n + si.q;
    if (dim == 0) return 0;
    HybChol tmp;
    HYB_CHECK(hyb_chol_alloc(tmp, dim));
    std::swap(tmp.F, si.fact.F);  // assemble into scratch storage of the same shape
    int rc = assemble(c, si);
    cudaError_t e = cudaSuccess;
    if (rc == 0)
        e = cudaMemcpy2DAsync(lhs, sizeof(double) * dim, si.fact.F, sizeof(double) * si.fact.ld, sizeof(double) * dim, dim,
                              cudaMemcpyDeviceToHost, c->st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->st);
    std::swap(tmp.F, si.fact.F);
    hyb_chol_free(tmp);
    if (e != cudaSuccess) {
        hyb_set_error("hyb_symindef_get: %s", cudaGetErrorString(e));
        return -1;
    }
    return rc;
}

}  // extern "C"
