// Large PosSemidefTri cones (side > 32) as sequences of dense device operations.
//
// Follows src/Cones/possemideftri.jl (real case):
//   update_feas :89-99 (svec_to_smat!, cholesky!), is_dual_feas :101-104, update_grad :106-116
//   (inverse from the factor), hess_prod! :135-147 (X^-1 A X^-1 by solves with the factor),
//   inv_hess_prod! :149-160 (X A X), sqrt_hess_prod! :162-174 (U^-T A U^-1),
//   inv_sqrt_hess_prod! :176-192 (U A U'), dder3 :194-204, and the generic check_numerics /
//   get_proxsqr of Cones.jl:267-304.
// The reference loops over the columns of `arr` through one scratch matrix; here all columns
// are unpacked side by side (side x side*m) so each step is ONE multi-RHS triangular solve or
// ONE DMMA contraction, with a batched block transpose in between (A symmetric =>
// (M A) M' = (M (M A)')').
#include "cones_dense.cuh"

#include "cones.cuh"

namespace {

hyb_counter g_dense_launches{0};
#define DLAUNCHED() (++g_dense_launches)
constexpr double RT2 = 1.4142135623730951;

__device__ __forceinline__ void svec_rc(int64_t k, int& r, int& c) {
    int cc = (int)((sqrt(8.0 * (double)k + 1.0) - 1.0) * 0.5);
    while ((int64_t)cc * (cc + 1) / 2 > k) --cc;
    while ((int64_t)(cc + 1) * (cc + 2) / 2 <= k) ++cc;
    c = cc;
    r = (int)(k - (int64_t)cc * (cc + 1) / 2);
}

// svec columns -> full symmetric blocks (arrayutilities.jl:236-251 + copytri!)
__global__ void svec_unpack_kernel(const double* __restrict__ v, int64_t vstride, int s, double* __restrict__ M,
                                   int64_t ld, int64_t blk_stride, double scale) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t j = blockIdx.y;
    const int64_t dim = (int64_t)s * (s + 1) / 2;
    if (k >= dim) return;
    int r, c;
    svec_rc(k, r, c);
    const double x = v[k + j * vstride] * ((r == c) ? scale : scale * (1.0 / RT2));
    double* B = M + j * blk_stride;
    B[r + (int64_t)c * ld] = x;
    B[c + (int64_t)r * ld] = x;
}

// upper triangles of blocks -> svec columns (arrayutilities.jl:183-197)
__global__ void svec_pack_kernel(const double* __restrict__ M, int64_t ld, int64_t blk_stride, int s,
                                 double* __restrict__ v, int64_t vstride, double scale) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t j = blockIdx.y;
    const int64_t dim = (int64_t)s * (s + 1) / 2;
    if (k >= dim) return;
    int r, c;
    svec_rc(k, r, c);
    const double x = M[j * blk_stride + r + (int64_t)c * ld];
    v[k + j * vstride] = scale * ((r == c) ? x : x * RT2);
}

// dst block j = (src block j)'
__global__ void block_transpose_kernel(const double* __restrict__ src, double* __restrict__ dst, int64_t ld,
                                       int64_t blk_stride, int s) {
    __shared__ double tile[32][33];
    const double* S = src + blockIdx.z * blk_stride;
    double* D = dst + blockIdx.z * blk_stride;
    const int r0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    for (int j = threadIdx.y; j < 32; j += 8) {
        const int r = r0 + threadIdx.x, c = c0 + j;
        tile[j][threadIdx.x] = (r < s && c < s) ? S[r + (int64_t)c * ld] : 0.0;
    }
    __syncthreads();
    for (int j = threadIdx.y; j < 32; j += 8) {
        const int c = c0 + threadIdx.x, r = r0 + j;
        if (r < s && c < s) D[c + (int64_t)r * ld] = tile[threadIdx.x][j];
    }
}

__global__ void set_identity_kernel(double* M, int64_t ld, int s) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (int64_t)s * s) return;
    const int r = (int)(e % s), c = (int)(e / s);
    M[r + (int64_t)c * ld] = (r == c) ? 1.0 : 0.0;
}

__global__ void copy_block_kernel(const double* __restrict__ src, double* __restrict__ dst, int64_t ld, int rows,
                                  int64_t cols) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (int64_t)rows * cols) return;
    const int64_t r = e % rows, c = e / rows;
    dst[r + c * ld] = src[r + c * ld];
}

__global__ void lincomb_kernel(int64_t n, double a, const double* __restrict__ x, double b, const double* __restrict__ y,
                               double* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = a * x[i] + (y ? b * y[i] : 0.0);
}

// fixed-order single-CTA reductions
__global__ void __launch_bounds__(1024) dot_kernel(int64_t n, const double* __restrict__ a, const double* __restrict__ b,
                                                   double* out) {
    __shared__ double sp[1024];
    double s = 0.0;
    for (int64_t i = threadIdx.x; i < n; i += 1024) s += a[i] * b[i];
    sp[threadIdx.x] = s;
    __syncthreads();
    for (int k = 512; k > 0; k >>= 1) {
        if (threadIdx.x < k) sp[threadIdx.x] += sp[threadIdx.x + k];
        __syncthreads();
    }
    if (threadIdx.x == 0) *out = sp[0];
}
__global__ void __launch_bounds__(1024) frob_dot_kernel(const double* __restrict__ A, const double* __restrict__ B,
                                                        int64_t ld, int s, double* out) {
    __shared__ double sp[1024];
    double acc = 0.0;
    for (int64_t e = threadIdx.x; e < (int64_t)s * s; e += 1024) {
        const int64_t r = e % s, c = e / s;
        acc += A[r + c * ld] * B[r + c * ld];
    }
    sp[threadIdx.x] = acc;
    __syncthreads();
    for (int k = 512; k > 0; k >>= 1) {
        if (threadIdx.x < k) sp[threadIdx.x] += sp[threadIdx.x + k];
        __syncthreads();
    }
    if (threadIdx.x == 0) *out = sp[0];
}

// status bits and proxsqr of the cone from the factorization infos and the three scalars
// scal[0] = <g, p>, scal[1] = <Hinv g, g>, scal[2] = psi' Hinv psi   (Cones.jl:267-304)
__global__ void psd_finalize_kernel(const int* info_x, const int* info_z, const double* scal, int s, int64_t dim,
                                    int* flag, double* prox) {
    int f = 0;
    double p = 0.0;
    if (*info_z == 0) f |= HYB_ST_DUAL_FEAS;
    if (*info_x == 0) {
        f |= HYB_ST_FEAS;
        const double nu = (double)s;
        bool ok = !(fabs(1.0 + scal[0] / nu) > 1.220703125e-4 * (double)dim);
        if (fabs(1.0 - scal[1] / nu) > 1.104854345603981e-1 * (double)dim) ok = false;
        if (ok) f |= HYB_ST_NUMERICS;
        p = (scal[2] < -1.4901161193847656e-08 * (double)dim) ? INFINITY : fabs(scal[2]);
    }
    *flag = f;
    *prox = p;
}

inline unsigned nb1(int64_t n, int per = 256) { return (unsigned)((n + per - 1) / per); }

}  // namespace

int64_t hybd_launch_count() { return g_dense_launches; }

struct HybDensePsd {
    int s = 0;
    int64_t dim = 0, ld = 0;
    double *X = nullptr, *Ux = nullptr, *Lx = nullptr, *Xinv = nullptr, *W1 = nullptr, *W2 = nullptr, *W3 = nullptr;
    double *vec1 = nullptr, *scal = nullptr;
    HybChol chX, chW;
    double *big1 = nullptr, *big2 = nullptr;
    int64_t big_cols = 0;
};

static int psd_ensure_big(HybDensePsd* c, int64_t m) {
    if (m <= c->big_cols) return 0;
    cudaFree(c->big1);
    cudaFree(c->big2);
    c->big1 = c->big2 = nullptr;
    HYB_CUDA(cudaMalloc(&c->big1, sizeof(double) * c->ld * c->s * m));
    HYB_CUDA(cudaMalloc(&c->big2, sizeof(double) * c->ld * c->s * m));
    c->big_cols = m;
    return 0;
}

int hybd_psd_create(HybDensePsd** out, int side) {
    HybDensePsd* c = new HybDensePsd();
    c->s = side;
    c->dim = (int64_t)side * (side + 1) / 2;
    c->ld = hyb_round_up(side, 16);
    const size_t bytes = sizeof(double) * c->ld * side;
    double** mats[] = {&c->X, &c->Ux, &c->Lx, &c->Xinv, &c->W1, &c->W2, &c->W3};
    for (double** m : mats) {
        HYB_CUDA(cudaMalloc(m, bytes));
        HYB_CUDA(cudaMemset(*m, 0, bytes));
    }
    HYB_CUDA(cudaMalloc(&c->vec1, sizeof(double) * c->dim));
    HYB_CUDA(cudaMalloc(&c->scal, sizeof(double) * 8));
    HYB_CUDA(cudaMemset(c->scal, 0, sizeof(double) * 8));
    HYB_CHECK(hyb_chol_attach(c->chX, c->Ux, c->ld, side));
    HYB_CHECK(hyb_chol_attach(c->chW, c->W1, c->ld, side));
    *out = c;
    return 0;
}

void hybd_psd_destroy(HybDensePsd* c) {
    if (!c) return;
    hyb_chol_free(c->chX);
    hyb_chol_free(c->chW);
    cudaFree(c->X);
    cudaFree(c->Ux);
    cudaFree(c->Lx);
    cudaFree(c->Xinv);
    cudaFree(c->W1);
    cudaFree(c->W2);
    cudaFree(c->W3);
    cudaFree(c->vec1);
    cudaFree(c->scal);
    cudaFree(c->big1);
    cudaFree(c->big2);
    delete c;
}

// out = X * smat(v) * X and *dot_out = <out, smat(v)>   (inv_hess_prod of one vector + inner product)
static int psd_xvx_dot(cudaStream_t st, HybDensePsd* c, const double* v, double* dot_out) {
    const int s = c->s;
    const int64_t ld = c->ld;
    svec_unpack_kernel<<<dim3(nb1(c->dim), 1), 256, 0, st>>>(v, 0, s, c->W1, ld, 0, 1.0);
    DLAUNCHED();
    HYB_CHECK(hyb_gemm_tn(st, s, s, s, 1.0, c->X, ld, c->W1, ld, 0.0, c->W2, ld, 0));           // X V
    HYB_CHECK(hyb_transpose(st, c->W2, ld, s, s, c->W3, ld, 0));                               // V X
    HYB_CHECK(hyb_gemm_tn(st, s, s, s, 1.0, c->X, ld, c->W3, ld, 0.0, c->W2, ld, 0));           // X V X
    frob_dot_kernel<<<1, 1024, 0, st>>>(c->W2, c->W1, ld, s, dot_out);
    DLAUNCHED();
    HYB_CUDA(cudaGetLastError());
    return 0;
}

int hybd_psd_update(cudaStream_t st, HybDensePsd* c, const double* primal, const double* dual_in, double scal,
                    double irtmu, double* point, double* dual_out, double* grad, int* flag, double* prox) {
    const int s = c->s;
    const int64_t ld = c->ld, dim = c->dim;
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, scal, primal, 0.0, nullptr, point);
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, 1.0, dual_in, 0.0, nullptr, dual_out);
    // dual feasibility: cholesky of smat(dual)
    svec_unpack_kernel<<<dim3(nb1(dim), 1), 256, 0, st>>>(dual_out, 0, s, c->W1, ld, 0, 1.0);
    HYB_CHECK(hyb_potrf(st, c->chW, nullptr));
    // primal feasibility: X = smat(point), Ux = chol(X)
    svec_unpack_kernel<<<dim3(nb1(dim), 1), 256, 0, st>>>(point, 0, s, c->X, ld, 0, 1.0);
    copy_block_kernel<<<nb1((int64_t)s * s), 256, 0, st>>>(c->X, c->Ux, ld, s, s);
    g_dense_launches += 5;
    HYB_CUDA(cudaGetLastError());
    int inf_x = 0;
    HYB_CHECK(hyb_potrf(st, c->chX, &inf_x));
    if (inf_x != 0) {  // infeasible candidate: rejected before grad / Hessian are needed (search.jl:146-150)
        psd_finalize_kernel<<<1, 1, 0, st>>>(c->chX.info, c->chW.info, c->scal, s, dim, flag, prox);
        DLAUNCHED();
        HYB_CUDA(cudaGetLastError());
        return 0;
    }
    HYB_CHECK(hyb_transpose(st, c->Ux, ld, s, s, c->Lx, ld, 1));
    // Xinv = X^-1 by two triangular solves on the identity (update_grad, possemideftri.jl:106-116)
    set_identity_kernel<<<nb1((int64_t)s * s), 256, 0, st>>>(c->Xinv, ld, s);
    DLAUNCHED();
    HYB_CHECK(hyb_trsm_fwd(st, c->Ux, ld, s, c->Xinv, ld, s));
    HYB_CHECK(hyb_trsm_bwd(st, c->Ux, c->Lx, ld, s, c->Xinv, ld, s));
    svec_pack_kernel<<<dim3(nb1(dim), 1), 256, 0, st>>>(c->Xinv, ld, 0, s, grad, 0, -1.0);
    // check_numerics and proxsqr (Cones.jl:267-304) with inv_hess_prod = X . X
    dot_kernel<<<1, 1024, 0, st>>>(dim, grad, point, c->scal + 0);
    g_dense_launches += 2;
    HYB_CHECK(psd_xvx_dot(st, c, grad, c->scal + 1));
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, irtmu, dual_out, 1.0, grad, c->vec1);
    DLAUNCHED();
    HYB_CHECK(psd_xvx_dot(st, c, c->vec1, c->scal + 2));
    psd_finalize_kernel<<<1, 1, 0, st>>>(c->chX.info, c->chW.info, c->scal, s, dim, flag, prox);
    DLAUNCHED();
    HYB_CUDA(cudaGetLastError());
    return 0;
}

int hybd_psd_prod(cudaStream_t st, HybDensePsd* c, int op, const double* arr, int64_t lda, double* prod, int64_t ldp,
                  int64_t m) {
    if (m <= 0) return 0;
    HYB_CHECK(psd_ensure_big(c, m));
    const int s = c->s;
    const int64_t ld = c->ld, dim = c->dim, blk = ld * s, cols = (int64_t)s * m;
    double *A = c->big1, *B = c->big2;
    svec_unpack_kernel<<<dim3(nb1(dim), (unsigned)m), 256, 0, st>>>(arr, lda, s, A, ld, blk, 1.0);
    DLAUNCHED();
    HYB_CUDA(cudaGetLastError());
    dim3 tgrid((unsigned)((s + 31) / 32), (unsigned)((s + 31) / 32), (unsigned)m);
    auto btrans = [&](const double* src, double* dst) -> int {
        block_transpose_kernel<<<tgrid, dim3(32, 8), 0, st>>>(src, dst, ld, blk, s);
        DLAUNCHED();
        HYB_CUDA(cudaGetLastError());
        return 0;
    };
    const double* result = nullptr;
    if (op == HYB_OP_HESS) {  // X^-1 A X^-1 : solves with the factor on both sides
        HYB_CHECK(hyb_trsm_fwd(st, c->Ux, ld, s, A, ld, cols));
        HYB_CHECK(hyb_trsm_bwd(st, c->Ux, c->Lx, ld, s, A, ld, cols));
        HYB_CHECK(btrans(A, B));
        HYB_CHECK(hyb_trsm_fwd(st, c->Ux, ld, s, B, ld, cols));
        HYB_CHECK(hyb_trsm_bwd(st, c->Ux, c->Lx, ld, s, B, ld, cols));
        result = B;
    } else if (op == HYB_OP_INV_HESS) {  // X A X
        HYB_CHECK(hyb_gemm_tn(st, s, cols, s, 1.0, c->X, ld, A, ld, 0.0, B, ld, 0));
        HYB_CHECK(btrans(B, A));
        HYB_CHECK(hyb_gemm_tn(st, s, cols, s, 1.0, c->X, ld, A, ld, 0.0, B, ld, 0));
        result = B;
    } else if (op == HYB_OP_SQRT_HESS) {  // U^-T A U^-1
        HYB_CHECK(hyb_trsm_fwd(st, c->Ux, ld, s, A, ld, cols));
        HYB_CHECK(btrans(A, B));
        HYB_CHECK(hyb_trsm_fwd(st, c->Ux, ld, s, B, ld, cols));
        result = B;
    } else {  // U A U' : U = Lx'
        HYB_CHECK(hyb_gemm_tn(st, s, cols, s, 1.0, c->Lx, ld, A, ld, 0.0, B, ld, 0));
        HYB_CHECK(btrans(B, A));
        HYB_CHECK(hyb_gemm_tn(st, s, cols, s, 1.0, c->Lx, ld, A, ld, 0.0, B, ld, 0));
        result = B;
    }
    svec_pack_kernel<<<dim3(nb1(dim), (unsigned)m), 256, 0, st>>>(result, ld, blk, s, prod, ldp, 1.0);
    DLAUNCHED();
    HYB_CUDA(cudaGetLastError());
    return 0;
}

// dder3 = svec(S S'), S = X^-1 D U^-1 (possemideftri.jl:194-204): with V = U^-T (X^-1 D)' = S',  S S' = V' V
int hybd_psd_dder3(cudaStream_t st, HybDensePsd* c, const double* dir, double* out) {
    const int s = c->s;
    const int64_t ld = c->ld, dim = c->dim;
    svec_unpack_kernel<<<dim3(nb1(dim), 1), 256, 0, st>>>(dir, 0, s, c->W2, ld, 0, 1.0);
    DLAUNCHED();
    HYB_CUDA(cudaGetLastError());
    HYB_CHECK(hyb_trsm_fwd(st, c->Ux, ld, s, c->W2, ld, s));
    HYB_CHECK(hyb_trsm_bwd(st, c->Ux, c->Lx, ld, s, c->W2, ld, s));  // X^-1 D
    HYB_CHECK(hyb_transpose(st, c->W2, ld, s, s, c->W3, ld, 0));      // D X^-1
    HYB_CHECK(hyb_trsm_fwd(st, c->Ux, ld, s, c->W3, ld, s));          // V = U^-T D X^-1
    HYB_CHECK(hyb_gemm_tn(st, s, s, s, 1.0, c->W3, ld, c->W3, ld, 0.0, c->W2, ld, 0));
    svec_pack_kernel<<<dim3(nb1(dim), 1), 256, 0, st>>>(c->W2, ld, 0, s, out, 0, 1.0);
    DLAUNCHED();
    HYB_CUDA(cudaGetLastError());
    return 0;
}

// =======================================================================================
// WSOSInterpNonnegative (real): src/Cones/wsosinterpnonnegative.jl
//   update_feas :99-128   Lambda_k = P_k' diag(s) P_k, Cholesky each
//   is_dual_feas :130-133 all(z > eps) (necessary condition only)
//   update_grad :135-149  F_k = R_k^-T P_k'  (Lambda_k = R_k' R_k),  g_j = -sum_k |F_k[:,j]|^2
//   update_hess :151-166  H = sum_k (F_k' F_k) .^ 2
//   generic products with H / chol(H): Cones.jl:108-125,202-214; update_hess_fact :233-245
//   hess_prod_slow! :168-191, dder3 :193-204, partial_lambda! :206-211
//   update_use_hess_prod_slow: Cones.jl:218-227
// Everything is evaluated eagerly at load time (the reference fills the same caches lazily
// during check_cone_points); if chol(H) fails the candidate is reported as failing
// check_numerics (the reference would continue with a Bunch-Kaufman factorization).
// =======================================================================================
namespace {

__global__ void rowscale_kernel(const double* __restrict__ in, const double* __restrict__ d, int64_t rows, int64_t cols,
                                int64_t ld, double* __restrict__ out) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= rows * cols) return;
    const int64_t r = e % rows, c = e / rows;
    out[r + c * ld] = d[r] * in[r + c * ld];
}
// out[j] = beta * out[j] + alpha * sum_r A[r,j] * B[r,j]   (one warp per column)
__global__ void coldot_kernel(const double* __restrict__ A, const double* __restrict__ B, int64_t rows, int64_t cols,
                              int64_t ld, double alpha, double beta, double* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t j = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (j >= cols) return;
    double s = 0.0;
    for (int64_t r = lane; r < rows; r += 32) s += A[r + j * ld] * B[r + j * ld];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) out[j] = (beta == 0.0 ? 0.0 : beta * out[j]) + alpha * s;
}
// H_upper (+)= UU_upper .^ 2
__global__ void sq_accum_kernel(const double* __restrict__ UU, double* __restrict__ H, int64_t n, int64_t ld, int first) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n * n) return;
    const int64_t r = e % n, c = e / n;
    if (r > c) return;
    const double u = UU[r + c * ld];
    H[r + c * ld] = (first ? 0.0 : H[r + c * ld]) + u * u;
}
__global__ void symmetrize_kernel(double* __restrict__ H, int64_t n, int64_t ld) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n * n) return;
    const int64_t r = e % n, c = e / n;
    if (r > c) H[r + c * ld] = H[c + r * ld];
}
__global__ void copy_upper_dense_kernel(const double* __restrict__ src, double* __restrict__ dst, int64_t n, int64_t ld) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n * n) return;
    const int64_t r = e % n, c = e / n;
    dst[r + c * ld] = (r <= c) ? src[r + c * ld] : 0.0;
}
__global__ void __launch_bounds__(1024) all_gt_eps_kernel(int64_t n, const double* __restrict__ v, int* out) {
    __shared__ int ok;
    if (threadIdx.x == 0) ok = 1;
    __syncthreads();
    int mine = 1;
    for (int64_t i = threadIdx.x; i < n; i += 1024)
        if (!(v[i] > HYB_EPS)) mine = 0;
    if (!mine) ok = 0;
    __syncthreads();
    if (threadIdx.x == 0) *out = ok;
}
__global__ void copy_mat_kernel(const double* __restrict__ src, int64_t lds_, double* __restrict__ dst, int64_t ldd,
                                int64_t rows, int64_t cols) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= rows * cols) return;
    const int64_t r = e % rows, c = e / rows;
    dst[r + c * ldd] = src[r + c * lds_];
}
// scal: [0] <g,p>  [1] <Hinv g, g>  [2] psi' Hinv psi  [3] p' H p ; infos: K Lambda infos then H info
__global__ void wsos_finalize_kernel(const int* const* lam_infos, int nPs, const int* info_h, const int* dual_ok,
                                     const double* scal, double nu, int64_t dim, int* flag, double* prox, int* slow,
                                     int* sqrt_ok) {
    int f = 0;
    double p = 0.0;
    bool feas = true;
    for (int k = 0; k < nPs; ++k)
        if (*lam_infos[k] != 0) feas = false;
    if (*dual_ok) f |= HYB_ST_DUAL_FEAS;
    *sqrt_ok = (feas && *info_h == 0) ? 1 : 0;
    *slow = 0;
    if (feas) {
        f |= HYB_ST_FEAS;
        bool ok = (*info_h == 0);
        if (fabs(1.0 + scal[0] / nu) > 1.220703125e-4 * (double)dim) ok = false;
        if (fabs(1.0 - scal[1] / nu) > 1.104854345603981e-1 * (double)dim) ok = false;
        if (ok) f |= HYB_ST_NUMERICS;
        p = (scal[2] < -1.4901161193847656e-08 * (double)dim) ? INFINITY : fabs(scal[2]);
        // update_use_hess_prod_slow (Cones.jl:218-227)
        if (fabs(1.0 - scal[3] / nu) > (double)dim * 1.4901161193847656e-08) *slow = 1;
    }
    *flag = f;
    *prox = p;
}

}  // namespace

struct HybDenseWsos {
    int64_t U = 0, ldU = 0;
    int nPs = 0;
    double nu = 0.0;
    std::vector<int64_t> L, ldL;
    std::vector<double*> P, Pt, F, Ft, Lam, M1, LU;  // per k
    std::vector<HybChol> chLam;
    double *PD = nullptr;                             // U x Lmax scratch
    double *H = nullptr, *Hf = nullptr, *Lh = nullptr, *UU = nullptr;
    // the K chains Lambda_k -> chol -> F_k -> F_k'F_k are independent of each other (wsosinterpnonnegative.jl:
    // 99-166 loops over k) and each is a string of latency-bound launches on small matrices: they run side by
    // side on one stream per k, with scratch of their own, and are joined in k order (fixed summation order)
    std::vector<double*> PDk, UUk;
    std::vector<cudaStream_t> sk;
    std::vector<cudaEvent_t> evk;
    cudaEvent_t ev_fork = nullptr;
    HybChol chH;
    double *v1 = nullptr, *v2 = nullptr, *v3 = nullptr, *scal = nullptr;
    int *dual_ok = nullptr, *slow = nullptr, *sqrt_ok = nullptr;
    const int** lam_infos_dev = nullptr;
    double *big1 = nullptr, *big2 = nullptr;
    int64_t big_cols = 0;
    double* gemv_scratch = nullptr;
};

static int wsos_ensure_big(HybDenseWsos* c, int64_t m) {
    if (m <= c->big_cols) return 0;
    cudaFree(c->big1);
    cudaFree(c->big2);
    c->big1 = c->big2 = nullptr;
    HYB_CUDA(cudaMalloc(&c->big1, sizeof(double) * c->ldU * m));
    HYB_CUDA(cudaMalloc(&c->big2, sizeof(double) * c->ldU * m));
    c->big_cols = m;
    return 0;
}

int hybd_wsos_create(HybDenseWsos** out, int64_t U, int nPs, const double* const* Ps_host, const int64_t* Ls) {
    HybDenseWsos* c = new HybDenseWsos();
    c->U = U;
    c->ldU = hyb_round_up(U, 16);
    c->nPs = nPs;
    int64_t Lmax = 1;
    c->chLam.resize(nPs);
    for (int k = 0; k < nPs; ++k) {
        const int64_t L = Ls[k], ldL = hyb_round_up(L, 16);
        c->L.push_back(L);
        c->ldL.push_back(ldL);
        c->nu += (double)L;
        Lmax = L > Lmax ? L : Lmax;
        double *P = nullptr, *Pt = nullptr, *F = nullptr, *Ft = nullptr, *Lam = nullptr, *M1 = nullptr, *LU = nullptr;
        HYB_CUDA(cudaMalloc(&P, sizeof(double) * c->ldU * L));
        HYB_CUDA(cudaMemset(P, 0, sizeof(double) * c->ldU * L));
        HYB_CUDA(cudaMemcpy2D(P, sizeof(double) * c->ldU, Ps_host[k], sizeof(double) * U, sizeof(double) * U, L,
                              cudaMemcpyHostToDevice));
        HYB_CUDA(cudaMalloc(&Pt, sizeof(double) * ldL * U));
        HYB_CUDA(cudaMemset(Pt, 0, sizeof(double) * ldL * U));
        HYB_CHECK(hyb_transpose(nullptr, P, c->ldU, U, L, Pt, ldL, 0));
        HYB_CUDA(cudaMalloc(&F, sizeof(double) * ldL * U));
        HYB_CUDA(cudaMemset(F, 0, sizeof(double) * ldL * U));
        HYB_CUDA(cudaMalloc(&LU, sizeof(double) * ldL * U));
        HYB_CUDA(cudaMalloc(&Ft, sizeof(double) * c->ldU * L));
        HYB_CUDA(cudaMemset(Ft, 0, sizeof(double) * c->ldU * L));
        HYB_CUDA(cudaMalloc(&Lam, sizeof(double) * ldL * L));
        HYB_CUDA(cudaMemset(Lam, 0, sizeof(double) * ldL * L));
        HYB_CUDA(cudaMalloc(&M1, sizeof(double) * ldL * L));
        c->P.push_back(P);
        c->Pt.push_back(Pt);
        c->F.push_back(F);
        c->Ft.push_back(Ft);
        c->Lam.push_back(Lam);
        c->M1.push_back(M1);
        c->LU.push_back(LU);
        HYB_CHECK(hyb_chol_attach(c->chLam[k], Lam, ldL, L));
    }
    HYB_CUDA(cudaMalloc(&c->PD, sizeof(double) * c->ldU * Lmax));
    const size_t ub = sizeof(double) * c->ldU * U;
    HYB_CUDA(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
    for (int k = 0; k < nPs; ++k) {
        double *pd = nullptr, *uu = nullptr;
        HYB_CUDA(cudaMalloc(&pd, sizeof(double) * c->ldU * c->L[k]));
        HYB_CUDA(cudaMalloc(&uu, ub));
        c->PDk.push_back(pd);
        c->UUk.push_back(uu);
        cudaStream_t s2 = nullptr;
        cudaEvent_t e2 = nullptr;
        HYB_CUDA(cudaStreamCreateWithFlags(&s2, cudaStreamNonBlocking));
        HYB_CUDA(cudaEventCreateWithFlags(&e2, cudaEventDisableTiming));
        c->sk.push_back(s2);
        c->evk.push_back(e2);
    }
    double** mats[] = {&c->H, &c->Hf, &c->Lh, &c->UU};
    for (double** m : mats) {
        HYB_CUDA(cudaMalloc(m, ub));
        HYB_CUDA(cudaMemset(*m, 0, ub));
    }
    HYB_CHECK(hyb_chol_attach(c->chH, c->Hf, c->ldU, U));
    HYB_CUDA(cudaMalloc(&c->v1, sizeof(double) * c->ldU));
    HYB_CUDA(cudaMalloc(&c->v2, sizeof(double) * c->ldU));
    HYB_CUDA(cudaMalloc(&c->v3, sizeof(double) * c->ldU));
    HYB_CUDA(cudaMalloc(&c->scal, sizeof(double) * 8));
    HYB_CUDA(cudaMemset(c->scal, 0, sizeof(double) * 8));
    HYB_CUDA(cudaMalloc(&c->dual_ok, sizeof(int)));
    HYB_CUDA(cudaMalloc(&c->slow, sizeof(int)));
    HYB_CUDA(cudaMalloc(&c->sqrt_ok, sizeof(int)));
    HYB_CUDA(cudaMemset(c->slow, 0, sizeof(int)));
    HYB_CUDA(cudaMemset(c->sqrt_ok, 0, sizeof(int)));
    std::vector<const int*> infos;
    for (int k = 0; k < nPs; ++k) infos.push_back(c->chLam[k].info);
    HYB_CUDA(cudaMalloc((void**)&c->lam_infos_dev, sizeof(int*) * (nPs > 0 ? nPs : 1)));
    HYB_CUDA(cudaMemcpy((void*)c->lam_infos_dev, infos.data(), sizeof(int*) * nPs, cudaMemcpyHostToDevice));
    HYB_CUDA(cudaMalloc(&c->gemv_scratch, sizeof(double) * U * HYB_GEMV_SPLITS));
    HYB_CUDA(cudaDeviceSynchronize());
    *out = c;
    return 0;
}

void hybd_wsos_destroy(HybDenseWsos* c) {
    if (!c) return;
    for (int k = 0; k < c->nPs; ++k) {
        hyb_chol_free(c->chLam[k]);
        cudaFree(c->P[k]);
        cudaFree(c->Pt[k]);
        cudaFree(c->F[k]);
        cudaFree(c->Ft[k]);
        cudaFree(c->Lam[k]);
        cudaFree(c->M1[k]);
        cudaFree(c->LU[k]);
    }
    hyb_chol_free(c->chH);
    for (double* b : c->PDk) cudaFree(b);
    for (double* b : c->UUk) cudaFree(b);
    for (cudaStream_t s2 : c->sk) cudaStreamDestroy(s2);
    for (cudaEvent_t e2 : c->evk) cudaEventDestroy(e2);
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    cudaFree(c->PD);
    cudaFree(c->H);
    cudaFree(c->Hf);
    cudaFree(c->Lh);
    cudaFree(c->UU);
    cudaFree(c->v1);
    cudaFree(c->v2);
    cudaFree(c->v3);
    cudaFree(c->scal);
    cudaFree(c->dual_ok);
    cudaFree(c->slow);
    cudaFree(c->sqrt_ok);
    cudaFree((void*)c->lam_infos_dev);
    cudaFree(c->big1);
    cudaFree(c->big2);
    cudaFree(c->gemv_scratch);
    delete c;
}

// LU_k = (F_k diag(d) F_k') F_k   (partial_lambda!, wsosinterpnonnegative.jl:206-211), d a device U-vector
static int wsos_partial_lambda(cudaStream_t st, HybDenseWsos* c, int k, const double* d) {
    const int64_t U = c->U, L = c->L[k], ldU = c->ldU, ldL = c->ldL[k];
    rowscale_kernel<<<nb1(U * L), 256, 0, st>>>(c->Ft[k], d, U, L, ldU, c->PD);
    DLAUNCHED();
    HYB_CUDA(cudaGetLastError());
    HYB_CHECK(hyb_gemm_tn(st, L, L, U, 1.0, c->Ft[k], ldU, c->PD, ldU, 0.0, c->M1[k], ldL, 0));
    HYB_CHECK(hyb_gemm_tn(st, L, U, L, 1.0, c->M1[k], ldL, c->F[k], ldL, 0.0, c->LU[k], ldL, 0));
    return 0;
}

int hybd_wsos_update(cudaStream_t st, HybDenseWsos* c, const double* primal, const double* dual_in, double scal,
                     double irtmu, double* point, double* dual_out, double* grad, int* flag, double* prox) {
    const int64_t U = c->U, ldU = c->ldU;
    lincomb_kernel<<<nb1(U), 256, 0, st>>>(U, scal, primal, 0.0, nullptr, point);
    lincomb_kernel<<<nb1(U), 256, 0, st>>>(U, 1.0, dual_in, 0.0, nullptr, dual_out);
    lincomb_kernel<<<nb1(U), 256, 0, st>>>(U, 1.0, point, 0.0, nullptr, c->v3);  // aligned copy of the point
    all_gt_eps_kernel<<<1, 1024, 0, st>>>(U, dual_out, c->dual_ok);
    g_dense_launches += 4;
    HYB_CUDA(cudaGetLastError());
    // fork: Lambda_k = P_k' diag(s) P_k (upper) and its Cholesky, one stream per k
    HYB_CUDA(cudaEventRecord(c->ev_fork, st));
    for (int k = 0; k < c->nPs; ++k) {
        const int64_t L = c->L[k], ldL = c->ldL[k];
        cudaStream_t s2 = c->sk[k];
        HYB_CUDA(cudaStreamWaitEvent(s2, c->ev_fork, 0));
        rowscale_kernel<<<nb1(U * L), 256, 0, s2>>>(c->P[k], c->v3, U, L, ldU, c->PDk[k]);
        DLAUNCHED();
        HYB_CHECK(hyb_gemm_tn(s2, L, L, U, 1.0, c->P[k], ldU, c->PDk[k], ldU, 0.0, c->Lam[k], ldL, HYB_GEMM_UPPER));
        HYB_CHECK(hyb_potrf(s2, c->chLam[k], nullptr));
        HYB_CUDA(cudaEventRecord(c->evk[k], s2));
        HYB_CUDA(cudaStreamWaitEvent(st, c->evk[k], 0));
    }
    {
        int infs[16] = {0};
        const int nk = c->nPs < 16 ? c->nPs : 16;
        for (int k = 0; k < nk; ++k)
            HYB_CUDA(cudaMemcpyAsync(&infs[k], c->chLam[k].info, sizeof(int), cudaMemcpyDeviceToHost, st));
        HYB_CUDA(cudaStreamSynchronize(st));
        bool infeasible = false;
        for (int k = 0; k < nk; ++k) infeasible = infeasible || infs[k] != 0;
        if (infeasible) {
            // not in the cone: update_feas reports the first Lambda_k that is not positive definite
            // (wsosinterpnonnegative.jl:99-128) and the line search rejects the candidate before any
            // gradient / Hessian oracle is asked for (search.jl:146-150) -- skip that work (3/4 of it)
            wsos_finalize_kernel<<<1, 1, 0, st>>>(c->lam_infos_dev, c->nPs, c->chH.info, c->dual_ok, c->scal, c->nu, U, flag,
                                                  prox, c->slow, c->sqrt_ok);
            DLAUNCHED();
            HYB_CUDA(cudaGetLastError());
            return 0;
        }
    }
    // fork again: F_k = R_k^-T P_k' and UU_k = F_k' F_k
    HYB_CUDA(cudaEventRecord(c->ev_fork, st));
    for (int k = 0; k < c->nPs; ++k) {
        const int64_t L = c->L[k], ldL = c->ldL[k];
        cudaStream_t s2 = c->sk[k];
        HYB_CUDA(cudaStreamWaitEvent(s2, c->ev_fork, 0));
        copy_mat_kernel<<<nb1(L * U), 256, 0, s2>>>(c->Pt[k], ldL, c->F[k], ldL, L, U);
        DLAUNCHED();
        HYB_CHECK(hyb_trsm_fwd(s2, c->Lam[k], ldL, L, c->F[k], ldL, U));
        HYB_CHECK(hyb_transpose(s2, c->F[k], ldL, L, U, c->Ft[k], ldU, 0));
        HYB_CHECK(hyb_gemm_tn(s2, U, U, L, 1.0, c->F[k], ldL, c->F[k], ldL, 0.0, c->UUk[k], ldU, HYB_GEMM_UPPER));
        HYB_CUDA(cudaEventRecord(c->evk[k], s2));
    }
    // join in k order: grad_j -= |F_k[:, j]|^2 and H (+)= (F_k' F_k) .^ 2, fixed summation order
    for (int k = 0; k < c->nPs; ++k) {
        const int64_t L = c->L[k], ldL = c->ldL[k];
        HYB_CUDA(cudaStreamWaitEvent(st, c->evk[k], 0));
        coldot_kernel<<<nb1(U, 8), 256, 0, st>>>(c->F[k], c->F[k], L, U, ldL, -1.0, k == 0 ? 0.0 : 1.0, grad);
        sq_accum_kernel<<<nb1(U * U), 256, 0, st>>>(c->UUk[k], c->H, U, ldU, k == 0);
        g_dense_launches += 2;
        HYB_CUDA(cudaGetLastError());
    }
    symmetrize_kernel<<<nb1(U * U), 256, 0, st>>>(c->H, U, ldU);
    copy_upper_dense_kernel<<<nb1(U * U), 256, 0, st>>>(c->H, c->Hf, U, ldU);
    g_dense_launches += 2;
    HYB_CHECK(hyb_potrf(st, c->chH, nullptr));
    HYB_CHECK(hyb_transpose(st, c->Hf, ldU, U, U, c->Lh, ldU, 1));
    // scalars for check_numerics / proxsqr / hess_prod_slow switch
    lincomb_kernel<<<nb1(U), 256, 0, st>>>(U, 1.0, grad, 0.0, nullptr, c->v1);
    dot_kernel<<<1, 1024, 0, st>>>(U, c->v1, c->v3, c->scal + 0);
    HYB_CHECK(hyb_potrs(st, c->chH, c->v1));  // Hinv g
    lincomb_kernel<<<nb1(U), 256, 0, st>>>(U, 1.0, grad, 0.0, nullptr, c->v2);
    dot_kernel<<<1, 1024, 0, st>>>(U, c->v1, c->v2, c->scal + 1);
    lincomb_kernel<<<nb1(U), 256, 0, st>>>(U, irtmu, dual_out, 1.0, grad, c->v2);  // psi
    lincomb_kernel<<<nb1(U), 256, 0, st>>>(U, 1.0, c->v2, 0.0, nullptr, c->v1);
    HYB_CHECK(hyb_potrs(st, c->chH, c->v1));
    dot_kernel<<<1, 1024, 0, st>>>(U, c->v1, c->v2, c->scal + 2);
    HYB_CHECK(hyb_gemv_t(st, U, U, c->H, ldU, c->v3, c->v1));  // H p (H symmetric)
    dot_kernel<<<1, 1024, 0, st>>>(U, c->v1, c->v3, c->scal + 3);
    wsos_finalize_kernel<<<1, 1, 0, st>>>(c->lam_infos_dev, c->nPs, c->chH.info, c->dual_ok, c->scal, c->nu, U, flag, prox,
                                          c->slow, c->sqrt_ok);
    g_dense_launches += 10;
    HYB_CUDA(cudaGetLastError());
    return 0;
}

int hybd_wsos_sqrt_ok(cudaStream_t st, HybDenseWsos* c, int* ok) {
    HYB_CUDA(cudaMemcpyAsync(ok, c->sqrt_ok, sizeof(int), cudaMemcpyDeviceToHost, st));
    HYB_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int hybd_wsos_prod(cudaStream_t st, HybDenseWsos* c, int op, const double* arr, int64_t lda, double* prod, int64_t ldp,
                   int64_t m) {
    if (m <= 0) return 0;
    const int64_t U = c->U, ldU = c->ldU;
    HYB_CHECK(wsos_ensure_big(c, m));
    double *A = c->big1, *B = c->big2;
    // aligned working copy of the input block (the cone's rows may start at an odd offset)
    copy_mat_kernel<<<nb1(U * m), 256, 0, st>>>(arr, lda, A, ldU, U, m);
    DLAUNCHED();
    HYB_CUDA(cudaGetLastError());
    const double* result = nullptr;
    if (op == 6) {
        int slow = 0;
        HYB_CUDA(cudaMemcpyAsync(&slow, c->slow, sizeof(int), cudaMemcpyDeviceToHost, st));
        HYB_CUDA(cudaStreamSynchronize(st));
        if (!slow) op = HYB_OP_HESS;
    }
    if (op == HYB_OP_HESS) {
        if (m == 1) HYB_CHECK(hyb_gemv_t(st, U, U, c->H, ldU, A, B));
        else HYB_CHECK(hyb_gemm_tn(st, U, m, U, 1.0, c->H, ldU, A, ldU, 0.0, B, ldU, 0));
        result = B;
    } else if (op == HYB_OP_INV_HESS) {
        if (m == 1) {
            HYB_CHECK(hyb_potrs(st, c->chH, A));
        } else {
            HYB_CHECK(hyb_trsm_fwd(st, c->Hf, ldU, U, A, ldU, m));
            HYB_CHECK(hyb_trsm_bwd(st, c->Hf, c->Lh, ldU, U, A, ldU, m));
        }
        result = A;
    } else if (op == HYB_OP_SQRT_HESS) {  // chol(H).U * arr = Lh' arr
        if (m == 1) HYB_CHECK(hyb_gemv_t(st, U, U, c->Lh, ldU, A, B));
        else HYB_CHECK(hyb_gemm_tn(st, U, m, U, 1.0, c->Lh, ldU, A, ldU, 0.0, B, ldU, 0));
        result = B;
    } else if (op == HYB_OP_INV_SQRT_HESS) {  // chol(H).U' \ arr
        HYB_CHECK(hyb_trsm_fwd(st, c->Hf, ldU, U, A, ldU, m));
        result = A;
    } else {  // hess_prod_slow!, factor-based (wsosinterpnonnegative.jl:168-191), column by column
        for (int64_t j = 0; j < m; ++j) {
            for (int k = 0; k < c->nPs; ++k) {
                HYB_CHECK(wsos_partial_lambda(st, c, k, A + j * ldU));
                coldot_kernel<<<nb1(U, 8), 256, 0, st>>>(c->F[k], c->LU[k], c->L[k], U, c->ldL[k], 1.0, k == 0 ? 0.0 : 1.0,
                                                        B + j * ldU);
                DLAUNCHED();
                HYB_CUDA(cudaGetLastError());
            }
        }
        result = B;
    }
    copy_mat_kernel<<<nb1(U * m), 256, 0, st>>>(result, ldU, prod, ldp, U, m);
    DLAUNCHED();
    HYB_CUDA(cudaGetLastError());
    return 0;
}

int hybd_wsos_dder3(cudaStream_t st, HybDenseWsos* c, const double* dir, double* out) {
    const int64_t U = c->U;
    lincomb_kernel<<<nb1(U), 256, 0, st>>>(U, 1.0, dir, 0.0, nullptr, c->v1);
    DLAUNCHED();
    for (int k = 0; k < c->nPs; ++k) {
        HYB_CHECK(wsos_partial_lambda(st, c, k, c->v1));
        coldot_kernel<<<nb1(U, 8), 256, 0, st>>>(c->LU[k], c->LU[k], c->L[k], U, c->ldL[k], 1.0, k == 0 ? 0.0 : 1.0, c->v2);
        DLAUNCHED();
        HYB_CUDA(cudaGetLastError());
    }
    lincomb_kernel<<<nb1(U), 256, 0, st>>>(U, 1.0, c->v2, 0.0, nullptr, out);
    DLAUNCHED();
    HYB_CUDA(cudaGetLastError());
    return 0;
}

// =======================================================================================
// WSOSInterpPosSemidefTri: src/Cones/wsosinterppossemideftri.jl
//   point = R(R+1)/2 blocks of length U in svec order of (p, q), p >= q (off-diagonal blocks
//   carry the sqrt2 scaling)                                                 :63-123
//   update_feas :125-158   Lambda_k (RL x RL) block (p,q) = P_k' diag(s_pq) P_k, Cholesky
//   is_dual_feas :160-169  diagonal blocks of the dual point > eps
//   update_grad :171-220   F_k = C_k^-T kron(I_R, P_k')  (Lambda_k = C_k' C_k), block_diag_prod!
//   update_hess :222-270   block Hadamard products of PLiP = F_k' F_k
//   hess_prod_slow! / dder3 = partial_prod! :272-358 ; block_diag_prod! :289-312
//   generic Hessian-based products and the slow switch as for WSOSInterpNonnegative
// =======================================================================================
namespace {

__device__ __forceinline__ void blk_pq(int b, int& p, int& q) {
    int pp = (int)((sqrt(8.0 * b + 1.0) - 1.0) * 0.5);
    while (pp * (pp + 1) / 2 > b) --pp;
    while ((pp + 1) * (pp + 2) / 2 <= b) ++pp;
    p = pp;
    q = b - pp * (pp + 1) / 2;
}

// out[r, c] = fac * d[r] * in[r, c]
__global__ void rowscale_fac_kernel(const double* __restrict__ in, const double* __restrict__ d, double fac, int64_t rows,
                                    int64_t cols, int64_t ld, double* __restrict__ out) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= rows * cols) return;
    const int64_t r = e % rows, c = e / rows;
    out[r + c * ld] = fac * d[r] * in[r + c * ld];
}

// block_diag_prod! (wsosinterppossemideftri.jl:289-312): one warp per (block (j,i), u):
// out[blk(j,i) * U + u] = beta * out + alpha * scale_ij * <M1[:, i*U+u], M2[:, j*U+u]>
__global__ void blockdiag_dot_kernel(const double* __restrict__ M1, const double* __restrict__ M2, int64_t rows, int64_t U,
                                     int R, int64_t ld, double alpha, double beta, double* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t item = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int nblk = R * (R + 1) / 2;
    if (item >= (int64_t)nblk * U) return;
    const int b = (int)(item / U);
    const int64_t u = item % U;
    int j, i;
    blk_pq(b, j, i);  // j >= i
    const double* a = M1 + ((int64_t)i * U + u) * ld;
    const double* bb = M2 + ((int64_t)j * U + u) * ld;
    double s = 0.0;
    for (int64_t r = lane; r < rows; r += 32) s += a[r] * bb[r];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) {
        const double v = alpha * ((i != j) ? RT2 : 1.0) * s;
        out[(int64_t)b * U + u] = (beta == 0.0 ? 0.0 : beta * out[(int64_t)b * U + u]) + v;
    }
}

// update_hess (wsosinterppossemideftri.jl:245-266): H[block, block2] (+)= PLP[p,p2] .* PLP[q,q2] * scal
//   (+ PLP[p,q2] .* PLP[q,p2] when both blocks are off-diagonal), for block2 >= block
__global__ void wsospsd_hess_kernel(const double* __restrict__ PLP, int64_t ldP, int64_t U, int R, double* __restrict__ H,
                                    int64_t ldH, int first) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int block = blockIdx.y, block2 = blockIdx.z;
    if (e >= U * U || block2 < block) return;
    const int64_t i = e % U, j = e / U;
    int p, q, p2, q2;
    blk_pq(block, p, q);
    blk_pq(block2, p2, q2);
    const double scal = ((p == q) != (p2 == q2)) ? RT2 : 1.0;
    auto el = [&](int a, int b) { return PLP[(int64_t)a * U + i + ((int64_t)b * U + j) * ldP]; };
    double v = el(p, p2) * el(q, q2) * scal;
    if (p != q && p2 != q2) v += el(p, q2) * el(q, p2);
    double* h = H + (int64_t)block * U + i + ((int64_t)block2 * U + j) * ldH;
    *h = (first ? 0.0 : *h) + v;
}

// dual feasibility: the diagonal blocks of the dual point are positive
__global__ void __launch_bounds__(1024) wsospsd_dualfeas_kernel(const double* __restrict__ z, int64_t U, int R, int* out) {
    __shared__ int ok;
    if (threadIdx.x == 0) ok = 1;
    __syncthreads();
    int mine = 1;
    for (int p = 0; p < R; ++p) {
        const int64_t base = (int64_t)(p * (p + 1) / 2 + p) * U;
        for (int64_t i = threadIdx.x; i < U; i += 1024)
            if (!(z[base + i] > HYB_EPS)) mine = 0;
    }
    if (!mine) ok = 0;
    __syncthreads();
    if (threadIdx.x == 0) *out = ok;
}

}  // namespace

struct HybDenseWsosPsd {
    int R = 1;
    int64_t U = 0, dim = 0, ldD = 0, ldU = 0, RU = 0, ldRU = 0;
    int nPs = 0;
    double nu = 0.0;
    std::vector<int64_t> L, RL, ldRL;
    std::vector<double*> P, Pt, F, Lam, M, Mt, LRUR;  // per k
    std::vector<HybChol> chLam;
    double *PD = nullptr, *PLP = nullptr;
    double *H = nullptr, *Hf = nullptr, *Lh = nullptr;
    HybChol chH;
    double *v1 = nullptr, *v2 = nullptr, *v3 = nullptr, *scal = nullptr;
    int *dual_ok = nullptr, *slow = nullptr, *sqrt_ok = nullptr;
    const int** lam_infos_dev = nullptr;
    double *big1 = nullptr, *big2 = nullptr;
    int64_t big_cols = 0;
};

int hybd_wsospsd_create(HybDenseWsosPsd** out, int R, int64_t U, int nPs, const double* const* Ps_host,
                        const int64_t* Ls) {
    HybDenseWsosPsd* c = new HybDenseWsosPsd();
    c->R = R;
    c->U = U;
    c->ldU = hyb_round_up(U, 16);
    c->dim = U * (R * (R + 1) / 2);
    c->ldD = hyb_round_up(c->dim, 16);
    c->RU = (int64_t)R * U;
    c->ldRU = hyb_round_up(c->RU, 16);
    c->nPs = nPs;
    c->chLam.resize(nPs);
    int64_t Lmax = 1;
    for (int k = 0; k < nPs; ++k) {
        const int64_t L = Ls[k], RL = (int64_t)R * L, ldRL = hyb_round_up(RL, 16), ldL = hyb_round_up(L, 16);
        c->L.push_back(L);
        c->RL.push_back(RL);
        c->ldRL.push_back(ldRL);
        c->nu += (double)RL;
        Lmax = L > Lmax ? L : Lmax;
        double *P = nullptr, *Pt = nullptr, *F = nullptr, *Lam = nullptr, *M = nullptr, *Mt = nullptr, *LRUR = nullptr;
        HYB_CUDA(cudaMalloc(&P, sizeof(double) * c->ldU * L));
        HYB_CUDA(cudaMemset(P, 0, sizeof(double) * c->ldU * L));
        HYB_CUDA(cudaMemcpy2D(P, sizeof(double) * c->ldU, Ps_host[k], sizeof(double) * U, sizeof(double) * U, L,
                              cudaMemcpyHostToDevice));
        HYB_CUDA(cudaMalloc(&Pt, sizeof(double) * ldL * U));
        HYB_CUDA(cudaMemset(Pt, 0, sizeof(double) * ldL * U));
        HYB_CHECK(hyb_transpose(nullptr, P, c->ldU, U, L, Pt, ldL, 0));
        HYB_CUDA(cudaMalloc(&F, sizeof(double) * ldRL * c->RU));
        HYB_CUDA(cudaMalloc(&LRUR, sizeof(double) * ldRL * c->RU));
        HYB_CUDA(cudaMalloc(&Lam, sizeof(double) * ldRL * RL));
        HYB_CUDA(cudaMemset(Lam, 0, sizeof(double) * ldRL * RL));
        HYB_CUDA(cudaMalloc(&M, sizeof(double) * ldRL * RL));
        HYB_CUDA(cudaMalloc(&Mt, sizeof(double) * ldRL * RL));
        HYB_CUDA(cudaMemset(M, 0, sizeof(double) * ldRL * RL));
        c->P.push_back(P);
        c->Pt.push_back(Pt);
        c->F.push_back(F);
        c->Lam.push_back(Lam);
        c->M.push_back(M);
        c->Mt.push_back(Mt);
        c->LRUR.push_back(LRUR);
        HYB_CHECK(hyb_chol_attach(c->chLam[k], Lam, ldRL, RL));
    }
    HYB_CUDA(cudaMalloc(&c->PD, sizeof(double) * c->ldU * Lmax));
    HYB_CUDA(cudaMalloc(&c->PLP, sizeof(double) * c->ldRU * c->RU));
    const size_t hb = sizeof(double) * c->ldD * c->dim;
    double** mats[] = {&c->H, &c->Hf, &c->Lh};
    for (double** m : mats) {
        HYB_CUDA(cudaMalloc(m, hb));
        HYB_CUDA(cudaMemset(*m, 0, hb));
    }
    HYB_CHECK(hyb_chol_attach(c->chH, c->Hf, c->ldD, c->dim));
    HYB_CUDA(cudaMalloc(&c->v1, sizeof(double) * c->ldD));
    HYB_CUDA(cudaMalloc(&c->v2, sizeof(double) * c->ldD));
    HYB_CUDA(cudaMalloc(&c->v3, sizeof(double) * c->ldD));
    HYB_CUDA(cudaMalloc(&c->scal, sizeof(double) * 8));
    HYB_CUDA(cudaMemset(c->scal, 0, sizeof(double) * 8));
    HYB_CUDA(cudaMalloc(&c->dual_ok, sizeof(int)));
    HYB_CUDA(cudaMalloc(&c->slow, sizeof(int)));
    HYB_CUDA(cudaMalloc(&c->sqrt_ok, sizeof(int)));
    HYB_CUDA(cudaMemset(c->slow, 0, sizeof(int)));
    HYB_CUDA(cudaMemset(c->sqrt_ok, 0, sizeof(int)));
    std::vector<const int*> infos;
    for (int k = 0; k < nPs; ++k) infos.push_back(c->chLam[k].info);
    HYB_CUDA(cudaMalloc((void**)&c->lam_infos_dev, sizeof(int*) * (nPs > 0 ? nPs : 1)));
    HYB_CUDA(cudaMemcpy((void*)c->lam_infos_dev, infos.data(), sizeof(int*) * nPs, cudaMemcpyHostToDevice));
    HYB_CUDA(cudaDeviceSynchronize());
    *out = c;
    return 0;
}

void hybd_wsospsd_destroy(HybDenseWsosPsd* c) {
    if (!c) return;
    for (int k = 0; k < c->nPs; ++k) {
        hyb_chol_free(c->chLam[k]);
        cudaFree(c->P[k]);
        cudaFree(c->Pt[k]);
        cudaFree(c->F[k]);
        cudaFree(c->Lam[k]);
        cudaFree(c->M[k]);
        cudaFree(c->Mt[k]);
        cudaFree(c->LRUR[k]);
    }
    hyb_chol_free(c->chH);
    cudaFree(c->PD);
    cudaFree(c->PLP);
    cudaFree(c->H);
    cudaFree(c->Hf);
    cudaFree(c->Lh);
    cudaFree(c->v1);
    cudaFree(c->v2);
    cudaFree(c->v3);
    cudaFree(c->scal);
    cudaFree(c->dual_ok);
    cudaFree(c->slow);
    cudaFree(c->sqrt_ok);
    cudaFree((void*)c->lam_infos_dev);
    cudaFree(c->big1);
    cudaFree(c->big2);
    delete c;
}

// dst (RL x RL, upper blocks then symmetrized if asked) : block (q,p), q <= p, = P_k' diag(vec_pq') P_k
static int wsospsd_build_lambda(cudaStream_t st, HybDenseWsosPsd* c, int k, const double* vec, double* dst, bool full) {
    const int64_t U = c->U, L = c->L[k], ldU = c->ldU, ldRL = c->ldRL[k];
    for (int p = 0; p < c->R; ++p)
        for (int q = 0; q <= p; ++q) {
            const double* d = vec + (int64_t)(p * (p + 1) / 2 + q) * U;
            rowscale_fac_kernel<<<nb1(U * L), 256, 0, st>>>(c->P[k], d, (p != q) ? 1.0 / RT2 : 1.0, U, L, ldU, c->PD);
            DLAUNCHED();
            HYB_CUDA(cudaGetLastError());
            HYB_CHECK(hyb_gemm_tn(st, L, L, U, 1.0, c->P[k], ldU, c->PD, ldU, 0.0, dst + (int64_t)q * L + (int64_t)p * L * ldRL,
                                  ldRL, 0));
        }
    if (full) {
        symmetrize_kernel<<<nb1(c->RL[k] * c->RL[k]), 256, 0, st>>>(dst, c->RL[k], ldRL);
        DLAUNCHED();
        HYB_CUDA(cudaGetLastError());
    }
    return 0;
}

int hybd_wsospsd_update(cudaStream_t st, HybDenseWsosPsd* c, const double* primal, const double* dual_in, double scal,
                        double irtmu, double* point, double* dual_out, double* grad, int* flag, double* prox) {
    const int64_t U = c->U, dim = c->dim, RU = c->RU, ldRU = c->ldRU, ldD = c->ldD;
    const int R = c->R;
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, scal, primal, 0.0, nullptr, point);
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, 1.0, dual_in, 0.0, nullptr, dual_out);
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, 1.0, point, 0.0, nullptr, c->v3);
    wsospsd_dualfeas_kernel<<<1, 1024, 0, st>>>(dual_out, U, R, c->dual_ok);
    g_dense_launches += 4;
    HYB_CUDA(cudaGetLastError());
    for (int k = 0; k < c->nPs; ++k) {
        const int64_t L = c->L[k], RL = c->RL[k], ldRL = c->ldRL[k], ldL = hyb_round_up(L, 16);
        HYB_CHECK(wsospsd_build_lambda(st, c, k, c->v3, c->Lam[k], false));
        HYB_CHECK(hyb_potrf(st, c->chLam[k], nullptr));
        // F_k = C_k^-T kron(I_R, P_k')
        HYB_CUDA(cudaMemsetAsync(c->F[k], 0, sizeof(double) * ldRL * RU, st));
        for (int p = 0; p < R; ++p) {
            copy_mat_kernel<<<nb1(L * U), 256, 0, st>>>(c->Pt[k], ldL, c->F[k] + (int64_t)p * L + (int64_t)p * U * ldRL, ldRL, L, U);
            DLAUNCHED();
        }
        HYB_CUDA(cudaGetLastError());
        HYB_CHECK(hyb_trsm_fwd(st, c->Lam[k], ldRL, RL, c->F[k], ldRL, RU));
        // grad -= block_diag_prod(F, F)
        blockdiag_dot_kernel<<<nb1((int64_t)(R * (R + 1) / 2) * U, 8), 256, 0, st>>>(c->F[k], c->F[k], RL, U, R, ldRL, -1.0,
                                                                                     k == 0 ? 0.0 : 1.0, grad);
        DLAUNCHED();
        // PLiP = F' F (full symmetric), then the block Hadamard update of H
        HYB_CHECK(hyb_gemm_tn(st, RU, RU, RL, 1.0, c->F[k], ldRL, c->F[k], ldRL, 0.0, c->PLP, ldRU, HYB_GEMM_UPPER));
        symmetrize_kernel<<<nb1(RU * RU), 256, 0, st>>>(c->PLP, RU, ldRU);
        const int nblk = R * (R + 1) / 2;
        wsospsd_hess_kernel<<<dim3(nb1(U * U), nblk, nblk), 256, 0, st>>>(c->PLP, ldRU, U, R, c->H, ldD, k == 0);
        g_dense_launches += 2;
        HYB_CUDA(cudaGetLastError());
    }
    symmetrize_kernel<<<nb1(dim * dim), 256, 0, st>>>(c->H, dim, ldD);
    copy_upper_dense_kernel<<<nb1(dim * dim), 256, 0, st>>>(c->H, c->Hf, dim, ldD);
    g_dense_launches += 2;
    HYB_CHECK(hyb_potrf(st, c->chH, nullptr));
    HYB_CHECK(hyb_transpose(st, c->Hf, ldD, dim, dim, c->Lh, ldD, 1));
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, 1.0, grad, 0.0, nullptr, c->v1);
    dot_kernel<<<1, 1024, 0, st>>>(dim, c->v1, c->v3, c->scal + 0);
    HYB_CHECK(hyb_potrs(st, c->chH, c->v1));
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, 1.0, grad, 0.0, nullptr, c->v2);
    dot_kernel<<<1, 1024, 0, st>>>(dim, c->v1, c->v2, c->scal + 1);
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, irtmu, dual_out, 1.0, grad, c->v2);
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, 1.0, c->v2, 0.0, nullptr, c->v1);
    HYB_CHECK(hyb_potrs(st, c->chH, c->v1));
    dot_kernel<<<1, 1024, 0, st>>>(dim, c->v1, c->v2, c->scal + 2);
    HYB_CHECK(hyb_gemv_t(st, dim, dim, c->H, ldD, c->v3, c->v1));
    dot_kernel<<<1, 1024, 0, st>>>(dim, c->v1, c->v3, c->scal + 3);
    wsos_finalize_kernel<<<1, 1, 0, st>>>(c->lam_infos_dev, c->nPs, c->chH.info, c->dual_ok, c->scal, c->nu, dim, flag, prox,
                                          c->slow, c->sqrt_ok);
    g_dense_launches += 10;
    HYB_CUDA(cudaGetLastError());
    return 0;
}

int hybd_wsospsd_sqrt_ok(cudaStream_t st, HybDenseWsosPsd* c, int* ok) {
    HYB_CUDA(cudaMemcpyAsync(ok, c->sqrt_ok, sizeof(int), cudaMemcpyDeviceToHost, st));
    HYB_CUDA(cudaStreamSynchronize(st));
    return 0;
}

// partial_prod! (wsosinterppossemideftri.jl:314-358) for one direction d (aligned device dim-vector):
// out (+)= block_diag_prod(left, LRUR) with LRUR = (C^-T Lambda(d) C^-1) F, left = LRUR (dder3) or F (slow product)
static int wsospsd_partial_prod(cudaStream_t st, HybDenseWsosPsd* c, const double* d, bool symm, double* out) {
    const int R = c->R;
    const int64_t U = c->U, RU = c->RU;
    for (int k = 0; k < c->nPs; ++k) {
        const int64_t RL = c->RL[k], ldRL = c->ldRL[k];
        HYB_CHECK(wsospsd_build_lambda(st, c, k, d, c->M[k], true));
        HYB_CHECK(hyb_trsm_fwd(st, c->Lam[k], ldRL, RL, c->M[k], ldRL, RL));
        HYB_CHECK(hyb_transpose(st, c->M[k], ldRL, RL, RL, c->Mt[k], ldRL, 0));
        HYB_CHECK(hyb_trsm_fwd(st, c->Lam[k], ldRL, RL, c->Mt[k], ldRL, RL));
        HYB_CHECK(hyb_gemm_tn(st, RL, RU, RL, 1.0, c->Mt[k], ldRL, c->F[k], ldRL, 0.0, c->LRUR[k], ldRL, 0));
        blockdiag_dot_kernel<<<nb1((int64_t)(R * (R + 1) / 2) * U, 8), 256, 0, st>>>(
            symm ? c->LRUR[k] : c->F[k], c->LRUR[k], RL, U, R, ldRL, 1.0, k == 0 ? 0.0 : 1.0, out);
        DLAUNCHED();
        HYB_CUDA(cudaGetLastError());
    }
    return 0;
}

int hybd_wsospsd_prod(cudaStream_t st, HybDenseWsosPsd* c, int op, const double* arr, int64_t lda, double* prod,
                      int64_t ldp, int64_t m) {
    if (m <= 0) return 0;
    const int64_t dim = c->dim, ldD = c->ldD;
    if (m > c->big_cols) {
        cudaFree(c->big1);
        cudaFree(c->big2);
        c->big1 = c->big2 = nullptr;
        HYB_CUDA(cudaMalloc(&c->big1, sizeof(double) * ldD * m));
        HYB_CUDA(cudaMalloc(&c->big2, sizeof(double) * ldD * m));
        c->big_cols = m;
    }
    double *A = c->big1, *B = c->big2;
    copy_mat_kernel<<<nb1(dim * m), 256, 0, st>>>(arr, lda, A, ldD, dim, m);
    DLAUNCHED();
    HYB_CUDA(cudaGetLastError());
    const double* result = nullptr;
    if (op == 6) {
        int slow = 0;
        HYB_CUDA(cudaMemcpyAsync(&slow, c->slow, sizeof(int), cudaMemcpyDeviceToHost, st));
        HYB_CUDA(cudaStreamSynchronize(st));
        if (!slow) op = HYB_OP_HESS;
    }
    if (op == HYB_OP_HESS) {
        if (m == 1) HYB_CHECK(hyb_gemv_t(st, dim, dim, c->H, ldD, A, B));
        else HYB_CHECK(hyb_gemm_tn(st, dim, m, dim, 1.0, c->H, ldD, A, ldD, 0.0, B, ldD, 0));
        result = B;
    } else if (op == HYB_OP_INV_HESS) {
        if (m == 1) {
            HYB_CHECK(hyb_potrs(st, c->chH, A));
        } else {
            HYB_CHECK(hyb_trsm_fwd(st, c->Hf, ldD, dim, A, ldD, m));
            HYB_CHECK(hyb_trsm_bwd(st, c->Hf, c->Lh, ldD, dim, A, ldD, m));
        }
        result = A;
    } else if (op == HYB_OP_SQRT_HESS) {
        if (m == 1) HYB_CHECK(hyb_gemv_t(st, dim, dim, c->Lh, ldD, A, B));
        else HYB_CHECK(hyb_gemm_tn(st, dim, m, dim, 1.0, c->Lh, ldD, A, ldD, 0.0, B, ldD, 0));
        result = B;
    } else if (op == HYB_OP_INV_SQRT_HESS) {
        HYB_CHECK(hyb_trsm_fwd(st, c->Hf, ldD, dim, A, ldD, m));
        result = A;
    } else {
        for (int64_t j = 0; j < m; ++j) HYB_CHECK(wsospsd_partial_prod(st, c, A + j * ldD, false, B + j * ldD));
        result = B;
    }
    copy_mat_kernel<<<nb1(dim * m), 256, 0, st>>>(result, ldD, prod, ldp, dim, m);
    DLAUNCHED();
    HYB_CUDA(cudaGetLastError());
    return 0;
}

int hybd_wsospsd_dder3(cudaStream_t st, HybDenseWsosPsd* c, const double* dir, double* out) {
    const int64_t dim = c->dim;
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, 1.0, dir, 0.0, nullptr, c->v1);
    DLAUNCHED();
    HYB_CHECK(wsospsd_partial_prod(st, c, c->v1, true, c->v2));
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, 1.0, c->v2, 0.0, nullptr, out);
    DLAUNCHED();
    HYB_CUDA(cudaGetLastError());
    return 0;
}

// =======================================================================================
// DoublyNonnegativeTri: src/Cones/doublynonnegativetri.jl  (barrier -logdet(S) - sum log(off-diagonal s))
//   update_feas :134-147   all(point > eps), then the Cholesky of smat(point)
//   is_dual_feas           true (Cones.jl:71)
//   update_grad :149-160   -svec(S^-1) - 1 ./ s_offdiag
//   update_hess :162-175   symm_kron(S^-1) (arrayutilities.jl:287-324) + diag(1 ./ s_offdiag.^2)
//   dder3       :198-212   svec(T T'), T = S^-1 D U^-1, plus (d ./ s).^2 ./ s on the off-diagonal entries
//   every Hessian product goes through the stored Hessian and its Cholesky (generic oracles of Cones.jl:
//   108-125, 202-214, 233-252); hess_prod!(:177-196) is the same linear map evaluated from S^-1 instead.
// Same eager evaluation at load time and the same engine as for the WSOS cones.
// =======================================================================================
namespace {

__global__ void dnn_grad_offdiag_kernel(int64_t dim, const double* __restrict__ point, double* __restrict__ grad) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= dim) return;
    int r, c;
    svec_rc(k, r, c);
    if (r != c) grad[k] -= 1.0 / point[k];
}

// H[a, b] for svec positions a = (i, j), b = (k, l), i <= j, k <= l (full symmetric matrix, both triangles)
__global__ void dnn_hess_kernel(int64_t dim, const double* __restrict__ Xi, int64_t ld, const double* __restrict__ point,
                                double* __restrict__ H, int64_t ldH) {
    const int64_t a = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t b = blockIdx.y;
    if (a >= dim) return;
    int i, j, k, l;
    svec_rc(a, i, j);
    svec_rc(b, k, l);
    double v;
    if (k == l) {
        v = (i == j) ? Xi[j + (int64_t)l * ld] * Xi[j + (int64_t)l * ld] : RT2 * Xi[i + (int64_t)l * ld] * Xi[j + (int64_t)l * ld];
    } else {
        v = (i == j) ? RT2 * Xi[j + (int64_t)k * ld] * Xi[j + (int64_t)l * ld]
                     : Xi[i + (int64_t)k * ld] * Xi[j + (int64_t)l * ld] + Xi[i + (int64_t)l * ld] * Xi[j + (int64_t)k * ld];
    }
    if (a == b && i != j) {
        const double inv = 1.0 / point[a];
        v += inv * inv;
    }
    H[a + b * ldH] = v;
}

__global__ void dnn_dder3_offdiag_kernel(int64_t dim, const double* __restrict__ point, const double* __restrict__ dir,
                                         double* __restrict__ out) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= dim) return;
    int r, c;
    svec_rc(k, r, c);
    if (r != c) {
        const double t = dir[k] / point[k];
        out[k] += t * t / point[k];
    }
}

// ---- PosSemidefTriSparse, dense implementation (possemideftrisparse/denseimpl.jl, real case): the same engine with
// pattern-indexed packing (svec_to_smat_sparse! :216-231, smat_to_svec_sparse! :251-267), the pattern-restricted
// Hessian (update_hess :61-88) and the hess_prod_slow! switch (:142-163, Cones.jl:218-227)
__global__ void spsd_unpack_kernel(int64_t dim, const double* __restrict__ v, int64_t vstride, const int* __restrict__ ri,
                                   const int* __restrict__ ci, double* __restrict__ M, int64_t ld, int64_t blk_stride) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t j = blockIdx.y;
    if (k >= dim) return;
    const int r = ri[k], c = ci[k];
    const double x = v[k + j * vstride] * ((r == c) ? 1.0 : (1.0 / RT2));
    double* B = M + j * blk_stride;
    B[r + (int64_t)c * ld] = x;
    B[c + (int64_t)r * ld] = x;
}
__global__ void spsd_pack_kernel(int64_t dim, const double* __restrict__ M, int64_t ld, int64_t blk_stride,
                                 const int* __restrict__ ri, const int* __restrict__ ci, double* __restrict__ v,
                                 int64_t vstride, double scale) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t j = blockIdx.y;
    if (k >= dim) return;
    const int r = ri[k], c = ci[k];
    const double x = M[j * blk_stride + r + (int64_t)c * ld];
    v[k + j * vstride] = scale * ((r == c) ? x : x * RT2);
}
__global__ void spsd_hess_kernel(int64_t dim, const double* __restrict__ Xi, int64_t ld, const int* __restrict__ ri,
                                 const int* __restrict__ ci, double* __restrict__ H, int64_t ldH) {
    const int64_t a = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t b = blockIdx.y;
    if (a >= dim) return;
    const int i1 = ri[a], j1 = ci[a], i2 = ri[b], j2 = ci[b];
    const bool d1 = (i1 == j1), d2 = (i2 == j2);
    double v;
    if (d1 && d2) v = Xi[i1 + (int64_t)i2 * ld] * Xi[i1 + (int64_t)i2 * ld];
    else if (d1 != d2) v = RT2 * Xi[i1 + (int64_t)i2 * ld] * Xi[j1 + (int64_t)j2 * ld];
    else v = Xi[i1 + (int64_t)i2 * ld] * Xi[j1 + (int64_t)j2 * ld] + Xi[i1 + (int64_t)j2 * ld] * Xi[j1 + (int64_t)i2 * ld];
    H[a + b * ldH] = v;
}
__global__ void set_int_kernel(int* p, int v) { *p = v; }

// scal: [0] <g,p>  [1] <Hinv g, g>  [2] psi' Hinv psi  [3] p' H p
__global__ void dnn_finalize_kernel(const int* pt_ok, const int* info_x, const int* info_h, const double* scal, double nu,
                                    int64_t dim, int with_slow, int* flag, double* prox, int* slow, int* sqrt_ok) {
    int f = HYB_ST_DUAL_FEAS;
    double p = 0.0;
    const bool feas = (*pt_ok != 0) && (*info_x == 0);
    *sqrt_ok = (feas && *info_h == 0) ? 1 : 0;
    *slow = 0;
    if (feas) {
        f |= HYB_ST_FEAS;
        bool ok = (*info_h == 0);
        if (fabs(1.0 + scal[0] / nu) > 1.220703125e-4 * (double)dim) ok = false;
        if (fabs(1.0 - scal[1] / nu) > 1.104854345603981e-1 * (double)dim) ok = false;
        if (ok) f |= HYB_ST_NUMERICS;
        p = (scal[2] < -1.4901161193847656e-08 * (double)dim) ? INFINITY : fabs(scal[2]);
        // update_use_hess_prod_slow (Cones.jl:218-227)
        if (with_slow && fabs(1.0 - scal[3] / nu) > (double)dim * 1.4901161193847656e-08) *slow = 1;
    }
    *flag = f;
    *prox = p;
}

}  // namespace

struct HybDenseDnn {
    int s = 0;
    int64_t dim = 0, ld = 0, ldD = 0;
    double nu = 0.0;
    double *X = nullptr, *Ux = nullptr, *Lx = nullptr, *Xinv = nullptr, *W2 = nullptr, *W3 = nullptr;  // s x s
    HybChol chX;
    double *H = nullptr, *Hf = nullptr, *Lh = nullptr;  // dim x dim
    HybChol chH;
    double *v1 = nullptr, *v2 = nullptr, *v3 = nullptr, *scal = nullptr;
    int *pt_ok = nullptr, *slow = nullptr, *sqrt_ok = nullptr;
    double *big1 = nullptr, *big2 = nullptr;
    int64_t big_cols = 0;
    // PosSemidefTriSparse: pattern positions (0-based, row >= col) of the point's coordinates
    bool sparse = false;
    int *ri = nullptr, *ci = nullptr;
    double *sbig1 = nullptr, *sbig2 = nullptr;  // side x side blocks for the factor-based hess_prod_slow!
    int64_t sbig_cols = 0;
};

// vector (dim) -> full symmetric side x side matrix, and back (scaled)
static int dnn_unpack(cudaStream_t st, HybDenseDnn* c, const double* v, double* M) {
    if (c->sparse) {
        HYB_CUDA(cudaMemsetAsync(M, 0, sizeof(double) * c->ld * c->s, st));
        spsd_unpack_kernel<<<dim3(nb1(c->dim), 1), 256, 0, st>>>(c->dim, v, 0, c->ri, c->ci, M, c->ld, 0);
    } else {
        svec_unpack_kernel<<<dim3(nb1(c->dim), 1), 256, 0, st>>>(v, 0, c->s, M, c->ld, 0, 1.0);
    }
    DLAUNCHED();
    HYB_CUDA(cudaGetLastError());
    return 0;
}
static int dnn_pack(cudaStream_t st, HybDenseDnn* c, const double* M, double* v, double scale) {
    if (c->sparse) spsd_pack_kernel<<<dim3(nb1(c->dim), 1), 256, 0, st>>>(c->dim, M, c->ld, 0, c->ri, c->ci, v, 0, scale);
    else svec_pack_kernel<<<dim3(nb1(c->dim), 1), 256, 0, st>>>(M, c->ld, 0, c->s, v, 0, scale);
    DLAUNCHED();
    HYB_CUDA(cudaGetLastError());
    return 0;
}

static int dnn_ensure_big(HybDenseDnn* c, int64_t m) {
    if (m <= c->big_cols) return 0;
    cudaFree(c->big1);
    cudaFree(c->big2);
    c->big1 = c->big2 = nullptr;
    HYB_CUDA(cudaMalloc(&c->big1, sizeof(double) * c->ldD * m));
    HYB_CUDA(cudaMalloc(&c->big2, sizeof(double) * c->ldD * m));
    c->big_cols = m;
    return 0;
}

static int dnn_create_common(HybDenseDnn** out, int side, int64_t dim, double nu, const int64_t* rows_cols);

int hybd_dnn_create(HybDenseDnn** out, int side) {
    return dnn_create_common(out, side, (int64_t)side * (side + 1) / 2, (double)((int64_t)side * (side + 1) / 2), nullptr);
}
// PosSemidefTriSparse: rows_cols = dim 0-based rows followed by dim 0-based columns of the pattern (row >= col)
int hybd_spsd_create(HybDenseDnn** out, int side, int64_t dim, const int64_t* rows_cols) {
    if (!rows_cols) {
        hyb_set_error("PosSemidefTriSparse needs its pattern (hyb_cone_desc.Ls)");
        return -3;
    }
    return dnn_create_common(out, side, dim, (double)side, rows_cols);
}

static int dnn_create_common(HybDenseDnn** out, int side, int64_t dim, double nu, const int64_t* rows_cols) {
    HybDenseDnn* c = new HybDenseDnn();
    c->s = side;
    c->dim = dim;
    c->nu = nu;
    c->ld = hyb_round_up(side, 16);
    c->ldD = hyb_round_up(c->dim, 16);
    const size_t sb = sizeof(double) * c->ld * side;
    double** smats[] = {&c->X, &c->Ux, &c->Lx, &c->Xinv, &c->W2, &c->W3};
    for (double** m : smats) {
        HYB_CUDA(cudaMalloc(m, sb));
        HYB_CUDA(cudaMemset(*m, 0, sb));
    }
    const size_t db = sizeof(double) * c->ldD * c->dim;
    double** dmats[] = {&c->H, &c->Hf, &c->Lh};
    for (double** m : dmats) {
        HYB_CUDA(cudaMalloc(m, db));
        HYB_CUDA(cudaMemset(*m, 0, db));
    }
    double** vecs[] = {&c->v1, &c->v2, &c->v3};
    for (double** v : vecs) HYB_CUDA(cudaMalloc(v, sizeof(double) * c->ldD));
    HYB_CUDA(cudaMalloc(&c->scal, sizeof(double) * 8));
    HYB_CUDA(cudaMemset(c->scal, 0, sizeof(double) * 8));
    HYB_CUDA(cudaMalloc(&c->pt_ok, sizeof(int)));
    HYB_CUDA(cudaMalloc(&c->slow, sizeof(int)));
    HYB_CUDA(cudaMalloc(&c->sqrt_ok, sizeof(int)));
    HYB_CUDA(cudaMemset(c->sqrt_ok, 0, sizeof(int)));
    HYB_CHECK(hyb_chol_attach(c->chX, c->Ux, c->ld, side));
    HYB_CHECK(hyb_chol_attach(c->chH, c->Hf, c->ldD, c->dim));
    if (rows_cols) {
        c->sparse = true;
        std::vector<int> r(dim), cc(dim);
        for (int64_t k = 0; k < dim; ++k) {
            r[k] = (int)rows_cols[k];
            cc[k] = (int)rows_cols[dim + k];
            if (cc[k] < 0 || cc[k] > r[k] || r[k] >= side) {
                hyb_set_error("PosSemidefTriSparse: pattern entry %lld = (%d, %d) outside the lower triangle of side %d",
                              (long long)k, r[k], cc[k], side);
                return -3;
            }
        }
        HYB_CUDA(cudaMalloc(&c->ri, sizeof(int) * dim));
        HYB_CUDA(cudaMalloc(&c->ci, sizeof(int) * dim));
        HYB_CUDA(cudaMemcpy(c->ri, r.data(), sizeof(int) * dim, cudaMemcpyHostToDevice));
        HYB_CUDA(cudaMemcpy(c->ci, cc.data(), sizeof(int) * dim, cudaMemcpyHostToDevice));
        set_int_kernel<<<1, 1>>>(c->pt_ok, 1);  // no entrywise condition on the point
        HYB_CUDA(cudaDeviceSynchronize());
    }
    *out = c;
    return 0;
}

void hybd_dnn_destroy(HybDenseDnn* c) {
    if (!c) return;
    cudaFree(c->ri);
    cudaFree(c->ci);
    cudaFree(c->sbig1);
    cudaFree(c->sbig2);
    hyb_chol_free(c->chX);
    hyb_chol_free(c->chH);
    double* ptrs[] = {c->X, c->Ux, c->Lx, c->Xinv, c->W2, c->W3, c->H, c->Hf, c->Lh, c->v1, c->v2, c->v3, c->scal, c->big1, c->big2};
    for (double* p : ptrs) cudaFree(p);
    cudaFree(c->pt_ok);
    cudaFree(c->slow);
    cudaFree(c->sqrt_ok);
    delete c;
}

int hybd_dnn_update(cudaStream_t st, HybDenseDnn* c, const double* primal, const double* dual_in, double scal, double irtmu,
                    double* point, double* dual_out, double* grad, int* flag, double* prox) {
    const int s = c->s;
    const int64_t ld = c->ld, dim = c->dim, ldD = c->ldD;
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, scal, primal, 0.0, nullptr, point);
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, 1.0, dual_in, 0.0, nullptr, dual_out);
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, 1.0, point, 0.0, nullptr, c->v3);  // aligned copy of the point
    if (!c->sparse) all_gt_eps_kernel<<<1, 1024, 0, st>>>(dim, c->v3, c->pt_ok);
    HYB_CHECK(dnn_unpack(st, c, c->v3, c->X));
    copy_block_kernel<<<nb1((int64_t)s * s), 256, 0, st>>>(c->X, c->Ux, ld, s, s);
    g_dense_launches += 5;
    HYB_CUDA(cudaGetLastError());
    int inf_x = 0, pt_ok = 0;
    HYB_CHECK(hyb_potrf(st, c->chX, nullptr));
    HYB_CUDA(cudaMemcpyAsync(&inf_x, c->chX.info, sizeof(int), cudaMemcpyDeviceToHost, st));
    HYB_CUDA(cudaMemcpyAsync(&pt_ok, c->pt_ok, sizeof(int), cudaMemcpyDeviceToHost, st));
    HYB_CUDA(cudaStreamSynchronize(st));
    if (inf_x != 0 || !pt_ok) {  // infeasible candidate: rejected before grad / Hessian are needed (search.jl:146-150)
        dnn_finalize_kernel<<<1, 1, 0, st>>>(c->pt_ok, c->chX.info, c->chH.info, c->scal, c->nu, dim, c->sparse ? 1 : 0, flag,
                                             prox, c->slow, c->sqrt_ok);
        DLAUNCHED();
        HYB_CUDA(cudaGetLastError());
        return 0;
    }
    HYB_CHECK(hyb_transpose(st, c->Ux, ld, s, s, c->Lx, ld, 1));
    set_identity_kernel<<<nb1((int64_t)s * s), 256, 0, st>>>(c->Xinv, ld, s);
    DLAUNCHED();
    HYB_CHECK(hyb_trsm_fwd(st, c->Ux, ld, s, c->Xinv, ld, s));
    HYB_CHECK(hyb_trsm_bwd(st, c->Ux, c->Lx, ld, s, c->Xinv, ld, s));
    HYB_CHECK(dnn_pack(st, c, c->Xinv, c->v1, -1.0));
    if (!c->sparse) dnn_grad_offdiag_kernel<<<nb1(dim), 256, 0, st>>>(dim, c->v3, c->v1);
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, 1.0, c->v1, 0.0, nullptr, grad);
    // Hessian, its Cholesky and the transposed factor for the backward solves
    if (c->sparse) spsd_hess_kernel<<<dim3(nb1(dim), (unsigned)dim), 256, 0, st>>>(dim, c->Xinv, ld, c->ri, c->ci, c->H, ldD);
    else dnn_hess_kernel<<<dim3(nb1(dim), (unsigned)dim), 256, 0, st>>>(dim, c->Xinv, ld, c->v3, c->H, ldD);
    copy_upper_dense_kernel<<<nb1(dim * dim), 256, 0, st>>>(c->H, c->Hf, dim, ldD);
    g_dense_launches += 5;
    HYB_CUDA(cudaGetLastError());
    HYB_CHECK(hyb_potrf(st, c->chH, nullptr));
    HYB_CHECK(hyb_transpose(st, c->Hf, ldD, dim, dim, c->Lh, ldD, 1));
    // scalars for check_numerics / proxsqr (Cones.jl:267-304); v1 still holds the gradient
    dot_kernel<<<1, 1024, 0, st>>>(dim, c->v1, c->v3, c->scal + 0);
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, 1.0, c->v1, 0.0, nullptr, c->v2);  // g
    HYB_CHECK(hyb_potrs(st, c->chH, c->v1));                                         // Hinv g
    dot_kernel<<<1, 1024, 0, st>>>(dim, c->v1, c->v2, c->scal + 1);
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, irtmu, dual_out, 1.0, grad, c->v2);  // psi
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, 1.0, c->v2, 0.0, nullptr, c->v1);
    HYB_CHECK(hyb_potrs(st, c->chH, c->v1));
    dot_kernel<<<1, 1024, 0, st>>>(dim, c->v1, c->v2, c->scal + 2);
    if (c->sparse) {  // p' H p for the hess_prod_slow! switch
        HYB_CHECK(hyb_gemv_t(st, dim, dim, c->H, ldD, c->v3, c->v1));
        dot_kernel<<<1, 1024, 0, st>>>(dim, c->v1, c->v3, c->scal + 3);
        DLAUNCHED();
    }
    dnn_finalize_kernel<<<1, 1, 0, st>>>(c->pt_ok, c->chX.info, c->chH.info, c->scal, c->nu, dim, c->sparse ? 1 : 0, flag, prox,
                                         c->slow, c->sqrt_ok);
    g_dense_launches += 7;
    HYB_CUDA(cudaGetLastError());
    return 0;
}

int hybd_dnn_sqrt_ok(cudaStream_t st, HybDenseDnn* c, int* ok) {
    HYB_CUDA(cudaMemcpyAsync(ok, c->sqrt_ok, sizeof(int), cudaMemcpyDeviceToHost, st));
    HYB_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int hybd_dnn_prod(cudaStream_t st, HybDenseDnn* c, int op, const double* arr, int64_t lda, double* prod, int64_t ldp,
                  int64_t m) {
    if (m <= 0) return 0;
    const int64_t U = c->dim, ldU = c->ldD;
    HYB_CHECK(dnn_ensure_big(c, m));
    double *A = c->big1, *B = c->big2;
    copy_mat_kernel<<<nb1(U * m), 256, 0, st>>>(arr, lda, A, ldU, U, m);
    DLAUNCHED();
    HYB_CUDA(cudaGetLastError());
    const double* result = nullptr;
    if (op == 6 && c->sparse) {  // hess_prod_slow! (denseimpl.jl:142-163): Lambda^-1 A Lambda^-1 through the factor
        int slow = 0;
        HYB_CUDA(cudaMemcpyAsync(&slow, c->slow, sizeof(int), cudaMemcpyDeviceToHost, st));
        HYB_CUDA(cudaStreamSynchronize(st));
        if (getenv("HYB_FORCE_HESS_PROD_SLOW")) slow = 1;  // test hook: take the factor-based branch at a benign point
        if (slow) {
            const int s = c->s;
            const int64_t ld = c->ld, blk = ld * s, cols = (int64_t)s * m;
            if (m > c->sbig_cols) {
                cudaFree(c->sbig1);
                cudaFree(c->sbig2);
                c->sbig1 = c->sbig2 = nullptr;
                HYB_CUDA(cudaMalloc(&c->sbig1, sizeof(double) * blk * m));
                HYB_CUDA(cudaMalloc(&c->sbig2, sizeof(double) * blk * m));
                c->sbig_cols = m;
            }
            double *SA = c->sbig1, *SB = c->sbig2;
            HYB_CUDA(cudaMemsetAsync(SA, 0, sizeof(double) * blk * m, st));
            spsd_unpack_kernel<<<dim3(nb1(U), (unsigned)m), 256, 0, st>>>(U, A, ldU, c->ri, c->ci, SA, ld, blk);
            DLAUNCHED();
            HYB_CHECK(hyb_trsm_fwd(st, c->Ux, ld, s, SA, ld, cols));
            HYB_CHECK(hyb_trsm_bwd(st, c->Ux, c->Lx, ld, s, SA, ld, cols));
            dim3 tgrid((unsigned)((s + 31) / 32), (unsigned)((s + 31) / 32), (unsigned)m);
            block_transpose_kernel<<<tgrid, dim3(32, 8), 0, st>>>(SA, SB, ld, blk, s);
            DLAUNCHED();
            HYB_CHECK(hyb_trsm_fwd(st, c->Ux, ld, s, SB, ld, cols));
            HYB_CHECK(hyb_trsm_bwd(st, c->Ux, c->Lx, ld, s, SB, ld, cols));
            spsd_pack_kernel<<<dim3(nb1(U), (unsigned)m), 256, 0, st>>>(U, SB, ld, blk, c->ri, c->ci, prod, ldp, 1.0);
            DLAUNCHED();
            HYB_CUDA(cudaGetLastError());
            return 0;
        }
    }
    if (op == HYB_OP_HESS || op == 6) {
        if (m == 1) HYB_CHECK(hyb_gemv_t(st, U, U, c->H, ldU, A, B));
        else HYB_CHECK(hyb_gemm_tn(st, U, m, U, 1.0, c->H, ldU, A, ldU, 0.0, B, ldU, 0));
        result = B;
    } else if (op == HYB_OP_INV_HESS) {
        if (m == 1) {
            HYB_CHECK(hyb_potrs(st, c->chH, A));
        } else {
            HYB_CHECK(hyb_trsm_fwd(st, c->Hf, ldU, U, A, ldU, m));
            HYB_CHECK(hyb_trsm_bwd(st, c->Hf, c->Lh, ldU, U, A, ldU, m));
        }
        result = A;
    } else if (op == HYB_OP_SQRT_HESS) {  // chol(H).U * arr = Lh' arr
        if (m == 1) HYB_CHECK(hyb_gemv_t(st, U, U, c->Lh, ldU, A, B));
        else HYB_CHECK(hyb_gemm_tn(st, U, m, U, 1.0, c->Lh, ldU, A, ldU, 0.0, B, ldU, 0));
        result = B;
    } else {  // chol(H).U' \ arr
        HYB_CHECK(hyb_trsm_fwd(st, c->Hf, ldU, U, A, ldU, m));
        result = A;
    }
    copy_mat_kernel<<<nb1(U * m), 256, 0, st>>>(result, ldU, prod, ldp, U, m);
    DLAUNCHED();
    HYB_CUDA(cudaGetLastError());
    return 0;
}

int hybd_dnn_dder3(cudaStream_t st, HybDenseDnn* c, const double* dir, double* out) {
    const int s = c->s;
    const int64_t ld = c->ld, dim = c->dim;
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, 1.0, dir, 0.0, nullptr, c->v1);
    DLAUNCHED();
    HYB_CHECK(dnn_unpack(st, c, c->v1, c->W2));
    HYB_CHECK(hyb_trsm_fwd(st, c->Ux, ld, s, c->W2, ld, s));
    HYB_CHECK(hyb_trsm_bwd(st, c->Ux, c->Lx, ld, s, c->W2, ld, s));  // S^-1 D
    HYB_CHECK(hyb_transpose(st, c->W2, ld, s, s, c->W3, ld, 0));      // D S^-1
    HYB_CHECK(hyb_trsm_fwd(st, c->Ux, ld, s, c->W3, ld, s));          // V = U^-T D S^-1 = T'
    HYB_CHECK(hyb_gemm_tn(st, s, s, s, 1.0, c->W3, ld, c->W3, ld, 0.0, c->W2, ld, 0));
    HYB_CHECK(dnn_pack(st, c, c->W2, c->v2, 1.0));
    if (!c->sparse) dnn_dder3_offdiag_kernel<<<nb1(dim), 256, 0, st>>>(dim, c->v3, c->v1, c->v2);
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, 1.0, c->v2, 0.0, nullptr, out);
    g_dense_launches += 2;
    HYB_CUDA(cudaGetLastError());
    return 0;
}

// =======================================================================================
// WSOSInterpEpiNormEucl: src/Cones/wsosinterpepinormeucl.jl   (point = R blocks of length U)
//   update_feas :131-178   per P_k: Lam11 = P' diag(p_1) P = R11' R11;  M_r = R11^-T (P' diag(p_r) P);
//                          mat = Lam11 - sum_r M_r' M_r = Rm' Rm
//   is_dual_feas :180-183  first block of the dual point > eps
//   update_grad :185-223   A = R11^-T P', B = Rm^-T P', E_r = Rm^-T (-M_r' A)
//   update_hess :225-296   U x U blocks of P Lam^-1 P from A, B, E_r, combined entrywise (wsoseuc_hess_kernel)
//   dder3       :298-396   the two parts restated with every product X Y' taken as (X')' (Y') on stored transposes
//   generic Hessian-based products as for WSOSInterpNonnegative (Cones.jl:108-125, 202-252)
// =======================================================================================
namespace {

// out[r, c] = d1[r] * X1[r, c] + (X2 ? d2[r] * X2[r, c] : 0)   (rows x cols, one leading dimension)
__global__ void rowscale2_kernel(const double* __restrict__ X1, const double* __restrict__ d1, const double* __restrict__ X2,
                                 const double* __restrict__ d2, int64_t rows, int64_t cols, int64_t ld, double beta,
                                 double* __restrict__ out) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= rows * cols) return;
    const int64_t r = e % rows, c = e / rows, o = r + c * ld;
    double v = d1[r] * X1[o];
    if (X2) v += d2[r] * X2[o];
    out[o] = (beta == 0.0 ? 0.0 : beta * out[o]) + v;
}

// One P_k's contribution to the upper triangle of H (update_hess :248-292) from its blocks
// Bk(r, s), r >= s, stored at blk[(r (r + 1) / 2 + s) * bstride] (U x U, ld), Bk(s, r) = Bk(r, s)'; P11 = A'A.
__global__ void wsoseuc_hess_kernel(int64_t U, int R, const double* __restrict__ blk, int64_t bstride, int64_t ld,
                                    const double* __restrict__ P11, double* __restrict__ H, int64_t ldH, int first) {
    const int64_t a = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t b = blockIdx.y;
    const int64_t dim = U * R;
    if (a >= dim || a > b) return;
    const int ra = (int)(a / U), rb = (int)(b / U);
    const int64_t i = a - ra * U, j = b - rb * U;
    auto B = [&](int r, int s, int64_t x, int64_t y) -> double {  // block (r, s) entry (x, y)
        return (r >= s) ? blk[((int64_t)r * (r + 1) / 2 + s) * bstride + x + y * ld]
                        : blk[((int64_t)s * (s + 1) / 2 + r) * bstride + y + x * ld];
    };
    double v = 0.0;
    if (ra == 0 && rb == 0) {
        const double p = P11[i + j * ld];
        v = -(double)(R - 2) * p * p;
        const double b00 = B(0, 0, i, j);
        v += b00 * b00;
        for (int r = 1; r < R; ++r) {
            for (int s = 0; s < r; ++s) {
                const double x = B(r, s, i, j), y = B(r, s, j, i);
                v += x * x + y * y;
            }
            const double d = B(r, r, i, j);
            v += d * d;
        }
    } else if (ra == 0) {
        for (int s = 0; s < R; ++s) v += B(s, 0, i, j) * B(s, rb, i, j);
        v *= 2.0;
    } else if (ra == rb) {
        v = 2.0 * (B(ra, 0, i, j) * B(ra, 0, j, i) + B(0, 0, i, j) * B(ra, ra, i, j));
    } else {  // 1 <= ra < rb
        v = 2.0 * (B(ra, 0, i, j) * B(rb, 0, j, i) + B(0, 0, i, j) * B(rb, ra, j, i));
    }
    H[a + b * ldH] = (first ? 0.0 : H[a + b * ldH]) + v;
}

__global__ void scale_vec_kernel(int64_t n, double a, double* __restrict__ x) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) x[i] *= a;
}

}  // namespace

struct HybDenseWsosEuc {
    int R = 0, nPs = 0;
    int64_t U = 0, ldU = 0, dim = 0, ldD = 0;
    double nu = 0.0;
    std::vector<int64_t> L, ldL;
    // per k: U x L (ld ldU): P, PD, At, Bmt, SPt, S1t, S2t;  L x U (ld ldL): A, Bm;  L x L (ld ldL): Lam, Mat, X, diagT
    std::vector<double*> P, PD, At, Bmt, SPt, S1t, S2t, A, Bm, Lam, Mat, X, diagT;
    std::vector<std::vector<double*>> M, E, Et, SRt, SAt, colT, rowT, half;  // [k][...]
    std::vector<HybChol> chLam, chMat;
    double *blk = nullptr, *P11 = nullptr;  // U x U blocks of the current k
    int64_t bstride = 0;
    double *H = nullptr, *Hf = nullptr, *Lh = nullptr;
    HybChol chH;
    double *v1 = nullptr, *v2 = nullptr, *v3 = nullptr, *scal = nullptr;
    int *dual_ok = nullptr, *slow = nullptr, *sqrt_ok = nullptr;
    const int** infos_dev = nullptr;
    double *big1 = nullptr, *big2 = nullptr;
    int64_t big_cols = 0;
    std::vector<double*> owned;  // everything cudaMalloc'd, for destroy
};

static int euc_alloc(HybDenseWsosEuc* c, double** p, size_t elems) {
    HYB_CUDA(cudaMalloc(p, sizeof(double) * (elems > 0 ? elems : 1)));
    HYB_CUDA(cudaMemset(*p, 0, sizeof(double) * (elems > 0 ? elems : 1)));
    c->owned.push_back(*p);
    return 0;
}

int hybd_wsoseuc_create(HybDenseWsosEuc** out, int R, int64_t U, int nPs, const double* const* Ps_host, const int64_t* Ls) {
    HybDenseWsosEuc* c = new HybDenseWsosEuc();
    c->R = R;
    c->U = U;
    c->nPs = nPs;
    c->ldU = hyb_round_up(U, 16);
    c->dim = U * R;
    c->ldD = hyb_round_up(c->dim, 16);
    const int R1 = R - 1;
    auto resize = [&](std::vector<double*>& v) { v.assign(nPs, nullptr); };
    for (auto* v : {&c->P, &c->PD, &c->At, &c->Bmt, &c->SPt, &c->S1t, &c->S2t, &c->A, &c->Bm, &c->Lam, &c->Mat, &c->X, &c->diagT})
        resize(*v);
    for (auto* v : {&c->M, &c->E, &c->Et, &c->SRt, &c->SAt, &c->colT, &c->rowT, &c->half}) v->assign(nPs, {});
    c->chLam.resize(nPs);
    c->chMat.resize(nPs);
    double sumL = 0;
    for (int k = 0; k < nPs; ++k) {
        const int64_t L = Ls[k], ldL = hyb_round_up(L, 16), ldU = c->ldU;
        c->L.push_back(L);
        c->ldL.push_back(ldL);
        sumL += (double)L;
        for (double** p : {&c->P[k], &c->PD[k], &c->At[k], &c->Bmt[k], &c->SPt[k], &c->S1t[k], &c->S2t[k]})
            HYB_CHECK(euc_alloc(c, p, ldU * L));
        for (double** p : {&c->A[k], &c->Bm[k]}) HYB_CHECK(euc_alloc(c, p, ldL * U));
        for (double** p : {&c->Lam[k], &c->Mat[k], &c->X[k], &c->diagT[k]}) HYB_CHECK(euc_alloc(c, p, ldL * L));
        auto many = [&](std::vector<double*>& v, int n, size_t elems) -> int {
            v.assign(n, nullptr);
            for (int i = 0; i < n; ++i) HYB_CHECK(euc_alloc(c, &v[i], elems));
            return 0;
        };
        HYB_CHECK(many(c->M[k], R1, ldL * L));
        HYB_CHECK(many(c->E[k], R1, ldL * U));
        HYB_CHECK(many(c->Et[k], R1, ldU * L));
        HYB_CHECK(many(c->SRt[k], R1, ldU * L));
        HYB_CHECK(many(c->SAt[k], R, ldU * L));
        HYB_CHECK(many(c->colT[k], R, ldL * L));
        HYB_CHECK(many(c->rowT[k], R1, ldL * L));
        HYB_CHECK(many(c->half[k], R * R, ldL * U));
        HYB_CUDA(cudaMemcpy2D(c->P[k], sizeof(double) * ldU, Ps_host[k], sizeof(double) * U, sizeof(double) * U, L,
                              cudaMemcpyHostToDevice));
        HYB_CHECK(hyb_chol_attach(c->chLam[k], c->Lam[k], ldL, L));
        HYB_CHECK(hyb_chol_attach(c->chMat[k], c->Mat[k], ldL, L));
    }
    c->nu = 2.0 * sumL;
    c->bstride = c->ldU * U;
    HYB_CHECK(euc_alloc(c, &c->blk, c->bstride * (R * (R + 1) / 2)));
    HYB_CHECK(euc_alloc(c, &c->P11, c->bstride));
    for (double** p : {&c->H, &c->Hf, &c->Lh}) HYB_CHECK(euc_alloc(c, p, c->ldD * c->dim));
    for (double** p : {&c->v1, &c->v2, &c->v3}) HYB_CHECK(euc_alloc(c, p, c->ldD));
    HYB_CHECK(euc_alloc(c, &c->scal, 8));
    HYB_CUDA(cudaMalloc(&c->dual_ok, sizeof(int)));
    HYB_CUDA(cudaMalloc(&c->slow, sizeof(int)));
    HYB_CUDA(cudaMalloc(&c->sqrt_ok, sizeof(int)));
    HYB_CUDA(cudaMemset(c->sqrt_ok, 0, sizeof(int)));
    HYB_CHECK(hyb_chol_attach(c->chH, c->Hf, c->ldD, c->dim));
    std::vector<const int*> infos;
    for (int k = 0; k < nPs; ++k) {
        infos.push_back(c->chLam[k].info);
        infos.push_back(c->chMat[k].info);
    }
    HYB_CUDA(cudaMalloc(&c->infos_dev, sizeof(int*) * infos.size()));
    HYB_CUDA(cudaMemcpy(c->infos_dev, infos.data(), sizeof(int*) * infos.size(), cudaMemcpyHostToDevice));
    *out = c;
    return 0;
}

void hybd_wsoseuc_destroy(HybDenseWsosEuc* c) {
    if (!c) return;
    for (auto& ch : c->chLam) hyb_chol_free(ch);
    for (auto& ch : c->chMat) hyb_chol_free(ch);
    hyb_chol_free(c->chH);
    for (double* p : c->owned) cudaFree(p);
    cudaFree(c->dual_ok);
    cudaFree(c->slow);
    cudaFree(c->sqrt_ok);
    cudaFree((void*)c->infos_dev);
    cudaFree(c->big1);
    cudaFree(c->big2);
    delete c;
}

int hybd_wsoseuc_update(cudaStream_t st, HybDenseWsosEuc* c, const double* primal, const double* dual_in, double scal,
                        double irtmu, double* point, double* dual_out, double* grad, int* flag, double* prox) {
    const int64_t U = c->U, ldU = c->ldU, dim = c->dim, ldD = c->ldD;
    const int R = c->R;
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, scal, primal, 0.0, nullptr, point);
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, 1.0, dual_in, 0.0, nullptr, dual_out);
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, 1.0, point, 0.0, nullptr, c->v3);
    all_gt_eps_kernel<<<1, 1024, 0, st>>>(U, dual_out, c->dual_ok);
    g_dense_launches += 4;
    HYB_CUDA(cudaGetLastError());
    // update_feas
    for (int k = 0; k < c->nPs; ++k) {
        const int64_t L = c->L[k], ldL = c->ldL[k];
        rowscale_kernel<<<nb1(U * L), 256, 0, st>>>(c->P[k], c->v3, U, L, ldU, c->PD[k]);
        DLAUNCHED();
        HYB_CHECK(hyb_gemm_tn(st, L, L, U, 1.0, c->P[k], ldU, c->PD[k], ldU, 0.0, c->Lam[k], ldL, HYB_GEMM_UPPER));
        copy_mat_kernel<<<nb1(L * L), 256, 0, st>>>(c->Lam[k], ldL, c->Mat[k], ldL, L, L);
        DLAUNCHED();
        HYB_CHECK(hyb_potrf(st, c->chLam[k], nullptr));
        for (int r = 1; r < R; ++r) {
            rowscale_kernel<<<nb1(U * L), 256, 0, st>>>(c->P[k], c->v3 + r * U, U, L, ldU, c->PD[k]);
            DLAUNCHED();
            HYB_CHECK(hyb_gemm_tn(st, L, L, U, 1.0, c->P[k], ldU, c->PD[k], ldU, 0.0, c->M[k][r - 1], ldL, 0));
            HYB_CHECK(hyb_trsm_fwd(st, c->Lam[k], ldL, L, c->M[k][r - 1], ldL, L));
            HYB_CHECK(hyb_gemm_tn(st, L, L, L, -1.0, c->M[k][r - 1], ldL, c->M[k][r - 1], ldL, 1.0, c->Mat[k], ldL, HYB_GEMM_UPPER));
        }
        HYB_CHECK(hyb_potrf(st, c->chMat[k], nullptr));
    }
    {
        std::vector<int> infs(2 * c->nPs, 0);
        for (int k = 0; k < c->nPs; ++k) {
            HYB_CUDA(cudaMemcpyAsync(&infs[2 * k], c->chLam[k].info, sizeof(int), cudaMemcpyDeviceToHost, st));
            HYB_CUDA(cudaMemcpyAsync(&infs[2 * k + 1], c->chMat[k].info, sizeof(int), cudaMemcpyDeviceToHost, st));
        }
        HYB_CUDA(cudaStreamSynchronize(st));
        bool infeasible = false;
        for (int v : infs) infeasible = infeasible || v != 0;
        if (infeasible) {  // rejected before any gradient / Hessian oracle is asked for (search.jl:146-150)
            wsos_finalize_kernel<<<1, 1, 0, st>>>(c->infos_dev, 2 * c->nPs, c->chH.info, c->dual_ok, c->scal, c->nu, dim, flag, prox,
                                                  c->slow, c->sqrt_ok);
            DLAUNCHED();
            HYB_CUDA(cudaGetLastError());
            return 0;
        }
    }
    // update_grad + update_hess
    for (int k = 0; k < c->nPs; ++k) {
        const int64_t L = c->L[k], ldL = c->ldL[k];
        HYB_CHECK(hyb_transpose(st, c->P[k], ldU, U, L, c->A[k], ldL, 0));  // P' (L x U)
        copy_mat_kernel<<<nb1(L * U), 256, 0, st>>>(c->A[k], ldL, c->Bm[k], ldL, L, U);
        DLAUNCHED();
        HYB_CHECK(hyb_trsm_fwd(st, c->Lam[k], ldL, L, c->A[k], ldL, U));   // A = R11^-T P'
        HYB_CHECK(hyb_trsm_fwd(st, c->Mat[k], ldL, L, c->Bm[k], ldL, U));  // B = Rm^-T P'
        HYB_CHECK(hyb_transpose(st, c->A[k], ldL, L, U, c->At[k], ldU, 0));
        HYB_CHECK(hyb_transpose(st, c->Bm[k], ldL, L, U, c->Bmt[k], ldU, 0));
        coldot_kernel<<<nb1(U, 8), 256, 0, st>>>(c->A[k], c->A[k], L, U, ldL, -1.0, k == 0 ? 0.0 : 1.0, grad);
        coldot_kernel<<<nb1(U, 8), 256, 0, st>>>(c->Bm[k], c->Bm[k], L, U, ldL, -1.0, 1.0, grad);
        g_dense_launches += 2;
        for (int r = 1; r < R; ++r) {
            double* E = c->E[k][r - 1];
            HYB_CHECK(hyb_gemm_tn(st, L, U, L, -1.0, c->M[k][r - 1], ldL, c->A[k], ldL, 0.0, E, ldL, 0));  // -M_r' A
            HYB_CHECK(hyb_trsm_fwd(st, c->Mat[k], ldL, L, E, ldL, U));
            HYB_CHECK(hyb_transpose(st, E, ldL, L, U, c->Et[k][r - 1], ldU, 0));
            coldot_kernel<<<nb1(U, 8), 256, 0, st>>>(E, c->Bm[k], L, U, ldL, -2.0, k == 0 ? 0.0 : 1.0, grad + r * U);
            coldot_kernel<<<nb1(U, 8), 256, 0, st>>>(E, E, L, U, ldL, -1.0, 1.0, grad);
            g_dense_launches += 2;
        }
        HYB_CUDA(cudaGetLastError());
        // U x U blocks of P Lam^-1 P (update_hess :236-247)
        auto B = [&](int r, int s) { return c->blk + ((int64_t)r * (r + 1) / 2 + s) * c->bstride; };
        HYB_CHECK(hyb_gemm_tn(st, U, U, L, 1.0, c->A[k], ldL, c->A[k], ldL, 0.0, c->P11, ldU, 0));
        HYB_CHECK(hyb_gemm_tn(st, U, U, L, 1.0, c->Bm[k], ldL, c->Bm[k], ldL, 0.0, B(0, 0), ldU, 0));
        for (int r = 1; r < R; ++r) {
            const double* Er = c->E[k][r - 1];
            HYB_CHECK(hyb_gemm_tn(st, U, U, L, 1.0, Er, ldL, c->Bm[k], ldL, 0.0, B(r, 0), ldU, 0));
            HYB_CHECK(hyb_gemm_tn(st, U, U, L, 1.0, Er, ldL, Er, ldL, 0.0, B(r, r), ldU, 0));
            lincomb_kernel<<<nb1(c->bstride), 256, 0, st>>>(c->bstride, 1.0, B(r, r), 1.0, c->P11, B(r, r));
            DLAUNCHED();
            for (int r2 = 1; r2 < r; ++r2)
                HYB_CHECK(hyb_gemm_tn(st, U, U, L, 1.0, Er, ldL, c->E[k][r2 - 1], ldL, 0.0, B(r, r2), ldU, 0));
        }
        wsoseuc_hess_kernel<<<dim3(nb1(dim), (unsigned)dim), 256, 0, st>>>(U, R, c->blk, c->bstride, ldU, c->P11, c->H, ldD, k == 0);
        DLAUNCHED();
        HYB_CUDA(cudaGetLastError());
    }
    symmetrize_kernel<<<nb1(dim * dim), 256, 0, st>>>(c->H, dim, ldD);
    copy_upper_dense_kernel<<<nb1(dim * dim), 256, 0, st>>>(c->H, c->Hf, dim, ldD);
    g_dense_launches += 2;
    HYB_CHECK(hyb_potrf(st, c->chH, nullptr));
    HYB_CHECK(hyb_transpose(st, c->Hf, ldD, dim, dim, c->Lh, ldD, 1));
    // scalars for check_numerics / proxsqr / the hess_prod_slow switch (Cones.jl:218-227, 267-304)
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, 1.0, grad, 0.0, nullptr, c->v1);
    dot_kernel<<<1, 1024, 0, st>>>(dim, c->v1, c->v3, c->scal + 0);
    HYB_CHECK(hyb_potrs(st, c->chH, c->v1));
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, 1.0, grad, 0.0, nullptr, c->v2);
    dot_kernel<<<1, 1024, 0, st>>>(dim, c->v1, c->v2, c->scal + 1);
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, irtmu, dual_out, 1.0, grad, c->v2);
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, 1.0, c->v2, 0.0, nullptr, c->v1);
    HYB_CHECK(hyb_potrs(st, c->chH, c->v1));
    dot_kernel<<<1, 1024, 0, st>>>(dim, c->v1, c->v2, c->scal + 2);
    HYB_CHECK(hyb_gemv_t(st, dim, dim, c->H, ldD, c->v3, c->v1));
    dot_kernel<<<1, 1024, 0, st>>>(dim, c->v1, c->v3, c->scal + 3);
    wsos_finalize_kernel<<<1, 1, 0, st>>>(c->infos_dev, 2 * c->nPs, c->chH.info, c->dual_ok, c->scal, c->nu, dim, flag, prox, c->slow,
                                          c->sqrt_ok);
    g_dense_launches += 10;
    HYB_CUDA(cudaGetLastError());
    return 0;
}

int hybd_wsoseuc_sqrt_ok(cudaStream_t st, HybDenseWsosEuc* c, int* ok) {
    HYB_CUDA(cudaMemcpyAsync(ok, c->sqrt_ok, sizeof(int), cudaMemcpyDeviceToHost, st));
    HYB_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int hybd_wsoseuc_prod(cudaStream_t st, HybDenseWsosEuc* c, int op, const double* arr, int64_t lda, double* prod, int64_t ldp,
                      int64_t m) {
    if (m <= 0) return 0;
    const int64_t D = c->dim, ldD = c->ldD;
    if (m > c->big_cols) {
        cudaFree(c->big1);
        cudaFree(c->big2);
        c->big1 = c->big2 = nullptr;
        HYB_CUDA(cudaMalloc(&c->big1, sizeof(double) * ldD * m));
        HYB_CUDA(cudaMalloc(&c->big2, sizeof(double) * ldD * m));
        c->big_cols = m;
    }
    double *A = c->big1, *B = c->big2;
    copy_mat_kernel<<<nb1(D * m), 256, 0, st>>>(arr, lda, A, ldD, D, m);
    DLAUNCHED();
    HYB_CUDA(cudaGetLastError());
    const double* result = nullptr;
    if (op == HYB_OP_HESS || op == 6) {  // no factor-based hess_prod_slow! for this cone: Cones.jl:229-231
        if (m == 1) HYB_CHECK(hyb_gemv_t(st, D, D, c->H, ldD, A, B));
        else HYB_CHECK(hyb_gemm_tn(st, D, m, D, 1.0, c->H, ldD, A, ldD, 0.0, B, ldD, 0));
        result = B;
    } else if (op == HYB_OP_INV_HESS) {
        if (m == 1) {
            HYB_CHECK(hyb_potrs(st, c->chH, A));
        } else {
            HYB_CHECK(hyb_trsm_fwd(st, c->Hf, ldD, D, A, ldD, m));
            HYB_CHECK(hyb_trsm_bwd(st, c->Hf, c->Lh, ldD, D, A, ldD, m));
        }
        result = A;
    } else if (op == HYB_OP_SQRT_HESS) {
        if (m == 1) HYB_CHECK(hyb_gemv_t(st, D, D, c->Lh, ldD, A, B));
        else HYB_CHECK(hyb_gemm_tn(st, D, m, D, 1.0, c->Lh, ldD, A, ldD, 0.0, B, ldD, 0));
        result = B;
    } else {
        HYB_CHECK(hyb_trsm_fwd(st, c->Hf, ldD, D, A, ldD, m));
        result = A;
    }
    copy_mat_kernel<<<nb1(D * m), 256, 0, st>>>(result, ldD, prod, ldp, D, m);
    DLAUNCHED();
    HYB_CUDA(cudaGetLastError());
    return 0;
}

int hybd_wsoseuc_dder3(cudaStream_t st, HybDenseWsosEuc* c, const double* dir, double* out) {
    const int64_t U = c->U, ldU = c->ldU, dim = c->dim;
    const int R = c->R;
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, 1.0, dir, 0.0, nullptr, c->v1);  // aligned copy of the direction
    DLAUNCHED();
    const double* d = c->v1;
    double* acc = c->v2;
    auto scale = [&](const double* X, const double* dv, int64_t L, double* dst) {  // dst = diag(dv) X  (U x L)
        rowscale2_kernel<<<nb1(U * L), 256, 0, st>>>(X, dv, nullptr, nullptr, U, L, ldU, 0.0, dst);
        DLAUNCHED();
    };
    // part 1 (:305-319): -(R - 2) * sum_k |Herm(A D0 A') A|^2 column norms
    for (int k = 0; k < c->nPs; ++k) {
        const int64_t L = c->L[k], ldL = c->ldL[k];
        scale(c->At[k], d, L, c->SAt[k][0]);
        HYB_CHECK(hyb_gemm_tn(st, L, L, U, 1.0, c->SAt[k][0], ldU, c->At[k], ldU, 0.0, c->X[k], ldL, 0));   // A D0 A'
        HYB_CHECK(hyb_gemm_tn(st, L, U, L, 1.0, c->X[k], ldL, c->A[k], ldL, 0.0, c->half[k][0], ldL, 0));  // (A D0 A')' A
        coldot_kernel<<<nb1(U, 8), 256, 0, st>>>(c->half[k][0], c->half[k][0], L, U, ldL, 1.0, k == 0 ? 0.0 : 1.0, acc);
        DLAUNCHED();
    }
    scale_vec_kernel<<<nb1(U), 256, 0, st>>>(U, (double)(2 - R), acc);
    for (int r = 1; r < R; ++r) HYB_CUDA(cudaMemsetAsync(acc + r * U, 0, sizeof(double) * U, st));
    DLAUNCHED();
    // part 2 (:321-393)
    for (int k = 0; k < c->nPs; ++k) {
        const int64_t L = c->L[k], ldL = c->ldL[k];
        double *At = c->At[k], *Bmt = c->Bmt[k];
        // transposes of scaled_pt, scaled_row_r, scaled_col_r, scaled_diag (U x L each)
        for (int r = 1; r < R; ++r) scale(At, d + r * U, L, c->SAt[k][r]);  // SAt[0] = diag(d0) A' from part 1
        rowscale2_kernel<<<nb1(U * L), 256, 0, st>>>(Bmt, d, nullptr, nullptr, U, L, ldU, 0.0, c->SPt[k]);
        DLAUNCHED();
        for (int r = 1; r < R; ++r) {
            rowscale2_kernel<<<nb1(U * L), 256, 0, st>>>(c->Et[k][r - 1], d + r * U, nullptr, nullptr, U, L, ldU, 1.0, c->SPt[k]);
            rowscale2_kernel<<<nb1(U * L), 256, 0, st>>>(Bmt, d + r * U, c->Et[k][r - 1], d, U, L, ldU, 0.0, c->SRt[k][r - 1]);
            g_dense_launches += 2;
        }
        HYB_CUDA(cudaGetLastError());
        // colT[0] = (scaled_pt B' + sum_r scaled_row_r E_r')' ; rowT_r = (scaled_row_r A')' ; colT[r] = (scaled_col_r B' +
        // scaled_diag E_r')' ; diagT = (scaled_diag A')'   -- X Y' = (X')' (Y'), its transpose = (Y')' (X')
        HYB_CHECK(hyb_gemm_tn(st, L, L, U, 1.0, Bmt, ldU, c->SPt[k], ldU, 0.0, c->colT[k][0], ldL, 0));
        HYB_CHECK(hyb_gemm_tn(st, L, L, U, 1.0, At, ldU, c->SAt[k][0], ldU, 0.0, c->diagT[k], ldL, 0));
        for (int r = 1; r < R; ++r) {
            HYB_CHECK(hyb_gemm_tn(st, L, L, U, 1.0, c->Et[k][r - 1], ldU, c->SRt[k][r - 1], ldU, 1.0, c->colT[k][0], ldL, 0));
            HYB_CHECK(hyb_gemm_tn(st, L, L, U, 1.0, At, ldU, c->SRt[k][r - 1], ldU, 0.0, c->rowT[k][r - 1], ldL, 0));
            HYB_CHECK(hyb_gemm_tn(st, L, L, U, 1.0, Bmt, ldU, c->SAt[k][r], ldU, 0.0, c->colT[k][r], ldL, 0));
            HYB_CHECK(hyb_gemm_tn(st, L, L, U, 1.0, c->Et[k][r - 1], ldU, c->SAt[k][0], ldU, 1.0, c->colT[k][r], ldL, 0));
        }
        // half[rho][sigma] = col[rho] * ctg[sigma] (+ row_sigma A at rho = 0, + diag A at rho = sigma), L x U each
        for (int rho = 0; rho < R; ++rho)
            for (int sg = 0; sg < R; ++sg) {
                const double* ctg = sg == 0 ? c->Bm[k] : c->E[k][sg - 1];
                double* hb = c->half[k][rho * R + sg];
                HYB_CHECK(hyb_gemm_tn(st, L, U, L, 1.0, c->colT[k][rho], ldL, ctg, ldL, 0.0, hb, ldL, 0));
                if (sg >= 1 && rho == 0)
                    HYB_CHECK(hyb_gemm_tn(st, L, U, L, 1.0, c->rowT[k][sg - 1], ldL, c->A[k], ldL, 1.0, hb, ldL, 0));
                if (sg >= 1 && rho == sg)
                    HYB_CHECK(hyb_gemm_tn(st, L, U, L, 1.0, c->diagT[k], ldL, c->A[k], ldL, 1.0, hb, ldL, 0));
            }
        for (int rho = 0; rho < R; ++rho) {
            const double* h0 = c->half[k][rho * R];
            coldot_kernel<<<nb1(U, 8), 256, 0, st>>>(h0, h0, L, U, ldL, 1.0, 1.0, acc);
            DLAUNCHED();
            for (int sg = 1; sg < R; ++sg) {
                const double* hs = c->half[k][rho * R + sg];
                coldot_kernel<<<nb1(U, 8), 256, 0, st>>>(hs, h0, L, U, ldL, 2.0, 1.0, acc + sg * U);
                coldot_kernel<<<nb1(U, 8), 256, 0, st>>>(hs, hs, L, U, ldL, 1.0, 1.0, acc);
                g_dense_launches += 2;
            }
        }
        HYB_CUDA(cudaGetLastError());
    }
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, 1.0, acc, 0.0, nullptr, out);
    DLAUNCHED();
    HYB_CUDA(cudaGetLastError());
    return 0;
}

// =======================================================================================
// WSOSInterpEpiNormOne: src/Cones/wsosinterpepinormone.jl   (point = R blocks of length U)
//   update_feas :161-218   per P_k: Lam11 = P' diag(p_1) P = R11' R11; per r: M_r = R11^-T (P' diag(p_r) P),
//                          mat_r = Lam11 - M_r' M_r = Rm_r' Rm_r
//   update_grad :226-265   A = R11^-T P', S12_r = Rm_r^-T (-M_r' A), S11_r = Rm_r^-T P'
//   update_hess(_prod) :267-330  diagonal and edge blocks of the arrow-shaped Hessian from U x U contractions
//   dder3 :429-521
//   hess_prod! / inv_hess_prod! (:332-427) exploit the arrow shape in the reference; here they are the generic
//   products with the stored Hessian and its Cholesky (the same linear maps)
// =======================================================================================
namespace {

// upper triangles (i <= j): d0 += uu, dr += uu, er += 4 P1 P2 with uu = 2 (P1^2 + P2^2)   (:316-323)
__global__ void wsosone_accum_kernel(int64_t U, int64_t ld, const double* __restrict__ P1, const double* __restrict__ P2,
                                     double* __restrict__ d0, double* __restrict__ dr, double* __restrict__ er, int first0,
                                     int firstr) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= U * U) return;
    const int64_t i = e % U, j = e / U, o = i + j * ld;
    if (i > j) return;
    const double a = P1[o], b = P2[o], uu = 2.0 * (a * a + b * b);
    d0[o] = (first0 ? 0.0 : d0[o]) + uu;
    dr[o] = (firstr ? 0.0 : dr[o]) + uu;
    er[o] = (firstr ? 0.0 : er[o]) + 4.0 * a * b;
}
__global__ void wsosone_sub11_kernel(int64_t U, int64_t ld, const double* __restrict__ P11, double fac, double* __restrict__ d0) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= U * U) return;
    const int64_t i = e % U, j = e / U, o = i + j * ld;
    if (i <= j) d0[o] -= fac * P11[o] * P11[o];
}
// upper triangle of H from the R diagonal blocks and the R - 1 edge blocks (update_hess :267-284)
__global__ void wsosone_hess_kernel(int64_t U, int R, const double* __restrict__ diag, const double* __restrict__ edge,
                                    int64_t bstride, int64_t ld, double* __restrict__ H, int64_t ldH) {
    const int64_t a = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t b = blockIdx.y;
    if (a >= U * R || a > b) return;
    const int ra = (int)(a / U), rb = (int)(b / U);
    const int64_t i = a - ra * U, j = b - rb * U;
    double v = 0.0;
    if (ra == rb) v = diag[ra * bstride + i + j * ld];
    else if (ra == 0) v = (i <= j) ? edge[(rb - 1) * bstride + i + j * ld] : edge[(rb - 1) * bstride + j + i * ld];
    H[a + b * ldH] = v;
}

}  // namespace

struct HybDenseWsosOne {
    int R = 0, nPs = 0;
    int64_t U = 0, ldU = 0, dim = 0, ldD = 0;
    double nu = 0.0;
    std::vector<int64_t> L, ldL;
    std::vector<double*> P, PD, At, D22t, A, LUlast, Lam, Lam0, XT, LLt;          // per k
    std::vector<std::vector<double*>> M, Mat, S11, S12, S11t, S12t, D11t, D12t, D21t, half;  // [k][r] (half: 4 per r)
    std::vector<HybChol> chLam;
    std::vector<std::vector<HybChol>> chMat;
    double *P11 = nullptr, *P1 = nullptr, *P2 = nullptr, *diag = nullptr, *edge = nullptr;
    int64_t bstride = 0;
    double *H = nullptr, *Hf = nullptr, *Lh = nullptr;
    HybChol chH;
    double *v1 = nullptr, *v2 = nullptr, *v3 = nullptr, *scal = nullptr;
    int *dual_ok = nullptr, *slow = nullptr, *sqrt_ok = nullptr;
    const int** infos_dev = nullptr;
    int ninfos = 0;
    double *big1 = nullptr, *big2 = nullptr;
    int64_t big_cols = 0;
    std::vector<double*> owned;
};

static int one_alloc(HybDenseWsosOne* c, double** p, size_t elems) {
    HYB_CUDA(cudaMalloc(p, sizeof(double) * (elems > 0 ? elems : 1)));
    HYB_CUDA(cudaMemset(*p, 0, sizeof(double) * (elems > 0 ? elems : 1)));
    c->owned.push_back(*p);
    return 0;
}

int hybd_wsosone_create(HybDenseWsosOne** out, int R, int64_t U, int nPs, const double* const* Ps_host, const int64_t* Ls) {
    HybDenseWsosOne* c = new HybDenseWsosOne();
    c->R = R;
    c->U = U;
    c->nPs = nPs;
    c->ldU = hyb_round_up(U, 16);
    c->dim = U * R;
    c->ldD = hyb_round_up(c->dim, 16);
    const int R1 = R - 1;
    for (auto* v : {&c->P, &c->PD, &c->At, &c->D22t, &c->A, &c->LUlast, &c->Lam, &c->Lam0, &c->XT, &c->LLt}) v->assign(nPs, nullptr);
    for (auto* v : {&c->M, &c->Mat, &c->S11, &c->S12, &c->S11t, &c->S12t, &c->D11t, &c->D12t, &c->D21t, &c->half}) v->assign(nPs, {});
    c->chLam.resize(nPs);
    c->chMat.resize(nPs);
    double sumL = 0;
    std::vector<const int*> infos;
    for (int k = 0; k < nPs; ++k) {
        const int64_t L = Ls[k], ldL = hyb_round_up(L, 16), ldU = c->ldU;
        c->L.push_back(L);
        c->ldL.push_back(ldL);
        sumL += (double)L;
        for (double** p : {&c->P[k], &c->PD[k], &c->At[k], &c->D22t[k]}) HYB_CHECK(one_alloc(c, p, ldU * L));
        for (double** p : {&c->A[k], &c->LUlast[k]}) HYB_CHECK(one_alloc(c, p, ldL * U));
        for (double** p : {&c->Lam[k], &c->Lam0[k], &c->XT[k], &c->LLt[k]}) HYB_CHECK(one_alloc(c, p, ldL * L));
        auto many = [&](std::vector<double*>& v, int n, size_t elems) -> int {
            v.assign(n, nullptr);
            for (int i = 0; i < n; ++i) HYB_CHECK(one_alloc(c, &v[i], elems));
            return 0;
        };
        HYB_CHECK(many(c->M[k], R1, ldL * L));
        HYB_CHECK(many(c->Mat[k], R1, ldL * L));
        for (auto* v : {&c->S11[k], &c->S12[k]}) HYB_CHECK(many(*v, R1, ldL * U));
        for (auto* v : {&c->S11t[k], &c->S12t[k], &c->D11t[k], &c->D12t[k], &c->D21t[k]}) HYB_CHECK(many(*v, R1, ldU * L));
        HYB_CHECK(many(c->half[k], 4 * R1, ldL * U));
        HYB_CUDA(cudaMemcpy2D(c->P[k], sizeof(double) * ldU, Ps_host[k], sizeof(double) * U, sizeof(double) * U, L,
                              cudaMemcpyHostToDevice));
        HYB_CHECK(hyb_chol_attach(c->chLam[k], c->Lam[k], ldL, L));
        infos.push_back(c->chLam[k].info);
        c->chMat[k].resize(R1);
        for (int r = 0; r < R1; ++r) {
            HYB_CHECK(hyb_chol_attach(c->chMat[k][r], c->Mat[k][r], ldL, L));
            infos.push_back(c->chMat[k][r].info);
        }
    }
    c->nu = (double)R * sumL;
    c->bstride = c->ldU * U;
    for (double** p : {&c->P11, &c->P1, &c->P2}) HYB_CHECK(one_alloc(c, p, c->bstride));
    HYB_CHECK(one_alloc(c, &c->diag, c->bstride * R));
    HYB_CHECK(one_alloc(c, &c->edge, c->bstride * (R1 > 0 ? R1 : 1)));
    for (double** p : {&c->H, &c->Hf, &c->Lh}) HYB_CHECK(one_alloc(c, p, c->ldD * c->dim));
    for (double** p : {&c->v1, &c->v2, &c->v3}) HYB_CHECK(one_alloc(c, p, c->ldD));
    HYB_CHECK(one_alloc(c, &c->scal, 8));
    HYB_CUDA(cudaMalloc(&c->dual_ok, sizeof(int)));
    HYB_CUDA(cudaMalloc(&c->slow, sizeof(int)));
    HYB_CUDA(cudaMalloc(&c->sqrt_ok, sizeof(int)));
    HYB_CUDA(cudaMemset(c->sqrt_ok, 0, sizeof(int)));
    HYB_CHECK(hyb_chol_attach(c->chH, c->Hf, c->ldD, c->dim));
    c->ninfos = (int)infos.size();
    HYB_CUDA(cudaMalloc(&c->infos_dev, sizeof(int*) * infos.size()));
    HYB_CUDA(cudaMemcpy(c->infos_dev, infos.data(), sizeof(int*) * infos.size(), cudaMemcpyHostToDevice));
    *out = c;
    return 0;
}

void hybd_wsosone_destroy(HybDenseWsosOne* c) {
    if (!c) return;
    for (auto& ch : c->chLam) hyb_chol_free(ch);
    for (auto& v : c->chMat)
        for (auto& ch : v) hyb_chol_free(ch);
    hyb_chol_free(c->chH);
    for (double* p : c->owned) cudaFree(p);
    cudaFree(c->dual_ok);
    cudaFree(c->slow);
    cudaFree(c->sqrt_ok);
    cudaFree((void*)c->infos_dev);
    cudaFree(c->big1);
    cudaFree(c->big2);
    delete c;
}

int hybd_wsosone_update(cudaStream_t st, HybDenseWsosOne* c, const double* primal, const double* dual_in, double scal,
                        double irtmu, double* point, double* dual_out, double* grad, int* flag, double* prox) {
    const int64_t U = c->U, ldU = c->ldU, dim = c->dim, ldD = c->ldD;
    const int R = c->R, R1 = c->R - 1;
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, scal, primal, 0.0, nullptr, point);
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, 1.0, dual_in, 0.0, nullptr, dual_out);
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, 1.0, point, 0.0, nullptr, c->v3);
    all_gt_eps_kernel<<<1, 1024, 0, st>>>(U, dual_out, c->dual_ok);
    g_dense_launches += 4;
    HYB_CUDA(cudaGetLastError());
    for (int k = 0; k < c->nPs; ++k) {  // update_feas
        const int64_t L = c->L[k], ldL = c->ldL[k];
        rowscale_kernel<<<nb1(U * L), 256, 0, st>>>(c->P[k], c->v3, U, L, ldU, c->PD[k]);
        DLAUNCHED();
        HYB_CHECK(hyb_gemm_tn(st, L, L, U, 1.0, c->P[k], ldU, c->PD[k], ldU, 0.0, c->Lam[k], ldL, HYB_GEMM_UPPER));
        copy_mat_kernel<<<nb1(L * L), 256, 0, st>>>(c->Lam[k], ldL, c->Lam0[k], ldL, L, L);
        DLAUNCHED();
        HYB_CHECK(hyb_potrf(st, c->chLam[k], nullptr));
        for (int r = 0; r < R1; ++r) {
            rowscale_kernel<<<nb1(U * L), 256, 0, st>>>(c->P[k], c->v3 + (r + 1) * U, U, L, ldU, c->PD[k]);
            DLAUNCHED();
            HYB_CHECK(hyb_gemm_tn(st, L, L, U, 1.0, c->P[k], ldU, c->PD[k], ldU, 0.0, c->M[k][r], ldL, 0));
            HYB_CHECK(hyb_trsm_fwd(st, c->Lam[k], ldL, L, c->M[k][r], ldL, L));
            copy_mat_kernel<<<nb1(L * L), 256, 0, st>>>(c->Lam0[k], ldL, c->Mat[k][r], ldL, L, L);
            DLAUNCHED();
            HYB_CHECK(hyb_gemm_tn(st, L, L, L, -1.0, c->M[k][r], ldL, c->M[k][r], ldL, 1.0, c->Mat[k][r], ldL, HYB_GEMM_UPPER));
            HYB_CHECK(hyb_potrf(st, c->chMat[k][r], nullptr));
        }
    }
    {
        std::vector<int> infs(c->ninfos, 0);
        int t = 0;
        for (int k = 0; k < c->nPs; ++k) {
            HYB_CUDA(cudaMemcpyAsync(&infs[t++], c->chLam[k].info, sizeof(int), cudaMemcpyDeviceToHost, st));
            for (int r = 0; r < R1; ++r)
                HYB_CUDA(cudaMemcpyAsync(&infs[t++], c->chMat[k][r].info, sizeof(int), cudaMemcpyDeviceToHost, st));
        }
        HYB_CUDA(cudaStreamSynchronize(st));
        bool infeasible = false;
        for (int v : infs) infeasible = infeasible || v != 0;
        if (infeasible) {
            wsos_finalize_kernel<<<1, 1, 0, st>>>(c->infos_dev, c->ninfos, c->chH.info, c->dual_ok, c->scal, c->nu, dim, flag, prox,
                                                  c->slow, c->sqrt_ok);
            DLAUNCHED();
            HYB_CUDA(cudaGetLastError());
            return 0;
        }
    }
    for (int k = 0; k < c->nPs; ++k) {  // update_grad + the Hessian blocks
        const int64_t L = c->L[k], ldL = c->ldL[k];
        HYB_CHECK(hyb_transpose(st, c->P[k], ldU, U, L, c->A[k], ldL, 0));
        for (int r = 0; r < R1; ++r) {
            copy_mat_kernel<<<nb1(L * U), 256, 0, st>>>(c->A[k], ldL, c->S11[k][r], ldL, L, U);  // P'
            DLAUNCHED();
        }
        HYB_CHECK(hyb_trsm_fwd(st, c->Lam[k], ldL, L, c->A[k], ldL, U));
        HYB_CHECK(hyb_transpose(st, c->A[k], ldL, L, U, c->At[k], ldU, 0));
        coldot_kernel<<<nb1(U, 8), 256, 0, st>>>(c->A[k], c->A[k], L, U, ldL, -(double)R, k == 0 ? 0.0 : 1.0, grad);
        DLAUNCHED();
        HYB_CHECK(hyb_gemm_tn(st, U, U, L, 1.0, c->A[k], ldL, c->A[k], ldL, 0.0, c->P11, ldU, 0));
        for (int r = 0; r < R1; ++r) {
            double *S12 = c->S12[k][r], *S11 = c->S11[k][r];
            HYB_CHECK(hyb_gemm_tn(st, L, U, L, -1.0, c->M[k][r], ldL, c->A[k], ldL, 0.0, S12, ldL, 0));
            HYB_CHECK(hyb_trsm_fwd(st, c->Mat[k][r], ldL, L, S12, ldL, U));
            HYB_CHECK(hyb_trsm_fwd(st, c->Mat[k][r], ldL, L, S11, ldL, U));
            HYB_CHECK(hyb_transpose(st, S12, ldL, L, U, c->S12t[k][r], ldU, 0));
            HYB_CHECK(hyb_transpose(st, S11, ldL, L, U, c->S11t[k][r], ldU, 0));
            coldot_kernel<<<nb1(U, 8), 256, 0, st>>>(S12, S12, L, U, ldL, -2.0, 1.0, grad);
            coldot_kernel<<<nb1(U, 8), 256, 0, st>>>(S12, S11, L, U, ldL, -2.0, k == 0 ? 0.0 : 1.0, grad + (r + 1) * U);
            g_dense_launches += 2;
            HYB_CHECK(hyb_gemm_tn(st, U, U, L, 1.0, S12, ldL, S11, ldL, 0.0, c->P2, ldU, 0));
            HYB_CHECK(hyb_gemm_tn(st, U, U, L, 1.0, S12, ldL, S12, ldL, 0.0, c->P1, ldU, 0));
            lincomb_kernel<<<nb1(c->bstride), 256, 0, st>>>(c->bstride, 1.0, c->P1, 1.0, c->P11, c->P1);
            wsosone_accum_kernel<<<nb1(U * U), 256, 0, st>>>(U, ldU, c->P1, c->P2, c->diag, c->diag + (r + 1) * c->bstride,
                                                            c->edge + r * c->bstride, (k == 0 && r == 0) ? 1 : 0, k == 0 ? 1 : 0);
            g_dense_launches += 2;
        }
        if (R1 == 0 && k == 0) HYB_CUDA(cudaMemsetAsync(c->diag, 0, sizeof(double) * c->bstride, st));
        wsosone_sub11_kernel<<<nb1(U * U), 256, 0, st>>>(U, ldU, c->P11, (double)(R - 2), c->diag);
        DLAUNCHED();
        HYB_CUDA(cudaGetLastError());
    }
    wsosone_hess_kernel<<<dim3(nb1(dim), (unsigned)dim), 256, 0, st>>>(U, R, c->diag, c->edge, c->bstride, ldU, c->H, ldD);
    symmetrize_kernel<<<nb1(dim * dim), 256, 0, st>>>(c->H, dim, ldD);
    copy_upper_dense_kernel<<<nb1(dim * dim), 256, 0, st>>>(c->H, c->Hf, dim, ldD);
    g_dense_launches += 3;
    HYB_CHECK(hyb_potrf(st, c->chH, nullptr));
    HYB_CHECK(hyb_transpose(st, c->Hf, ldD, dim, dim, c->Lh, ldD, 1));
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, 1.0, grad, 0.0, nullptr, c->v1);
    dot_kernel<<<1, 1024, 0, st>>>(dim, c->v1, c->v3, c->scal + 0);
    HYB_CHECK(hyb_potrs(st, c->chH, c->v1));
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, 1.0, grad, 0.0, nullptr, c->v2);
    dot_kernel<<<1, 1024, 0, st>>>(dim, c->v1, c->v2, c->scal + 1);
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, irtmu, dual_out, 1.0, grad, c->v2);
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, 1.0, c->v2, 0.0, nullptr, c->v1);
    HYB_CHECK(hyb_potrs(st, c->chH, c->v1));
    dot_kernel<<<1, 1024, 0, st>>>(dim, c->v1, c->v2, c->scal + 2);
    HYB_CHECK(hyb_gemv_t(st, dim, dim, c->H, ldD, c->v3, c->v1));
    dot_kernel<<<1, 1024, 0, st>>>(dim, c->v1, c->v3, c->scal + 3);
    wsos_finalize_kernel<<<1, 1, 0, st>>>(c->infos_dev, c->ninfos, c->chH.info, c->dual_ok, c->scal, c->nu, dim, flag, prox, c->slow,
                                          c->sqrt_ok);
    g_dense_launches += 10;
    HYB_CUDA(cudaGetLastError());
    return 0;
}

int hybd_wsosone_sqrt_ok(cudaStream_t st, HybDenseWsosOne* c, int* ok) {
    HYB_CUDA(cudaMemcpyAsync(ok, c->sqrt_ok, sizeof(int), cudaMemcpyDeviceToHost, st));
    HYB_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int hybd_wsosone_prod(cudaStream_t st, HybDenseWsosOne* c, int op, const double* arr, int64_t lda, double* prod, int64_t ldp,
                      int64_t m) {
    if (m <= 0) return 0;
    const int64_t D = c->dim, ldD = c->ldD;
    if (m > c->big_cols) {
        cudaFree(c->big1);
        cudaFree(c->big2);
        c->big1 = c->big2 = nullptr;
        HYB_CUDA(cudaMalloc(&c->big1, sizeof(double) * ldD * m));
        HYB_CUDA(cudaMalloc(&c->big2, sizeof(double) * ldD * m));
        c->big_cols = m;
    }
    double *A = c->big1, *B = c->big2;
    copy_mat_kernel<<<nb1(D * m), 256, 0, st>>>(arr, lda, A, ldD, D, m);
    DLAUNCHED();
    HYB_CUDA(cudaGetLastError());
    const double* result = nullptr;
    if (op == HYB_OP_HESS || op == 6) {
        if (m == 1) HYB_CHECK(hyb_gemv_t(st, D, D, c->H, ldD, A, B));
        else HYB_CHECK(hyb_gemm_tn(st, D, m, D, 1.0, c->H, ldD, A, ldD, 0.0, B, ldD, 0));
        result = B;
    } else if (op == HYB_OP_INV_HESS) {
        if (m == 1) {
            HYB_CHECK(hyb_potrs(st, c->chH, A));
        } else {
            HYB_CHECK(hyb_trsm_fwd(st, c->Hf, ldD, D, A, ldD, m));
            HYB_CHECK(hyb_trsm_bwd(st, c->Hf, c->Lh, ldD, D, A, ldD, m));
        }
        result = A;
    } else if (op == HYB_OP_SQRT_HESS) {
        if (m == 1) HYB_CHECK(hyb_gemv_t(st, D, D, c->Lh, ldD, A, B));
        else HYB_CHECK(hyb_gemm_tn(st, D, m, D, 1.0, c->Lh, ldD, A, ldD, 0.0, B, ldD, 0));
        result = B;
    } else {
        HYB_CHECK(hyb_trsm_fwd(st, c->Hf, ldD, D, A, ldD, m));
        result = A;
    }
    copy_mat_kernel<<<nb1(D * m), 256, 0, st>>>(result, ldD, prod, ldp, D, m);
    DLAUNCHED();
    HYB_CUDA(cudaGetLastError());
    return 0;
}

int hybd_wsosone_dder3(cudaStream_t st, HybDenseWsosOne* c, const double* dir, double* out) {
    const int64_t U = c->U, ldU = c->ldU, dim = c->dim;
    const int R = c->R, R1 = c->R - 1;
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, 1.0, dir, 0.0, nullptr, c->v1);
    DLAUNCHED();
    const double* d = c->v1;
    double* acc = c->v2;
    // part 1 (:437-450) and the pieces of part 2 that depend on Lam11 only
    for (int k = 0; k < c->nPs; ++k) {
        const int64_t L = c->L[k], ldL = c->ldL[k];
        rowscale2_kernel<<<nb1(U * L), 256, 0, st>>>(c->At[k], d, nullptr, nullptr, U, L, ldU, 0.0, c->D22t[k]);
        DLAUNCHED();
        HYB_CHECK(hyb_gemm_tn(st, L, L, U, 1.0, c->At[k], ldU, c->D22t[k], ldU, 0.0, c->XT[k], ldL, 0));   // (A D0 A')'
        HYB_CHECK(hyb_gemm_tn(st, L, U, L, 1.0, c->XT[k], ldL, c->A[k], ldL, 0.0, c->LUlast[k], ldL, 0));  // A D0 A' A
        coldot_kernel<<<nb1(U, 8), 256, 0, st>>>(c->LUlast[k], c->LUlast[k], L, U, ldL, 1.0, k == 0 ? 0.0 : 1.0, acc);
        DLAUNCHED();
    }
    scale_vec_kernel<<<nb1(U), 256, 0, st>>>(U, (double)(2 - R), acc);
    DLAUNCHED();
    for (int r = 1; r < R; ++r) HYB_CUDA(cudaMemsetAsync(acc + r * U, 0, sizeof(double) * U, st));
    // part 2 (:452-518)
    for (int k = 0; k < c->nPs; ++k) {
        const int64_t L = c->L[k], ldL = c->ldL[k];
        double *At = c->At[k], *A = c->A[k], *LLt = c->LLt[k], *D22t = c->D22t[k];
        for (int s = 0; s < R1; ++s) {
            const double* ds = d + (s + 1) * U;
            double *S11 = c->S11[k][s], *S12 = c->S12[k][s], *S11t = c->S11t[k][s], *S12t = c->S12t[k][s];
            double *D11t = c->D11t[k][s], *D12t = c->D12t[k][s], *D21t = c->D21t[k][s];
            double *h00 = c->half[k][4 * s], *h01 = c->half[k][4 * s + 1], *h10 = c->half[k][4 * s + 2], *h11 = c->half[k][4 * s + 3];
            rowscale2_kernel<<<nb1(U * L), 256, 0, st>>>(S11t, d, S12t, ds, U, L, ldU, 0.0, D11t);
            rowscale2_kernel<<<nb1(U * L), 256, 0, st>>>(S11t, ds, S12t, d, U, L, ldU, 0.0, D12t);
            rowscale2_kernel<<<nb1(U * L), 256, 0, st>>>(At, ds, nullptr, nullptr, U, L, ldU, 0.0, D21t);
            g_dense_launches += 3;
            HYB_CUDA(cudaGetLastError());
            auto llt = [&](const double* Yt, const double* Xt) {  // LLt = (X Y')' = (Y')' (X')
                return hyb_gemm_tn(st, L, L, U, 1.0, Yt, ldU, Xt, ldU, 0.0, LLt, ldL, 0);
            };
            auto mul = [&](const double* Z, double beta, double* hb) {  // hb (+)= LL Z
                return hyb_gemm_tn(st, L, U, L, 1.0, LLt, ldL, Z, ldL, beta, hb, ldL, 0);
            };
            HYB_CHECK(llt(S12t, D11t));
            HYB_CHECK(mul(S12, 0.0, h00));
            HYB_CHECK(mul(S11, 0.0, h01));
            HYB_CHECK(llt(S12t, D12t));
            HYB_CHECK(mul(S11, 1.0, h00));
            HYB_CHECK(mul(S12, 1.0, h01));
            HYB_CHECK(llt(S12t, D21t));
            HYB_CHECK(mul(S12, 0.0, h10));
            HYB_CHECK(mul(S11, 0.0, h11));
            HYB_CHECK(llt(S12t, D22t));
            HYB_CHECK(mul(S11, 1.0, h10));
            HYB_CHECK(mul(S12, 1.0, h11));
            HYB_CHECK(llt(At, D11t));
            HYB_CHECK(mul(A, 1.0, h00));
            HYB_CHECK(llt(At, D12t));
            HYB_CHECK(mul(A, 1.0, h01));
            HYB_CHECK(llt(At, D21t));
            HYB_CHECK(mul(A, 1.0, h10));
            lincomb_kernel<<<nb1(ldL * U), 256, 0, st>>>(ldL * U, 1.0, h11, 1.0, c->LUlast[k], h11);
            coldot_kernel<<<nb1(U, 8), 256, 0, st>>>(h00, h00, L, U, ldL, 1.0, 1.0, acc);
            coldot_kernel<<<nb1(U, 8), 256, 0, st>>>(h10, h10, L, U, ldL, 1.0, 1.0, acc);
            coldot_kernel<<<nb1(U, 8), 256, 0, st>>>(h01, h01, L, U, ldL, 1.0, 1.0, acc);
            coldot_kernel<<<nb1(U, 8), 256, 0, st>>>(h11, h11, L, U, ldL, 1.0, 1.0, acc);
            coldot_kernel<<<nb1(U, 8), 256, 0, st>>>(h01, h00, L, U, ldL, 2.0, 1.0, acc + (s + 1) * U);
            coldot_kernel<<<nb1(U, 8), 256, 0, st>>>(h11, h10, L, U, ldL, 2.0, 1.0, acc + (s + 1) * U);
            g_dense_launches += 7;
            HYB_CUDA(cudaGetLastError());
        }
    }
    lincomb_kernel<<<nb1(dim), 256, 0, st>>>(dim, 1.0, acc, 0.0, nullptr, out);
    DLAUNCHED();
    HYB_CUDA(cudaGetLastError());
    return 0;
}
