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

int64_t g_dense_launches = 0;
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
    HYB_CHECK(hyb_potrf(st, c->chX, nullptr));
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
