// HBM-bound matrix-vector products with the resident q x N matrix G (= G*Q2).
//
// Replaces the dgemv call sites of the reference's solve path:
// qrchol.jl:59,80 (G'z, Gx in solve_subsystem3), systemsolvers/common.jl:95,98,151
// (apply_lhs, solve_system) and Solvers.jl:638,656 (residuals).
// Algorithmic traffic: 8*q*N bytes per product.  Fixed-order reductions (no atomics).
#include "common.cuh"

namespace {

// y = G x : thread per row (coalesced down the rows), columns split over blockIdx.y,
// partial sums to scratch[split][q]; second kernel adds the splits in fixed order.
__global__ void __launch_bounds__(128)
gemv_n_partial_kernel(int64_t q, int64_t N, const double* __restrict__ G, int64_t ld, const double* __restrict__ x,
                      double* __restrict__ scratch) {
    __shared__ double xs[512];
    const int64_t row0 = (int64_t)blockIdx.x * 256 + threadIdx.x;  // each thread owns rows row0 and row0 + 128
    const int split = blockIdx.y, nsplit = gridDim.y;
    const int64_t cols_per = (N + nsplit - 1) / nsplit;
    const int64_t c_begin = split * cols_per;
    const int64_t c_end = (c_begin + cols_per < N) ? c_begin + cols_per : N;
    double a0 = 0.0, a1 = 0.0;
    const bool ok0 = row0 < q, ok1 = row0 + 128 < q;
    for (int64_t cb = c_begin; cb < c_end; cb += 512) {
        const int nc = (int)((c_end - cb < 512) ? (c_end - cb) : 512);
        __syncthreads();
        for (int i = threadIdx.x; i < nc; i += 128) xs[i] = x[cb + i];
        __syncthreads();
        const double* p = G + row0 + cb * ld;
        int c = 0;
        for (; c + 3 < nc; c += 4) {
            double v00 = 0, v01 = 0, v02 = 0, v03 = 0, v10 = 0, v11 = 0, v12 = 0, v13 = 0;
            if (ok0) {
                v00 = __ldg(p + (int64_t)c * ld);
                v01 = __ldg(p + (int64_t)(c + 1) * ld);
                v02 = __ldg(p + (int64_t)(c + 2) * ld);
                v03 = __ldg(p + (int64_t)(c + 3) * ld);
            }
            if (ok1) {
                v10 = __ldg(p + 128 + (int64_t)c * ld);
                v11 = __ldg(p + 128 + (int64_t)(c + 1) * ld);
                v12 = __ldg(p + 128 + (int64_t)(c + 2) * ld);
                v13 = __ldg(p + 128 + (int64_t)(c + 3) * ld);
            }
            a0 += v00 * xs[c] + v01 * xs[c + 1] + v02 * xs[c + 2] + v03 * xs[c + 3];
            a1 += v10 * xs[c] + v11 * xs[c + 1] + v12 * xs[c + 2] + v13 * xs[c + 3];
        }
        for (; c < nc; ++c) {
            if (ok0) a0 += __ldg(p + (int64_t)c * ld) * xs[c];
            if (ok1) a1 += __ldg(p + 128 + (int64_t)c * ld) * xs[c];
        }
    }
    if (ok0) scratch[(int64_t)split * q + row0] = a0;
    if (ok1) scratch[(int64_t)split * q + row0 + 128] = a1;
}

__global__ void gemv_n_reduce_kernel(int64_t q, int nsplit, const double* __restrict__ scratch, double* __restrict__ y) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= q) return;
    double s = 0.0;
    for (int k = 0; k < nsplit; ++k) s += scratch[(int64_t)k * q + r];
    y[r] = s;
}

// y = G' z : one warp per column, lanes stride down the column with double2 loads.
__global__ void __launch_bounds__(256)
gemv_t_kernel(int64_t q, int64_t N, const double* __restrict__ G, int64_t ld, const double* __restrict__ z,
              double* __restrict__ y) {
    const int lane = threadIdx.x & 31;
    const int64_t col = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (col >= N) return;
    const double* p = G + col * ld;
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
    const int64_t q2 = q & ~(int64_t)1;
    int64_t r = 2 * lane;
    for (; r + 192 < q2; r += 256) {
        const double2 g0 = __ldg((const double2*)(p + r)), z0 = __ldg((const double2*)(z + r));
        const double2 g1 = __ldg((const double2*)(p + r + 64)), z1 = __ldg((const double2*)(z + r + 64));
        const double2 g2 = __ldg((const double2*)(p + r + 128)), z2 = __ldg((const double2*)(z + r + 128));
        const double2 g3 = __ldg((const double2*)(p + r + 192)), z3 = __ldg((const double2*)(z + r + 192));
        a0 += g0.x * z0.x + g0.y * z0.y;
        a1 += g1.x * z1.x + g1.y * z1.y;
        a2 += g2.x * z2.x + g2.y * z2.y;
        a3 += g3.x * z3.x + g3.y * z3.y;
    }
    for (; r < q2; r += 64) {
        const double2 g0 = __ldg((const double2*)(p + r)), z0 = __ldg((const double2*)(z + r));
        a0 += g0.x * z0.x + g0.y * z0.y;
    }
    if (lane == 0 && (q & 1)) a1 += p[q - 1] * z[q - 1];
    double v = (a0 + a1) + (a2 + a3);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) y[col] = v;
}

// ---- the same two products for a few right-hand sides at once (the batched direction solves of
// dirs.cu): G is read ONCE for all M columns, which is the whole point -- the products are bound by
// the 8*q*N bytes of G, not by the M vectors.
template <int M>
__global__ void __launch_bounds__(128)
gemm_n_partial_kernel(int64_t q, int64_t N, const double* __restrict__ G, int64_t ld, const double* __restrict__ X,
                      int64_t ldx, double* __restrict__ scratch) {
    __shared__ double xs[M][256];
    const int64_t row0 = (int64_t)blockIdx.x * 256 + threadIdx.x;
    const int split = blockIdx.y, nsplit = gridDim.y;
    const int64_t cols_per = (N + nsplit - 1) / nsplit;
    const int64_t c_begin = split * cols_per;
    const int64_t c_end = (c_begin + cols_per < N) ? c_begin + cols_per : N;
    double a0[M], a1[M];
#pragma unroll
    for (int m = 0; m < M; ++m) a0[m] = a1[m] = 0.0;
    const bool ok0 = row0 < q, ok1 = row0 + 128 < q;
    for (int64_t cb = c_begin; cb < c_end; cb += 256) {
        const int nc = (int)((c_end - cb < 256) ? (c_end - cb) : 256);
        __syncthreads();
#pragma unroll
        for (int m = 0; m < M; ++m)
            for (int i = threadIdx.x; i < nc; i += 128) xs[m][i] = X[cb + i + (int64_t)m * ldx];
        __syncthreads();
        const double* p = G + row0 + cb * ld;
        int c = 0;
        for (; c + 3 < nc; c += 4) {
            double v0[4], v1[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                v0[u] = ok0 ? __ldg(p + (int64_t)(c + u) * ld) : 0.0;
                v1[u] = ok1 ? __ldg(p + 128 + (int64_t)(c + u) * ld) : 0.0;
            }
#pragma unroll
            for (int m = 0; m < M; ++m) {
                a0[m] += v0[0] * xs[m][c] + v0[1] * xs[m][c + 1] + v0[2] * xs[m][c + 2] + v0[3] * xs[m][c + 3];
                a1[m] += v1[0] * xs[m][c] + v1[1] * xs[m][c + 1] + v1[2] * xs[m][c + 2] + v1[3] * xs[m][c + 3];
            }
        }
        for (; c < nc; ++c) {
            const double v0 = ok0 ? __ldg(p + (int64_t)c * ld) : 0.0;
            const double v1 = ok1 ? __ldg(p + 128 + (int64_t)c * ld) : 0.0;
#pragma unroll
            for (int m = 0; m < M; ++m) {
                a0[m] += v0 * xs[m][c];
                a1[m] += v1 * xs[m][c];
            }
        }
    }
#pragma unroll
    for (int m = 0; m < M; ++m) {
        double* sc = scratch + ((int64_t)m * nsplit + split) * q;
        if (ok0) sc[row0] = a0[m];
        if (ok1) sc[row0 + 128] = a1[m];
    }
}

__global__ void gemm_n_reduce_kernel(int64_t q, int nsplit, const double* __restrict__ scratch, double* __restrict__ Y,
                                     int64_t ldy) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int m = blockIdx.y;
    if (r >= q) return;
    double s = 0.0;
    for (int k = 0; k < nsplit; ++k) s += scratch[((int64_t)m * nsplit + k) * q + r];
    Y[r + (int64_t)m * ldy] = s;
}

template <int M>
__global__ void __launch_bounds__(256)
gemm_t_kernel(int64_t q, int64_t N, const double* __restrict__ G, int64_t ld, const double* __restrict__ Z, int64_t ldz,
              double* __restrict__ Y, int64_t ldy) {
    const int lane = threadIdx.x & 31;
    const int64_t col = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (col >= N) return;
    const double* p = G + col * ld;
    double a0[M], a1[M];
#pragma unroll
    for (int m = 0; m < M; ++m) a0[m] = a1[m] = 0.0;
    const int64_t q2 = q & ~(int64_t)1;
    int64_t r = 2 * lane;
    for (; r + 64 < q2; r += 128) {
        const double2 g0 = __ldg((const double2*)(p + r)), g1 = __ldg((const double2*)(p + r + 64));
#pragma unroll
        for (int m = 0; m < M; ++m) {
            const double2 z0 = __ldg((const double2*)(Z + r + (int64_t)m * ldz));
            const double2 z1 = __ldg((const double2*)(Z + r + 64 + (int64_t)m * ldz));
            a0[m] += g0.x * z0.x + g0.y * z0.y;
            a1[m] += g1.x * z1.x + g1.y * z1.y;
        }
    }
    for (; r < q2; r += 64) {
        const double2 g0 = __ldg((const double2*)(p + r));
#pragma unroll
        for (int m = 0; m < M; ++m) {
            const double2 z0 = __ldg((const double2*)(Z + r + (int64_t)m * ldz));
            a0[m] += g0.x * z0.x + g0.y * z0.y;
        }
    }
    if (lane == 0 && (q & 1)) {
#pragma unroll
        for (int m = 0; m < M; ++m) a1[m] += p[q - 1] * Z[q - 1 + (int64_t)m * ldz];
    }
#pragma unroll
    for (int m = 0; m < M; ++m) {
        double v = a0[m] + a1[m];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) Y[col + (int64_t)m * ldy] = v;
    }
}

}  // namespace

int hyb_gemv_n(cudaStream_t st, int64_t q, int64_t N, const double* G, int64_t ld, const double* x, double* y,
               double* scratch) {
    if (q == 0) return 0;
    if (N == 0) {
        HYB_CUDA(cudaMemsetAsync(y, 0, sizeof(double) * q, st));
        return 0;
    }
    dim3 grid((unsigned)((q + 255) / 256), HYB_GEMV_SPLITS);
    gemv_n_partial_kernel<<<grid, 128, 0, st>>>(q, N, G, ld, x, scratch);
    HYB_CUDA(cudaGetLastError());
    gemv_n_reduce_kernel<<<(unsigned)((q + 255) / 256), 256, 0, st>>>(q, HYB_GEMV_SPLITS, scratch, y);
    HYB_CUDA(cudaGetLastError());
    return 0;
}

int hyb_gemv_t(cudaStream_t st, int64_t q, int64_t N, const double* G, int64_t ld, const double* z, double* y) {
    if (N == 0) return 0;
    if (q == 0) {
        HYB_CUDA(cudaMemsetAsync(y, 0, sizeof(double) * N, st));
        return 0;
    }
    gemv_t_kernel<<<(unsigned)((N + 7) / 8), 256, 0, st>>>(q, N, G, ld, z, y);
    HYB_CUDA(cudaGetLastError());
    return 0;
}


// Y (q x M) = G X (N x M), M <= HYB_MAX_RHS ; scratch must hold M * HYB_GEMV_SPLITS * q doubles
int hyb_gemm_n_small(cudaStream_t st, int64_t q, int64_t N, const double* G, int64_t ld, const double* X, int64_t ldx,
                     int M, double* Y, int64_t ldy, double* scratch) {
    if (q == 0 || M <= 0) return 0;
    if (N == 0) {
        HYB_CUDA(cudaMemset2DAsync(Y, sizeof(double) * ldy, 0, sizeof(double) * q, M, st));
        return 0;
    }
    dim3 grid((unsigned)((q + 255) / 256), HYB_GEMV_SPLITS);
    switch (M) {
        case 1: gemm_n_partial_kernel<1><<<grid, 128, 0, st>>>(q, N, G, ld, X, ldx, scratch); break;
        case 2: gemm_n_partial_kernel<2><<<grid, 128, 0, st>>>(q, N, G, ld, X, ldx, scratch); break;
        case 3: gemm_n_partial_kernel<3><<<grid, 128, 0, st>>>(q, N, G, ld, X, ldx, scratch); break;
        case 4: gemm_n_partial_kernel<4><<<grid, 128, 0, st>>>(q, N, G, ld, X, ldx, scratch); break;
        default: hyb_set_error("hyb_gemm_n_small: M = %d out of range", M); return -3;
    }
    HYB_CUDA(cudaGetLastError());
    gemm_n_reduce_kernel<<<dim3((unsigned)((q + 255) / 256), M), 256, 0, st>>>(q, HYB_GEMV_SPLITS, scratch, Y, ldy);
    HYB_CUDA(cudaGetLastError());
    return 0;
}

// Y (N x M) = G' Z (q x M) ; Z columns 16-byte aligned (even ldz)
int hyb_gemm_t_small(cudaStream_t st, int64_t q, int64_t N, const double* G, int64_t ld, const double* Z, int64_t ldz,
                     int M, double* Y, int64_t ldy) {
    if (N == 0 || M <= 0) return 0;
    if (q == 0) {
        HYB_CUDA(cudaMemset2DAsync(Y, sizeof(double) * ldy, 0, sizeof(double) * N, M, st));
        return 0;
    }
    const unsigned grid = (unsigned)((N + 7) / 8);
    switch (M) {
        case 1: gemm_t_kernel<1><<<grid, 256, 0, st>>>(q, N, G, ld, Z, ldz, Y, ldy); break;
        case 2: gemm_t_kernel<2><<<grid, 256, 0, st>>>(q, N, G, ld, Z, ldz, Y, ldy); break;
        case 3: gemm_t_kernel<3><<<grid, 256, 0, st>>>(q, N, G, ld, Z, ldz, Y, ldy); break;
        case 4: gemm_t_kernel<4><<<grid, 256, 0, st>>>(q, N, G, ld, Z, ldz, Y, ldy); break;
        default: hyb_set_error("hyb_gemm_t_small: M = %d out of range", M); return -3;
    }
    HYB_CUDA(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------
// y = G x without any allocation: the split-sum scratch of the gemv above lives in a
// fixed, library-owned device array instead of a buffer from cudaMalloc.
//
// Who needs this: a host whose device memory is entirely managed by another allocator
// (for instance a Julia session in which CUDA.jl's memory pool, or a PyTorch process whose
// caching allocator, has already claimed the GPU, so that a foreign cudaMalloc may fail or
// fragment the pool), and callers inside a latency-critical section, where cudaMalloc and
// cudaFree are unwelcome because they synchronise the device.  Such a caller owns G, x and y
// (device pointers, on the current device) and gets the product with no allocation made by
// this library, by torch, or by any routine this function calls.
//
// The scratch is HYB_GEMV_SPLITS * HYB_NOALLOC_MAX_Q doubles (4 MiB).  Every GPU that loads
// the library holds that array until the process exits.  All calls on a device share it, so
// two calls that overlap on the GPU (different streams, or different host threads) race on
// it: serialise them.  It keeps whatever the previous call left behind; that is harmless
// here because each call writes every scratch entry it later reads (all HYB_GEMV_SPLITS
// partial sums of rows 0..q-1), so nothing has to start at zero and nothing is cleared.
//
// Set-up: the array's name is only a host-side handle; the device address is looked up with
// cudaGetSymbolAddress once per device, by the first call made while that device is current.
// Make that first call outside any stream capture.
// ---------------------------------------------------------------------------------------
#define HYB_NOALLOC_MAX_Q 65536
#define HYB_NOALLOC_MAX_DEVICES 64

static __device__ __align__(256) double g_noalloc_scratch[(size_t)HYB_GEMV_SPLITS * HYB_NOALLOC_MAX_Q];
static double* g_noalloc_ptr[HYB_NOALLOC_MAX_DEVICES] = {nullptr};

extern "C" int hyb_gemv_n_noalloc(const double* G_dev, int64_t q, int64_t N, int64_t ld, const double* x_dev,
                                  double* y_dev, void* stream) {
    if (q > HYB_NOALLOC_MAX_Q) {
        hyb_set_error("hyb_gemv_n_noalloc: q = %lld exceeds the fixed scratch capacity of %d rows; use hyb_G_mul instead",
                      (long long)q, HYB_NOALLOC_MAX_Q);
        return -4;
    }
    if (q < 0 || N < 0 || ld < q || (q > 0 && N > 0 && (!G_dev || !x_dev)) || (q > 0 && !y_dev)) {
        hyb_set_error("hyb_gemv_n_noalloc: bad arguments (q=%lld N=%lld ld=%lld)", (long long)q, (long long)N, (long long)ld);
        return -3;
    }
    if (q == 0) return 0;
    int dev = -1;
    HYB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= HYB_NOALLOC_MAX_DEVICES) {
        hyb_set_error("hyb_gemv_n_noalloc: device index %d out of range", dev);
        return -3;
    }
    if (!g_noalloc_ptr[dev]) {
        double* ptr = nullptr;
        HYB_CUDA(cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch));
        g_noalloc_ptr[dev] = ptr;
    }
    cudaStream_t st = (cudaStream_t)stream;
    if (N == 0) {
        HYB_CUDA(cudaMemsetAsync(y_dev, 0, sizeof(double) * q, st));
        return 0;
    }
    double* scratch = g_noalloc_ptr[dev];
    dim3 grid((unsigned)((q + 255) / 256), HYB_GEMV_SPLITS);
    gemv_n_partial_kernel<<<grid, 128, 0, st>>>(q, N, G_dev, ld, x_dev, scratch);
    HYB_CUDA(cudaGetLastError());
    gemv_n_reduce_kernel<<<(unsigned)((q + 255) / 256), 256, 0, st>>>(q, HYB_GEMV_SPLITS, scratch, y_dev);
    HYB_CUDA(cudaGetLastError());
    return 0;
}
