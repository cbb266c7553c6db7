// Dense FP64 Cholesky (upper, column-major) and triangular solves for sm_100a.
//
// Replaces cholesky!(Symmetric(:U), check=false) -> dpotrf('U') of the reference
// (src/linearalgebra/dense.jl:226, called from qrchol.jl:257-258) and
// ldiv!(x, fact, rhs) -> dpotrs (qrchol.jl:75).
//
// Blocked right-looking factorization A = U'U with NB = 128:
//   panel  : one CTA factors the 128x128 diagonal block in shared memory and also
//            writes inv(U11) (kept: the solves reuse it),
//   row    : U12 = inv(U11)' * A12          -> DMMA gemm_tn (in place, M = 128),
//   update : A22 -= U12' * U12 (upper only) -> DMMA gemm_tn (SYRK form).
// Every contraction is K-major on both sides because U is stored column-major and
// the contraction runs down columns, so the same TMA/DMMA tile engine serves all.
//
// Solves: two persistent "wavefront" kernels (forward U'y = b, backward Ux = y),
// one CTA per 128-column block, cooperative launch so all CTAs are co-resident;
// a CTA consumes the x-blocks it depends on as their ready flags are published,
// multiplies by the stored inverse diagonal block, and publishes its own block.
// Fixed summation order -> bitwise reproducible.
#include <cooperative_groups.h>

#include "common.cuh"

namespace {

constexpr int NB = 128;
constexpr int LDS = NB + 1;  // padded shared-memory leading dimension

// --------------------------------------------------------------------------------
// panel: in-place upper Cholesky of an nb x nb block (nb <= 128) + its inverse
// --------------------------------------------------------------------------------
__global__ void __launch_bounds__(512, 1)
potf2_trtri_kernel(double* __restrict__ A, int64_t lda, int nb, double* __restrict__ Tinv, int* __restrict__ info,
                   int k0) {
    // S[r * LDS + c]: upper triangle (r <= c) holds A then U; once U is final the strictly
    // lower triangle receives inv(U) transposed (S[c][r] = inv(U)[r][c], r < c) and Td its diagonal.
    extern __shared__ double sm[];
    double* S = sm;
    double* Td = sm + NB * LDS;
    __shared__ int fail;
    const int tid = threadIdx.x, nth = blockDim.x;
    if (*info != 0) return;  // an earlier panel failed: nothing meaningful to do
    if (tid == 0) fail = 0;
    for (int e = tid; e < NB * NB; e += nth) {
        const int r = e % NB, c = e / NB;
        S[r * LDS + c] = (r < nb && c < nb && r <= c) ? A[r + (int64_t)c * lda] : 0.0;
    }
    __syncthreads();

    // right-looking unblocked factorization, rows of U
    for (int j = 0; j < nb; ++j) {
        const double d = S[j * LDS + j];
        if (!(d > 0.0)) {  // also catches NaN
            if (tid == 0) {
                fail = 1;
                *info = k0 + j + 1;
            }
            break;  // uniform: every thread read the same d
        }
        const double r = sqrt(d);
        const double rinv = 1.0 / r;
        __syncthreads();
        for (int c = j + tid; c < nb; c += nth) S[j * LDS + c] = (c == j) ? r : S[j * LDS + c] * rinv;
        __syncthreads();
        // trailing update: S[i][c] -= U[j][i] * U[j][c], j < i <= c
        const int m = nb - j - 1;
        for (int e = tid; e < m * m; e += nth) {
            const int i = j + 1 + e / m, c = j + 1 + e % m;
            if (c >= i) S[i * LDS + c] -= S[j * LDS + i] * S[j * LDS + c];
        }
        __syncthreads();
    }
    __syncthreads();
    if (fail) return;

    // write U11 back (upper triangle; the strictly lower part of the block is left untouched)
    for (int e = tid; e < nb * nb; e += nth) {
        const int r = e % nb, c = e / nb;
        if (r <= c) A[r + (int64_t)c * lda] = S[r * LDS + c];
    }
    if (tid < NB) Td[tid] = (tid < nb) ? 1.0 / S[tid * LDS + tid] : 0.0;
    __syncthreads();

    // inv(U): thread c back-substitutes column c, t[i] = -(sum_{k=i+1..c} U[i][k] t[k]) / U[i][i].
    // Threads walk i in lockstep: U[i][k] is a broadcast read, t[k] = S[c][k] is private to thread c.
    if (tid < nb) {
        const int c = tid;
        const double tcc = Td[c];
        for (int i = c - 1; i >= 0; --i) {
            double s0 = S[i * LDS + c] * tcc, s1 = 0.0, s2 = 0.0, s3 = 0.0;
            int k = i + 1;
            for (; k + 3 < c; k += 4) {
                s0 += S[i * LDS + k] * S[c * LDS + k];
                s1 += S[i * LDS + k + 1] * S[c * LDS + k + 1];
                s2 += S[i * LDS + k + 2] * S[c * LDS + k + 2];
                s3 += S[i * LDS + k + 3] * S[c * LDS + k + 3];
            }
            for (; k < c; ++k) s0 += S[i * LDS + k] * S[c * LDS + k];
            S[c * LDS + i] = -((s0 + s1) + (s2 + s3)) * Td[i];
        }
    }
    __syncthreads();
    // Tinv block: dense 128 x 128 column-major tile, zero outside the nb x nb upper triangle
    for (int e = tid; e < NB * NB; e += nth) {
        const int r = e % NB, c = e / NB;
        double v = 0.0;
        if (r < nb && c < nb) v = (r < c) ? S[c * LDS + r] : ((r == c) ? Td[c] : 0.0);
        Tinv[r + c * NB] = v;
    }
}

__global__ void copy_upper_kernel(const double* __restrict__ src, int64_t lds_, double* __restrict__ dst, int64_t ldd,
                                  int64_t N) {
    const int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    const int64_t c = blockIdx.y;
    if (r < N && r <= c) dst[r + c * ldd] = src[r + c * lds_];
}

// --------------------------------------------------------------------------------
// wavefront triangular solves
// --------------------------------------------------------------------------------
__device__ __forceinline__ int ld_acquire(const int* p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release(int* p, int v) {
    asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// forward: solve U' y = b in place (x holds b on entry, y on exit)
__global__ void __launch_bounds__(256, 1)
trsv_fwd_kernel(const double* __restrict__ U, int64_t ld, const double* __restrict__ Tinv, double* x, int N, int nblk,
                int* flags, int epoch) {
    __shared__ double ys[NB];
    __shared__ double rs[NB];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int ib = blockIdx.x; ib < nblk; ib += gridDim.x) {
        double acc[16];
#pragma unroll
        for (int c = 0; c < 16; ++c) acc[c] = 0.0;
        const int col0 = ib * NB + warp * 16;
        for (int jb = 0; jb < ib; ++jb) {
            if (tid == 0) {
                while (ld_acquire(&flags[jb]) != epoch) {
                }
            }
            __syncthreads();
            if (tid < NB) ys[tid] = __ldcg(&x[jb * NB + tid]);
            __syncthreads();
            const double y0 = ys[2 * lane], y1 = ys[2 * lane + 1], y2 = ys[64 + 2 * lane], y3 = ys[65 + 2 * lane];
            const double* base = U + (int64_t)jb * NB + 2 * lane;
#pragma unroll
            for (int c = 0; c < 16; ++c) {
                const int col = col0 + c;
                if (col < N) {
                    const double2 v0 = __ldg((const double2*)(base + (int64_t)col * ld));
                    const double2 v1 = __ldg((const double2*)(base + (int64_t)col * ld + 64));
                    acc[c] += (v0.x * y0 + v0.y * y1) + (v1.x * y2 + v1.y * y3);
                }
            }
        }
#pragma unroll
        for (int c = 0; c < 16; ++c) {
            double v = acc[c];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            acc[c] = v;
        }
        __syncthreads();  // everyone is done with ys
        if (lane == 0) {
#pragma unroll
            for (int c = 0; c < 16; ++c) {
                const int col = col0 + c;
                rs[warp * 16 + c] = (col < N) ? (x[col] - acc[c]) : 0.0;
            }
        }
        __syncthreads();
        // y_ib = Tinv_ib' * r : column dots with the (upper, zero-padded) inverse block
        {
            const double r0 = rs[2 * lane], r1 = rs[2 * lane + 1], r2 = rs[64 + 2 * lane], r3 = rs[65 + 2 * lane];
            const double* tb = Tinv + (int64_t)ib * NB * NB + 2 * lane;
#pragma unroll
            for (int c = 0; c < 16; ++c) {
                const int lc = warp * 16 + c;
                const double2 v0 = __ldg((const double2*)(tb + lc * NB));
                const double2 v1 = __ldg((const double2*)(tb + lc * NB + 64));
                double v = (v0.x * r0 + v0.y * r1) + (v1.x * r2 + v1.y * r3);
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                const int col = col0 + c;
                if (lane == 0 && col < N) x[col] = v;
            }
        }
        __syncthreads();
        if (tid == 0) {
            __threadfence();
            st_release(&flags[ib], epoch);
        }
    }
}

// backward: solve U x = y in place
__global__ void __launch_bounds__(256, 1)
trsv_bwd_kernel(const double* __restrict__ U, int64_t ld, const double* __restrict__ Tinv, double* x, int N, int nblk,
                int* flags, int epoch) {
    __shared__ double xs[NB];
    __shared__ double part[2][NB];
    const int tid = threadIdx.x, r = tid & (NB - 1), half = tid >> 7;
    for (int ib = nblk - 1 - (int)blockIdx.x; ib >= 0; ib -= gridDim.x) {
        const int row = ib * NB + r;
        const bool row_ok = row < N;
        double acc = 0.0;
        for (int jb = nblk - 1; jb > ib; --jb) {
            if (tid == 0) {
                while (ld_acquire(&flags[jb]) != epoch) {
                }
            }
            __syncthreads();
            if (tid < NB) xs[tid] = (jb * NB + tid < N) ? __ldcg(&x[jb * NB + tid]) : 0.0;
            __syncthreads();
            if (row_ok) {
                const int ncol = min(NB, N - jb * NB);
                const double* p = U + row + (int64_t)(jb * NB + half * 64) * ld;
                const int cend = min(64, ncol - half * 64);
                double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
                int c = 0;
                for (; c + 3 < cend; c += 4) {
                    a0 += __ldg(p + (int64_t)c * ld) * xs[half * 64 + c];
                    a1 += __ldg(p + (int64_t)(c + 1) * ld) * xs[half * 64 + c + 1];
                    a2 += __ldg(p + (int64_t)(c + 2) * ld) * xs[half * 64 + c + 2];
                    a3 += __ldg(p + (int64_t)(c + 3) * ld) * xs[half * 64 + c + 3];
                }
                for (; c < cend; ++c) a0 += __ldg(p + (int64_t)c * ld) * xs[half * 64 + c];
                acc += (a0 + a1) + (a2 + a3);
            }
        }
        __syncthreads();
        part[half][r] = acc;
        __syncthreads();
        if (tid < NB) xs[tid] = row_ok ? (x[row] - (part[0][tid] + part[1][tid])) : 0.0;  // rhs block
        __syncthreads();
        // x_ib = Tinv_ib * rhs (row form over the upper-triangular inverse block)
        {
            const double* tb = Tinv + (int64_t)ib * NB * NB + r + (int64_t)(half * 64) * NB;
            double a0 = 0.0, a1 = 0.0;
            for (int c = 0; c < 64; c += 2) {
                a0 += __ldg(tb + c * NB) * xs[half * 64 + c];
                a1 += __ldg(tb + (c + 1) * NB) * xs[half * 64 + c + 1];
            }
            part[half][r] = a0 + a1;
        }
        __syncthreads();
        if (tid < NB && row_ok) x[row] = part[0][tid] + part[1][tid];
        __syncthreads();
        if (tid == 0) {
            __threadfence();
            st_release(&flags[ib], epoch);
        }
    }
}

int g_num_sms = 0;

}  // namespace

int hyb_chol_alloc(HybChol& ch, int64_t N) {
    ch.N = N;
    ch.ld = hyb_round_up(N > 0 ? N : 1, 16);
    ch.nblk = (N + NB - 1) / NB;
    HYB_CUDA(cudaMalloc(&ch.F, sizeof(double) * ch.ld * (N > 0 ? N : 1)));
    HYB_CUDA(cudaMalloc(&ch.Tinv, sizeof(double) * NB * NB * (ch.nblk > 0 ? ch.nblk : 1)));
    HYB_CUDA(cudaMalloc(&ch.flags, sizeof(int) * (ch.nblk > 0 ? ch.nblk : 1)));
    HYB_CUDA(cudaMemset(ch.flags, 0, sizeof(int) * (ch.nblk > 0 ? ch.nblk : 1)));
    HYB_CUDA(cudaMalloc(&ch.info, sizeof(int)));
    HYB_CUDA(cudaMalloc(&ch.work, sizeof(double) * (N > 0 ? N : 1)));
    ch.epoch = 0;
    return 0;
}

void hyb_chol_free(HybChol& ch) {
    cudaFree(ch.F);
    cudaFree(ch.Tinv);
    cudaFree(ch.flags);
    cudaFree(ch.info);
    cudaFree(ch.work);
    ch = HybChol();
}

int hyb_potrf(cudaStream_t st, HybChol& ch, int* info_out) {
    const int64_t N = ch.N, ld = ch.ld;
    static bool attr_set = false;
    const int panel_smem = (NB * LDS + NB) * (int)sizeof(double);
    if (!attr_set) {
        HYB_CUDA(cudaFuncSetAttribute(potf2_trtri_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, panel_smem));
        attr_set = true;
    }
    HYB_CUDA(cudaMemsetAsync(ch.info, 0, sizeof(int), st));
    for (int64_t k0 = 0, b = 0; k0 < N; k0 += NB, ++b) {
        const int nb = (int)((N - k0 < NB) ? (N - k0) : NB);
        double* A11 = ch.F + k0 + k0 * ld;
        double* Tb = ch.Tinv + b * NB * NB;
        potf2_trtri_kernel<<<1, 512, panel_smem, st>>>(A11, ld, nb, Tb, ch.info, (int)k0);
        HYB_CUDA(cudaGetLastError());
        const int64_t n2 = N - k0 - nb;
        if (n2 > 0) {
            double* A12 = ch.F + k0 + (k0 + nb) * ld;
            double* A22 = ch.F + (k0 + nb) + (k0 + nb) * ld;
            // U12 = inv(U11)' A12  (in place: one 128-row tile, each CTA owns its columns)
            HYB_CHECK(hyb_gemm_tn(st, nb, n2, nb, 1.0, Tb, NB, A12, ld, 0.0, A12, ld, 0));
            // A22 -= U12' U12  (upper)
            HYB_CHECK(hyb_gemm_tn(st, n2, n2, nb, -1.0, A12, ld, A12, ld, 1.0, A22, ld, HYB_GEMM_UPPER));
        }
    }
    if (info_out) {
        HYB_CUDA(cudaMemcpyAsync(info_out, ch.info, sizeof(int), cudaMemcpyDeviceToHost, st));
        HYB_CUDA(cudaStreamSynchronize(st));
    }
    return 0;
}

int hyb_potrs(cudaStream_t st, HybChol& ch, double* x) {
    if (ch.N == 0) return 0;
    if (g_num_sms == 0) {
        int dev = 0;
        HYB_CUDA(cudaGetDevice(&dev));
        HYB_CUDA(cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev));
    }
    int N = (int)ch.N, nblk = (int)ch.nblk;
    int grid = nblk < g_num_sms ? nblk : g_num_sms;
    const double* U = ch.F;
    const double* Tinv = ch.Tinv;
    int64_t ld = ch.ld;
    int* flags = ch.flags;
    int epoch = ++ch.epoch;
    void* args[] = {(void*)&U, (void*)&ld, (void*)&Tinv, (void*)&x, (void*)&N, (void*)&nblk, (void*)&flags, (void*)&epoch};
    HYB_CUDA(cudaLaunchCooperativeKernel((void*)trsv_fwd_kernel, dim3(grid), dim3(256), args, 0, st));
    int epoch2 = ++ch.epoch;
    void* args2[] = {(void*)&U, (void*)&ld, (void*)&Tinv, (void*)&x, (void*)&N, (void*)&nblk, (void*)&flags, (void*)&epoch2};
    HYB_CUDA(cudaLaunchCooperativeKernel((void*)trsv_bwd_kernel, dim3(grid), dim3(256), args2, 0, st));
    return 0;
}

// copy the upper triangle of an N x N matrix (posdef_fact_copy!'s copyto!, dense.jl:232)
int hyb_copy_upper(cudaStream_t st, const double* src, int64_t lds_, double* dst, int64_t ldd, int64_t N) {
    if (N == 0) return 0;
    dim3 grid((unsigned)((N + 255) / 256), (unsigned)N);
    copy_upper_kernel<<<grid, 256, 0, st>>>(src, lds_, dst, ldd, N);
    HYB_CUDA(cudaGetLastError());
    return 0;
}
