// Column-pivoted Householder QR on the device: the `qr!(M, ColumnNorm())` of the reference's
// preprocessing (src/Solvers/process.jl:109 for [A; G], :233 for A'), i.e. LAPACK dgeqp3.
//
// Why it is here: at Schur dimension 8192 the reference-style preprocessing of the 16368 x 8192
// matrix G takes about a minute on the host cores while the whole interior-point loop takes under a
// second on the GPU ("the dominant setup cost").  The factorization is bound by the two
// passes over the trailing matrix that every Householder step makes (w = v'A, A -= tau v w'): HBM
// traffic 16 (m - i)(n - i) bytes at step i, about 0.83 * 16 * n^3 bytes in total for m = 2n.
//
// Algorithm = LAPACK dlaqp2 (the unblocked kernel of dgeqp3), same conventions so the output drops
// into dormqr / the host code unchanged: R on and above the diagonal, Householder vectors below it
// (v[0] = 1 implicit), tau, and the pivot vector; pivot = first column of maximal partial norm;
// partial norms are down-dated with the dlaqp2 safeguard (recomputed when cancellation is detected).
// Two launches per step, nothing goes back to the host until the end.
#include "ctx.cuh"

namespace {

constexpr int QR_THREADS = 1024;
constexpr int QR_WARPS = 8;  // columns per CTA in the trailing update

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_max_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
// 2-norm of x[0:len) by one warp, scaled like dnrm2 (no overflow / underflow of the squares)
__device__ __forceinline__ double warp_nrm2(const double* x, int64_t len, int lane) {
    double amax = 0.0;
    for (int64_t r = lane; r < len; r += 32) amax = fmax(amax, fabs(x[r]));
    amax = warp_max_d(amax);
    if (!(amax > 0.0)) return 0.0;
    const double inv = 1.0 / amax;
    double ssq = 0.0;
    for (int64_t r = lane; r < len; r += 32) {
        const double t = x[r] * inv;
        ssq += t * t;
    }
    return amax * sqrt(warp_sum_d(ssq));
}

// initial column norms (dgeqp3: vn1 = vn2 = dnrm2 of every column), identity pivots
__global__ void __launch_bounds__(QR_WARPS * 32)
qr_init_kernel(int64_t m, int64_t n, const double* __restrict__ A, int64_t ld, double* vn1, double* vn2, int* jpvt) {
    const int lane = threadIdx.x & 31;
    const int64_t j = (int64_t)blockIdx.x * QR_WARPS + (threadIdx.x >> 5);
    if (j >= n) return;
    const double nr = warp_nrm2(A + j * ld, m, lane);
    if (lane == 0) {
        vn1[j] = nr;
        vn2[j] = nr;
        jpvt[j] = (int)j;
    }
}

// step i, part 1 (one CTA): pivot search, column swap, Householder vector of column i (dlarfg)
__global__ void __launch_bounds__(QR_THREADS)
qr_pivot_house_kernel(int64_t m, int64_t n, int64_t i, double* __restrict__ A, int64_t ld, double* vn1, double* vn2,
                      int* jpvt, double* tau, double* __restrict__ vbuf) {
    __shared__ double sv[32];
    __shared__ int si[32];
    __shared__ double sred[32];
    __shared__ int pvt_s;
    __shared__ double bcast[2];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // (1) idamax over vn1[i:n): first index of the maximum
    double best = -1.0;
    int bidx = 0x7fffffff;
    for (int64_t j = i + tid; j < n; j += QR_THREADS) {
        const double v = vn1[j];
        if (v > best) {
            best = v;
            bidx = (int)j;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bidx, o);
        if (ov > best || (ov == best && oi < bidx)) {
            best = ov;
            bidx = oi;
        }
    }
    if (lane == 0) {
        sv[warp] = best;
        si[warp] = bidx;
    }
    __syncthreads();
    if (warp == 0) {
        best = sv[lane];
        bidx = si[lane];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double ov = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bidx, o);
            if (ov > best || (ov == best && oi < bidx)) {
                best = ov;
                bidx = oi;
            }
        }
        if (lane == 0) pvt_s = (bidx == 0x7fffffff) ? (int)i : bidx;
    }
    __syncthreads();
    const int64_t pvt = pvt_s;
    // (2) swap columns i and pvt (all m rows, like dswap in dlaqp2), pivots and norms
    double* ci = A + i * ld;
    if (pvt != i) {
        double* cp = A + pvt * ld;
        for (int64_t r = tid; r < m; r += QR_THREADS) {
            const double t = ci[r];
            ci[r] = cp[r];
            cp[r] = t;
        }
        if (tid == 0) {
            const int t = jpvt[pvt];
            jpvt[pvt] = jpvt[i];
            jpvt[i] = t;
            vn1[pvt] = vn1[i];
            vn2[pvt] = vn2[i];
        }
    }
    __syncthreads();
    // (3) dlarfg on ci[i:m): beta = -sign(alpha) * ||(alpha, x)||, tau = (beta - alpha) / beta, v = x / (alpha - beta)
    const int64_t len = m - i - 1;  // entries below the diagonal
    double amax = 0.0;
    for (int64_t r = tid; r < len; r += QR_THREADS) amax = fmax(amax, fabs(ci[i + 1 + r]));
    amax = warp_max_d(amax);
    if (lane == 0) sred[warp] = amax;
    __syncthreads();
    if (warp == 0) {
        amax = warp_max_d(sred[lane]);
        if (lane == 0) bcast[0] = amax;
    }
    __syncthreads();
    amax = bcast[0];
    double xnorm = 0.0;
    if (amax > 0.0) {
        const double inv = 1.0 / amax;
        double ssq = 0.0;
        for (int64_t r = tid; r < len; r += QR_THREADS) {
            const double t = ci[i + 1 + r] * inv;
            ssq += t * t;
        }
        ssq = warp_sum_d(ssq);
        __syncthreads();
        if (lane == 0) sred[warp] = ssq;
        __syncthreads();
        if (warp == 0) {
            ssq = warp_sum_d(sred[lane]);
            if (lane == 0) bcast[1] = amax * sqrt(ssq);
        }
        __syncthreads();
        xnorm = bcast[1];
    }
    const double alpha = ci[i];
    __syncthreads();
    if (xnorm == 0.0) {
        if (tid == 0) {
            tau[i] = 0.0;
            vbuf[0] = 1.0;
        }
        for (int64_t r = tid; r < len; r += QR_THREADS) vbuf[1 + r] = 0.0;
        return;
    }
    const double beta = -copysign(hypot(alpha, xnorm), alpha);
    const double scal = 1.0 / (alpha - beta);
    for (int64_t r = tid; r < len; r += QR_THREADS) {
        const double v = ci[i + 1 + r] * scal;
        ci[i + 1 + r] = v;
        vbuf[1 + r] = v;
    }
    if (tid == 0) {
        tau[i] = (beta - alpha) / beta;
        ci[i] = beta;
        vbuf[0] = 1.0;
    }
}

// step i, part 2: apply H = I - tau v v' to the trailing columns (dlarf) and down-date their partial
// norms (dlaqp2's safeguarded update).  One warp per column.
__global__ void __launch_bounds__(QR_WARPS * 32)
qr_apply_kernel(int64_t m, int64_t n, int64_t i, double* __restrict__ A, int64_t ld, double* vn1, double* vn2,
                const double* __restrict__ tau, const double* __restrict__ vbuf) {
    const int lane = threadIdx.x & 31;
    const int64_t j = i + 1 + (int64_t)blockIdx.x * QR_WARPS + (threadIdx.x >> 5);
    if (j >= n) return;
    double* col = A + j * ld + i;
    const int64_t len = m - i;
    const double t = tau[i];
    if (t != 0.0) {
        double w = 0.0;
        for (int64_t r = lane; r < len; r += 32) w += vbuf[r] * col[r];
        w = warp_sum_d(w) * t;
        for (int64_t r = lane; r < len; r += 32) col[r] -= w * vbuf[r];
    }
    __syncwarp();
    const double v1 = vn1[j];
    if (v1 != 0.0) {
        double temp = fabs(col[0]) / v1;
        temp = 1.0 - temp * temp;
        temp = temp > 0.0 ? temp : 0.0;
        const double ratio = v1 / vn2[j];
        const double temp2 = temp * ratio * ratio;
        if (temp2 <= 1.4901161193847656e-08) {  // tol3z = sqrt(eps): cancellation, recompute
            const double nr = (len > 1) ? warp_nrm2(col + 1, len - 1, lane) : 0.0;
            if (lane == 0) {
                vn1[j] = nr;
                vn2[j] = nr;
            }
        } else if (lane == 0) {
            vn1[j] = v1 * sqrt(temp);
        }
    }
}

}  // namespace

extern "C" int hyb_geqp3(hyb_ctx* c, int64_t m, int64_t n, double* A, int64_t lda, int* jpvt, double* tau) {
    if (!c) return -1;
    if (m < 0 || n < 0 || lda < (m > 1 ? m : 1)) {
        hyb_set_error("hyb_geqp3: bad dimensions m=%lld n=%lld lda=%lld", (long long)m, (long long)n, (long long)lda);
        return -3;
    }
    const int64_t k = m < n ? m : n;
    if (k == 0) {
        for (int64_t j = 0; j < n; ++j) jpvt[j] = (int)j;
        return 0;
    }
    HYB_CUDA(cudaSetDevice(c->device));
    cudaStream_t st = c->st;
    const int64_t ld = hyb_round_up(m, 16);
    double *dA = nullptr, *dv = nullptr, *dtau = nullptr, *vn1 = nullptr, *vn2 = nullptr;
    int* dp = nullptr;
    int rc = 0;
    auto cleanup = [&]() {
        cudaFree(dA);
        cudaFree(dv);
        cudaFree(dtau);
        cudaFree(vn1);
        cudaFree(vn2);
        cudaFree(dp);
    };
#define QR_CUDA(call)                                                                        \
    do {                                                                                     \
        cudaError_t _e = (call);                                                             \
        if (_e != cudaSuccess) {                                                             \
            hyb_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(_e)); \
            cleanup();                                                                       \
            return -1;                                                                       \
        }                                                                                    \
    } while (0)
    QR_CUDA(cudaMalloc(&dA, sizeof(double) * ld * n));
    QR_CUDA(cudaMalloc(&dv, sizeof(double) * (m + 1)));
    QR_CUDA(cudaMalloc(&dtau, sizeof(double) * k));
    QR_CUDA(cudaMalloc(&vn1, sizeof(double) * n));
    QR_CUDA(cudaMalloc(&vn2, sizeof(double) * n));
    QR_CUDA(cudaMalloc(&dp, sizeof(int) * n));
    QR_CUDA(cudaMemcpy2DAsync(dA, sizeof(double) * ld, A, sizeof(double) * lda, sizeof(double) * m, n,
                              cudaMemcpyHostToDevice, st));
    c->h2d_bytes += (int64_t)sizeof(double) * m * n;
    KBEGIN(c);
    qr_init_kernel<<<(unsigned)((n + QR_WARPS - 1) / QR_WARPS), QR_WARPS * 32, 0, st>>>(m, n, dA, ld, vn1, vn2, dp);
    ++g_misc_launches;
    for (int64_t i = 0; i < k; ++i) {
        qr_pivot_house_kernel<<<1, QR_THREADS, 0, st>>>(m, n, i, dA, ld, vn1, vn2, dp, dtau, dv);
        const int64_t ntr = n - i - 1;
        if (ntr > 0)
            qr_apply_kernel<<<(unsigned)((ntr + QR_WARPS - 1) / QR_WARPS), QR_WARPS * 32, 0, st>>>(m, n, i, dA, ld, vn1, vn2,
                                                                                                 dtau, dv);
        g_misc_launches += ntr > 0 ? 2 : 1;
    }
    QR_CUDA(cudaGetLastError());
    KEND(c);
    QR_CUDA(cudaMemcpy2DAsync(A, sizeof(double) * lda, dA, sizeof(double) * ld, sizeof(double) * m, n, cudaMemcpyDeviceToHost,
                              st));
    QR_CUDA(cudaMemcpyAsync(tau, dtau, sizeof(double) * k, cudaMemcpyDeviceToHost, st));
    QR_CUDA(cudaMemcpyAsync(jpvt, dp, sizeof(int) * n, cudaMemcpyDeviceToHost, st));
    QR_CUDA(cudaStreamSynchronize(st));
    c->d2h_bytes += (int64_t)sizeof(double) * (m * n + k) + (int64_t)sizeof(int) * n;
    ksync(c);
#undef QR_CUDA
    cleanup();
    return rc;
}

// --------------------------------------------------------------------------------
// C <- C Q for the Q of a Householder QR in factored form (LAPACK dormqr 'R', 'N'): the G * Ap_Q of the reduce
// (src/Solvers/process.jl:310-338, `rmul!(G, Ap_fact.Q)`).  Q = H_0 H_1 ... H_{k-1}, H_j = I - tau_j v_j v_j' with
// v_j = [0 .. 0, 1, V[j+1:n, j]].  The rows of C are independent of each other under a right multiplication, so one
// thread owns one row and walks through ALL reflectors without any grid-wide step: per reflector w = C[r, j:] v_j,
// then C[r, j:] -= tau_j w v_j'; consecutive threads own consecutive rows, so every access of the column-major C is
// coalesced; v_j passes through shared memory in chunks.
// --------------------------------------------------------------------------------
namespace {
constexpr int ORM_THREADS = 256, ORM_CHUNK = 2048;

__global__ void __launch_bounds__(ORM_THREADS)
ormqr_right_kernel(int64_t m, int64_t n, int64_t k, const double* __restrict__ V, int64_t ldv, const double* __restrict__ tau,
                   double* __restrict__ C, int64_t ldc) {
    __shared__ double vs[ORM_CHUNK];
    const int64_t r = (int64_t)blockIdx.x * ORM_THREADS + threadIdx.x;
    const bool live = r < m;
    for (int64_t j = 0; j < k; ++j) {
        const int64_t len = n - j;  // v_j[0] = 1, v_j[i] = V[j + i, j]
        const double tj = tau[j];
        if (tj == 0.0) continue;  // H_j = I (uniform over the grid)
        double w = 0.0;
        for (int64_t c0 = 0; c0 < len; c0 += ORM_CHUNK) {
            const int64_t cl = len - c0 < ORM_CHUNK ? len - c0 : ORM_CHUNK;
            __syncthreads();
            for (int64_t i = threadIdx.x; i < cl; i += ORM_THREADS) vs[i] = (c0 + i == 0) ? 1.0 : V[j + c0 + i + j * ldv];
            __syncthreads();
            if (live) {
                const double* row = C + r + (j + c0) * ldc;
                for (int64_t i = 0; i < cl; ++i) w += row[i * ldc] * vs[i];
            }
        }
        const double tw = tj * w;
        for (int64_t c0 = 0; c0 < len; c0 += ORM_CHUNK) {
            const int64_t cl = len - c0 < ORM_CHUNK ? len - c0 : ORM_CHUNK;
            __syncthreads();
            for (int64_t i = threadIdx.x; i < cl; i += ORM_THREADS) vs[i] = (c0 + i == 0) ? 1.0 : V[j + c0 + i + j * ldv];
            __syncthreads();
            if (live) {
                double* row = C + r + (j + c0) * ldc;
                for (int64_t i = 0; i < cl; ++i) row[i * ldc] -= tw * vs[i];
            }
        }
    }
}
}  // namespace

extern "C" int hyb_ormqr_right(hyb_ctx* c, int64_t m, int64_t n, int64_t k, const double* V, int64_t ldv, const double* tau,
                               double* C, int64_t ldc) {
    if (!c) return -1;
    if (m < 0 || n < 0 || k < 0 || k > n || ldv < (n > 1 ? n : 1) || ldc < (m > 1 ? m : 1)) {
        hyb_set_error("hyb_ormqr_right: bad dimensions m=%lld n=%lld k=%lld ldv=%lld ldc=%lld", (long long)m, (long long)n,
                      (long long)k, (long long)ldv, (long long)ldc);
        return -3;
    }
    if (m == 0 || n == 0 || k == 0) return 0;
    HYB_CUDA(cudaSetDevice(c->device));
    cudaStream_t st = c->st;
    const int64_t ldV = hyb_round_up(n, 16), ldC = hyb_round_up(m, 16);
    double *dV = nullptr, *dtau = nullptr, *dC = nullptr;
    cudaError_t e = cudaMalloc(&dV, sizeof(double) * ldV * k);
    if (e == cudaSuccess) e = cudaMalloc(&dtau, sizeof(double) * k);
    if (e == cudaSuccess) e = cudaMalloc(&dC, sizeof(double) * ldC * n);
    if (e == cudaSuccess)
        e = cudaMemcpy2DAsync(dV, sizeof(double) * ldV, V, sizeof(double) * ldv, sizeof(double) * n, k, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dtau, tau, sizeof(double) * k, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess)
        e = cudaMemcpy2DAsync(dC, sizeof(double) * ldC, C, sizeof(double) * ldc, sizeof(double) * m, n, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) {
        c->h2d_bytes += (int64_t)sizeof(double) * (n * k + k + m * n);
        KBEGIN(c);
        ormqr_right_kernel<<<(unsigned)((m + ORM_THREADS - 1) / ORM_THREADS), ORM_THREADS, 0, st>>>(m, n, k, dV, ldV, dtau, dC, ldC);
        ++g_misc_launches;
        e = cudaGetLastError();
        KEND(c);
    }
    if (e == cudaSuccess)
        e = cudaMemcpy2DAsync(C, sizeof(double) * ldc, dC, sizeof(double) * ldC, sizeof(double) * m, n, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e == cudaSuccess) {
        c->d2h_bytes += (int64_t)sizeof(double) * m * n;
        ksync(c);
    }
    cudaFree(dV);
    cudaFree(dtau);
    cudaFree(dC);
    if (e != cudaSuccess) {
        hyb_set_error("hyb_ormqr_right: %s", cudaGetErrorString(e));
        return -1;
    }
    return 0;
}
