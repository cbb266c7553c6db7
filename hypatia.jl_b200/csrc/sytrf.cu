// Symmetric-indefinite factorization and solves on the device: rook-pivoted Bunch-Kaufman,
// A = U D U' with 1 x 1 and 2 x 2 diagonal blocks (upper triangle referenced).
//
// Replaces the fallback of the reference's posdef_fact_copy! (src/linearalgebra/dense.jl:228-249):
// when cholesky! fails, symm_fact! = bunchkaufman!(A, true) -> LAPACK dsytrf_rook (dense.jl:199-203)
// factors the SAME matrix, and only if that fails too is the diagonal shifted (increase_diag!,
// dense.jl:100-107) and Bunch-Kaufman tried again.  ldiv! with that factorization is dsytrs_rook.
//
// This is a rare path (a handful of late, ill-conditioned iterations at most), so it is built for
// exact LAPACK semantics, not for speed: the unblocked algorithm dsytf2_rook, same pivot search
// (alpha = (1 + sqrt(17)) / 8, first index of the maximum like idamax), same interchanges, same
// storage (multipliers in place, ipiv > 0: 1 x 1 pivot with interchange row; ipiv(k), ipiv(k-1) < 0:
// 2 x 2 pivot with one interchange each) and the same order of floating-point operations inside an
// entry's update -- checked against LAPACK's output (pivot vectors identical).  Per elimination step
// two launches: one CTA searches the pivot, swaps, scales the pivot column(s) and parks the
// unscaled copies in a workspace; a grid then applies the rank-1 / rank-2 update to the leading
// triangle (HBM-bound: 16 bytes per entry).  The step index lives on the device (a step retires one
// or two columns, which the host cannot know): the host issues N step pairs, spare ones return.
#include <cfloat>

#include "common.cuh"

namespace {

constexpr int PT = 1024;        // threads of the pivot / solve kernels
constexpr int UR = 128;         // rows per CTA of the update kernel (one per thread)
constexpr int UC = 32;          // columns per CTA of the update kernel

struct SyState {
    int k;      // next column to eliminate (0-based, counts down); < 0: done
    int info;   // LAPACK info: 1-based index of the first exactly zero pivot column, 0 = none
    int kind;   // update of this step: 0 none, 1 rank-1, 2 rank-2
    int kdim;   // order of the leading block the update touches
    double a;   // rank-1: alpha of dsyr (A += u (-alpha u)'); rank-2: d12
};

struct ValIdx {
    double v;
    int i;
};

// maximum of |x| with the FIRST index attaining it (idamax), over a strided segment, whole CTA;
// result broadcast to every thread.  Empty segment: {0, -1}.
__device__ ValIdx block_iamax(const double* x, int64_t stride, int len, ValIdx* sh) {
    ValIdx b{-1.0, -1};
    for (int i = threadIdx.x; i < len; i += PT) {
        const double v = fabs(x[(int64_t)i * stride]);
        if (v > b.v) {  // strict: the thread's indices ascend, so its first maximum is kept
            b.v = v;
            b.i = i;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double ov = __shfl_xor_sync(0xffffffffu, b.v, o);
        const int oi = __shfl_xor_sync(0xffffffffu, b.i, o);
        if (oi >= 0 && (ov > b.v || (ov == b.v && oi < b.i) || b.i < 0)) {
            b.v = ov;
            b.i = oi;
        }
    }
    __syncthreads();  // sh may still be read from a previous call
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = b;
    __syncthreads();
    ValIdx r = sh[0];
    for (int w = 1; w < PT / 32; ++w) {
        const ValIdx o = sh[w];
        if (o.i >= 0 && (o.v > r.v || (o.v == r.v && o.i < r.i) || r.i < 0)) r = o;
    }
    if (r.i < 0) r.v = 0.0;
    return r;
}

__device__ void block_swap(double* a, int64_t sa, double* b, int64_t sb, int len) {
    for (int i = threadIdx.x; i < len; i += PT) {
        const double t = a[(int64_t)i * sa];
        a[(int64_t)i * sa] = b[(int64_t)i * sb];
        b[(int64_t)i * sb] = t;
    }
}

#define AT(i, j) A[(int64_t)(i) + (int64_t)(j) * ld]

// one elimination step of dsytf2_rook (UPLO = 'U'): pivot search, interchanges, pivot-column scaling
__global__ void __launch_bounds__(PT)
sytf2_pivot_kernel(double* __restrict__ A, int64_t ld, int n, SyState* __restrict__ st, int* __restrict__ ipiv,
                   double* __restrict__ ws) {
    __shared__ ValIdx sh[PT / 32];
    const int tid = threadIdx.x;
    const int k = st->k;
    if (k < 0) {
        if (tid == 0) st->kind = 0;
        return;
    }
    const double ALPHA = (1.0 + sqrt(17.0)) / 8.0;
    int kstep = 1, p = k, kp = k, kind = 0, kdim = 0;
    double sa = 0.0;
    const double absakk = fabs(AT(k, k));
    ValIdx cm = block_iamax(&AT(0, k), 1, k, sh);
    int imax = cm.i;
    double colmax = cm.v;
    bool zero_col = false;
    if (fmax(absakk, colmax) == 0.0) {
        zero_col = true;  // column is exactly zero: record it, no elimination
    } else {
        if (absakk < ALPHA * colmax) {
            while (true) {  // rook search: every thread follows the same (broadcast) decisions
                double rowmax = 0.0;
                int jmax = imax;
                if (imax != k) {
                    const ValIdx r = block_iamax(&AT(imax, imax + 1), ld, k - imax, sh);
                    jmax = imax + 1 + r.i;
                    rowmax = r.v;
                }
                if (imax > 0) {
                    const ValIdx c = block_iamax(&AT(0, imax), 1, imax, sh);
                    if (c.v > rowmax) {
                        rowmax = c.v;
                        jmax = c.i;
                    }
                }
                if (!(fabs(AT(imax, imax)) < ALPHA * rowmax)) {
                    kp = imax;  // 1 x 1 pivot with interchange
                    break;
                } else if (p == jmax || rowmax <= colmax) {
                    kp = imax;  // 2 x 2 pivot
                    kstep = 2;
                    break;
                } else {
                    p = imax;
                    colmax = rowmax;
                    imax = jmax;
                }
            }
        }
        const int kk = k - kstep + 1;
        __syncthreads();
        if (kstep == 2 && p != k) {  // first interchange of a 2 x 2 pivot: rows / columns k and p
            if (p > 0) block_swap(&AT(0, k), 1, &AT(0, p), 1, p);
            if (p < k - 1) block_swap(&AT(p + 1, k), 1, &AT(p, p + 1), ld, k - p - 1);
            if (tid == 0) {
                const double t = AT(k, k);
                AT(k, k) = AT(p, p);
                AT(p, p) = t;
            }
            __syncthreads();
        }
        if (kp != kk) {  // interchange rows / columns kk and kp
            if (kp > 0) block_swap(&AT(0, kk), 1, &AT(0, kp), 1, kp);
            if (kk > 0 && kp < kk - 1) block_swap(&AT(kp + 1, kk), 1, &AT(kp, kp + 1), ld, kk - kp - 1);
            if (tid == 0) {
                double t = AT(kk, kk);
                AT(kk, kk) = AT(kp, kp);
                AT(kp, kp) = t;
                if (kstep == 2) {
                    t = AT(k - 1, k);
                    AT(k - 1, k) = AT(kp, k);
                    AT(kp, k) = t;
                }
            }
            __syncthreads();
        }
        if (kstep == 1) {
            if (k > 0) {
                const double akk = AT(k, k);
                kind = 1;
                kdim = k;
                if (fabs(akk) >= DBL_MIN) {
                    const double d11 = 1.0 / akk;
                    sa = d11;
                    for (int i = tid; i < k; i += PT) {
                        const double x = AT(i, k);
                        ws[i] = x;
                        AT(i, k) = d11 * x;
                    }
                } else {
                    sa = akk;
                    for (int i = tid; i < k; i += PT) {
                        const double u = AT(i, k) / akk;
                        ws[i] = u;
                        AT(i, k) = u;
                    }
                }
            }
        } else if (k > 1) {
            const double d12 = AT(k - 1, k);
            const double d22 = AT(k - 1, k - 1) / d12;
            const double d11 = AT(k, k) / d12;
            const double t = 1.0 / (d11 * d22 - 1.0);
            kind = 2;
            kdim = k - 1;
            sa = d12;
            for (int j = tid; j < k - 1; j += PT) {
                const double xk = AT(j, k), xm = AT(j, k - 1);
                const double wkm1 = t * (d11 * xm - xk);
                const double wk = t * (d22 * xk - xm);
                ws[j] = xk / d12;
                ws[n + j] = xm / d12;
                ws[2 * (int64_t)n + j] = wk;
                ws[3 * (int64_t)n + j] = wkm1;
                AT(j, k) = wk / d12;
                AT(j, k - 1) = wkm1 / d12;
            }
        }
    }
    __syncthreads();
    if (tid == 0) {
        if (zero_col && st->info == 0) st->info = k + 1;
        if (kstep == 1) {
            ipiv[k] = kp + 1;
        } else {
            ipiv[k] = -(p + 1);
            ipiv[k - 1] = -(kp + 1);
        }
        st->kind = kind;
        st->kdim = kdim;
        st->a = sa;
        st->k = k - kstep;
    }
}

// rank-1 / rank-2 update of the leading kdim x kdim upper triangle (dsyr / the 2 x 2 loop of dsytf2_rook)
__global__ void __launch_bounds__(UR)
sytf2_update_kernel(double* __restrict__ A, int64_t ld, int n, const SyState* __restrict__ st, const double* __restrict__ ws) {
    const int kind = st->kind;
    if (kind == 0) return;
    const int kdim = st->kdim;
    const int i = blockIdx.x * UR + threadIdx.x;
    const int j0 = blockIdx.y * UC;
    if (j0 >= kdim || (int)(blockIdx.x * UR) > j0 + UC - 1) return;  // outside the block / below the diagonal
    if (i >= kdim) return;
    const int j1 = min(j0 + UC, kdim);
    if (kind == 1) {
        const double alpha = st->a, ui = ws[i];
        for (int j = max(j0, i); j < j1; ++j) {
            const double temp = -alpha * ws[j];
            AT(i, j) = AT(i, j) + ui * temp;
        }
    } else {
        const double a1 = ws[i], a2 = ws[n + i];
        const double* wk = ws + 2 * (int64_t)n;
        const double* wkm1 = ws + 3 * (int64_t)n;
        for (int j = max(j0, i); j < j1; ++j) AT(i, j) = AT(i, j) - a1 * wk[j] - a2 * wkm1[j];
    }
}

// dsytrs_rook (UPLO = 'U') for R right-hand sides (columns of X), one CTA: first U D X = B going up the
// columns, then U' X = B going down, interchanges interleaved exactly as LAPACK does.
__global__ void __launch_bounds__(PT)
sytrs_rook_kernel(const double* __restrict__ A, int64_t ld, int n, const int* __restrict__ ipiv, double* X, int64_t ldx, int R) {
    __shared__ double red[4][PT / 32];
    __shared__ double piv[8];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
#define XR(i, r) X[(int64_t)(i) + (int64_t)(r) * ldx]
    auto swap_rows = [&](int a, int b) {
        if (a != b && tid < R) {
            const double t = XR(a, tid);
            XR(a, tid) = XR(b, tid);
            XR(b, tid) = t;
        }
    };
    int k = n - 1;
    while (k >= 0) {
        if (ipiv[k] > 0) {
            swap_rows(k, ipiv[k] - 1);
            __syncthreads();
            for (int r = 0; r < R; ++r) {
                const double xk = XR(k, r);
                for (int i = tid; i < k; i += PT) XR(i, r) -= AT(i, k) * xk;
            }
            __syncthreads();
            if (tid < R) XR(k, tid) = XR(k, tid) * (1.0 / AT(k, k));
            k -= 1;
        } else {
            swap_rows(k, -ipiv[k] - 1);
            __syncthreads();
            swap_rows(k - 1, -ipiv[k - 1] - 1);
            __syncthreads();
            for (int r = 0; r < R; ++r) {
                const double xk = XR(k, r), xm = XR(k - 1, r);
                for (int i = tid; i < k - 1; i += PT) {
                    double v = XR(i, r);
                    v -= AT(i, k) * xk;
                    v -= AT(i, k - 1) * xm;
                    XR(i, r) = v;
                }
            }
            __syncthreads();
            if (tid < R) {
                const double akm1k = AT(k - 1, k);
                const double akm1 = AT(k - 1, k - 1) / akm1k;
                const double ak = AT(k, k) / akm1k;
                const double denom = akm1 * ak - 1.0;
                const double bkm1 = XR(k - 1, tid) / akm1k;
                const double bk = XR(k, tid) / akm1k;
                XR(k - 1, tid) = (ak * bkm1 - bk) / denom;
                XR(k, tid) = (akm1 * bk - bkm1) / denom;
            }
            k -= 2;
        }
        __syncthreads();
    }
    // U' X = B
    k = 0;
    while (k < n) {
        const int ncol = (ipiv[k] > 0) ? 1 : 2;
        if (k > 0) {
            for (int c = 0; c < ncol; ++c)
                for (int r = 0; r < R; ++r) {
                    double s = 0.0;
                    for (int i = tid; i < k; i += PT) s += AT(i, k + c) * XR(i, r);
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
                    if (lane == 0) red[r][warp] = s;
                    __syncthreads();
                    if (tid == 0) {
                        double t = 0.0;
                        for (int w = 0; w < PT / 32; ++w) t += red[r][w];
                        piv[c * 4 + r] = t;
                    }
                    __syncthreads();
                }
            if (tid < R) {
                XR(k, tid) -= piv[tid];
                if (ncol == 2) XR(k + 1, tid) -= piv[4 + tid];
            }
            __syncthreads();
        }
        if (ncol == 1) {
            swap_rows(k, ipiv[k] - 1);
        } else {
            swap_rows(k, -ipiv[k] - 1);
            __syncthreads();
            swap_rows(k + 1, -ipiv[k + 1] - 1);
        }
        __syncthreads();
        k += ncol;
    }
#undef XR
}
#undef AT

__global__ void sy_init_kernel(SyState* st, int n) {
    st->k = n - 1;
    st->info = 0;
    st->kind = 0;
    st->kdim = 0;
    st->a = 0.0;
}

hyb_counter g_sytrf_launches{0};

}  // namespace

int64_t hyb_sytrf_launch_count() { return g_sytrf_launches; }

// In-place rook-pivoted Bunch-Kaufman of ch.F (upper triangle); ch.ipiv receives LAPACK's pivot vector.
// *info_out = LAPACK info (0: D has no exactly zero block).  Synchronizes the stream.
int hyb_sytrf_rook(cudaStream_t st, HybChol& ch, int* info_out) {
    const int n = (int)ch.N;
    if (info_out) *info_out = 0;
    if (n == 0) {
        ch.kind = 1;
        return 0;
    }
    if (!ch.ipiv) {
        HYB_CUDA(cudaMalloc(&ch.ipiv, sizeof(int) * n));
        HYB_CUDA(cudaMalloc(&ch.bk_ws, sizeof(double) * 4 * (size_t)n));
        HYB_CUDA(cudaMalloc(&ch.bk_state, sizeof(SyState)));
    }
    SyState* state = (SyState*)ch.bk_state;
    sy_init_kernel<<<1, 1, 0, st>>>(state, n);
    for (int s = 0; s < n; ++s) {
        const int kmax = n - s;  // at most this many columns are left at host step s
        sytf2_pivot_kernel<<<1, PT, 0, st>>>(ch.F, ch.ld, n, state, ch.ipiv, ch.bk_ws);
        if (kmax > 1) {
            dim3 grid((unsigned)((kmax + UR - 1) / UR), (unsigned)((kmax + UC - 1) / UC));
            sytf2_update_kernel<<<grid, UR, 0, st>>>(ch.F, ch.ld, n, state, ch.bk_ws);
        }
        g_sytrf_launches += 2;
    }
    HYB_CUDA(cudaGetLastError());
    SyState h;
    HYB_CUDA(cudaMemcpyAsync(&h, state, sizeof(SyState), cudaMemcpyDeviceToHost, st));
    HYB_CUDA(cudaStreamSynchronize(st));
    if (info_out) *info_out = h.info;
    ch.kind = 1;
    return 0;
}

// X <- A^-1 X for R <= HYB_MAX_RHS columns with the factorization of hyb_sytrf_rook
int hyb_sytrs_rook(cudaStream_t st, HybChol& ch, double* X, int64_t ldx, int R) {
    if (ch.N == 0 || R <= 0) return 0;
    if (R > HYB_MAX_RHS) {
        hyb_set_error("hyb_sytrs_rook: %d right-hand sides (max %d)", R, HYB_MAX_RHS);
        return -3;
    }
    sytrs_rook_kernel<<<1, PT, 0, st>>>(ch.F, ch.ld, (int)ch.N, ch.ipiv, X, ldx, R);
    ++g_sytrf_launches;
    HYB_CUDA(cudaGetLastError());
    return 0;
}
