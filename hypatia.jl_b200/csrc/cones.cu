// Batched cone barrier oracles (sm_100a), one launch per block of identical cones.
//
// Follows the reference oracles formula for formula:
//   Nonnegative   src/Cones/nonnegative.jl:55-141
//   EpiNormEucl   src/Cones/epinormeucl.jl:62-228
//   EpiPerSquare  src/Cones/epipersquare.jl:73-280
//   PosSemidefTri src/Cones/possemideftri.jl:89-204 (real, side <= 32: warp-per-cone in shared memory)
//   check_numerics / get_proxsqr  src/Cones/Cones.jl:267-304
// HBM-bound: a product on a d x m block moves 16*d*m bytes; vector oracles <= 32*d.
// Small cones (dim <= 32): thread per (cone, column), adjacent threads on adjacent cones of
// the same column so a warp covers contiguous rows.  Larger cones: warp per (cone, column)
// with lanes striding down the rows.  All reductions are fixed-order.
#include "cones.cuh"

#include "cones_dense.cuh"

namespace {

hyb_counter g_cone_launches{0};
#define LAUNCHED() (++g_cone_launches)

constexpr double RT2 = 1.4142135623730951;

template <bool WARP>
__device__ __forceinline__ double red_sum(double v) {
    if (WARP) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    }
    return v;
}

// ------------------------------------------------------------------ Nonnegative
__global__ void nonneg_update_kernel(int64_t n, const double* __restrict__ primal, const double* __restrict__ dual_in,
                                     double scal, double irtmu, double* __restrict__ point, double* __restrict__ dual,
                                     double* __restrict__ grad, int* __restrict__ flag, double* __restrict__ prox,
                                     const double* __restrict__ ctl) {
    if (ctl) {  // speculative line search (dirs.cu): a decided sequence skips; the scaling comes from the device
        if (ctl[0] != 0.0) return;
        scal = irtmu = ctl[1];
    }
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double p = scal * primal[i], d = dual_in[i];
    point[i] = p;
    dual[i] = d;
    int f = 0;
    if (p > HYB_EPS) f |= HYB_ST_FEAS;
    if (d > HYB_EPS) f |= HYB_ST_DUAL_FEAS;
    const double g = -1.0 / p;
    grad[i] = g;
    // check_numerics per coordinate: 1 + g*p == 0 and 1 - (g*p*p)*g == 0 up to rounding
    if (f & HYB_ST_FEAS) {
        const double t1 = fabs(1.0 + g * p), t2 = fabs(1.0 - (g * p * p) * g);
        if (t1 <= 1.220703125e-4 && t2 <= 1.104854345603981e-1) f |= HYB_ST_NUMERICS;
    }
    flag[i] = f;
    const double v = p * d * irtmu - 1.0;
    prox[i] = v * v;
}

__global__ void nonneg_prod_kernel(int64_t n, int64_t m, int op, const double* __restrict__ point,
                                   const double* __restrict__ arr, int64_t lda, double* __restrict__ prod, int64_t ldp) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double p = point[i];
    for (int64_t j = blockIdx.y; j < m; j += gridDim.y) {
        const double a = arr[i + j * lda];
        double r;
        if (op == HYB_OP_HESS) r = a / p / p;
        else if (op == HYB_OP_INV_HESS) r = a * p * p;
        else if (op == HYB_OP_SQRT_HESS) r = a / p;
        else r = a * p;
        prod[i + j * ldp] = r;
    }
}

__global__ void nonneg_dder3_kernel(int64_t n, const double* __restrict__ point, const double* __restrict__ dir,
                                    double* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double p = point[i], t = dir[i] / p;
    out[i] = t * t / p;
}

// ------------------------------------------------------------------ EpiNormEucl / EpiPerSquare
// TYPE 1: (u, w), dist = (u^2 - |w|^2)/2.   TYPE 2: (u, v, w), dist = u v - |w|^2/2.
template <int TYPE>
struct Soc {
    static constexpr int K0 = TYPE;  // number of leading "epigraph" coordinates
};

// dist-like quantity of a vector (without the feasibility logic)
template <int TYPE, bool WARP>
__device__ __forceinline__ double soc_wnorm2(const double* v, int d, int i0, int stride) {
    double s = 0.0;
    for (int i = TYPE + i0; i < d; i += stride) s += v[i] * v[i];
    return red_sum<WARP>(s);
}

template <bool WARP>
__device__ __forceinline__ double dot_strided(const double* a, const double* b, int lo, int d, int i0, int stride) {
    double s = 0.0;
    for (int i = lo + i0; i < d; i += stride) s += a[i] * b[i];
    return red_sum<WARP>(s);
}

// One item = one cone.  WARP: lanes stride over coordinates; else one thread loops.
template <int TYPE, bool WARP>
__global__ void soc_update_kernel(int64_t count, int d, int64_t offset, int64_t cone0, const double* __restrict__ primal,
                                  const double* __restrict__ dual_in, double scal, double irtmu,
                                  double* __restrict__ point, double* __restrict__ dual, double* __restrict__ grad,
                                  double* __restrict__ aux, int* __restrict__ flag, double* __restrict__ prox,
                                  const double* __restrict__ ctl) {
    if (ctl) {  // speculative line search (dirs.cu): a decided sequence skips; the scaling comes from the device
        if (ctl[0] != 0.0) return;
        scal = irtmu = ctl[1];
    }
    const int64_t gt = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t c = WARP ? (gt >> 5) : gt;
    const int i0 = WARP ? (int)(gt & 31) : 0;
    const int stride = WARP ? 32 : 1;
    if (c >= count) return;
    const int64_t base = offset + c * d;
    double* p = point + base;
    double* z = dual + base;
    double* g = grad + base;
    for (int i = i0; i < d; i += stride) {
        p[i] = scal * primal[base + i];
        z[i] = dual_in[base + i];
    }
    if (WARP) __syncwarp();
    int f = 0;
    double dist = 0.0;
    // primal feasibility (epinormeucl.jl:62-76, epipersquare.jl:73-88)
    {
        const double w2 = soc_wnorm2<TYPE, WARP>(p, d, i0, stride);
        if (TYPE == 1) {
            const double u = p[0];
            if (u > HYB_EPS) {
                dist = (u * u - w2) / 2;
                if (dist > HYB_EPS) f |= HYB_ST_FEAS;
            }
        } else {
            const double u = p[0], v = p[1];
            if (u > HYB_EPS && v > HYB_EPS) {
                dist = u * v - w2 / 2;
                if (dist > HYB_EPS) f |= HYB_ST_FEAS;
            }
        }
    }
    // dual feasibility (epinormeucl.jl:78-88, epipersquare.jl:90-100)
    {
        const double w2 = soc_wnorm2<TYPE, WARP>(z, d, i0, stride);
        if (TYPE == 1) {
            const double u = z[0];
            if (u > HYB_EPS && (u * u - w2) > 2 * HYB_EPS) f |= HYB_ST_DUAL_FEAS;
        } else {
            const double u = z[0], v = z[1];
            if (u > HYB_EPS && v > HYB_EPS && (u * v - w2 / 2) > HYB_EPS) f |= HYB_ST_DUAL_FEAS;
        }
    }
    double proxsqr = 0.0;
    if (f & HYB_ST_FEAS) {
        // grad (epinormeucl.jl:90-98, epipersquare.jl:102-112)
        for (int i = TYPE + i0; i < d; i += stride) g[i] = p[i] / dist;
        if (i0 == 0) {
            if (TYPE == 1) {
                g[0] = -(p[0] / dist);
            } else {
                g[0] = -(p[1] / dist);
                g[1] = -(p[0] / dist);
            }
        }
        if (WARP) __syncwarp();
        // check_numerics (Cones.jl:267-284): nu = 2
        const double gp = dot_strided<WARP>(g, p, 0, d, i0, stride);
        bool ok = !(fabs(1.0 + gp / 2.0) > 1.220703125e-4 * d);
        // Hig = inv_hess_prod(g) = (p.g) p + dist * J g ;  <Hig, g>
        double s = 0.0;
        for (int i = i0; i < d; i += stride) {
            double hi = gp * p[i];
            if (TYPE == 1) hi += (i == 0) ? -dist * g[0] : dist * g[i];
            else hi += (i == 0) ? -dist * g[1] : ((i == 1) ? -dist * g[0] : dist * g[i]);
            s += hi * g[i];
        }
        s = red_sum<WARP>(s);
        if (fabs(1.0 - s / 2.0) > 1.104854345603981e-1 * d) ok = false;
        if (ok) f |= HYB_ST_NUMERICS;
        // proxsqr (Cones.jl:288-304): psi = irtmu * z + g ; psi' Hinv psi
        double ppsi = 0.0;
        for (int i = i0; i < d; i += stride) ppsi += p[i] * (irtmu * z[i] + g[i]);
        ppsi = red_sum<WARP>(ppsi);
        double acc = 0.0;
        for (int i = i0; i < d; i += stride) {
            const double psi = irtmu * z[i] + g[i];
            double hi = ppsi * p[i];
            if (TYPE == 1) {
                hi += (i == 0) ? -dist * psi : dist * psi;
            } else {
                if (i == 0) hi -= dist * (irtmu * z[1] + g[1]);
                else if (i == 1) hi -= dist * (irtmu * z[0] + g[0]);
                else hi += dist * psi;
            }
            acc += hi * psi;
        }
        acc = red_sum<WARP>(acc);
        proxsqr = (acc < -1.4901161193847656e-08 * d) ? INFINITY : fabs(acc);
    }
    if (i0 == 0) {
        aux[cone0 + c] = dist;
        flag[base] = f;
        prox[base] = proxsqr;
    }
}

// products on a q_block x m array.  Item = (cone c, column j); thread mapping puts adjacent
// threads on adjacent cones of the same column.
// one (cone, column) item: r = op(H) a for the cone with point p and dist; r may alias a (every entry of a
// that is needed after the first store has been read into a register by then)
template <int TYPE, bool WARP>
__device__ __forceinline__ void soc_prod_item(const double* p, const double* a, double* r, double dist, int d, int op,
                                              int i0, int stride) {
    if (op == HYB_OP_HESS) {
        // epinormeucl.jl:129-144, epipersquare.jl:145-162
        const double wwa = dot_strided<WARP>(p, a, TYPE, d, i0, stride);
        double ga;
        if (TYPE == 1) ga = (wwa - p[0] * a[0]) / dist;
        else ga = (wwa - p[1] * a[0] - p[0] * a[1]) / dist;
        const double a0 = a[0], a1 = a[TYPE == 2 ? 1 : 0];
        if (WARP) __syncwarp();
        for (int i = TYPE + i0; i < d; i += stride) r[i] = (ga * p[i] + a[i]) / dist;
        if (i0 == 0) {
            if (TYPE == 1) {
                r[0] = (-ga * p[0] - a0) / dist;
            } else {
                r[0] = (-ga * p[1] - a1) / dist;
                r[1] = (-ga * p[0] - a0) / dist;
            }
        }
    } else if (op == HYB_OP_INV_HESS) {
        // epinormeucl.jl:146-157, epipersquare.jl:164-176
        const double pa = dot_strided<WARP>(p, a, 0, d, i0, stride);
        const double a0 = a[0], a1 = a[TYPE == 2 ? 1 : 0];
        if (WARP) __syncwarp();
        for (int i = TYPE + i0; i < d; i += stride) r[i] = pa * p[i] + dist * a[i];
        if (i0 == 0) {
            if (TYPE == 1) {
                r[0] = pa * p[0] - dist * a0;
            } else {
                r[0] = pa * p[0] - dist * a1;
                r[1] = pa * p[1] - dist * a0;
            }
        }
    } else if (TYPE == 1) {
        // epinormeucl.jl:159-206
        const double u = p[0];
        const double rtdist = sqrt(dist);
        const double urtdist = u + rtdist * RT2;
        const double dotwwj = dot_strided<WARP>(p, a, 1, d, i0, stride);
        const double uj = a[0];
        if (WARP) __syncwarp();
        if (op == HYB_OP_SQRT_HESS) {
            const double distrt2 = dist * RT2;
            const double wmulj = (dotwwj / urtdist - uj) / distrt2;
            for (int i = 1 + i0; i < d; i += stride) r[i] = p[i] * wmulj + a[i] / rtdist;
            if (i0 == 0) r[0] = (u * uj - dotwwj) / distrt2;
        } else {
            const double wmulj = (dotwwj / urtdist + uj) / RT2;
            for (int i = 1 + i0; i < d; i += stride) r[i] = p[i] * wmulj + a[i] * rtdist;
            if (i0 == 0) r[0] = (u * uj + dotwwj) / RT2;
        }
    } else {
        // epipersquare.jl:178-255
        const double rtdist = sqrt(dist);
        const double denom = 2 * rtdist + p[0] + p[1];
        const double a0 = a[0], a1 = a[1];
        if (op == HYB_OP_SQRT_HESS) {
            const double v0 = -p[1] / rtdist - 1, v1 = -p[0] / rtdist - 1;
            double s = 0.0;
            for (int i = 2 + i0; i < d; i += stride) s += (p[i] / rtdist) * a[i];
            s = red_sum<WARP>(s);
            const double dotj = (v0 * a0 + v1 * a1 + s) / denom;
            if (WARP) __syncwarp();
            for (int i = 2 + i0; i < d; i += stride) r[i] = dotj * (p[i] / rtdist) + a[i] / rtdist;
            if (i0 == 0) {
                r[0] = dotj * v0 - a1 / rtdist;
                r[1] = dotj * v1 - a0 / rtdist;
            }
        } else {
            const double v0 = p[0] + rtdist, v1 = p[1] + rtdist;
            double s = 0.0;
            for (int i = 2 + i0; i < d; i += stride) s += p[i] * a[i];
            s = red_sum<WARP>(s);
            const double dotj = (v0 * a0 + v1 * a1 + s) / denom;
            if (WARP) __syncwarp();
            for (int i = 2 + i0; i < d; i += stride) r[i] = dotj * p[i] + a[i] * rtdist;
            if (i0 == 0) {
                r[0] = dotj * v0 - a1 * rtdist;
                r[1] = dotj * v1 - a0 * rtdist;
            }
        }
    }
}

template <int TYPE, bool WARP>
__global__ void soc_prod_kernel(int64_t count, int d, int64_t offset, int64_t cone0, int64_t m, int op,
                                const double* __restrict__ point, const double* __restrict__ aux,
                                const double* __restrict__ arr, int64_t lda, double* __restrict__ prod, int64_t ldp) {
    const int64_t gt = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t item = WARP ? (gt >> 5) : gt;
    const int i0 = WARP ? (int)(gt & 31) : 0;
    const int stride = WARP ? 32 : 1;
    if (item >= count * m) return;
    const int64_t c = item % count, j = item / count;
    soc_prod_item<TYPE, WARP>(point + offset + c * d, arr + offset + c * d + j * lda, prod + offset + c * d + j * ldp,
                              aux[cone0 + c], d, op, i0, stride);
}

// Wide products (m >= 32 columns, the W * GQ2 product of the Schur assembly) of small cones: a thread that walks
// its cone's d consecutive rows makes the warp touch 32 segments d * 8 bytes apart per load instruction (measured:
// 25 % of the HBM copy rate at d = 8).  Here a CTA moves a tile of 256 cones x 1 column (256 d consecutive
// doubles) between global and shared memory with fully coalesced accesses, and the per-cone arithmetic runs
// out of shared memory (row stride d + 1: conflict-free).  The cones' points are staged once per CTA and
// reused for CPB columns.
constexpr int SOC_TILE = 256, SOC_CPB = 8;
template <int TYPE>
__global__ void __launch_bounds__(SOC_TILE)
soc_prod_tile_kernel(int64_t count, int d, int64_t offset, int64_t cone0, int64_t m, int op,
                     const double* __restrict__ point, const double* __restrict__ aux, const double* __restrict__ arr,
                     int64_t lda, double* __restrict__ prod, int64_t ldp) {
    extern __shared__ double sm[];
    const int ds = d + 1;
    double* P = sm;
    double* A = sm + SOC_TILE * ds;
    const int tid = threadIdx.x;
    const int64_t c0 = (int64_t)blockIdx.x * SOC_TILE;
    const int nc = (int)((count - c0 < SOC_TILE) ? (count - c0) : SOC_TILE);
    const int ne = nc * d;
    const int64_t base = offset + c0 * d;
    for (int e = tid; e < ne; e += SOC_TILE) P[(e / d) * ds + (e % d)] = point[base + e];
    const double dist = (tid < nc) ? aux[cone0 + c0 + tid] : 1.0;
    const int64_t j0 = (int64_t)blockIdx.y * SOC_CPB;
    const int64_t j1 = (j0 + SOC_CPB < m) ? (j0 + SOC_CPB) : m;
    for (int64_t j = j0; j < j1; ++j) {
        const double* src = arr + base + j * lda;
        for (int e = tid; e < ne; e += SOC_TILE) A[(e / d) * ds + (e % d)] = src[e];
        __syncthreads();
        if (tid < nc) soc_prod_item<TYPE, false>(P + tid * ds, A + tid * ds, A + tid * ds, dist, d, op, 0, 1);
        __syncthreads();
        double* dst = prod + base + j * ldp;
        for (int e = tid; e < ne; e += SOC_TILE) dst[e] = A[(e / d) * ds + (e % d)];
        __syncthreads();
    }
}

// dder3 (epinormeucl.jl:208-228, epipersquare.jl:257-280)
template <int TYPE, bool WARP>
__global__ void soc_dder3_kernel(int64_t count, int d, int64_t offset, int64_t cone0, const double* __restrict__ point,
                                 const double* __restrict__ aux, const double* __restrict__ dir,
                                 double* __restrict__ out) {
    const int64_t gt = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t c = WARP ? (gt >> 5) : gt;
    const int i0 = WARP ? (int)(gt & 31) : 0;
    const int stride = WARP ? 32 : 1;
    if (c >= count) return;
    const double* p = point + offset + c * d;
    const double* dr = dir + offset + c * d;
    double* o = out + offset + c * d;
    const double dist = aux[cone0 + c];
    const double wwd = dot_strided<WARP>(p, dr, TYPE, d, i0, stride);
    double jdotpd, ga;
    if (TYPE == 1) {
        jdotpd = p[0] * dr[0] - wwd;
        ga = (wwd - p[0] * dr[0]) / dist;
    } else {
        jdotpd = p[0] * dr[1] + p[1] * dr[0] - wwd;
        ga = (wwd - p[1] * dr[0] - p[0] * dr[1]) / dist;
    }
    // h = hess_prod(dir); accumulate <dir,h>, <p,h>
    double sdh = 0.0, sph = 0.0;
    for (int i = i0; i < d; i += stride) {
        double h;
        if (TYPE == 1) h = (i == 0) ? (-ga * p[0] - dr[0]) / dist : (ga * p[i] + dr[i]) / dist;
        else h = (i == 0) ? (-ga * p[1] - dr[1]) / dist : ((i == 1) ? (-ga * p[0] - dr[0]) / dist : (ga * p[i] + dr[i]) / dist);
        sdh += dr[i] * h;
        sph += p[i] * h;
    }
    const double dotdHd = -red_sum<WARP>(sdh);
    const double dotpHd = red_sum<WARP>(sph);
    const double den = 2 * dist;
    for (int i = i0; i < d; i += stride) {
        double h, add;
        if (TYPE == 1) {
            if (i == 0) {
                h = (-ga * p[0] - dr[0]) / dist;
                add = -dotdHd * p[0] - dotpHd * dr[0];
            } else {
                h = (ga * p[i] + dr[i]) / dist;
                add = dotdHd * p[i] + dotpHd * dr[i];
            }
        } else {
            if (i == 0) {
                h = (-ga * p[1] - dr[1]) / dist;
                add = -dotdHd * p[1] - dotpHd * dr[1];
            } else if (i == 1) {
                h = (-ga * p[0] - dr[0]) / dist;
                add = -dotdHd * p[0] - dotpHd * dr[0];
            } else {
                h = (ga * p[i] + dr[i]) / dist;
                add = dotdHd * p[i] + dotpHd * dr[i];
            }
        }
        o[i] = (h * jdotpd + add) / den;
    }
}

// ------------------------------------------------------------------ EpiNormInf (real)
// src/Cones/epinorminf.jl:100-392.  point = (u, w), d = dim - 1, nu = 1 + d.  Per coordinate
//   mu_i = w_i / u, zeta_i = (u - mu_i w_i) / 2, cu = (d - 1) / u
// are recomputed from the point (two flops) instead of stored; the one scalar worth keeping per cone is
//   Zu = -cu + sum_i 1 / (u - zeta_i)                 (update_inv_hess_aux, :257-273)
// in aux[cone].  No sqrt oracles (use_sqrt_hess_oracles = false, :73): the cone takes the gemm branch of
// the Schur assembly (qrchol.jl:248-254).  Same thread mapping as the second-order cones above.
template <bool WARP>
__device__ __forceinline__ double red_max(double v) {
    if (WARP) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    }
    return v;
}

// H^-1 a . b for the loaded point (inv_hess_prod!, :314-335): c1 = u (a_0 + wumzi . a_w) / Zu,
// (H^-1 a)_0 = c1, (H^-1 a)_i = c1 wumzi_i + zeta_i zti_i a_i, wumzi = w / (u - zeta), zti = u - wumzi w
template <bool WARP, class FA, class FB>
__device__ __forceinline__ double ninf_hinv_dot(const double* p, int d, int i0, int stride, double Zu, FA a, FB b) {
    const double u = p[0];
    double s = 0.0;
    for (int i = 1 + i0; i < d; i += stride) {
        const double w = p[i], zeta = 0.5 * (u - (w / u) * w);
        s += (w / (u - zeta)) * a(i);
    }
    s = red_sum<WARP>(s);
    const double c1 = u * (a(0) + s) / Zu;
    double acc = (i0 == 0) ? c1 * b(0) : 0.0;
    for (int i = 1 + i0; i < d; i += stride) {
        const double w = p[i], zeta = 0.5 * (u - (w / u) * w);
        const double wumzi = w / (u - zeta), zti = u - wumzi * w;
        acc += (c1 * wumzi + zeta * a(i) * zti) * b(i);
    }
    return red_sum<WARP>(acc);
}

template <bool WARP>
__global__ void ninf_update_kernel(int64_t count, int d, int64_t offset, int64_t cone0, const double* __restrict__ primal,
                                   const double* __restrict__ dual_in, double scal, double irtmu,
                                   double* __restrict__ point, double* __restrict__ dual, double* __restrict__ grad,
                                   double* __restrict__ aux, int* __restrict__ flag, double* __restrict__ prox,
                                   const double* __restrict__ ctl) {
    if (ctl) {  // speculative line search (dirs.cu): a decided sequence skips; the scaling comes from the device
        if (ctl[0] != 0.0) return;
        scal = irtmu = ctl[1];
    }
    const int64_t gt = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t c = WARP ? (gt >> 5) : gt;
    const int i0 = WARP ? (int)(gt & 31) : 0;
    const int stride = WARP ? 32 : 1;
    if (c >= count) return;
    const int64_t base = offset + c * d;
    double* p = point + base;
    double* z = dual + base;
    double* g = grad + base;
    for (int i = i0; i < d; i += stride) {
        p[i] = scal * primal[base + i];
        z[i] = dual_in[base + i];
    }
    if (WARP) __syncwarp();
    int f = 0;
    const double u = p[0];
    {  // update_feas (:100-113): u > eps and u - max|w| > eps
        double m = 0.0;
        for (int i = 1 + i0; i < d; i += stride) m = fmax(m, fabs(p[i]));
        m = red_max<WARP>(m);
        if (u > HYB_EPS && (u - m) > HYB_EPS) f |= HYB_ST_FEAS;
    }
    {  // is_dual_feas (:115-128): u > eps and u - sum|w| > eps
        double s1 = 0.0;
        for (int i = 1 + i0; i < d; i += stride) s1 += fabs(z[i]);
        s1 = red_sum<WARP>(s1);
        if (z[0] > HYB_EPS && (z[0] - s1) > HYB_EPS) f |= HYB_ST_DUAL_FEAS;
    }
    double proxsqr = 0.0, Zu = 0.0;
    if (f & HYB_ST_FEAS) {
        const int dw = d - 1;
        const double nu = (double)d, cu = (double)(dw - 1) / u;
        // update_grad (:136-162)
        double sz = 0.0, su = 0.0;
        for (int i = 1 + i0; i < d; i += stride) {
            const double w = p[i], mu = w / u, zeta = 0.5 * (u - mu * w);
            sz += 1.0 / zeta;
            su += 1.0 / (u - zeta);
            g[i] = mu / zeta;
        }
        sz = red_sum<WARP>(sz);
        su = red_sum<WARP>(su);
        Zu = -cu + su;
        if (i0 == 0) g[0] = cu - sz;
        if (WARP) __syncwarp();
        // check_numerics (Cones.jl:267-284)
        double gp = 0.0;
        for (int i = i0; i < d; i += stride) gp += g[i] * p[i];
        gp = red_sum<WARP>(gp);
        bool ok = !(fabs(1.0 + gp / nu) > 1.220703125e-4 * d);
        auto gv = [&](int i) { return g[i]; };
        const double hgg = ninf_hinv_dot<WARP>(p, d, i0, stride, Zu, gv, gv);
        if (fabs(1.0 - hgg / nu) > 1.104854345603981e-1 * d) ok = false;
        if (ok) f |= HYB_ST_NUMERICS;
        // get_proxsqr (Cones.jl:288-304): psi = irtmu z + g ; psi' H^-1 psi
        auto psi = [&](int i) { return irtmu * z[i] + g[i]; };
        const double acc = ninf_hinv_dot<WARP>(p, d, i0, stride, Zu, psi, psi);
        proxsqr = (acc < -1.4901161193847656e-08 * d) ? INFINITY : fabs(acc);
    }
    if (i0 == 0) {
        aux[cone0 + c] = Zu;
        flag[base] = f;
        prox[base] = proxsqr;
    }
}

// hess_prod! (:205-229) / inv_hess_prod! (:314-335) on a q_block x m array; item = (cone, column)
template <bool WARP>
__device__ __forceinline__ void ninf_prod_item(const double* p, const double* a, double* r, double Zu, int d, int op, int i0,
                                               int stride) {
    const double u = p[0], a0 = a[0];
    if (op == HYB_OP_HESS) {
        const double pui = a0 / u, cu = (double)(d - 2) / u;
        double s = 0.0;
        for (int i = 1 + i0; i < d; i += stride) {
            const double w = p[i], mu = w / u, zeta = 0.5 * (u - mu * w);
            const double ai = a[i];
            const double s1 = (a0 - mu * ai) / zeta;
            s += (s1 - pui) / zeta;
            r[i] = (ai / u - s1 * mu) / zeta;
        }
        s = red_sum<WARP>(s);
        if (i0 == 0) r[0] = s - cu * pui;
    } else {
        double s = 0.0;
        for (int i = 1 + i0; i < d; i += stride) {
            const double w = p[i], zeta = 0.5 * (u - (w / u) * w);
            s += (w / (u - zeta)) * a[i];
        }
        s = red_sum<WARP>(s);
        const double c1 = u * (a0 + s) / Zu;
        if (WARP) __syncwarp();
        for (int i = 1 + i0; i < d; i += stride) {
            const double w = p[i], zeta = 0.5 * (u - (w / u) * w);
            const double wumzi = w / (u - zeta), zti = u - wumzi * w;
            r[i] = c1 * wumzi + zeta * a[i] * zti;
        }
        if (i0 == 0) r[0] = c1;
    }
}

template <bool WARP>
__global__ void ninf_prod_kernel(int64_t count, int d, int64_t offset, int64_t cone0, int64_t m, int op,
                                 const double* __restrict__ point, const double* __restrict__ aux,
                                 const double* __restrict__ arr, int64_t lda, double* __restrict__ prod, int64_t ldp) {
    const int64_t gt = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t item = WARP ? (gt >> 5) : gt;
    const int i0 = WARP ? (int)(gt & 31) : 0;
    const int stride = WARP ? 32 : 1;
    if (item >= count * m) return;
    const int64_t c = item % count, j = item / count;
    ninf_prod_item<WARP>(point + offset + c * d, arr + offset + c * d + j * lda, prod + offset + c * d + j * ldp, aux[cone0 + c],
                         d, op, i0, stride);
}

// coalesced-tile form for wide products, as soc_prod_tile_kernel
__global__ void __launch_bounds__(SOC_TILE)
ninf_prod_tile_kernel(int64_t count, int d, int64_t offset, int64_t cone0, int64_t m, int op, const double* __restrict__ point,
                      const double* __restrict__ aux, const double* __restrict__ arr, int64_t lda, double* __restrict__ prod,
                      int64_t ldp) {
    extern __shared__ double sm[];
    const int ds = d + 1;
    double* P = sm;
    double* A = sm + SOC_TILE * ds;
    const int tid = threadIdx.x;
    const int64_t c0 = (int64_t)blockIdx.x * SOC_TILE;
    const int nc = (int)((count - c0 < SOC_TILE) ? (count - c0) : SOC_TILE);
    const int ne = nc * d;
    const int64_t base = offset + c0 * d;
    for (int e = tid; e < ne; e += SOC_TILE) P[(e / d) * ds + (e % d)] = point[base + e];
    const double Zu = (tid < nc) ? aux[cone0 + c0 + tid] : 1.0;
    const int64_t j0 = (int64_t)blockIdx.y * SOC_CPB;
    const int64_t j1 = (j0 + SOC_CPB < m) ? (j0 + SOC_CPB) : m;
    for (int64_t j = j0; j < j1; ++j) {
        const double* src = arr + base + j * lda;
        for (int e = tid; e < ne; e += SOC_TILE) A[(e / d) * ds + (e % d)] = src[e];
        __syncthreads();
        if (tid < nc) ninf_prod_item<false>(P + tid * ds, A + tid * ds, A + tid * ds, Zu, d, op, 0, 1);
        __syncthreads();
        double* dst = prod + base + j * ldp;
        for (int e = tid; e < ne; e += SOC_TILE) dst[e] = A[(e / d) * ds + (e % d)];
        __syncthreads();
    }
}

// dder3 (:366-392)
template <bool WARP>
__global__ void ninf_dder3_kernel(int64_t count, int d, int64_t offset, const double* __restrict__ point,
                                  const double* __restrict__ dir, double* __restrict__ out) {
    const int64_t gt = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t c = WARP ? (gt >> 5) : gt;
    const int i0 = WARP ? (int)(gt & 31) : 0;
    const int stride = WARP ? 32 : 1;
    if (c >= count) return;
    const double* p = point + offset + c * d;
    const double* dr = dir + offset + c * d;
    double* o = out + offset + c * d;
    const double u = p[0], pp = dr[0], pui = pp / u, cu = (double)(d - 2) / u;
    double c1 = 0.0;
    for (int i = 1 + i0; i < d; i += stride) {
        const double w = p[i], mu = w / u, zeta = 0.5 * (u - mu * w);
        const double ri = dr[i], rui = ri / u;
        const double s1 = (pp - mu * ri) / zeta;
        const double s2 = 0.5 * (pp * pui - rui * ri) / zeta - s1 * s1;
        c1 += (s1 * pui + s2) / zeta;
        o[i] = (s1 * rui + s2 * mu) / zeta;
    }
    c1 = red_sum<WARP>(c1);
    if (i0 == 0) o[0] = -c1 - cu * pui * pui;
}

// ------------------------------------------------------------------ PosSemidefTri, side <= 32
// Per-cone matrices in global memory: [X | U | Uinv | Xinv], each side*side column-major.
constexpr int PSD_MAX_SIDE = 32;
constexpr int PSD_LD = PSD_MAX_SIDE + 1;

// svec (column-major upper triangle, off-diagonals * sqrt2) -> full symmetric matrix in shared memory
__device__ __forceinline__ void svec_to_smem(const double* v, int side, double* M, int lane) {
    const int dim = side * (side + 1) / 2;
    for (int k = lane; k < dim; k += 32) {
        int c = (int)((sqrt(8.0 * k + 1.0) - 1.0) * 0.5);
        while (c * (c + 1) / 2 > k) --c;
        while ((c + 1) * (c + 2) / 2 <= k) ++c;
        const int r = k - c * (c + 1) / 2;
        const double x = (r == c) ? v[k] : v[k] * (1.0 / RT2);
        M[r * PSD_LD + c] = x;
        M[c * PSD_LD + r] = x;
    }
}
__device__ __forceinline__ void smem_to_svec(const double* M, int side, double* v, int lane, double scale) {
    const int dim = side * (side + 1) / 2;
    for (int k = lane; k < dim; k += 32) {
        int c = (int)((sqrt(8.0 * k + 1.0) - 1.0) * 0.5);
        while (c * (c + 1) / 2 > k) --c;
        while ((c + 1) * (c + 2) / 2 <= k) ++c;
        const int r = k - c * (c + 1) / 2;
        v[k] = scale * ((r == c) ? M[r * PSD_LD + c] : M[r * PSD_LD + c] * RT2);
    }
}
// C = A * B (side x side, shared memory, row stride PSD_LD); transposes selectable
template <bool TA, bool TB>
__device__ __forceinline__ void warp_matmul(const double* A, const double* B, double* C, int side, int lane) {
    for (int e = lane; e < side * side; e += 32) {
        const int r = e / side, c = e % side;
        double s = 0.0;
        for (int k = 0; k < side; ++k) {
            const double a = TA ? A[k * PSD_LD + r] : A[r * PSD_LD + k];
            const double b = TB ? B[c * PSD_LD + k] : B[k * PSD_LD + c];
            s += a * b;
        }
        C[r * PSD_LD + c] = s;
    }
}
// in-place upper Cholesky of the matrix in M (upper triangle referenced); returns success.
// On exit M holds U in its upper triangle and zeros strictly below.
__device__ __forceinline__ bool warp_chol_upper(double* M, int side, int lane) {
    bool ok = true;
    for (int j = 0; j < side; ++j) {
        const double d = M[j * PSD_LD + j];
        if (!(d > 0.0)) {
            ok = false;
            break;
        }
        const double r = sqrt(d);
        __syncwarp();
        for (int c = j + lane; c < side; c += 32) M[j * PSD_LD + c] = (c == j) ? r : M[j * PSD_LD + c] / r;
        __syncwarp();
        for (int e = lane; e < (side - j - 1) * (side - j - 1); e += 32) {
            const int i = j + 1 + e / (side - j - 1), c = j + 1 + e % (side - j - 1);
            if (c >= i) M[i * PSD_LD + c] -= M[j * PSD_LD + i] * M[j * PSD_LD + c];
        }
        __syncwarp();
    }
    __syncwarp();
    for (int e = lane; e < side * side; e += 32) {
        const int r = e / side, c = e % side;
        if (r > c) M[r * PSD_LD + c] = 0.0;
    }
    __syncwarp();
    return ok;
}
// T = inv(U) for upper-triangular U (zeros below the diagonal in both)
__device__ __forceinline__ void warp_trtri_upper(const double* U, double* T, int side, int lane) {
    for (int e = lane; e < side * side; e += 32) T[(e / side) * PSD_LD + (e % side)] = 0.0;
    __syncwarp();
    for (int c = lane; c < side; c += 32) {
        T[c * PSD_LD + c] = 1.0 / U[c * PSD_LD + c];
        for (int i = c - 1; i >= 0; --i) {
            double s = 0.0;
            for (int k = i + 1; k <= c; ++k) s += U[i * PSD_LD + k] * T[k * PSD_LD + c];
            T[i * PSD_LD + c] = -s / U[i * PSD_LD + i];
        }
    }
    __syncwarp();
}
__device__ __forceinline__ void smem_to_global(const double* M, double* G, int side, int lane) {
    for (int e = lane; e < side * side; e += 32) {
        const int r = e % side, c = e / side;
        G[r + c * side] = M[r * PSD_LD + c];
    }
}
__device__ __forceinline__ void global_to_smem(const double* G, double* M, int side, int lane, bool trans) {
    for (int e = lane; e < side * side; e += 32) {
        const int r = e % side, c = e / side;
        if (trans) M[c * PSD_LD + r] = G[r + c * side];
        else M[r * PSD_LD + c] = G[r + c * side];
    }
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double smem_frob_dot(const double* A, const double* B, int side, int lane) {
    double s = 0.0;
    for (int e = lane; e < side * side; e += 32) s += A[(e / side) * PSD_LD + (e % side)] * B[(e / side) * PSD_LD + (e % side)];
    return warp_sum(s);
}

constexpr int PSD_WARPS = 4;
constexpr int PSD_MAT = PSD_MAX_SIDE * PSD_LD;  // doubles per shared matrix

// one warp per cone: feas, dual feas, grad, check_numerics, proxsqr; stores X, U, Uinv, Xinv
__global__ void __launch_bounds__(PSD_WARPS * 32)
psd_update_kernel(int64_t count, int side, int64_t offset, int64_t mat_off, const double* __restrict__ primal,
                  const double* __restrict__ dual_in, double scal, double irtmu, double* __restrict__ point,
                  double* __restrict__ dual, double* __restrict__ grad, double* __restrict__ mats,
                  int* __restrict__ flag, double* __restrict__ prox, const double* __restrict__ ctl) {
    extern __shared__ double sm[];
    if (ctl) {  // speculative line search (dirs.cu): a decided sequence skips; the scaling comes from the device
        if (ctl[0] != 0.0) return;
        scal = irtmu = ctl[1];
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t c = (int64_t)blockIdx.x * PSD_WARPS + warp;
    if (c >= count) return;
    double* X = sm + warp * 4 * PSD_MAT;
    double* W1 = X + PSD_MAT;
    double* W2 = W1 + PSD_MAT;
    double* W3 = W2 + PSD_MAT;
    const int dim = side * (side + 1) / 2;
    const int64_t base = offset + c * dim;
    double* p = point + base;
    double* z = dual + base;
    double* g = grad + base;
    double* cm = mats + mat_off + c * 4 * side * side;
    for (int i = lane; i < dim; i += 32) {
        p[i] = scal * primal[base + i];
        z[i] = dual_in[base + i];
    }
    __syncwarp();
    int f = 0;
    double proxsqr = 0.0;
    // dual feasibility: chol of Z (possemideftri.jl:101-104)
    svec_to_smem(z, side, W1, lane);
    __syncwarp();
    if (warp_chol_upper(W1, side, lane)) f |= HYB_ST_DUAL_FEAS;
    // primal feasibility: chol of X (possemideftri.jl:89-99)
    svec_to_smem(p, side, X, lane);
    __syncwarp();
    for (int e = lane; e < side * side; e += 32) W1[(e / side) * PSD_LD + (e % side)] = X[(e / side) * PSD_LD + (e % side)];
    __syncwarp();
    if (warp_chol_upper(W1, side, lane)) {
        f |= HYB_ST_FEAS;
        // W1 = U ; W2 = Uinv ; W3 = Xinv = Uinv Uinv'
        warp_trtri_upper(W1, W2, side, lane);
        warp_matmul<false, true>(W2, W2, W3, side, lane);
        __syncwarp();
        smem_to_global(X, cm, side, lane);
        smem_to_global(W1, cm + side * side, side, lane);
        smem_to_global(W2, cm + 2 * side * side, side, lane);
        smem_to_global(W3, cm + 3 * side * side, side, lane);
        // grad = -svec(Xinv) (possemideftri.jl:106-116)
        smem_to_svec(W3, side, g, lane, -1.0);
        __syncwarp();
        // check_numerics (Cones.jl:267-284), nu = side
        double gp = 0.0;
        for (int i = lane; i < dim; i += 32) gp += g[i] * p[i];
        gp = warp_sum(gp);
        bool ok = !(fabs(1.0 + gp / side) > 1.220703125e-4 * dim);
        // Hig = X G X with G = smat(g) = -Xinv ; <Hig, g> = <X G X, G>
        for (int e = lane; e < side * side; e += 32) W3[(e / side) * PSD_LD + (e % side)] = -W3[(e / side) * PSD_LD + (e % side)];
        __syncwarp();
        warp_matmul<false, false>(W3, X, W1, side, lane);  // G X
        __syncwarp();
        warp_matmul<false, false>(X, W1, W2, side, lane);  // X G X
        __syncwarp();
        const double hgg = smem_frob_dot(W2, W3, side, lane);
        if (fabs(1.0 - hgg / side) > 1.104854345603981e-1 * dim) ok = false;
        if (ok) f |= HYB_ST_NUMERICS;
        // proxsqr (Cones.jl:288-304): V = smat(irtmu z + g) ; <X V X, V>
        __syncwarp();
        svec_to_smem(z, side, W1, lane);
        __syncwarp();
        for (int e = lane; e < side * side; e += 32) {
            const int idx = (e / side) * PSD_LD + (e % side);
            W1[idx] = irtmu * W1[idx] + W3[idx];
        }
        __syncwarp();
        warp_matmul<false, false>(W1, X, W2, side, lane);  // V X
        __syncwarp();
        warp_matmul<false, false>(X, W2, W3, side, lane);  // X V X
        __syncwarp();
        const double acc = smem_frob_dot(W3, W1, side, lane);
        proxsqr = (acc < -1.4901161193847656e-08 * dim) ? INFINITY : fabs(acc);
    }
    if (lane == 0) {
        flag[base] = f;
        prox[base] = proxsqr;
    }
}

// prod column = svec( op(M) * smat(arr column) * op(M)' ), one warp per (cone, column)
//   HESS: M = Xinv ; INV_HESS: M = X ; SQRT: M = Uinv' ; INV_SQRT: M = U
__global__ void __launch_bounds__(PSD_WARPS * 32)
psd_prod_kernel(int64_t count, int side, int64_t offset, int64_t mat_off, int64_t m, int op,
                const double* __restrict__ mats, const double* __restrict__ arr, int64_t lda,
                double* __restrict__ prod, int64_t ldp) {
    extern __shared__ double sm[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t item = (int64_t)blockIdx.x * PSD_WARPS + warp;
    if (item >= count * m) return;
    const int64_t c = item % count, j = item / count;
    double* Ms = sm + warp * 3 * PSD_MAT;
    double* A = Ms + PSD_MAT;
    double* T = A + PSD_MAT;
    const int dim = side * (side + 1) / 2;
    const double* cm = mats + mat_off + c * 4 * side * side;
    const int which = (op == HYB_OP_HESS) ? 3 : (op == HYB_OP_INV_HESS) ? 0 : (op == HYB_OP_SQRT_HESS) ? 2 : 1;
    global_to_smem(cm + which * side * side, Ms, side, lane, op == HYB_OP_SQRT_HESS);
    svec_to_smem(arr + offset + c * dim + j * lda, side, A, lane);
    __syncwarp();
    warp_matmul<false, false>(Ms, A, T, side, lane);  // M A
    __syncwarp();
    warp_matmul<false, true>(T, Ms, A, side, lane);   // (M A) M'
    __syncwarp();
    smem_to_svec(A, side, prod + offset + c * dim + j * ldp, lane, 1.0);
}

// Wide products (many columns, e.g. the q x N matrix G*Q2 in the Schur assembly): the map
// a -> svec(M smat(a) M') is linear, so per cone it is one dim x dim matrix, the symmetric Kronecker
// product of M with itself (the reference builds the same matrix for its explicit Hessian,
// arrayutilities.jl:287-324 symm_kron!), and the product for all columns is a block-diagonal GEMM
// on the DMMA engine instead of count * m little 16 x 16 products.  With s_ij = 1 (i == j) or sqrt 2:
//   H[(i,j),(k,l)] = s_ij / s_kl * (k == l ? M_ik M_jk : M_ik M_jl + M_il M_jk)
// Stored K-major for hyb_gemm_tn_blockdiag: out[(c * dim + row) * ldh + col], row = (i,j), col = (k,l).
__global__ void psd_skron_kernel(int64_t count, int side, int64_t mat_off, int op, const double* __restrict__ mats,
                                 double* __restrict__ out, int ldh) {
    const int dim = side * (side + 1) / 2;
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= count * dim * dim) return;
    const int64_t c = e / ((int64_t)dim * dim);
    const int rem = (int)(e - c * dim * dim);
    const int row = rem / dim, col = rem - row * dim;
    auto unpack = [](int k, int& r, int& cc) {
        cc = (int)((sqrt(8.0 * k + 1.0) - 1.0) * 0.5);
        while (cc * (cc + 1) / 2 > k) --cc;
        while ((cc + 1) * (cc + 2) / 2 <= k) ++cc;
        r = k - cc * (cc + 1) / 2;
    };
    int i, j, k, l;
    unpack(row, i, j);
    unpack(col, k, l);
    const int which = (op == HYB_OP_HESS) ? 3 : (op == HYB_OP_INV_HESS) ? 0 : (op == HYB_OP_SQRT_HESS) ? 2 : 1;
    const double* G = mats + mat_off + c * 4 * side * side + which * side * side;
    const bool tr = (op == HYB_OP_SQRT_HESS);  // M = Uinv'
    auto Mel = [&](int r, int cc) { return tr ? G[cc + r * side] : G[r + cc * side]; };
    double v = (k == l) ? Mel(i, k) * Mel(j, k) : Mel(i, k) * Mel(j, l) + Mel(i, l) * Mel(j, k);
    if (i != j) v *= RT2;
    if (k != l) v *= (1.0 / RT2);
    out[(c * dim + row) * ldh + col] = v;
}

// dder3 = svec(S S'), S = Xinv D Uinv  (possemideftri.jl:194-204)
__global__ void __launch_bounds__(PSD_WARPS * 32)
psd_dder3_kernel(int64_t count, int side, int64_t offset, int64_t mat_off, const double* __restrict__ mats,
                 const double* __restrict__ dir, double* __restrict__ out) {
    extern __shared__ double sm[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t c = (int64_t)blockIdx.x * PSD_WARPS + warp;
    if (c >= count) return;
    double* Ms = sm + warp * 3 * PSD_MAT;
    double* A = Ms + PSD_MAT;
    double* T = A + PSD_MAT;
    const int dim = side * (side + 1) / 2;
    const double* cm = mats + mat_off + c * 4 * side * side;
    global_to_smem(cm + 3 * side * side, Ms, side, lane, false);  // Xinv
    svec_to_smem(dir + offset + c * dim, side, A, lane);
    __syncwarp();
    warp_matmul<false, false>(Ms, A, T, side, lane);  // Xinv D
    __syncwarp();
    global_to_smem(cm + 2 * side * side, Ms, side, lane, false);  // Uinv
    __syncwarp();
    warp_matmul<false, false>(T, Ms, A, side, lane);  // S = Xinv D Uinv
    __syncwarp();
    warp_matmul<false, true>(A, A, T, side, lane);    // S S'
    __syncwarp();
    smem_to_svec(T, side, out + offset + c * dim, lane, 1.0);
}

// ------------------------------------------------------------------ reduction over items
__global__ void __launch_bounds__(1024)
reduce_items_kernel(int64_t q, const int* __restrict__ flag, const double* __restrict__ prox, int use_max,
                    int* __restrict__ out_flag, double* __restrict__ out_prox, const double* __restrict__ ctl) {
    if (ctl && ctl[0] != 0.0) return;
    __shared__ int sf[1024];
    __shared__ double sp[1024];
    int f = HYB_ST_ALL;
    double a = 0.0;
    for (int64_t i = threadIdx.x; i < q; i += 1024) {
        f &= flag[i];
        const double v = prox[i];
        a = use_max ? fmax(a, v) : a + v;
        if (v != v) a = v;  // propagate NaN (fmax would drop it)
    }
    sf[threadIdx.x] = f;
    sp[threadIdx.x] = a;
    __syncthreads();
    for (int s = 512; s > 0; s >>= 1) {
        if (threadIdx.x < s) {
            sf[threadIdx.x] &= sf[threadIdx.x + s];
            const double x = sp[threadIdx.x], y = sp[threadIdx.x + s];
            double r = use_max ? fmax(x, y) : x + y;
            if (x != x || y != y) r = NAN;
            sp[threadIdx.x] = r;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        *out_flag = sf[0];
        *out_prox = sp[0];
    }
}

__global__ void fill_items_kernel(int64_t q, int* flag, double* prox) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < q) {
        flag[i] = HYB_ST_ALL;
        prox[i] = 0.0;
    }
}

inline unsigned nblocks(int64_t n, int per) { return (unsigned)((n + per - 1) / per); }

}  // namespace

int64_t hybi_cones_launch_count() { return g_cone_launches + hybd_launch_count(); }

int hybi_cones_alloc(HybCones& cs) {
    const int64_t q = cs.q > 0 ? cs.q : 1, K = cs.K > 0 ? cs.K : 1;
    HYB_CUDA(cudaMalloc(&cs.point, sizeof(double) * q));
    HYB_CUDA(cudaMalloc(&cs.dual, sizeof(double) * q));
    HYB_CUDA(cudaMalloc(&cs.grad, sizeof(double) * q));
    HYB_CUDA(cudaMalloc(&cs.tmp, sizeof(double) * q));
    HYB_CUDA(cudaMalloc(&cs.aux, sizeof(double) * K));
    HYB_CUDA(cudaMalloc(&cs.item_flag, sizeof(int) * q));
    HYB_CUDA(cudaMalloc(&cs.item_prox, sizeof(double) * q));
    HYB_CUDA(cudaMalloc(&cs.red_flag, sizeof(int)));
    HYB_CUDA(cudaMalloc(&cs.red_prox, sizeof(double)));
    HYB_CUDA(cudaMemset(cs.point, 0, sizeof(double) * q));
    HYB_CUDA(cudaMemset(cs.dual, 0, sizeof(double) * q));
    HYB_CUDA(cudaMemset(cs.grad, 0, sizeof(double) * q));
    HYB_CUDA(cudaMemset(cs.aux, 0, sizeof(double) * K));
    int64_t mat_total = 0;
    for (auto& b : cs.blocks) {
        if (b.type == HYB_POSSEMIDEFTRI) {
            if (b.side > PSD_MAX_SIDE) {  // large cone: dense engine (one object per cone)
                HybDensePsd* d = nullptr;
                HYB_CHECK(hybd_psd_create(&d, b.side));
                b.dense = d;
                continue;
            }
            b.mat_off = mat_total;
            mat_total += b.count * 4 * b.side * b.side;
        } else if (b.type == HYB_WSOSINTERPNONNEGATIVE) {
            HybDenseWsos* d = nullptr;
            HYB_CHECK(hybd_wsos_create(&d, b.U, b.nPs, b.Ps, b.Ls));
            b.dense = d;
        } else if (b.type == HYB_WSOSINTERPPOSSEMIDEFTRI) {
            HybDenseWsosPsd* d = nullptr;
            HYB_CHECK(hybd_wsospsd_create(&d, b.side, b.U, b.nPs, b.Ps, b.Ls));
            b.dense = d;
        } else if (b.type == HYB_WSOSINTERPEPINORMONE) {
            HybDenseWsosOne* d = nullptr;
            HYB_CHECK(hybd_wsosone_create(&d, b.side, b.U, b.nPs, b.Ps, b.Ls));
            b.dense = d;
        } else if (b.type == HYB_WSOSINTERPEPINORMEUCL) {
            HybDenseWsosEuc* d = nullptr;
            HYB_CHECK(hybd_wsoseuc_create(&d, b.side, b.U, b.nPs, b.Ps, b.Ls));
            b.dense = d;
        } else if (b.type == HYB_DOUBLYNONNEGATIVETRI) {
            HybDenseDnn* d = nullptr;
            HYB_CHECK(hybd_dnn_create(&d, b.side));
            b.dense = d;
        } else if (b.type == HYB_POSSEMIDEFTRISPARSE) {
            HybDenseDnn* d = nullptr;
            HYB_CHECK(hybd_spsd_create(&d, b.side, b.dim, b.Ls));
            b.dense = d;
        } else if (b.type != HYB_NONNEGATIVE && b.type != HYB_EPINORMEUCL && b.type != HYB_EPIPERSQUARE &&
                   b.type != HYB_EPINORMINF) {
            hyb_set_error("cone type %d is not supported by the GPU path yet", b.type);
            return -2;
        }
    }
    HYB_CUDA(cudaMalloc(&cs.mats, sizeof(double) * (mat_total > 0 ? mat_total : 1)));
    fill_items_kernel<<<nblocks(q, 256), 256>>>(cs.q, cs.item_flag, cs.item_prox);
    HYB_CUDA(cudaGetLastError());
    HYB_CUDA(cudaDeviceSynchronize());
    static HybDevOnce attr_once;
    if (HybDevOnce::Guard g_(attr_once); g_.first) {
        HYB_CUDA(cudaFuncSetAttribute(psd_update_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      PSD_WARPS * 4 * PSD_MAT * (int)sizeof(double)));
        HYB_CUDA(cudaFuncSetAttribute(psd_prod_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      PSD_WARPS * 3 * PSD_MAT * (int)sizeof(double)));
        HYB_CUDA(cudaFuncSetAttribute(psd_dder3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      PSD_WARPS * 3 * PSD_MAT * (int)sizeof(double)));
        HYB_CUDA(cudaFuncSetAttribute(soc_prod_tile_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      2 * SOC_TILE * 33 * (int)sizeof(double)));
        HYB_CUDA(cudaFuncSetAttribute(soc_prod_tile_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      2 * SOC_TILE * 33 * (int)sizeof(double)));
        HYB_CUDA(cudaFuncSetAttribute(ninf_prod_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      2 * SOC_TILE * 33 * (int)sizeof(double)));
    }
    return 0;
}

void hybi_cones_free(HybCones& cs) {
    for (auto& b : cs.blocks)
        if (b.dense) {
            if (b.type == HYB_WSOSINTERPNONNEGATIVE) hybd_wsos_destroy((HybDenseWsos*)b.dense);
            else if (b.type == HYB_WSOSINTERPPOSSEMIDEFTRI) hybd_wsospsd_destroy((HybDenseWsosPsd*)b.dense);
            else if (b.type == HYB_WSOSINTERPEPINORMEUCL) hybd_wsoseuc_destroy((HybDenseWsosEuc*)b.dense);
            else if (b.type == HYB_WSOSINTERPEPINORMONE) hybd_wsosone_destroy((HybDenseWsosOne*)b.dense);
            else if (b.type == HYB_DOUBLYNONNEGATIVETRI || b.type == HYB_POSSEMIDEFTRISPARSE) hybd_dnn_destroy((HybDenseDnn*)b.dense);
            else hybd_psd_destroy((HybDensePsd*)b.dense);
            b.dense = nullptr;
        }
    cudaFree(cs.point);
    cudaFree(cs.dual);
    cudaFree(cs.grad);
    cudaFree(cs.tmp);
    cudaFree(cs.aux);
    cudaFree(cs.mats);
    cudaFree(cs.skron);
    cudaFree(cs.item_flag);
    cudaFree(cs.item_prox);
    cudaFree(cs.red_flag);
    cudaFree(cs.red_prox);
    cs = HybCones();
}

// every block is served by one closed-form kernel family (no host-sequenced dense engine)
bool hybi_cones_all_batched(const HybCones& cs) {
    for (const auto& b : cs.blocks)
        if (b.dense || b.type == HYB_WSOSINTERPNONNEGATIVE || b.type == HYB_WSOSINTERPPOSSEMIDEFTRI) return false;
    return true;
}

int hybi_cones_update(cudaStream_t st, HybCones& cs, const double* primal, const double* dual_in, double scal,
                     double irtmu, int use_max_prox, const double* ctl) {
    (void)use_max_prox;
    if (ctl && !hybi_cones_all_batched(cs)) {
        hyb_set_error("device-controlled cone update: the model has cones of the dense engines");
        return -2;
    }
    for (const auto& b : cs.blocks) {
        const int64_t rows = b.count * b.dim;
        switch (b.type) {
            case HYB_NONNEGATIVE:
                nonneg_update_kernel<<<nblocks(rows, 256), 256, 0, st>>>(rows, primal + b.offset, dual_in + b.offset, scal,
                                                                         irtmu, cs.point + b.offset, cs.dual + b.offset,
                                                                         cs.grad + b.offset, cs.item_flag + b.offset,
                                                                         cs.item_prox + b.offset, ctl);
                break;
            case HYB_EPINORMEUCL:
            case HYB_EPIPERSQUARE: {
                const bool warp = b.dim > 32;
                const int64_t threads = warp ? b.count * 32 : b.count;
#define SOC_UPD(T, W)                                                                                              \
    soc_update_kernel<T, W><<<nblocks(threads, 128), 128, 0, st>>>(b.count, b.dim, b.offset, b.cone0, primal, dual_in, \
                                                                   scal, irtmu, cs.point, cs.dual, cs.grad, cs.aux,   \
                                                                   cs.item_flag, cs.item_prox, ctl)
                if (b.type == HYB_EPINORMEUCL) {
                    if (warp) SOC_UPD(1, true); else SOC_UPD(1, false);
                } else {
                    if (warp) SOC_UPD(2, true); else SOC_UPD(2, false);
                }
#undef SOC_UPD
                break;
            }
            case HYB_EPINORMINF: {
                const bool warp = b.dim > 32;
                const int64_t threads = warp ? b.count * 32 : b.count;
                if (warp)
                    ninf_update_kernel<true><<<nblocks(threads, 128), 128, 0, st>>>(b.count, b.dim, b.offset, b.cone0, primal,
                                                                                  dual_in, scal, irtmu, cs.point, cs.dual,
                                                                                  cs.grad, cs.aux, cs.item_flag, cs.item_prox, ctl);
                else
                    ninf_update_kernel<false><<<nblocks(threads, 128), 128, 0, st>>>(b.count, b.dim, b.offset, b.cone0, primal,
                                                                                   dual_in, scal, irtmu, cs.point, cs.dual,
                                                                                   cs.grad, cs.aux, cs.item_flag, cs.item_prox, ctl);
                break;
            }
            case HYB_POSSEMIDEFTRI:
                if (b.dense) {
                    HYB_CHECK(hybd_psd_update(st, (HybDensePsd*)b.dense, primal + b.offset, dual_in + b.offset, scal, irtmu,
                                              cs.point + b.offset, cs.dual + b.offset, cs.grad + b.offset,
                                              cs.item_flag + b.offset, cs.item_prox + b.offset));
                    break;
                }
                psd_update_kernel<<<nblocks(b.count, PSD_WARPS), PSD_WARPS * 32,
                                    PSD_WARPS * 4 * PSD_MAT * sizeof(double), st>>>(
                    b.count, b.side, b.offset, b.mat_off, primal, dual_in, scal, irtmu, cs.point, cs.dual, cs.grad,
                    cs.mats, cs.item_flag, cs.item_prox, ctl);
                break;
            case HYB_WSOSINTERPNONNEGATIVE:
                HYB_CHECK(hybd_wsos_update(st, (HybDenseWsos*)b.dense, primal + b.offset, dual_in + b.offset, scal, irtmu,
                                           cs.point + b.offset, cs.dual + b.offset, cs.grad + b.offset,
                                           cs.item_flag + b.offset, cs.item_prox + b.offset));
                break;
            case HYB_WSOSINTERPPOSSEMIDEFTRI:
                HYB_CHECK(hybd_wsospsd_update(st, (HybDenseWsosPsd*)b.dense, primal + b.offset, dual_in + b.offset, scal,
                                              irtmu, cs.point + b.offset, cs.dual + b.offset, cs.grad + b.offset,
                                              cs.item_flag + b.offset, cs.item_prox + b.offset));
                break;
            case HYB_WSOSINTERPEPINORMONE:
                HYB_CHECK(hybd_wsosone_update(st, (HybDenseWsosOne*)b.dense, primal + b.offset, dual_in + b.offset, scal, irtmu,
                                              cs.point + b.offset, cs.dual + b.offset, cs.grad + b.offset,
                                              cs.item_flag + b.offset, cs.item_prox + b.offset));
                break;
            case HYB_WSOSINTERPEPINORMEUCL:
                HYB_CHECK(hybd_wsoseuc_update(st, (HybDenseWsosEuc*)b.dense, primal + b.offset, dual_in + b.offset, scal, irtmu,
                                              cs.point + b.offset, cs.dual + b.offset, cs.grad + b.offset,
                                              cs.item_flag + b.offset, cs.item_prox + b.offset));
                break;
            case HYB_DOUBLYNONNEGATIVETRI:
            case HYB_POSSEMIDEFTRISPARSE:
                HYB_CHECK(hybd_dnn_update(st, (HybDenseDnn*)b.dense, primal + b.offset, dual_in + b.offset, scal, irtmu,
                                          cs.point + b.offset, cs.dual + b.offset, cs.grad + b.offset,
                                          cs.item_flag + b.offset, cs.item_prox + b.offset));
                break;
            default:
                hyb_set_error("cone type %d unsupported", b.type);
                return -2;
        }
        LAUNCHED();
        HYB_CUDA(cudaGetLastError());
    }
    return 0;
}

// device-side form of hybi_cones_reduce: red_flag / red_prox stay on the device (skipped when ctl[0] != 0)
int hybi_cones_reduce_dev(cudaStream_t st, HybCones& cs, int use_max_prox, const double* ctl) {
    reduce_items_kernel<<<1, 1024, 0, st>>>(cs.q, cs.item_flag, cs.item_prox, use_max_prox, cs.red_flag, cs.red_prox, ctl);
    LAUNCHED();
    HYB_CUDA(cudaGetLastError());
    return 0;
}

int hybi_cones_reduce(cudaStream_t st, HybCones& cs, int use_max_prox, int* all_flags, double* agg) {
    reduce_items_kernel<<<1, 1024, 0, st>>>(cs.q, cs.item_flag, cs.item_prox, use_max_prox, cs.red_flag, cs.red_prox, nullptr);
    LAUNCHED();
    HYB_CUDA(cudaGetLastError());
    HYB_CUDA(cudaMemcpyAsync(all_flags, cs.red_flag, sizeof(int), cudaMemcpyDeviceToHost, st));
    HYB_CUDA(cudaMemcpyAsync(agg, cs.red_prox, sizeof(double), cudaMemcpyDeviceToHost, st));
    HYB_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int hybi_cones_prod(cudaStream_t st, HybCones& cs, int which, const double* arr, int64_t lda, double* prod, int64_t ldp,
                   int64_t m) {
    if (m <= 0) return 0;
    for (size_t bi = 0; bi < cs.blocks.size(); ++bi) {
        const auto& b = cs.blocks[bi];
        if (!cs.block_mask.empty() && !cs.block_mask[bi]) continue;
        int op;
        if (which == 4) op = b.use_dual ? HYB_OP_INV_HESS : HYB_OP_HESS;
        else if (which == 5 || (which == 7 && cs.block_sqrt[bi])) op = b.use_dual ? HYB_OP_INV_SQRT_HESS : HYB_OP_SQRT_HESS;
        else if (which == 7) op = b.use_dual ? HYB_OP_INV_HESS : HYB_OP_HESS;
        else if (which == 8) op = b.use_dual ? HYB_OP_HESS : HYB_OP_INV_HESS;  // z-z block of the symmetric-indefinite system
        else if (which == 6)  // hess_prod_slow! is hess_prod! except for the WSOS cones (Cones.jl:229-231)
            op = (b.type == HYB_WSOSINTERPNONNEGATIVE || b.type == HYB_WSOSINTERPPOSSEMIDEFTRI ||
                  b.type == HYB_POSSEMIDEFTRISPARSE) ? 6 : HYB_OP_HESS;
        else op = which;
        const int64_t rows = b.count * b.dim;
        switch (b.type) {
            case HYB_NONNEGATIVE: {
                dim3 grid(nblocks(rows, 256), (unsigned)(m < 1024 ? m : 1024));
                nonneg_prod_kernel<<<grid, 256, 0, st>>>(rows, m, op, cs.point + b.offset, arr + b.offset, lda,
                                                         prod + b.offset, ldp);
                break;
            }
            case HYB_EPINORMEUCL:
            case HYB_EPIPERSQUARE: {
                const bool warp = b.dim > 32;
                const int64_t threads = (warp ? 32 : 1) * b.count * m;
#define SOC_PROD(T, W)                                                                                               \
    soc_prod_kernel<T, W><<<nblocks(threads, 128), 128, 0, st>>>(b.count, b.dim, b.offset, b.cone0, m, op, cs.point, \
                                                                 cs.aux, arr, lda, prod, ldp)
                if (!warp && m >= 32 && b.count >= 32) {  // wide product of many small cones: coalesced tiles
                    dim3 grid(nblocks(b.count, SOC_TILE), (unsigned)((m + SOC_CPB - 1) / SOC_CPB));
                    const size_t smem = sizeof(double) * 2 * SOC_TILE * (b.dim + 1);
                    if (b.type == HYB_EPINORMEUCL)
                        soc_prod_tile_kernel<1><<<grid, SOC_TILE, smem, st>>>(b.count, b.dim, b.offset, b.cone0, m, op, cs.point,
                                                                             cs.aux, arr, lda, prod, ldp);
                    else
                        soc_prod_tile_kernel<2><<<grid, SOC_TILE, smem, st>>>(b.count, b.dim, b.offset, b.cone0, m, op, cs.point,
                                                                             cs.aux, arr, lda, prod, ldp);
                    break;
                }
                if (b.type == HYB_EPINORMEUCL) {
                    if (warp) SOC_PROD(1, true); else SOC_PROD(1, false);
                } else {
                    if (warp) SOC_PROD(2, true); else SOC_PROD(2, false);
                }
#undef SOC_PROD
                break;
            }
            case HYB_EPINORMINF: {
                if (op != HYB_OP_HESS && op != HYB_OP_INV_HESS) {
                    hyb_set_error("EpiNormInf has no sqrt Hessian oracles (epinorminf.jl:73)");
                    return -3;
                }
                const bool warp = b.dim > 32;
                const int64_t threads = (warp ? 32 : 1) * b.count * m;
                if (!warp && m >= 32 && b.count >= 32) {
                    dim3 grid(nblocks(b.count, SOC_TILE), (unsigned)((m + SOC_CPB - 1) / SOC_CPB));
                    ninf_prod_tile_kernel<<<grid, SOC_TILE, sizeof(double) * 2 * SOC_TILE * (b.dim + 1), st>>>(
                        b.count, b.dim, b.offset, b.cone0, m, op, cs.point, cs.aux, arr, lda, prod, ldp);
                    break;
                }
                if (warp)
                    ninf_prod_kernel<true><<<nblocks(threads, 128), 128, 0, st>>>(b.count, b.dim, b.offset, b.cone0, m, op,
                                                                                cs.point, cs.aux, arr, lda, prod, ldp);
                else
                    ninf_prod_kernel<false><<<nblocks(threads, 128), 128, 0, st>>>(b.count, b.dim, b.offset, b.cone0, m, op,
                                                                                 cs.point, cs.aux, arr, lda, prod, ldp);
                break;
            }
            case HYB_POSSEMIDEFTRI:
                if (b.dense) {
                    HYB_CHECK(hybd_psd_prod(st, (HybDensePsd*)b.dense, op, arr + b.offset, lda, prod + b.offset, ldp, m));
                    break;
                }
                {
                    // many columns: block-diagonal DMMA GEMM with the symmetric-Kronecker matrices
                    const int64_t ldh = (b.dim + 1) & ~(int64_t)1;
                    const int64_t need = b.count * b.dim * ldh;
                    const bool aligned = !(b.offset & 1) && !(b.dim & 1) && !(lda & 1) && !((uintptr_t)arr & 15);
                    if (m >= 64 && b.dim >= 28 && aligned && b.count <= 65535 && need <= ((int64_t)1 << 28)) {
                        if (cs.skron_cap < need) {
                            HYB_CUDA(cudaStreamSynchronize(st));
                            cudaFree(cs.skron);
                            cs.skron = nullptr;
                            cs.skron_cap = 0;
                            HYB_CUDA(cudaMalloc(&cs.skron, sizeof(double) * need));
                            cs.skron_cap = need;
                        }
                        if (ldh != b.dim) HYB_CUDA(cudaMemsetAsync(cs.skron, 0, sizeof(double) * need, st));
                        psd_skron_kernel<<<nblocks(b.count * b.dim * b.dim, 256), 256, 0, st>>>(b.count, b.side, b.mat_off, op,
                                                                                              cs.mats, cs.skron, (int)ldh);
                        LAUNCHED();
                        HYB_CUDA(cudaGetLastError());
                        HYB_CHECK(hyb_gemm_tn_blockdiag(st, b.dim, m, b.dim, b.dim, b.count, 1.0, cs.skron, ldh,
                                                        arr + b.offset, lda, 0.0, prod + b.offset, ldp));
                        break;
                    }
                }
                psd_prod_kernel<<<nblocks(b.count * m, PSD_WARPS), PSD_WARPS * 32,
                                  PSD_WARPS * 3 * PSD_MAT * sizeof(double), st>>>(b.count, b.side, b.offset, b.mat_off, m,
                                                                                 op, cs.mats, arr, lda, prod, ldp);
                break;
            case HYB_WSOSINTERPNONNEGATIVE:
                HYB_CHECK(hybd_wsos_prod(st, (HybDenseWsos*)b.dense, op, arr + b.offset, lda, prod + b.offset, ldp, m));
                break;
            case HYB_WSOSINTERPPOSSEMIDEFTRI:
                HYB_CHECK(hybd_wsospsd_prod(st, (HybDenseWsosPsd*)b.dense, op, arr + b.offset, lda, prod + b.offset, ldp, m));
                break;
            case HYB_WSOSINTERPEPINORMONE:
                HYB_CHECK(hybd_wsosone_prod(st, (HybDenseWsosOne*)b.dense, op, arr + b.offset, lda, prod + b.offset, ldp, m));
                break;
            case HYB_WSOSINTERPEPINORMEUCL:
                HYB_CHECK(hybd_wsoseuc_prod(st, (HybDenseWsosEuc*)b.dense, op, arr + b.offset, lda, prod + b.offset, ldp, m));
                break;
            case HYB_DOUBLYNONNEGATIVETRI:
            case HYB_POSSEMIDEFTRISPARSE:
                HYB_CHECK(hybd_dnn_prod(st, (HybDenseDnn*)b.dense, op, arr + b.offset, lda, prod + b.offset, ldp, m));
                break;
            default:
                hyb_set_error("cone type %d unsupported", b.type);
                return -2;
        }
        LAUNCHED();
        HYB_CUDA(cudaGetLastError());
    }
    return 0;
}

int hybi_cones_dder3(cudaStream_t st, HybCones& cs, const double* dir, double* out) {
    for (const auto& b : cs.blocks) {
        const int64_t rows = b.count * b.dim;
        switch (b.type) {
            case HYB_NONNEGATIVE:
                nonneg_dder3_kernel<<<nblocks(rows, 256), 256, 0, st>>>(rows, cs.point + b.offset, dir + b.offset,
                                                                        out + b.offset);
                break;
            case HYB_EPINORMEUCL:
            case HYB_EPIPERSQUARE: {
                const bool warp = b.dim > 32;
                const int64_t threads = warp ? b.count * 32 : b.count;
#define SOC_D3(T, W)                                                                                            \
    soc_dder3_kernel<T, W><<<nblocks(threads, 128), 128, 0, st>>>(b.count, b.dim, b.offset, b.cone0, cs.point, cs.aux, \
                                                                  dir, out)
                if (b.type == HYB_EPINORMEUCL) {
                    if (warp) SOC_D3(1, true); else SOC_D3(1, false);
                } else {
                    if (warp) SOC_D3(2, true); else SOC_D3(2, false);
                }
#undef SOC_D3
                break;
            }
            case HYB_EPINORMINF: {
                const bool warp = b.dim > 32;
                const int64_t threads = warp ? b.count * 32 : b.count;
                if (warp)
                    ninf_dder3_kernel<true><<<nblocks(threads, 128), 128, 0, st>>>(b.count, b.dim, b.offset, cs.point, dir, out);
                else
                    ninf_dder3_kernel<false><<<nblocks(threads, 128), 128, 0, st>>>(b.count, b.dim, b.offset, cs.point, dir, out);
                break;
            }
            case HYB_POSSEMIDEFTRI:
                if (b.dense) {
                    HYB_CHECK(hybd_psd_dder3(st, (HybDensePsd*)b.dense, dir + b.offset, out + b.offset));
                    break;
                }
                psd_dder3_kernel<<<nblocks(b.count, PSD_WARPS), PSD_WARPS * 32, PSD_WARPS * 3 * PSD_MAT * sizeof(double),
                                   st>>>(b.count, b.side, b.offset, b.mat_off, cs.mats, dir, out);
                break;
            case HYB_WSOSINTERPNONNEGATIVE:
                HYB_CHECK(hybd_wsos_dder3(st, (HybDenseWsos*)b.dense, dir + b.offset, out + b.offset));
                break;
            case HYB_WSOSINTERPPOSSEMIDEFTRI:
                HYB_CHECK(hybd_wsospsd_dder3(st, (HybDenseWsosPsd*)b.dense, dir + b.offset, out + b.offset));
                break;
            case HYB_WSOSINTERPEPINORMONE:
                HYB_CHECK(hybd_wsosone_dder3(st, (HybDenseWsosOne*)b.dense, dir + b.offset, out + b.offset));
                break;
            case HYB_WSOSINTERPEPINORMEUCL:
                HYB_CHECK(hybd_wsoseuc_dder3(st, (HybDenseWsosEuc*)b.dense, dir + b.offset, out + b.offset));
                break;
            case HYB_DOUBLYNONNEGATIVETRI:
            case HYB_POSSEMIDEFTRISPARSE:
                HYB_CHECK(hybd_dnn_dder3(st, (HybDenseDnn*)b.dense, dir + b.offset, out + b.offset));
                break;
            default:
                hyb_set_error("cone type %d unsupported", b.type);
                return -2;
        }
        LAUNCHED();
        HYB_CUDA(cudaGetLastError());
    }
    return 0;
}
