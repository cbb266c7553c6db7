// Device-resident stepper inner loop: the iterate (x, z, s) lives on the device for the
// whole solve; the host keeps tau, kap, mu and sees a handful of scalars per iteration.
//
//   hyb_step_point_set / _get      the iterate goes up once (initialize_point, Solvers.jl:499-516) and comes down
//                                  once (postprocess, process.jl:415-469)
//   hyb_step_convergence           calc_convergence_params (Solvers.jl:631-693): residual vectors kept on the
//                                  device (they ARE the prediction right-hand side), norms / objective pieces out
//   hyb_step_rhs                   update_rhs_pred / _cent / _predadj / _centadj (steppers/common.jl:14-119) into a
//                                  column of the direction engine's right-hand sides
//   hyb_step_directions            get_directions (systemsolvers/common.jl:22-83) on those columns; only the
//                                  (tau, kap) entries of the directions and the residual norms come back
//   hyb_step_point_update          update_stepper_points(alpha, point, false) (combined.jl:129-177)
//
// Together with hyb_search_gates / hyb_search_check (dirs.cu) an interior-point iteration is a fixed
// sequence of about ten C-ABI calls with no q- or N-vector crossing PCIe.
#include "dirs.cuh"

namespace {

constexpr int RT = 1024;

__device__ __forceinline__ double blk_sum(double v, double* sh) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = (threadIdx.x < RT / 32) ? sh[threadIdx.x] : 0.0;
    if (threadIdx.x < 32) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    }
    return t;  // valid in warp 0
}
// maximum that propagates NaN (fmax would drop it; the reference's norm(., Inf) returns NaN)
__device__ __forceinline__ double nanmax(double a, double b) { return (a != a || b != b) ? NAN : fmax(a, b); }
__device__ __forceinline__ double blk_max(double v, double* sh) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = nanmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    __syncthreads();
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = (threadIdx.x < RT / 32) ? sh[threadIdx.x] : 0.0;
    if (threadIdx.x < 32) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t = nanmax(t, __shfl_xor_sync(0xffffffffu, t, o));
    }
    return t;
}

// calc_convergence_params (Solvers.jl:631-693, p = 0) from T = G'z and Gx = G x:
//   x residual: xr = G'z (norm -> out[0]); xr += c tau (norm -> out[1]); x_residual = -xr
//   z residual: zr = G x + s (norm -> out[2]); zr -= h tau (norm -> out[3]); z_residual = zr
//   out[4] = c'x, out[5] = h'z, out[6] = z's
__global__ void __launch_bounds__(RT)
conv_kernel(int64_t N, int64_t q, double tau, const double* __restrict__ c, const double* __restrict__ h,
            const double* __restrict__ T, const double* __restrict__ Gx, const double* __restrict__ px,
            const double* __restrict__ pz, const double* __restrict__ ps, double* __restrict__ resx,
            double* __restrict__ resz, double* __restrict__ out) {
    __shared__ double sh[RT / 32];
    double m0 = 0.0, m1 = 0.0, m2 = 0.0, m3 = 0.0, cx = 0.0, hz = 0.0, zs = 0.0;
    for (int64_t i = threadIdx.x; i < N; i += RT) {
        const double t = T[i];
        m0 = nanmax(m0, fabs(t));
        const double xr = t + c[i] * tau;
        m1 = nanmax(m1, fabs(xr));
        resx[i] = -xr;
        cx += c[i] * px[i];
    }
    for (int64_t i = threadIdx.x; i < q; i += RT) {
        const double zr0 = Gx[i] + ps[i];
        m2 = nanmax(m2, fabs(zr0));
        const double zr = zr0 - h[i] * tau;
        m3 = nanmax(m3, fabs(zr));
        resz[i] = zr;
        hz += h[i] * pz[i];
        zs += pz[i] * ps[i];
    }
    const double r0 = blk_max(m0, sh), r1 = blk_max(m1, sh), r2 = blk_max(m2, sh), r3 = blk_max(m3, sh);
    const double s0 = blk_sum(cx, sh), s1 = blk_sum(hz, sh), s2 = blk_sum(zs, sh);
    if (threadIdx.x == 0) {
        out[0] = r0;
        out[1] = r1;
        out[2] = r2;
        out[3] = r3;
        out[4] = s0;
        out[5] = s1;
        out[6] = s2;
    }
}

// update_rhs_pred (steppers/common.jl:14-28) and update_rhs_cent (:66-83), elementwise part:
//   pred: x = x_residual, z = z_residual, s = -dual            cent: x = 0, z = 0, s = -dual - rtmu * grad
// (dual = the cone's dual slice of the iterate: s for dual-barrier cones, z otherwise, point.jl:52-59)
__global__ void rhs_lin_kernel(int64_t N, int64_t q, int pred, double rtmu, const double* __restrict__ resx,
                               const double* __restrict__ resz, const double* __restrict__ pz, const double* __restrict__ ps,
                               const unsigned char* __restrict__ dual_row, const double* __restrict__ grad,
                               double* __restrict__ Rx, double* __restrict__ Rz, double* __restrict__ Rs) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < N) Rx[i] = pred ? resx[i] : 0.0;
    if (i < q) {
        Rz[i] = pred ? resz[i] : 0.0;
        const double dual = dual_row[i] ? ps[i] : pz[i];
        Rs[i] = pred ? -dual : (-dual - rtmu * grad[i]);
    }
}

// update_rhs_predadj / update_rhs_centadj (steppers/common.jl:30-63, :86-119), first half: the arguments of the
// two oracle calls.  prim_dir = the cone's primal slice of the kept direction; prim_scal = prim_dir / mu^(1/4).
//   A0 = argument of hess_prod_slow! (prim_dir for predadj, prim_scal for centadj), A1 = prim_scal (dder3)
__global__ void adj_args_kernel(int64_t q, int pred, double irtrtmu, const double* __restrict__ kz,
                                const double* __restrict__ ks, const unsigned char* __restrict__ dual_row,
                                double* __restrict__ A0, double* __restrict__ A1) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= q) return;
    const double pd = dual_row[i] ? kz[i] : ks[i];
    const double sc = irtrtmu * pd;
    A0[i] = pred ? pd : sc;
    A1[i] = sc;
}

// second half, one warp per cone: the identity gate
//   dot1 = <dder3_k, point_k>,  dot2 = <prim_scal_k, H_k> (x 1/mu^(1/4) for predadj)
//   keep iff |dot1 - dot2| / (sqrt(eps) + |dot2|) < 1e-4            (steppers/common.jl:49-55, :105-111)
// and rhs.s_k = keep ? (predadj: H_k + dder3_k, centadj: dder3_k) : 0;  x, z parts of the column are zero.
__global__ void __launch_bounds__(256)
adj_gate_kernel(int64_t K, int pred, double irtrtmu, const int64_t* __restrict__ off, const int* __restrict__ dim,
                const double* __restrict__ point, const double* __restrict__ A1, const double* __restrict__ H,
                const double* __restrict__ D3, double* __restrict__ Rs) {
    const int lane = threadIdx.x & 31;
    const int64_t k = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (k >= K) return;
    const int64_t o = off[k];
    const int n = dim[k];
    double d1 = 0.0, d2 = 0.0;
    for (int r = lane; r < n; r += 32) {
        d1 += D3[o + r] * point[o + r];
        d2 += A1[o + r] * H[o + r];
    }
#pragma unroll
    for (int w = 16; w > 0; w >>= 1) {
        d1 += __shfl_xor_sync(0xffffffffu, d1, w);
        d2 += __shfl_xor_sync(0xffffffffu, d2, w);
    }
    if (pred) d2 = irtrtmu * d2;
    const double viol = fabs(d1 - d2) / (1.4901161193847656e-08 + fabs(d2));
    const bool keep = viol < 1e-4;  // a NaN comparison is false: the cone is skipped, as in the reference
    for (int r = lane; r < n; r += 32) Rs[o + r] = keep ? (pred ? H[o + r] + D3[o + r] : D3[o + r]) : 0.0;
}

// update_stepper_points(alpha, point, false) (combined.jl:129-177) on x, z, s: point += sum_k coef_k dir_k with the
// same left-to-right order of additions as the host (pred, predadj, cent, centadj = slots 2, 3, 0, 1)
__global__ void point_update_kernel(int64_t N, int64_t q, int64_t ldn, int64_t ldq, double c0, double c1, double c2,
                                    double c3, const double* __restrict__ kx, const double* __restrict__ kz,
                                    const double* __restrict__ ks, double* __restrict__ px, double* __restrict__ pz,
                                    double* __restrict__ ps) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const double cf[4] = {c0, c1, c2, c3};
    const int order[4] = {2, 3, 0, 1};
    auto comb = [&](const double* k, int64_t ld, int64_t idx) {
        double acc = 0.0;
        bool first = true;
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int s = order[t];
            if (cf[s] != 0.0) {
                const double v = cf[s] * k[idx + (int64_t)s * ld];
                acc = first ? v : acc + v;
                first = false;
            }
        }
        return acc;
    };
    if (i < N) px[i] += comb(kx, ldn, i);
    if (i < q) {
        pz[i] += comb(kz, ldq, i);
        ps[i] += comb(ks, ldq, i);
    }
}

__global__ void __launch_bounds__(RT)
dot_zs_kernel(int64_t q, const double* __restrict__ pz, const double* __restrict__ ps, double* __restrict__ out) {
    __shared__ double sh[RT / 32];
    double s = 0.0;
    bool bad = false;
    for (int64_t i = threadIdx.x; i < q; i += RT) {
        const double a = pz[i], b = ps[i];
        bad |= (a != a) || (b != b);
        s += a * b;
    }
    const double t = blk_sum(bad ? NAN : s, sh);
    if (threadIdx.x == 0) out[0] = t;
}

}  // namespace

static inline unsigned nb256(int64_t n) { return (unsigned)(((n > 0 ? n : 1) + 255) / 256); }

extern "C" {

int hyb_step_point_set(hyb_ctx* c, const double* x, const double* z, const double* s) {
    if (!c || !c->loaded || !c->dirs) {
        hyb_set_error("hyb_step_point_set: call hyb_dirs_setup first");
        return -3;
    }
    HybDirs& d = *c->dirs;
    HYB_CUDA(cudaSetDevice(c->device));
    if (d.N > 0) HYB_CUDA(h2d(c, d.px, x, sizeof(double) * d.N));
    if (d.q > 0) {
        HYB_CUDA(h2d(c, d.pz, z, sizeof(double) * d.q));
        HYB_CUDA(h2d(c, d.ps, s, sizeof(double) * d.q));
    }
    HYB_CUDA(cudaStreamSynchronize(c->st));
    return 0;
}

int hyb_step_point_get(hyb_ctx* c, double* x, double* z, double* s) {
    if (!c || !c->loaded || !c->dirs) return -1;
    HybDirs& d = *c->dirs;
    HYB_CUDA(cudaSetDevice(c->device));
    if (d.N > 0 && x) HYB_CUDA(d2h(c, x, d.px, sizeof(double) * d.N));
    if (d.q > 0) {
        if (z) HYB_CUDA(d2h(c, z, d.pz, sizeof(double) * d.q));
        if (s) HYB_CUDA(d2h(c, s, d.ps, sizeof(double) * d.q));
    }
    HYB_CUDA(cudaStreamSynchronize(c->st));
    return 0;
}

int hyb_step_convergence(hyb_ctx* c, double tau, double* out7) {
    if (!c || !c->loaded || !c->dirs) return -1;
    HybDirs& d = *c->dirs;
    HYB_CUDA(cudaSetDevice(c->device));
    cudaStream_t st = c->st;
    const int64_t N = d.N, q = d.q;
    KBEGIN(c);
    if (N > 0 && q > 0) {
        HYB_CHECK(hyb_gemv_t(st, q, N, c->GQ2, c->ldq, d.pz, d.T));                       // G'z  (Solvers.jl:638)
        HYB_CHECK(hyb_gemv_n(st, q, N, c->GQ2, c->ldq, d.px, d.Gx, d.gemm_scratch));      // G x  (Solvers.jl:656)
        g_misc_launches += 3;
    } else {
        if (N > 0) HYB_CUDA(cudaMemsetAsync(d.T, 0, sizeof(double) * N, st));
        if (q > 0) HYB_CUDA(cudaMemsetAsync(d.Gx, 0, sizeof(double) * q, st));
    }
    conv_kernel<<<1, RT, 0, st>>>(N, q, tau, d.c, d.h, d.T, d.Gx, d.px, d.pz, d.ps, d.resx, d.resz, d.sout);
    HYB_CUDA(cudaGetLastError());
    ++g_misc_launches;
    KEND(c);
    HYB_CUDA(d2h(c, out7, d.sout, sizeof(double) * 7));
    HYB_CUDA(cudaStreamSynchronize(st));
    ksync(c);
    return 0;
}

// kind: 0 cent, 1 centadj, 2 pred, 3 predadj (the slot numbering of hyb_dirs_keep).  col: column of the
// direction engine's right-hand sides to fill.  The adjustments read the kept unadjusted direction of their
// kind (slot 0 for centadj, 2 for predadj) and its tau entry `dir_tau`.
int hyb_step_rhs(hyb_ctx* c, int kind, int col, double mu, double tau, double kap, double tau_residual, double dir_tau) {
    if (!c || !c->loaded || !c->dirs) return -1;
    HybDirs& d = *c->dirs;
    if (kind < 0 || kind > 3 || col < 0 || col >= HYB_MAX_RHS) {
        hyb_set_error("hyb_step_rhs: kind %d / column %d out of range", kind, col);
        return -3;
    }
    HYB_CUDA(cudaSetDevice(c->device));
    if (c->updated_irtmu != c->updated_irtmu) HYB_CHECK(hyb_ensure_updated(c, c->loaded_scal));
    cudaStream_t st = c->st;
    const int64_t N = d.N, q = d.q, nq = N > q ? N : q;
    double* Rx = d.rhs.x + (int64_t)col * d.ldn;
    double* Rz = d.rhs.z + (int64_t)col * d.ldq;
    double* Rs = d.rhs.s + (int64_t)col * d.ldq;
    double tk[2];
    KBEGIN(c);
    if (kind == 0 || kind == 2) {
        const int pred = kind == 2;
        rhs_lin_kernel<<<nb256(nq), 256, 0, st>>>(N, q, pred, sqrt(mu), d.resx, d.resz, d.pz, d.ps, d.dual_row, c->cones.grad,
                                                  Rx, Rz, Rs);
        HYB_CUDA(cudaGetLastError());
        ++g_misc_launches;
        tk[0] = pred ? tau_residual : 0.0;
        tk[1] = pred ? -kap : (-kap + mu / tau);
    } else {
        const int pred = kind == 3;
        const int src = pred ? 2 : 0;
        const double irtrtmu = 1.0 / sqrt(sqrt(mu));
        if (N > 0) HYB_CUDA(cudaMemsetAsync(Rx, 0, sizeof(double) * N, st));
        if (q > 0) {
            HYB_CUDA(cudaMemsetAsync(Rz, 0, sizeof(double) * q, st));
            double *A0 = d.A, *A1 = d.A + d.ldq, *B0 = d.B, *B1 = d.B + d.ldq;
            adj_args_kernel<<<nb256(q), 256, 0, st>>>(q, pred, irtrtmu, d.kz + (int64_t)src * d.ldq, d.ks + (int64_t)src * d.ldq,
                                                      d.dual_row, A0, A1);
            HYB_CUDA(cudaGetLastError());
            HYB_CHECK(hybi_cones_prod(st, c->cones, 6, A0, d.ldq, B0, d.ldq, 1));  // hess_prod_slow!
            HYB_CHECK(hybi_cones_dder3(st, c->cones, A1, B1));
            adj_gate_kernel<<<(unsigned)((d.K + 7) / 8), 256, 0, st>>>(d.K, pred, irtrtmu, d.cone_off, d.cone_dim, c->cones.point,
                                                                      A1, B0, B1, Rs);
            HYB_CUDA(cudaGetLastError());
            g_misc_launches += 2;
        }
        const double tdt = dir_tau / tau;
        tk[0] = 0.0;
        tk[1] = pred ? tdt * mu / tau * (1.0 + tdt) : tdt * mu / tau * tdt;
    }
    KEND(c);
    HYB_CUDA(h2d(c, d.rhs.tk + 2 * col, tk, sizeof(double) * 2));
    HYB_CUDA(cudaStreamSynchronize(st));
    ksync(c);
    return 0;
}

int hyb_step_directions(hyb_ctx* c, int nrhs, int refresh_const, double tau_bar, double mu, double res_norm_cutoff,
                        int max_ref_steps, double min_impr_tol, int* num_ref, double* res_norm, double* dir_taukap) {
    if (!c || !c->loaded || !c->dirs) return -1;
    HybDirs& d = *c->dirs;
    if (nrhs < 0 || nrhs + (refresh_const ? 1 : 0) > HYB_MAX_RHS) {
        hyb_set_error("hyb_step_directions: %d right-hand sides (+%d) exceed the batch limit %d", nrhs, refresh_const ? 1 : 0,
                      HYB_MAX_RHS);
        return -3;
    }
    HYB_CUDA(cudaSetDevice(c->device));
    if (c->updated_irtmu != c->updated_irtmu) HYB_CHECK(hyb_ensure_updated(c, c->loaded_scal));
    KBEGIN(c);
    HYB_CHECK(hyb_directions_core(c, nrhs, refresh_const != 0, mu / tau_bar / tau_bar, res_norm_cutoff, max_ref_steps,
                                  min_impr_tol, num_ref, res_norm));
    KEND(c);
    if (nrhs > 0 && dir_taukap) HYB_CUDA(d2h(c, dir_taukap, d.dir.tk, sizeof(double) * 2 * nrhs));
    HYB_CUDA(cudaStreamSynchronize(c->st));
    ksync(c);
    return 0;
}

int hyb_step_point_update(hyb_ctx* c, const double* coef, double* zs) {
    if (!c || !c->loaded || !c->dirs) return -1;
    HybDirs& d = *c->dirs;
    HYB_CUDA(cudaSetDevice(c->device));
    cudaStream_t st = c->st;
    const int64_t N = d.N, q = d.q, nq = N > q ? N : q;
    KBEGIN(c);
    point_update_kernel<<<nb256(nq), 256, 0, st>>>(N, q, d.ldn, d.ldq, coef[0], coef[1], coef[2], coef[3], d.kx, d.kz, d.ks, d.px,
                                                   d.pz, d.ps);
    dot_zs_kernel<<<1, RT, 0, st>>>(q, d.pz, d.ps, d.sout);
    HYB_CUDA(cudaGetLastError());
    g_misc_launches += 2;
    KEND(c);
    double v = 0.0;
    HYB_CUDA(d2h(c, &v, d.sout, sizeof(double)));
    HYB_CUDA(cudaStreamSynchronize(st));
    ksync(c);
    if (zs) *zs = v;
    return 0;
}

}  // extern "C"
