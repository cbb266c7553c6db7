// Device-resident direction solves: get_directions with iterative refinement
// (systemsolvers/common.jl:22-83), apply_lhs (:86-125), solve_system (:133-158),
// solve_subsystem4 (:161-189) and the QRChol setup_rhs3 / solve_subsystem3 (qrchol.jl:23-92,
// reduced model: p == 0) for up to HYB_MAX_RHS - 1 right-hand sides at once.
//
// Why batched: the four directions of a CombinedStepper iteration (combined.jl:60-125) come in two
// independent pairs -- (cent, pred), then (centadj, predadj), each adjustment needing only its own
// unadjusted direction -- and the constant solve of update_lhs (qrchol.jl:199-206) is independent
// of the first pair.  Every solve is bound by one pass over G for G'z, one for Gx (8 q N bytes
// each) and the latency of the two wavefront triangular solves; solving a pair together reads G and
// the Cholesky factor once for both and runs the two substitutions side by side.
//
// Vector layout at the boundary (point.jl:31-62 with p == 0): [x(N); z(q); tau; s(q); kap].
#include <cstdlib>

#include "dirs.cuh"

namespace {

constexpr int RED_THREADS = 512;

// deterministic block reduction (fixed tree)
__device__ __forceinline__ double block_sum(double v, double* sh) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) sh[w] = v;
    __syncthreads();
    double t = (threadIdx.x < (blockDim.x >> 5)) ? sh[threadIdx.x] : 0.0;
    if (w == 0) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    }
    return t;  // valid in warp 0
}
__device__ __forceinline__ double block_max(double v, double* sh) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) sh[w] = v;
    __syncthreads();
    double t = (threadIdx.x < (blockDim.x >> 5)) ? sh[threadIdx.x] : 0.0;
    if (w == 0) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t = fmax(t, __shfl_xor_sync(0xffffffffu, t, o));
    }
    return t;
}

// setup_rhs3 part 1 (qrchol.jl:23-44): the argument of the block Hessian product.
// Column `const_col` (if >= 0) is the constant system: argument h (qrchol.jl:199-203).
__global__ void rhs3_pre_kernel(int64_t q, int64_t ldq, int m, int const_col, const double* __restrict__ Rz,
                                const double* __restrict__ Rs, const double* __restrict__ h,
                                const unsigned char* __restrict__ dual_row, double* __restrict__ A) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= q || j >= m) return;
    const int64_t o = i + (int64_t)j * ldq;
    if (j == const_col)
        A[o] = h[i];
    else
        A[o] = dual_row[i] ? (-Rz[o] - Rs[o]) : Rz[o];
}
// part 2: rhs_sub.z from the product B = Htilde * A
__global__ void rhs3_post_kernel(int64_t q, int64_t ldq, int m, int const_col, const double* __restrict__ B,
                                 const double* __restrict__ Rs, const unsigned char* __restrict__ dual_row,
                                 double* __restrict__ Vz) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= q || j >= m) return;
    const int64_t o = i + (int64_t)j * ldq;
    if (j == const_col)
        Vz[o] = B[o];
    else
        Vz[o] = dual_row[i] ? B[o] : (-B[o] - Rs[o]);
}
// r = rhs_sub.x + G' rhs_sub.z (qrchol.jl:59-60); rhs_sub.x = rhs.x, or -c for the constant system
__global__ void subx_kernel(int64_t N, int64_t ldn, int m, int const_col, const double* __restrict__ Rx,
                            const double* __restrict__ c, const double* __restrict__ T, double* __restrict__ Vx) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= N || j >= m) return;
    const int64_t o = i + (int64_t)j * ldn;
    Vx[o] = ((j == const_col) ? -c[i] : Rx[o]) + T[o];
}
// sol_sub.z = Htilde (G x) - rhs_sub.z (qrchol.jl:84-91)
__global__ void subz_kernel(int64_t q, int64_t ldq, int m, const double* __restrict__ B, double* __restrict__ Vz) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= q || j >= m) return;
    const int64_t o = i + (int64_t)j * ldq;
    Vz[o] = B[o] - Vz[o];
}
// dot_obj (common.jl:222-224) of column j: c'x + h'z ; one block per column
__global__ void __launch_bounds__(RED_THREADS)
dot_obj_kernel(int64_t N, int64_t q, int64_t ldn, int64_t ldq, const double* __restrict__ c, const double* __restrict__ h,
               const double* __restrict__ X, const double* __restrict__ Z, double* __restrict__ out) {
    __shared__ double sh[32];
    const int j = blockIdx.x;
    double a = 0.0, b = 0.0;
    for (int64_t i = threadIdx.x; i < N; i += RED_THREADS) a += c[i] * X[i + (int64_t)j * ldn];
    for (int64_t i = threadIdx.x; i < q; i += RED_THREADS) b += h[i] * Z[i + (int64_t)j * ldq];
    const double sa = block_sum(a, sh);
    const double sb = block_sum(b, sh);
    if (threadIdx.x == 0) out[j] = sa + sb;
}
// solve_subsystem4's tau elimination and the lift of solve_system (common.jl:133-189):
//   tau = (rhs.tau + rhs.kap + dot_obj(sol_sub)) / (mu / taubar^2 - dot_obj(sol_const))
//   x = sol_sub.x + tau xc ; z = sol_sub.z + tau zc ; kap = -mu/taubar^2 tau + rhs.kap
// (s follows in lift_s_kernel from a fresh product G x, as common.jl:151-153 computes it: forming it
// from G sol_sub.x + tau G xc instead loses digits when the two terms cancel near convergence)
__global__ void lift_kernel(int64_t N, int64_t q, int64_t ldn, int64_t ldq, int m, double mu_tt,
                            const double* __restrict__ scal, const double* __restrict__ Rtk,
                            const double* __restrict__ Vx, const double* __restrict__ Vz,
                            const double* __restrict__ xc, const double* __restrict__ zc,
                            double* __restrict__ Dx, double* __restrict__ Dz, double* __restrict__ Dtk) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y;
    if (j >= m) return;
    const double rtau = Rtk[2 * j], rkap = Rtk[2 * j + 1];
    const double tau = (rtau + rkap + scal[1 + j]) / (mu_tt - scal[0]);
    if (i < N) Dx[i + (int64_t)j * ldn] = Vx[i + (int64_t)j * ldn] + tau * xc[i];
    if (i < q) {
        const int64_t o = i + (int64_t)j * ldq;
        Dz[o] = Vz[o] + tau * zc[i];
    }
    if (i == 0) {
        Dtk[2 * j] = tau;
        Dtk[2 * j + 1] = -mu_tt * tau + rkap;
    }
}
//   s = h tau - rhs.z - G x (common.jl:151-153)
__global__ void lift_s_kernel(int64_t q, int64_t ldq, int m, const double* __restrict__ h, const double* __restrict__ Rz,
                              const double* __restrict__ Dtk, const double* __restrict__ Gx, double* __restrict__ Ds) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= q || j >= m) return;
    const int64_t o = i + (int64_t)j * ldq;
    Ds[o] = h[i] * Dtk[2 * j] - Rz[o] - Gx[o];
}
// apply_lhs (common.jl:86-125) minus the right-hand side, elementwise parts:
//   res.x = G'z + tau c - rhs.x ; res.z = h tau - s - G x - rhs.z ; primal / dual direction split
__global__ void reslin_kernel(int64_t N, int64_t q, int64_t ldn, int64_t ldq, int m, const double* __restrict__ c,
                              const double* __restrict__ h, const unsigned char* __restrict__ dual_row,
                              const double* __restrict__ Dz, const double* __restrict__ Ds, const double* __restrict__ Dtk,
                              const double* __restrict__ T, const double* __restrict__ Gx, const double* __restrict__ Rx,
                              const double* __restrict__ Rz, double* __restrict__ Ex, double* __restrict__ Ez,
                              double* __restrict__ A) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y;
    if (j >= m) return;
    const double tau = Dtk[2 * j];
    if (i < N) {
        const int64_t o = i + (int64_t)j * ldn;
        Ex[o] = T[o] + tau * c[i] - Rx[o];
    }
    if (i < q) {
        const int64_t o = i + (int64_t)j * ldq;
        Ez[o] = h[i] * tau - Ds[o] - Gx[o] - Rz[o];
        A[o] = dual_row[i] ? Dz[o] : Ds[o];  // primal direction of the cone
    }
}
//   res.s = H_slow(prim_dir) + dual_dir - rhs.s
__global__ void ress_kernel(int64_t q, int64_t ldq, int m, const unsigned char* __restrict__ dual_row,
                            const double* __restrict__ B, const double* __restrict__ Dz, const double* __restrict__ Ds,
                            const double* __restrict__ Rs, double* __restrict__ Es) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= q || j >= m) return;
    const int64_t o = i + (int64_t)j * ldq;
    Es[o] = B[o] + (dual_row[i] ? Ds[o] : Dz[o]) - Rs[o];
}
//   res.tau = -c'x - h'z - kap - rhs.tau ; res.kap = mu/taubar^2 tau + kap - rhs.kap ; inf-norm of the column
__global__ void __launch_bounds__(RED_THREADS)
resnorm_kernel(int64_t N, int64_t q, int64_t ldn, int64_t ldq, double mu_tt, const double* __restrict__ c,
               const double* __restrict__ h, const double* __restrict__ Dx, const double* __restrict__ Dz,
               const double* __restrict__ Dtk, const double* __restrict__ Rtk, const double* __restrict__ Ex,
               const double* __restrict__ Ez, const double* __restrict__ Es, double* __restrict__ Etk,
               double* __restrict__ norms) {
    __shared__ double sh[32];
    const int j = blockIdx.x;
    double a = 0.0, b = 0.0, mx = 0.0;
    bool bad = false;
    for (int64_t i = threadIdx.x; i < N; i += RED_THREADS) {
        a += c[i] * Dx[i + (int64_t)j * ldn];
        const double e = fabs(Ex[i + (int64_t)j * ldn]);
        bad |= !(e == e);
        mx = fmax(mx, e);
    }
    for (int64_t i = threadIdx.x; i < q; i += RED_THREADS) {
        const int64_t o = i + (int64_t)j * ldq;
        b += h[i] * Dz[o];
        const double e1 = fabs(Ez[o]), e2 = fabs(Es[o]);
        bad |= !(e1 == e1) || !(e2 == e2);
        mx = fmax(mx, fmax(e1, e2));
    }
    const double sa = block_sum(a, sh);
    const double sb = block_sum(b, sh);
    double m2 = block_max(bad ? INFINITY : mx, sh);
    if (threadIdx.x == 0) {
        const double tau = Dtk[2 * j], kap = Dtk[2 * j + 1];
        const double et = -(sa + sb) - kap - Rtk[2 * j];
        const double ek = mu_tt * tau + kap - Rtk[2 * j + 1];
        Etk[2 * j] = et;
        Etk[2 * j + 1] = ek;
        m2 = fmax(m2, fmax(fabs(et), fabs(ek)));
        if (!(et == et) || !(ek == ek) || m2 == INFINITY) m2 = NAN;  // a NaN residual is reported as NaN
        norms[j] = m2;
    }
}
// dir = dir_temp - dir (common.jl:52-53), all pieces of one column
__global__ void refine_combine_kernel(int64_t N, int64_t q, const double* __restrict__ Tx, const double* __restrict__ Tz,
                                      const double* __restrict__ Ts, const double* __restrict__ Ttk, double* __restrict__ Dx,
                                      double* __restrict__ Dz, double* __restrict__ Ds, double* __restrict__ Dtk) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < N) Dx[i] = Tx[i] - Dx[i];
    if (i < q) {
        Dz[i] = Tz[i] - Dz[i];
        Ds[i] = Ts[i] - Ds[i];
    }
    if (i < 2) Dtk[i] = Ttk[i] - Dtk[i];
}

}  // namespace

// ----------------------------------------------------------------------------- host side
static void free_vecs(HybVecs& v) {
    cudaFree(v.x);
    cudaFree(v.z);
    cudaFree(v.s);
    cudaFree(v.tk);
    v = HybVecs();
}
static int alloc_vecs(HybVecs& v, const HybDirs& d, int m) {
    HYB_CUDA(cudaMalloc(&v.x, sizeof(double) * d.ldn * m));
    HYB_CUDA(cudaMalloc(&v.z, sizeof(double) * d.ldq * m));
    HYB_CUDA(cudaMalloc(&v.s, sizeof(double) * d.ldq * m));
    HYB_CUDA(cudaMalloc(&v.tk, sizeof(double) * 2 * m));
    HYB_CUDA(cudaMemset(v.x, 0, sizeof(double) * d.ldn * m));
    HYB_CUDA(cudaMemset(v.z, 0, sizeof(double) * d.ldq * m));
    HYB_CUDA(cudaMemset(v.s, 0, sizeof(double) * d.ldq * m));
    HYB_CUDA(cudaMemset(v.tk, 0, sizeof(double) * 2 * m));
    return 0;
}
static HybVecs column(const HybVecs& v, const HybDirs& d, int j) {
    HybVecs o;
    o.x = v.x + (int64_t)j * d.ldn;
    o.z = v.z + (int64_t)j * d.ldq;
    o.s = v.s + (int64_t)j * d.ldq;
    o.tk = v.tk + 2 * j;
    return o;
}

void hyb_dirs_free(hyb_ctx* c) {
    HybDirs* d = c->dirs;
    if (!d) return;
    cudaFree(d->c);
    cudaFree(d->h);
    cudaFree(d->dual_row);
    free_vecs(d->rhs);
    free_vecs(d->dir);
    free_vecs(d->res);
    free_vecs(d->tmp);
    free_vecs(d->rrhs);
    free_vecs(d->rdir);
    cudaFree(d->A);
    cudaFree(d->B);
    cudaFree(d->Vx);
    cudaFree(d->Vz);
    cudaFree(d->Gx);
    cudaFree(d->T);
    cudaFree(d->xc);
    cudaFree(d->zc);
    cudaFree(d->scal);
    cudaFree(d->gemm_scratch);
    cudaFree(d->pz);
    cudaFree(d->px);
    cudaFree(d->kx);
    cudaFree(d->resx);
    cudaFree(d->resz);
    cudaFree(d->sout);
    cudaFree(d->ps);
    cudaFree(d->kz);
    cudaFree(d->ks);
    cudaFree(d->szk);
    cudaFree(d->nus);
    cudaFree(d->cone_off);
    cudaFree(d->cone_dim);
    cudaFree(d->sred);
    cudaFree(d->sflag);
    cudaFree(d->sctl);
    delete d;
    c->dirs = nullptr;
}

static inline dim3 grid2(int64_t n, int m) { return dim3((unsigned)((n + 255) / 256), (unsigned)m); }

// solve_system (common.jl:133-158) for m columns of `rhs` into `out`; with_const appends the
// constant system of update_lhs (qrchol.jl:199-206) as column m and refreshes sol_const from it
static int solve_system_dev(hyb_ctx* c, HybDirs& d, int m, bool with_const, const HybVecs& rhs, const HybVecs& out,
                            double mu_tt) {
    cudaStream_t st = c->st;
    const int64_t N = d.N, q = d.q, ldn = d.ldn, ldq = d.ldq;
    const int mm = m + (with_const ? 1 : 0), cc = with_const ? m : -1;
    const int64_t nq = N > q ? N : q;
    if (q > 0) {
        rhs3_pre_kernel<<<grid2(q, mm), 256, 0, st>>>(q, ldq, mm, cc, rhs.z, rhs.s, d.h, d.dual_row, d.A);
        HYB_CUDA(cudaGetLastError());
        HYB_CHECK(hybi_cones_prod(st, c->cones, 4, d.A, ldq, d.B, ldq, mm));
        rhs3_post_kernel<<<grid2(q, mm), 256, 0, st>>>(q, ldq, mm, cc, d.B, rhs.s, d.dual_row, d.Vz);
        HYB_CUDA(cudaGetLastError());
        g_misc_launches += 2;
    }
    if (N > 0) {
        HYB_CHECK(hyb_gemm_t_small(st, q, N, c->GQ2, c->ldq, d.Vz, ldq, mm, d.T, ldn));
        subx_kernel<<<grid2(N, mm), 256, 0, st>>>(N, ldn, mm, cc, rhs.x, d.c, d.T, d.Vx);
        HYB_CUDA(cudaGetLastError());
        HYB_CHECK(hyb_potrs_multi(st, c->chol, d.Vx, ldn, mm));
        HYB_CHECK(hyb_gemm_n_small(st, q, N, c->GQ2, c->ldq, d.Vx, ldn, mm, d.Gx, ldq, d.gemm_scratch));
        g_misc_launches += 4;
    } else if (q > 0) {
        HYB_CUDA(cudaMemsetAsync(d.Gx, 0, sizeof(double) * ldq * mm, st));
    }
    if (q > 0) {
        HYB_CHECK(hybi_cones_prod(st, c->cones, 4, d.Gx, ldq, d.B, ldq, mm));
        subz_kernel<<<grid2(q, mm), 256, 0, st>>>(q, ldq, mm, d.B, d.Vz);
        HYB_CUDA(cudaGetLastError());
        ++g_misc_launches;
    }
    if (with_const) {
        if (N > 0) HYB_CUDA(cudaMemcpyAsync(d.xc, d.Vx + (int64_t)m * ldn, sizeof(double) * N, cudaMemcpyDeviceToDevice, st));
        if (q > 0) {
            HYB_CUDA(cudaMemcpyAsync(d.zc, d.Vz + (int64_t)m * ldq, sizeof(double) * q, cudaMemcpyDeviceToDevice, st));
        }
        dot_obj_kernel<<<1, RED_THREADS, 0, st>>>(N, q, ldn, ldq, d.c, d.h, d.xc, d.zc, d.scal);
        HYB_CUDA(cudaGetLastError());
        ++g_misc_launches;
        d.have_const = true;
    }
    if (m > 0) {
        if (!d.have_const) {
            hyb_set_error("hyb_get_directions: no constant solution yet (call with refresh_const after each hyb_schur_factor)");
            return -3;
        }
        dot_obj_kernel<<<m, RED_THREADS, 0, st>>>(N, q, ldn, ldq, d.c, d.h, d.Vx, d.Vz, d.scal + 1);
        HYB_CUDA(cudaGetLastError());
        lift_kernel<<<grid2(nq > 0 ? nq : 1, m), 256, 0, st>>>(N, q, ldn, ldq, m, mu_tt, d.scal, rhs.tk, d.Vx, d.Vz, d.xc, d.zc,
                                                             out.x, out.z, out.tk);
        HYB_CUDA(cudaGetLastError());
        g_misc_launches += 2;
        if (q > 0) {
            // d.Gx <- G x for the lifted x; residual_dev reuses it (gx_ready) for the same direction
            if (N > 0) {
                HYB_CHECK(hyb_gemm_n_small(st, q, N, c->GQ2, c->ldq, out.x, ldn, m, d.Gx, ldq, d.gemm_scratch));
                g_misc_launches += 2;
            }
            lift_s_kernel<<<grid2(q, m), 256, 0, st>>>(q, ldq, m, d.h, rhs.z, out.tk, d.Gx, out.s);
            HYB_CUDA(cudaGetLastError());
            ++g_misc_launches;
        }
    }
    return 0;
}

// res = LHS * dir - rhs (apply_lhs, common.jl:86-125, then common.jl:42) and the inf-norms into d.scal[8..]
// gx_ready: d.Gx already holds G * dir.x (left there by solve_system_dev for this very direction)
static int residual_dev(hyb_ctx* c, HybDirs& d, int m, const HybVecs& dir, const HybVecs& rhs, const HybVecs& res,
                        double mu_tt, double* norms_dev, bool gx_ready) {
    cudaStream_t st = c->st;
    const int64_t N = d.N, q = d.q, ldn = d.ldn, ldq = d.ldq;
    const int64_t nq = N > q ? N : q;
    if (N > 0) {
        HYB_CHECK(hyb_gemm_t_small(st, q, N, c->GQ2, c->ldq, dir.z, ldq, m, d.T, ldn));
        if (!gx_ready) HYB_CHECK(hyb_gemm_n_small(st, q, N, c->GQ2, c->ldq, dir.x, ldn, m, d.Gx, ldq, d.gemm_scratch));
        g_misc_launches += gx_ready ? 1 : 3;
    } else if (q > 0) {
        HYB_CUDA(cudaMemsetAsync(d.Gx, 0, sizeof(double) * ldq * m, st));
    }
    reslin_kernel<<<grid2(nq > 0 ? nq : 1, m), 256, 0, st>>>(N, q, ldn, ldq, m, d.c, d.h, d.dual_row, dir.z, dir.s, dir.tk, d.T,
                                                           d.Gx, rhs.x, rhs.z, res.x, res.z, d.A);
    HYB_CUDA(cudaGetLastError());
    if (q > 0) {
        HYB_CHECK(hybi_cones_prod(st, c->cones, 6, d.A, ldq, d.B, ldq, m));
        ress_kernel<<<grid2(q, m), 256, 0, st>>>(q, ldq, m, d.dual_row, d.B, dir.z, dir.s, rhs.s, res.s);
        HYB_CUDA(cudaGetLastError());
    }
    resnorm_kernel<<<m, RED_THREADS, 0, st>>>(N, q, ldn, ldq, mu_tt, d.c, d.h, dir.x, dir.z, dir.tk, rhs.tk, res.x, res.z,
                                              res.s, res.tk, norms_dev);
    HYB_CUDA(cudaGetLastError());
    g_misc_launches += 3;
    return 0;
}

static int copy_vecs(hyb_ctx* c, const HybDirs& d, const HybVecs& dst, const HybVecs& src) {
    cudaStream_t st = c->st;
    if (d.N > 0) HYB_CUDA(cudaMemcpyAsync(dst.x, src.x, sizeof(double) * d.N, cudaMemcpyDeviceToDevice, st));
    if (d.q > 0) {
        HYB_CUDA(cudaMemcpyAsync(dst.z, src.z, sizeof(double) * d.q, cudaMemcpyDeviceToDevice, st));
        HYB_CUDA(cudaMemcpyAsync(dst.s, src.s, sizeof(double) * d.q, cudaMemcpyDeviceToDevice, st));
    }
    HYB_CUDA(cudaMemcpyAsync(dst.tk, src.tk, sizeof(double) * 2, cudaMemcpyDeviceToDevice, st));
    return 0;
}

// get_directions (common.jl:22-83) on the m right-hand sides in d.rhs; the directions end up in d.dir
int hyb_directions_core(hyb_ctx* c, int m, bool refresh_const, double mu_tt, double res_norm_cutoff, int max_ref_steps,
                        double min_impr_tol, int* num_ref, double* res_norm) {
    HybDirs& d = *c->dirs;
    cudaStream_t st = c->st;
    const int64_t N = d.N, q = d.q;
    HYB_CHECK(solve_system_dev(c, d, m, refresh_const, d.rhs, d.dir, mu_tt));
    static const bool dbg = getenv("HYB_DEBUG_DIRS") != nullptr;
    double norms[HYB_MAX_RHS] = {0, 0, 0, 0};
    for (int j = 0; j < m; ++j) {
        if (num_ref) num_ref[j] = 0;
        if (res_norm) res_norm[j] = 0.0;
    }
    if (max_ref_steps > 0 && m > 0) {
        HYB_CHECK(residual_dev(c, d, m, d.dir, d.rhs, d.res, mu_tt, d.scal + 8, true));
        HYB_CUDA(cudaMemcpyAsync(norms, d.scal + 8, sizeof(double) * m, cudaMemcpyDeviceToHost, st));
        HYB_CUDA(cudaStreamSynchronize(st));
        for (int j = 0; j < m; ++j) {
            double rn = norms[j];
            int nref = 0;
            if (dbg) fprintf(stderr, "  dirs: col %d res_norm %.3e cutoff %.3e\n", j, rn, res_norm_cutoff);
            if (rn > res_norm_cutoff) {  // iterative refinement of column j (common.jl:46-79)
                const HybVecs dj = column(d.dir, d, j), rj = column(d.rhs, d, j), ej = column(d.res, d, j);
                HYB_CHECK(copy_vecs(c, d, d.tmp, dj));
                bool prev_slow = false;
                double prev = rn;
                for (int it = 0; it < max_ref_steps; ++it) {
                    ++nref;
                    HYB_CHECK(copy_vecs(c, d, d.rrhs, ej));
                    HYB_CHECK(solve_system_dev(c, d, 1, false, d.rrhs, d.rdir, mu_tt));
                    const int64_t nq = N > q ? N : q;
                    refine_combine_kernel<<<(unsigned)(((nq > 2 ? nq : 2) + 255) / 256), 256, 0, st>>>(
                        N, q, d.tmp.x, d.tmp.z, d.tmp.s, d.tmp.tk, d.rdir.x, d.rdir.z, d.rdir.s, d.rdir.tk);
                    HYB_CUDA(cudaGetLastError());
                    ++g_misc_launches;
                    HYB_CHECK(residual_dev(c, d, 1, d.rdir, rj, ej, mu_tt, d.scal + 12, false));
                    double rn_new = 0.0;
                    HYB_CUDA(cudaMemcpyAsync(&rn_new, d.scal + 12, sizeof(double), cudaMemcpyDeviceToHost, st));
                    HYB_CUDA(cudaStreamSynchronize(st));
                    if (dbg) fprintf(stderr, "    refine: %.3e -> %.3e\n", rn, rn_new);
                    if (rn_new != rn_new) {  // NaN: reported, the caller asserts like common.jl:81
                        rn = rn_new;
                        break;
                    }
                    if (rn_new >= rn) break;  // worse: keep the previous direction (still in column j)
                    HYB_CHECK(copy_vecs(c, d, dj, d.rdir));
                    HYB_CHECK(copy_vecs(c, d, d.tmp, d.rdir));
                    rn = rn_new;
                    if (rn < res_norm_cutoff) break;
                    const bool curr_slow = rn > min_impr_tol * prev;
                    if (prev_slow && curr_slow) break;
                    prev = rn;
                    prev_slow = curr_slow;
                }
            }
            if (num_ref) num_ref[j] = nref;
            if (res_norm) res_norm[j] = rn;
        }
    }
    return 0;
}

extern "C" {

int hyb_dirs_setup(hyb_ctx* c, const double* cvec, const double* hvec) {
    if (!c || !c->loaded) return -1;
    HYB_CUDA(cudaSetDevice(c->device));
    hyb_dirs_free(c);
    HybDirs* d = new HybDirs();
    c->dirs = d;
    const int64_t N = d->N = c->N, q = d->q = c->q;
    d->ldn = hyb_round_up(N > 0 ? N : 1, 16);
    d->ldq = c->ldq;
    const int M = HYB_MAX_RHS;
    HYB_CUDA(cudaMalloc(&d->c, sizeof(double) * d->ldn));
    HYB_CUDA(cudaMalloc(&d->h, sizeof(double) * d->ldq));
    HYB_CUDA(cudaMalloc(&d->dual_row, (size_t)d->ldq));
    if (N > 0) HYB_CUDA(cudaMemcpy(d->c, cvec, sizeof(double) * N, cudaMemcpyHostToDevice));
    if (q > 0) HYB_CUDA(cudaMemcpy(d->h, hvec, sizeof(double) * q, cudaMemcpyHostToDevice));
    std::vector<unsigned char> mask((size_t)d->ldq, 0);
    for (const HybConeBlock& b : c->cones.blocks) {
        if (!b.use_dual) continue;
        const int64_t rows = (int64_t)b.dim * b.count;
        for (int64_t i = 0; i < rows; ++i) mask[(size_t)(b.offset + i)] = 1;
    }
    HYB_CUDA(cudaMemcpy(d->dual_row, mask.data(), mask.size(), cudaMemcpyHostToDevice));
    HYB_CHECK(alloc_vecs(d->rhs, *d, M));
    HYB_CHECK(alloc_vecs(d->dir, *d, M));
    HYB_CHECK(alloc_vecs(d->res, *d, M));
    HYB_CHECK(alloc_vecs(d->tmp, *d, 1));
    HYB_CHECK(alloc_vecs(d->rrhs, *d, 1));
    HYB_CHECK(alloc_vecs(d->rdir, *d, 1));
    HYB_CUDA(cudaMalloc(&d->A, sizeof(double) * d->ldq * M));
    HYB_CUDA(cudaMalloc(&d->B, sizeof(double) * d->ldq * M));
    HYB_CUDA(cudaMalloc(&d->Vz, sizeof(double) * d->ldq * M));
    HYB_CUDA(cudaMalloc(&d->Gx, sizeof(double) * d->ldq * M));
    HYB_CUDA(cudaMalloc(&d->Vx, sizeof(double) * d->ldn * M));
    HYB_CUDA(cudaMalloc(&d->T, sizeof(double) * d->ldn * M));
    HYB_CUDA(cudaMalloc(&d->xc, sizeof(double) * d->ldn));
    HYB_CUDA(cudaMalloc(&d->zc, sizeof(double) * d->ldq));
    HYB_CUDA(cudaMalloc(&d->scal, sizeof(double) * 16));
    HYB_CUDA(cudaMalloc(&d->gemm_scratch, sizeof(double) * (q > 0 ? q : 1) * HYB_GEMV_SPLITS * M));
    HYB_CUDA(cudaMemset(d->Vx, 0, sizeof(double) * d->ldn * M));
    HYB_CUDA(cudaMemset(d->T, 0, sizeof(double) * d->ldn * M));
    HYB_CUDA(cudaMemset(d->scal, 0, sizeof(double) * 16));
    // line-search state
    const int64_t K = d->K = c->cones.K;
    const int64_t Ka = K > 0 ? K : 1;
    HYB_CUDA(cudaMalloc(&d->pz, sizeof(double) * d->ldq));
    HYB_CUDA(cudaMalloc(&d->ps, sizeof(double) * d->ldq));
    HYB_CUDA(cudaMalloc(&d->px, sizeof(double) * d->ldn));
    HYB_CUDA(cudaMalloc(&d->kx, sizeof(double) * d->ldn * 4));
    HYB_CUDA(cudaMalloc(&d->resx, sizeof(double) * d->ldn));
    HYB_CUDA(cudaMalloc(&d->resz, sizeof(double) * d->ldq));
    HYB_CUDA(cudaMalloc(&d->sout, sizeof(double) * 8));
    HYB_CUDA(cudaMemset(d->px, 0, sizeof(double) * d->ldn));
    HYB_CUDA(cudaMemset(d->kx, 0, sizeof(double) * d->ldn * 4));
    HYB_CUDA(cudaMalloc(&d->kz, sizeof(double) * d->ldq * 4));
    HYB_CUDA(cudaMalloc(&d->ks, sizeof(double) * d->ldq * 4));
    HYB_CUDA(cudaMemset(d->kz, 0, sizeof(double) * d->ldq * 4));
    HYB_CUDA(cudaMemset(d->ks, 0, sizeof(double) * d->ldq * 4));
    HYB_CUDA(cudaMalloc(&d->szk, sizeof(double) * Ka));
    HYB_CUDA(cudaMalloc(&d->nus, sizeof(double) * Ka));
    HYB_CUDA(cudaMalloc(&d->cone_off, sizeof(int64_t) * Ka));
    HYB_CUDA(cudaMalloc(&d->cone_dim, sizeof(int) * Ka));
    HYB_CUDA(cudaMalloc(&d->sred, sizeof(double) * 2));
    HYB_CUDA(cudaMalloc(&d->sflag, sizeof(int)));
    HYB_CUDA(cudaMalloc(&d->sctl, sizeof(double) * 8));
    std::vector<int64_t> off((size_t)Ka, 0);
    std::vector<int> dm((size_t)Ka, 0);
    for (const HybConeBlock& b : c->cones.blocks)
        for (int64_t i = 0; i < b.count; ++i) {
            off[(size_t)(b.cone0 + i)] = b.offset + i * b.dim;
            dm[(size_t)(b.cone0 + i)] = b.dim;
        }
    HYB_CUDA(cudaMemcpy(d->cone_off, off.data(), sizeof(int64_t) * Ka, cudaMemcpyHostToDevice));
    HYB_CUDA(cudaMemcpy(d->cone_dim, dm.data(), sizeof(int) * Ka, cudaMemcpyHostToDevice));
    d->search_ready = false;
    return 0;
}

int hyb_get_directions(hyb_ctx* c, int nrhs, const double* rhs, double* dir, int64_t ldv, int refresh_const,
                       double tau_bar, double mu, double res_norm_cutoff, int max_ref_steps, double min_impr_tol,
                       int* num_ref, double* res_norm, double* sol_const) {
    if (!c || !c->loaded) return -1;
    HybDirs* dp = c->dirs;
    if (!dp) {
        hyb_set_error("hyb_get_directions: call hyb_dirs_setup first");
        return -3;
    }
    HybDirs& d = *dp;
    const int m = nrhs;
    if (m < 0 || m + (refresh_const ? 1 : 0) > HYB_MAX_RHS) {
        hyb_set_error("hyb_get_directions: %d right-hand sides (+%d) exceed the batch limit %d", m, refresh_const ? 1 : 0,
                      HYB_MAX_RHS);
        return -3;
    }
    HYB_CUDA(cudaSetDevice(c->device));
    if (c->updated_irtmu != c->updated_irtmu) HYB_CHECK(hyb_ensure_updated(c, c->loaded_scal));
    cudaStream_t st = c->st;
    const int64_t N = d.N, q = d.q, L = N + 2 * q + 2;
    if (m > 0 && ldv < L) {
        hyb_set_error("hyb_get_directions: ldv %lld < N + 2q + 2 = %lld", (long long)ldv, (long long)L);
        return -3;
    }
    const size_t hp = sizeof(double) * (size_t)ldv;
    auto up = [&](double* dst, int64_t ldd, int64_t off, int64_t len) -> cudaError_t {
        if (len == 0 || m == 0) return cudaSuccess;
        c->h2d_bytes += (int64_t)sizeof(double) * len * m;
        return cudaMemcpy2DAsync(dst, sizeof(double) * ldd, rhs + off, hp, sizeof(double) * len, m, cudaMemcpyHostToDevice, st);
    };
    auto down = [&](double* src, int64_t lds_, int64_t off, int64_t len) -> cudaError_t {
        if (len == 0 || m == 0) return cudaSuccess;
        c->d2h_bytes += (int64_t)sizeof(double) * len * m;
        return cudaMemcpy2DAsync(dir + off, hp, src, sizeof(double) * lds_, sizeof(double) * len, m, cudaMemcpyDeviceToHost, st);
    };
    HYB_CUDA(up(d.rhs.x, d.ldn, 0, N));
    HYB_CUDA(up(d.rhs.z, d.ldq, N, q));
    HYB_CUDA(up(d.rhs.tk, 2, N + q, 1));
    HYB_CUDA(up(d.rhs.s, d.ldq, N + q + 1, q));
    HYB_CUDA(up(d.rhs.tk + 1, 2, N + 2 * q + 1, 1));
    const double mu_tt = mu / tau_bar / tau_bar;
    KBEGIN(c);
    HYB_CHECK(hyb_directions_core(c, m, refresh_const != 0, mu_tt, res_norm_cutoff, max_ref_steps, min_impr_tol, num_ref, res_norm));
    KEND(c);
    HYB_CUDA(down(d.dir.x, d.ldn, 0, N));
    HYB_CUDA(down(d.dir.z, d.ldq, N, q));
    HYB_CUDA(down(d.dir.tk, 2, N + q, 1));
    HYB_CUDA(down(d.dir.s, d.ldq, N + q + 1, q));
    HYB_CUDA(down(d.dir.tk + 1, 2, N + 2 * q + 1, 1));
    if (sol_const && d.have_const) {
        if (N > 0) HYB_CUDA(d2h(c, sol_const, d.xc, sizeof(double) * N));
        if (q > 0) HYB_CUDA(d2h(c, sol_const + N, d.zc, sizeof(double) * q));
    }
    HYB_CUDA(cudaStreamSynchronize(st));
    ksync(c);
    return 0;
}

}  // extern "C"

// =============================================================================================
// Device-resident line search: check_cone_points (search.jl:100-162) on a candidate
//   cand.ztsk = point.ztsk + sum_k coef_k dir_k.ztsk            (update_stepper_points, combined.jl:129-177)
// formed on the device from the point and the four directions kept after hyb_get_directions, so a
// candidate costs a few small kernels and two or three scalars over PCIe instead of two q-vectors
// and O(q) host arithmetic.  tau / kap (scalars) stay with the host.
// =============================================================================================
namespace {

// candidate primal / dual slices of every cone (point.jl:50-59: dual-barrier cones take z as primal)
__global__ void cand_kernel(int64_t q, int64_t ldq, double c0, double c1, double c2, double c3,
                            const double* __restrict__ pz, const double* __restrict__ ps, const double* __restrict__ kz,
                            const double* __restrict__ ks, const unsigned char* __restrict__ dual_row,
                            double* __restrict__ primal, double* __restrict__ dual, const double* __restrict__ ctl) {
    if (ctl && ctl[2] >= 0.0) return;  // speculative search: an earlier candidate was accepted
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= q) return;
    double z = pz[i], s = ps[i];
    // same order of additions as the host's `cand += a*d1 + b*d2 + ...` (left to right)
    double dz = 0.0, ds = 0.0;
    bool first = true;
    const double cf[4] = {c0, c1, c2, c3};
    // order used by update_stepper_points: pred, predadj, cent, centadj  (slots 2, 3, 0, 1)
    const int order[4] = {2, 3, 0, 1};
#pragma unroll
    for (int t = 0; t < 4; ++t) {
        const int k = order[t];
        if (cf[k] != 0.0) {
            const double az = cf[k] * kz[i + (int64_t)k * ldq], as = cf[k] * ks[i + (int64_t)k * ldq];
            dz = first ? az : dz + az;
            ds = first ? as : ds + as;
            first = false;
        }
    }
    z += dz;
    s += ds;
    const bool dr = dual_row[i] != 0;
    primal[i] = dr ? z : s;
    dual[i] = dr ? s : z;
}
// <primal_k, dual_k> per cone, one warp per cone
__global__ void __launch_bounds__(256)
cone_dots_kernel(int64_t K, const int64_t* __restrict__ off, const int* __restrict__ dim, const double* __restrict__ a,
                 const double* __restrict__ b, double* __restrict__ out, const double* __restrict__ ctl) {
    if (ctl && ctl[2] >= 0.0) return;
    const int lane = threadIdx.x & 31;
    const int64_t k = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (k >= K) return;
    const int64_t o = off[k];
    const int n = dim[k];
    double s = 0.0;
    for (int r = lane; r < n; r += 32) s += a[o + r] * b[o + r];
#pragma unroll
    for (int w = 16; w > 0; w >>= 1) s += __shfl_xor_sync(0xffffffffu, s, w);
    if (lane == 0) out[k] = s;
}
// min and sum of szk (one CTA of 1024 threads, fixed order); valid in thread 0
__device__ __forceinline__ void szk_reduce_block(int64_t K, const double* __restrict__ szk, double& mn_out, double& sm_out) {
    __shared__ double smin[32], ssum[32];
    double mn = INFINITY, sm = 0.0;
    bool nan = false;
    for (int64_t k = threadIdx.x; k < K; k += 1024) {
        const double v = szk[k];
        nan |= (v != v);
        mn = fmin(mn, v);
        sm += v;
    }
    if (nan) mn = NAN;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double om = __shfl_xor_sync(0xffffffffu, mn, o);
        mn = (mn != mn || om != om) ? NAN : fmin(mn, om);
        sm += __shfl_xor_sync(0xffffffffu, sm, o);
    }
    if (lane == 0) {
        smin[w] = mn;
        ssum[w] = sm;
    }
    __syncthreads();
    if (w == 0) {
        mn = smin[lane];
        sm = ssum[lane];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double om = __shfl_xor_sync(0xffffffffu, mn, o);
            mn = (mn != mn || om != om) ? NAN : fmin(mn, om);
            sm += __shfl_xor_sync(0xffffffffu, sm, o);
        }
    }
    mn_out = mn;
    sm_out = sm;
}
__global__ void __launch_bounds__(1024)
szk_reduce_kernel(int64_t K, const double* __restrict__ szk, double* __restrict__ out) {
    double mn, sm;
    szk_reduce_block(K, szk, mn, sm);
    if (threadIdx.x == 0) {
        out[0] = mn;
        out[1] = sm;
    }
}
// Speculative line search, first decision of candidate `idx` on the device: the scalar gates of search.jl:108-130
// that the sequential path takes on the host between hyb_search_gates and hyb_search_check, with the same
// arithmetic.  ctl[0] <- 1 to skip the oracle update of this candidate (rejected, or an earlier one accepted).
__global__ void __launch_bounds__(1024)
search_decide1_kernel(int64_t K, const double* __restrict__ szk, const double* __restrict__ nus, double taukap, double nup1,
                      double min_prox, double bound, double* __restrict__ ctl) {
    __shared__ int any;
    __shared__ double mu_s;
    if (ctl[2] >= 0.0) {  // uniform: written only by one thread of an earlier kernel
        if (threadIdx.x == 0) ctl[0] = 1.0;
        return;
    }
    double mn, sm;
    szk_reduce_block(K, szk, mn, sm);
    if (threadIdx.x == 0) {
        any = 0;
        const double eps = 2.220446049250313e-16;
        const double mu = (sm + taukap) / nup1;
        mu_s = mu;
        int reject = 0;
        if (mn < eps) reject = 1;
        if (mu < eps) reject = 1;
        const double rel = taukap / mu;
        if (rel < min_prox) reject = 1;
        const double tp = (rel - 1.0) * (rel - 1.0);
        if (tp > bound) reject = 1;
        any = reject;
        ctl[1] = 1.0 / sqrt(mu);
        ctl[3] = mu;
        ctl[5] = tp;
        ctl[6] += 1.0;
    }
    __syncthreads();
    const double mu = mu_s;
    int bad = 0;
    for (int64_t k = threadIdx.x; k < K; k += 1024) {
        const double nu = nus[k];
        const double rel = szk[k] / (mu * nu);
        const double dv = rel - 1.0;
        if (rel < min_prox || nu * (dv * dv) > bound) bad = 1;
    }
    if (bad) atomicOr(&any, 1);
    __syncthreads();
    if (threadIdx.x == 0) ctl[0] = any ? 1.0 : 0.0;
}
// ... and the last one (search.jl:141-158 after the oracle checks): accept candidate `idx` or go on
__global__ void search_decide2_kernel(int idx, const int* __restrict__ red_flag, const double* __restrict__ red_prox,
                                      int use_max, double bound, double* __restrict__ ctl) {
    if (ctl[0] != 0.0) return;
    const double a = *red_prox, agg0 = ctl[5];
    const double total = use_max ? fmax(agg0, a) : agg0 + a;
    if (*red_flag == HYB_ST_ALL && total < bound) {
        ctl[2] = (double)idx;
        ctl[4] = total;
        ctl[0] = 1.0;
    }
}
// per-cone gates of search.jl:124-130: sz_rel < min_prox or nu (sz_rel - 1)^2 > bound
__global__ void __launch_bounds__(1024)
szk_gate_kernel(int64_t K, const double* __restrict__ szk, const double* __restrict__ nus, double mu, double min_prox,
                double bound, int* __restrict__ flag) {
    __shared__ int any;
    if (threadIdx.x == 0) any = 0;
    __syncthreads();
    int bad = 0;
    for (int64_t k = threadIdx.x; k < K; k += 1024) {
        const double nu = nus[k];
        const double rel = szk[k] / (mu * nu);
        const double dv = rel - 1.0;
        if (rel < min_prox || nu * (dv * dv) > bound) bad = 1;
    }
    if (bad) atomicOr(&any, 1);
    __syncthreads();
    if (threadIdx.x == 0) *flag = any;
}

}  // namespace

extern "C" {

int hyb_search_setup(hyb_ctx* c, const double* nus) {
    if (!c || !c->loaded || !c->dirs) {
        hyb_set_error("hyb_search_setup: call hyb_dirs_setup first");
        return -3;
    }
    HybDirs& d = *c->dirs;
    HYB_CUDA(cudaSetDevice(c->device));
    if (d.K > 0) HYB_CUDA(cudaMemcpy(d.nus, nus, sizeof(double) * d.K, cudaMemcpyHostToDevice));
    d.search_ready = true;
    return 0;
}

int hyb_dirs_keep(hyb_ctx* c, int col, int slot) {
    if (!c || !c->loaded || !c->dirs) return -1;
    HybDirs& d = *c->dirs;
    if (col < 0 || col >= HYB_MAX_RHS || slot < 0 || slot > 3) {
        hyb_set_error("hyb_dirs_keep: column %d / slot %d out of range", col, slot);
        return -3;
    }
    HYB_CUDA(cudaSetDevice(c->device));
    if (d.N > 0)
        HYB_CUDA(cudaMemcpyAsync(d.kx + (int64_t)slot * d.ldn, d.dir.x + (int64_t)col * d.ldn, sizeof(double) * d.N,
                                 cudaMemcpyDeviceToDevice, c->st));
    if (d.q > 0) {
        HYB_CUDA(cudaMemcpyAsync(d.kz + (int64_t)slot * d.ldq, d.dir.z + (int64_t)col * d.ldq, sizeof(double) * d.q,
                                 cudaMemcpyDeviceToDevice, c->st));
        HYB_CUDA(cudaMemcpyAsync(d.ks + (int64_t)slot * d.ldq, d.dir.s + (int64_t)col * d.ldq, sizeof(double) * d.q,
                                 cudaMemcpyDeviceToDevice, c->st));
    }
    return 0;
}

int hyb_search_load_point(hyb_ctx* c, const double* z, const double* s) {
    if (!c || !c->loaded || !c->dirs) return -1;
    HybDirs& d = *c->dirs;
    HYB_CUDA(cudaSetDevice(c->device));
    if (d.q > 0) {
        HYB_CUDA(h2d(c, d.pz, z, sizeof(double) * d.q));
        HYB_CUDA(h2d(c, d.ps, s, sizeof(double) * d.q));
    }
    return 0;
}

int hyb_search_gates(hyb_ctx* c, const double* coef, double* min_szk, double* sum_szk) {
    if (!c || !c->loaded || !c->dirs || !c->dirs->search_ready) {
        hyb_set_error("hyb_search_gates: call hyb_search_setup first");
        return -3;
    }
    HybDirs& d = *c->dirs;
    HYB_CUDA(cudaSetDevice(c->device));
    cudaStream_t st = c->st;
    const int64_t q = d.q, K = d.K;
    *min_szk = INFINITY;
    *sum_szk = 0.0;
    if (q == 0 || K == 0) return 0;
    KBEGIN(c);
    cand_kernel<<<(unsigned)((q + 255) / 256), 256, 0, st>>>(q, d.ldq, coef[0], coef[1], coef[2], coef[3], d.pz, d.ps, d.kz, d.ks,
                                                           d.dual_row, c->in_primal, c->in_dual, nullptr);
    cone_dots_kernel<<<(unsigned)((K + 7) / 8), 256, 0, st>>>(K, d.cone_off, d.cone_dim, c->in_primal, c->in_dual, d.szk, nullptr);
    szk_reduce_kernel<<<1, 1024, 0, st>>>(K, d.szk, d.sred);
    HYB_CUDA(cudaGetLastError());
    g_misc_launches += 3;
    KEND(c);
    // the loaded cone state no longer matches in_primal / in_dual until hyb_search_check runs
    c->point_loaded = false;
    c->updated_irtmu = NAN;
    double out[2];
    HYB_CUDA(d2h(c, out, d.sred, sizeof(double) * 2));
    HYB_CUDA(cudaStreamSynchronize(st));
    ksync(c);
    *min_szk = out[0];
    *sum_szk = out[1];
    return 0;
}

int hyb_search_alpha_supported(hyb_ctx* c) {
    if (!c || !c->loaded) return 0;
    return hybi_cones_all_batched(c->cones) ? 1 : 0;
}

int hyb_search_alpha(hyb_ctx* c, int ncand, const double* coefs, const double* taukap, double nup1, double min_prox,
                     double proxsqr_bound, int use_max_prox, int* accepted, double* mu, double* agg) {
    if (!c || !c->loaded || !c->dirs || !c->dirs->search_ready) {
        hyb_set_error("hyb_search_alpha: call hyb_search_setup first");
        return -3;
    }
    HybDirs& d = *c->dirs;
    HYB_CUDA(cudaSetDevice(c->device));
    cudaStream_t st = c->st;
    const int64_t q = d.q, K = d.K;
    *accepted = -1;
    *mu = *agg = NAN;
    if (ncand <= 0) return 0;
    if (q == 0 || K == 0 || !hybi_cones_all_batched(c->cones)) {
        hyb_set_error("hyb_search_alpha: needs a model whose cones are all served by the closed-form kernel families");
        return -3;
    }
    const double init[8] = {0.0, 0.0, -1.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    HYB_CUDA(h2d(c, d.sctl, init, sizeof(init)));
    KBEGIN(c);
    for (int i = 0; i < ncand; ++i) {
        const double* cf = coefs + 4 * i;
        cand_kernel<<<(unsigned)((q + 255) / 256), 256, 0, st>>>(q, d.ldq, cf[0], cf[1], cf[2], cf[3], d.pz, d.ps, d.kz, d.ks,
                                                               d.dual_row, c->in_primal, c->in_dual, d.sctl);
        cone_dots_kernel<<<(unsigned)((K + 7) / 8), 256, 0, st>>>(K, d.cone_off, d.cone_dim, c->in_primal, c->in_dual, d.szk,
                                                                d.sctl);
        search_decide1_kernel<<<1, 1024, 0, st>>>(K, d.szk, d.nus, taukap[i], nup1, min_prox, proxsqr_bound, d.sctl);
        HYB_CUDA(cudaGetLastError());
        g_misc_launches += 3;
        // load_point(cone, primal, irtmu), load_dual_point, reset_data, the oracle checks (search.jl:141-158)
        HYB_CHECK(hybi_cones_update(st, c->cones, c->in_primal, c->in_dual, 0.0, 0.0, 1, d.sctl));
        HYB_CHECK(hybi_cones_reduce_dev(st, c->cones, use_max_prox, d.sctl));
        search_decide2_kernel<<<1, 1, 0, st>>>(i, c->cones.red_flag, c->cones.red_prox, use_max_prox, proxsqr_bound, d.sctl);
        HYB_CUDA(cudaGetLastError());
        ++g_misc_launches;
    }
    KEND(c);
    double out[8];
    HYB_CUDA(d2h(c, out, d.sctl, sizeof(out)));
    HYB_CUDA(cudaStreamSynchronize(st));
    ksync(c);
    if (out[2] >= 0.0) {  // the device holds the accepted candidate, oracles updated at its scaling
        *accepted = (int)out[2];
        *mu = out[3];
        *agg = out[4];
        c->loaded_scal = out[1];
        c->point_loaded = true;
        c->updated_irtmu = out[1];
    } else {  // the loaded cone state is some rejected candidate's
        c->point_loaded = false;
        c->updated_irtmu = NAN;
    }
    return 0;
}

int hyb_search_check(hyb_ctx* c, double mu, double min_prox, double proxsqr_bound, int use_max_prox, double agg0, int* ok,
                     double* agg) {
    if (!c || !c->loaded || !c->dirs || !c->dirs->search_ready) return -1;
    HybDirs& d = *c->dirs;
    HYB_CUDA(cudaSetDevice(c->device));
    cudaStream_t st = c->st;
    *ok = 0;
    *agg = agg0;
    const double irtmu = 1.0 / sqrt(mu);
    if (d.K > 0) {
        KBEGIN(c);
        szk_gate_kernel<<<1, 1024, 0, st>>>(d.K, d.szk, d.nus, mu, min_prox, proxsqr_bound, d.sflag);
        HYB_CUDA(cudaGetLastError());
        ++g_misc_launches;
        KEND(c);
        int viol = 0;
        HYB_CUDA(d2h(c, &viol, d.sflag, sizeof(int)));
        HYB_CUDA(cudaStreamSynchronize(st));
        ksync(c);
        if (viol) return 0;
    }
    // load_point(cone, primal, irtmu), load_dual_point, reset_data, then the oracle checks (search.jl:141-158)
    c->loaded_scal = irtmu;
    c->point_loaded = true;
    c->updated_irtmu = NAN;
    KBEGIN(c);
    HYB_CHECK(hyb_ensure_updated(c, irtmu));
    KEND(c);
    int flags = 0;
    double a = 0.0;
    HYB_CHECK(hybi_cones_reduce(st, c->cones, use_max_prox, &flags, &a));
    ksync(c);
    const double total = use_max_prox ? fmax(agg0, a) : agg0 + a;
    *agg = total;
    *ok = (flags == HYB_ST_ALL) && (total < proxsqr_bound) ? 1 : 0;
    return 0;
}

}  // extern "C"
