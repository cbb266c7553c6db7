// Dense FP64 Cholesky (upper, column-major) and triangular solves for sm_100a.
//
// Replaces cholesky!(Symmetric(:U), check=false) -> dpotrf('U') of the reference
// (src/linearalgebra/dense.jl:226, called from qrchol.jl:257-258) and
// ldiv!(x, fact, rhs) -> dpotrs (qrchol.jl:75).
//
// Blocked right-looking factorization A = U'U with NB = 128:
//   panel  : one CTA factors the 128x128 diagonal block in shared memory (blocked by 16),
//   row    : U12 = U11^-T A12 by forward substitution, one thread per column,
//   update : A22 -= U12' * U12 (upper only) -> DMMA gemm_tn (SYRK form; K-major on both
//            sides because U is column-major and the contraction runs down columns).
// No explicit inverses anywhere: inverse-based TRSM/TRSV has backward error
// cond(U11)*eps, which late interior-point iterations do not tolerate.
//
// Solves: two persistent "wavefront" kernels (forward U'y = b, backward Ux = y),
// one CTA per 128-column block, cooperative launch so all CTAs are co-resident;
// a CTA consumes the x-blocks it depends on as their ready flags are published,
// solves its diagonal block by substitution in shared memory, and publishes its block.
// Fixed summation order -> bitwise reproducible.
#include <cooperative_groups.h>
#include <cstdlib>

#include "common.cuh"

namespace {

constexpr int NB = 128;
constexpr int LDS = NB + 1;  // padded shared-memory leading dimension
constexpr int IB = 16;       // inner block of the panel factorization
constexpr int PLD = NB + 4;  // row stride of the panel kernel's block (conflict-free DMMA fragment loads)

// --------------------------------------------------------------------------------
// panel: in-place upper Cholesky of an nb x nb diagonal block (nb <= 128) in shared
// memory, blocked by IB = 16 columns: (1) one warp factors the 16 x 16 diagonal block in
// registers (lane c & 15 owns column c; pivots and multipliers travel by shuffle), (2)
// thread-per-column forward substitution gives the 16 x m row panel, (3) all threads apply
// the rank-16 update to the trailing triangle in 4 x 4 register tiles.  Substitution only
// (no explicit inverses), so backward stable like LAPACK dpotf2/dpotrf.
//
// The serial chain of the whole panel is 128 x (pivot -> reciprocal square root -> scale ->
// rank-1 update of the next pivot).  IB = 16 keeps the per-pivot shuffle/FMA work of the one
// warp on that chain short (15 instead of 31 steps).  Pivot arithmetic, two variants:
//   EXACT (default): r = sqrt(d) (IEEE), rinv = 1 / r (IEEE) -- what LAPACK's dpotf2 computes;
//   FAST : rinv = rsqrt(d) polished by one Newton step (<= 0.5 ulp + 2^-70 from 1/sqrt(d)),
//          r = d * rinv corrected by one Newton step off the chain (the IEEE sqrt in all but
//          near-tie cases): no divide and no sqrt->divide dependency on the chain.
// The strictly lower triangle of the block in global memory is zeroed on exit.
// --------------------------------------------------------------------------------
template <bool FAST>
__device__ __noinline__ void potf2_diag_block(double* S, double* rdiag, int jb, int lane, int* fail, int* info,
                                                 int k0) {
    double col[IB];
    const int c = lane & (IB - 1);  // lanes 16..31 mirror lanes 0..15
#pragma unroll
    for (int i = 0; i < IB; ++i) col[i] = S[(jb + i) * PLD + jb + c];
    bool bad = false;
    int badj = 0;
#pragma unroll
    for (int j = 0; j < IB; ++j) {
        const double d = __shfl_sync(0xffffffffu, col[j], j);
        if (!(d > 0.0) && !bad) {  // also catches NaN
            bad = true;
            badj = j;
        }
        if (FAST) {
            double r, rinv;
            double y = rsqrt(d);
            const double e = fma(-d * y, y, 1.0);
            rinv = fma(0.5 * y, e, y);
            r = d * rinv;
            r = fma(fma(-r, r, d), 0.5 * rinv, r);
            const double ujc = (c == j) ? r : col[j] * rinv;
            if (c >= j) col[j] = ujc;
            if (lane == j) rdiag[j] = rinv;
#pragma unroll
            for (int i = j + 1; i < IB; ++i) {
                const double uji = __shfl_sync(0xffffffffu, ujc, i);
                if (c >= i) col[i] -= uji * ujc;
            }
        } else {
            // The recurrence from one pivot to the next only needs a_ji a_jc / d, one reciprocal per pivot:
            // (a_ji * (1/d)) * a_jc.  The square roots and their reciprocals, which scale the rows of the
            // factor (u_jc = a_jc * (1 / sqrt(d_j)): IEEE sqrt, IEEE reciprocal, one multiplication -- what
            // LAPACK's dpotf2 computes for row j), are taken for all 16 pivots at once after the loop.
            const double dinv = __drcp_rn(d);
            const double ajc = col[j];
            const double ljc = ajc * dinv;
#pragma unroll
            for (int i = j + 1; i < IB; ++i) {
                const double lji = __shfl_sync(0xffffffffu, ljc, i);
                if (c >= i) col[i] -= lji * ajc;
            }
        }
    }
    if (!FAST) {
        // lane c holds pivot d_c in col[c]
        double dc = col[0];
#pragma unroll
        for (int i = 1; i < IB; ++i)
            if (c == i) dc = col[i];
        const double rc = sqrt(dc), rinvc = 1.0 / rc;
        if (lane < IB) rdiag[c] = rinvc;
#pragma unroll
        for (int j = 0; j < IB; ++j) {
            const double rj = __shfl_sync(0xffffffffu, rinvc, j);
            if (c > j) col[j] *= rj;
            if (c == j) col[j] = rc;
        }
    }
    if (bad) {
        if (lane == 0) {
            *fail = 1;
            *info = k0 + jb + badj + 1;
        }
    } else if (lane < IB) {
#pragma unroll
        for (int i = 0; i < IB; ++i)
            if (i <= c) S[(jb + i) * PLD + jb + c] = col[i];
    }
}

__device__ __noinline__ void potf2_row_panel(double* S, const double* rdiag, int jb, int c) {
    double y[IB];
#pragma unroll
    for (int i = 0; i < IB; ++i) y[i] = S[(jb + i) * PLD + c];
#pragma unroll
    for (int k = 0; k < IB; ++k) {
        y[k] *= rdiag[k];
#pragma unroll
        for (int i = k + 1; i < IB; ++i) y[i] -= S[(jb + k) * PLD + jb + i] * y[k];
    }
#pragma unroll
    for (int i = 0; i < IB; ++i) S[(jb + i) * PLD + c] = y[i];
}

// S[i][c] -= sum_k U[jb+k][i] U[jb+k][c] for t0 <= i, c < nbp (block upper triangle) on the FP64 tensor
// pipe: every warp takes 16 x 16 blocks of the trailing matrix (2 x 2 m8n8k4 tiles, K = IB = 16 in four
// k-steps).  Per 1024 FMAs a warp issues 4 shared-memory loads instead of the 16 a 4 x 4 register tile
// of scalar FMAs needs -- the scalar form was bound by shared-memory bandwidth (8.3k cycles for the first
// update of a panel against 1.8k here).  With the row stride PLD = 132 the fragment loads
// (word = 4 t + g mod 16 per half-warp) are conflict-free.  Entries below the diagonal inside the
// diagonal blocks are updated too; nothing ever reads them.
__device__ __forceinline__ void dmma884_sub(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}
__device__ __forceinline__ void potf2_trailing_update(double* S, int jb, int t0, int nbp, int first, int lane, int stride) {
    const int nt = (nbp - t0) >> 4;  // 16-wide blocks per side
    const int cnt = nt * (nt + 1) / 2;
    const int g = lane >> 2, t = lane & 3;
    for (int idx = first; idx < cnt; idx += stride) {
        int bj = (int)((sqrtf(8.0f * idx + 1.0f) - 1.0f) * 0.5f);
        while (bj * (bj + 1) / 2 > idx) --bj;
        while ((bj + 1) * (bj + 2) / 2 <= idx) ++bj;
        const int bi = idx - bj * (bj + 1) / 2;
        const int i0 = t0 + 16 * bi, c0 = t0 + 16 * bj;
        double acc[2][2][2] = {{{0.0, 0.0}, {0.0, 0.0}}, {{0.0, 0.0}, {0.0, 0.0}}};
#pragma unroll
        for (int kk = 0; kk < IB / 4; ++kk) {
            const double* row = S + (jb + 4 * kk + t) * PLD;
            const double a0 = row[i0 + g], a1 = row[i0 + 8 + g];
            const double b0 = row[c0 + g], b1 = row[c0 + 8 + g];
            dmma884_sub(acc[0][0][0], acc[0][0][1], a0, b0);
            dmma884_sub(acc[0][1][0], acc[0][1][1], a0, b1);
            dmma884_sub(acc[1][0][0], acc[1][0][1], a1, b0);
            dmma884_sub(acc[1][1][0], acc[1][1][1], a1, b1);
        }
#pragma unroll
        for (int mi = 0; mi < 2; ++mi)
#pragma unroll
            for (int ni = 0; ni < 2; ++ni) {
                double2* pc = (double2*)(S + (i0 + 8 * mi + g) * PLD + c0 + 8 * ni + 2 * t);
                double2 v = *pc;
                v.x -= acc[mi][ni][0];
                v.y -= acc[mi][ni][1];
                *pc = v;
            }
    }
}

// Prologue of the panel kernel: S -= T' T on the block upper triangle, T = the 128 x nb block right above the
// diagonal block in global memory (the rows of the factor the previous panel's row solve just produced).  With it
// the rank-128 update of the NEXT diagonal block is part of the panel kernel, so that update of the rest of the
// block row (a separate launch on another stream) runs beside the panel factorization instead of before it.
// T passes through shared memory in chunks of 32 rows, transposed (TcT[col * TLD + k], TLD = 36: the staging
// stores and the m8n8k4 fragment loads are both conflict-free), the next chunk prefetched into registers while
// the current one feeds the tensor pipe; every warp keeps the accumulators of its (at most three) 16 x 16
// blocks in registers across the four chunks.
constexpr int PTLD = 36;
constexpr int TCHUNK = 32;
__device__ __forceinline__ void potf2_block_of(int idx, int& bi, int& bj) {
    bj = (int)((sqrtf(8.0f * idx + 1.0f) - 1.0f) * 0.5f);
    while (bj * (bj + 1) / 2 > idx) --bj;
    while ((bj + 1) * (bj + 2) / 2 <= idx) ++bj;
    bi = idx - bj * (bj + 1) / 2;
}
__device__ __noinline__ void potf2_prologue(double* S, double* TcT, const double* __restrict__ T, int64_t ldt, int nb,
                                            int tid, int warp, int lane) {
    constexpr int PER = TCHUNK * NB / 512;  // elements of a chunk per thread
    constexpr int NBLK = (NB / 16) * (NB / 16 + 1) / 2;
    const int g = lane >> 2, t = lane & 3;
    double acc[3][2][2][2];
#pragma unroll
    for (int b = 0; b < 3; ++b)
#pragma unroll
        for (int mi = 0; mi < 2; ++mi)
#pragma unroll
            for (int ni = 0; ni < 2; ++ni) acc[b][mi][ni][0] = acc[b][mi][ni][1] = 0.0;
    int i0[3], c0[3];
#pragma unroll
    for (int b = 0; b < 3; ++b) {
        const int idx = warp + 16 * b;
        int bi = 0, bj = 0;
        if (idx < NBLK) potf2_block_of(idx, bi, bj);
        i0[b] = 16 * bi;
        c0[b] = 16 * bj;
    }
    double v[PER];
    auto fetch = [&](int chunk) {
#pragma unroll
        for (int u = 0; u < PER; ++u) {
            const int e = tid + u * 512, k = e & (TCHUNK - 1), c = e >> 5;
            v[u] = (c < nb) ? __ldg(T + chunk * TCHUNK + k + (int64_t)c * ldt) : 0.0;
        }
    };
    fetch(0);
    for (int chunk = 0; chunk < NB / TCHUNK; ++chunk) {
#pragma unroll
        for (int u = 0; u < PER; ++u) {
            const int e = tid + u * 512, k = e & (TCHUNK - 1), c = e >> 5;
            TcT[c * PTLD + k] = v[u];
        }
        __syncthreads();
        if (chunk + 1 < NB / TCHUNK) fetch(chunk + 1);
#pragma unroll
        for (int b = 0; b < 3; ++b) {
            if (warp + 16 * b >= NBLK) continue;
            const double* pa0 = TcT + (i0[b] + g) * PTLD + t;
            const double* pa1 = pa0 + 8 * PTLD;
            const double* pb0 = TcT + (c0[b] + g) * PTLD + t;
            const double* pb1 = pb0 + 8 * PTLD;
#pragma unroll
            for (int kk = 0; kk < TCHUNK / 4; ++kk) {
                const double a0 = pa0[4 * kk], a1 = pa1[4 * kk], b0 = pb0[4 * kk], b1 = pb1[4 * kk];
                dmma884_sub(acc[b][0][0][0], acc[b][0][0][1], a0, b0);
                dmma884_sub(acc[b][0][1][0], acc[b][0][1][1], a0, b1);
                dmma884_sub(acc[b][1][0][0], acc[b][1][0][1], a1, b0);
                dmma884_sub(acc[b][1][1][0], acc[b][1][1][1], a1, b1);
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int b = 0; b < 3; ++b) {
        if (warp + 16 * b >= NBLK) continue;
#pragma unroll
        for (int mi = 0; mi < 2; ++mi)
#pragma unroll
            for (int ni = 0; ni < 2; ++ni) {
                double2* pc = (double2*)(S + (i0[b] + 8 * mi + g) * PLD + c0[b] + 8 * ni + 2 * t);
                double2 w = *pc;
                w.x -= acc[b][mi][ni][0];
                w.y -= acc[b][mi][ni][1];
                *pc = w;
            }
    }
    __syncthreads();
}

#ifdef HYB_PROFILE_POTF2
__device__ long long g_potf2_ts[64];
__device__ long long g_trsm_ts[16];
#define POTF2_TS(i) do { if (tid == 0) g_potf2_ts[i] = clock64(); } while (0)
#else
#define POTF2_TS(i)
#endif

template <bool FAST>
__global__ void __launch_bounds__(512, 1)
potf2_kernel(double* __restrict__ A, int64_t lda, int nb, int* __restrict__ info, int k0, int pending) {
    extern __shared__ double sm[];
    double* S = sm;               // S[r * PLD + c]
    double* rdiag = sm + NB * PLD;  // reciprocals of the current diagonal block's pivots
    double* TcT = rdiag + IB;       // staging of the prologue (NB x TLD)
    __shared__ int fail;
    const int tid = threadIdx.x, nth = blockDim.x, lane = tid & 31, warp = tid >> 5;
    if (*info != 0) return;  // an earlier panel failed: nothing meaningful to do
    if (tid == 0) fail = 0;
    POTF2_TS(0);
    // stage the block: every thread issues all 32 of its loads before the first store, so a full
    // bandwidth-latency product of requests is in flight (a load->store loop leaves one per thread)
    {
        double v[NB * NB / 512];
#pragma unroll
        for (int t = 0; t < NB * NB / 512; ++t) {
            const int e = tid + t * 512, r = e & (NB - 1), c = e >> 7;
            v[t] = (r < nb && c < nb && r <= c) ? __ldg(A + r + (int64_t)c * lda) : ((r == c) ? 1.0 : 0.0);
        }
#pragma unroll
        for (int t = 0; t < NB * NB / 512; ++t) {
            const int e = tid + t * 512, r = e & (NB - 1), c = e >> 7;
            S[r * PLD + c] = v[t];
        }
    }
    __syncthreads();
    // rank-128 update from the previous panel's rows, when the caller left this block out of that update
    if (pending) potf2_prologue(S, TcT, A - NB, lda, nb, tid, warp, lane);

    POTF2_TS(1);
    // identity padding makes every IB x IB step full-size: no tail cases below
    const int nbp = (nb + IB - 1) / IB * IB;
    // Pipelined over the 16-column blocks: warp 0 owns the serial part.  After the row panel of block b it
    // updates the next diagonal block alone and factors it at once, while the other 15 warps apply the rest of
    // the rank-16 update; one barrier later the next row panel finds both done.
    if (warp == 0) potf2_diag_block<FAST>(S, rdiag, 0, lane, &fail, info, k0);
    __syncthreads();
    POTF2_TS(2);
    if (fail) return;
    for (int jb = 0; jb < nbp; jb += IB) {
        const int t0 = jb + IB;   // first trailing column
        const int m = nbp - t0;   // trailing size
        if (m <= 0) break;
        // row panel: U[jb:jb+IB, c] = U11^-T A[jb:jb+IB, c], thread per trailing column
        if (tid < m) potf2_row_panel(S, rdiag, jb, t0 + tid);
        __syncthreads();
        POTF2_TS(3 + 3 * (jb / IB));
        if (warp == 0) {
            potf2_trailing_update(S, jb, t0, nbp, 0, lane, 1 << 30);  // block (t0, t0) only
            __syncwarp();
            POTF2_TS(4 + 3 * (jb / IB));
            potf2_diag_block<FAST>(S, rdiag, t0, lane, &fail, info, k0);
        } else {
            potf2_trailing_update(S, jb, t0, nbp, warp, lane, (nth >> 5) - 1);
        }
        __syncthreads();
        POTF2_TS(5 + 3 * (jb / IB));
        if (fail) return;
    }
    POTF2_TS(40);
    // write back: U in the upper triangle, zeros strictly below (the block is workspace)
    for (int e = tid; e < NB * NB; e += nth) {
        const int r = e & (NB - 1), c = e >> 7;
        if (r < nb && c < nb) A[r + (int64_t)c * lda] = (r <= c) ? S[r * PLD + c] : 0.0;
    }
    POTF2_TS(41);
}

#ifdef HYB_PROFILE_POTF2
void hyb_potf2_dump_ts() {
    long long ts[64];
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(ts, g_potf2_ts, sizeof(ts));
    fprintf(stderr, "potf2 clocks: stage+prologue %lld | diag0 %lld", ts[1] - ts[0], ts[2] - ts[1]);
    for (int b = 0; b + 1 < NB / IB; ++b)
        fprintf(stderr, " | panel %lld upd0 %lld diag+join %lld", ts[3 + 3 * b] - ts[2 + 3 * b], ts[4 + 3 * b] - ts[3 + 3 * b],
                ts[5 + 3 * b] - ts[4 + 3 * b]);
    fprintf(stderr, " | writeback %lld | total %lld\n", ts[41] - ts[40], ts[41] - ts[0]);
    long long tt[16];
    cudaMemcpyFromSymbol(tt, g_trsm_ts, sizeof(tt));
    fprintf(stderr, "trsm clocks: stage %lld | s0 %lld u1 %lld s1 %lld u2 %lld s2 %lld u3 %lld s3 %lld | writeback %lld | total %lld\n",
            tt[1] - tt[0], tt[2] - tt[1], tt[3] - tt[2], tt[4] - tt[3], tt[5] - tt[4], tt[6] - tt[5], tt[7] - tt[6],
            tt[8] - tt[7], tt[9] - tt[8], tt[9] - tt[0]);
}
#endif

// --------------------------------------------------------------------------------
// row panel of the blocked factorization: A12 <- U11^-T A12 (U11 128 x 128 upper),
// forward substitution, one thread per column of A12 in 32-row register chunks.
// U11 is read straight from global memory with warp-uniform (broadcast) loads that
// stay in L1; solved chunks of y are parked in shared memory for the later chunks.
// Backward stable (LAPACK dtrsm semantics), no explicit inverse.
// --------------------------------------------------------------------------------
constexpr int TRSM_THREADS = 128;
constexpr int TRSM_COLS = 64;          // columns of the right-hand side per CTA (widest instance)
// The slab kernels are templates on COLS (64, 32 or 16 columns per CTA): the work of a CTA is
// 128 x 128 / 2 x COLS FP64 FMAs on one SM's FP64 pipe, so narrower slabs on more SMs shorten the
// kernel whenever the launch still fits one wave -- which is what matters for the row panel of the
// Cholesky factorization, where this kernel sits on the serial chain.
template <int COLS> struct TrsmCfg {
    static constexpr int TLD = COLS + 1;      // padded slab leading dimension
    static constexpr int CPT = COLS / 16;     // columns per thread in the register-tiled updates
    static constexpr int SMEM = (NB * LDS + NB * (COLS + 1) + NB) * (int)sizeof(double);
};

// substitution of one 32-row chunk, one thread per column, chunk in registers
// (template instances called in sequence: a register array declared inside a runtime
//  chunk loop gets demoted to local memory by nvcc)
template <int CB, int COLS>
__device__ __forceinline__ void trsm_solve_chunk(const double* Uc, const double* dinv, double* T, int c) {
    constexpr int TLD = TrsmCfg<COLS>::TLD;
    double v[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = T[(CB * 32 + i) * TLD + c];
#pragma unroll
    for (int k = 0; k < 32; ++k) {
        v[k] *= dinv[CB * 32 + k];
#pragma unroll
        for (int i = k + 1; i < 32; ++i) v[i] -= Uc[(CB * 32 + i) * LDS + CB * 32 + k] * v[k];
    }
#pragma unroll
    for (int i = 0; i < 32; ++i) T[(CB * 32 + i) * TLD + c] = v[i];
}

// rows R0..R0+31 of the slab -= U[0:K, R0:R0+32]' * Y[0:K, :]  (K = CB*32), 4 x CPT register tiles
template <int CB, int COLS>
__device__ __forceinline__ void trsm_update_chunk(const double* Uc, double* T, int tid) {
    if (CB == 0) return;
    constexpr int TLD = TrsmCfg<COLS>::TLD, CPT = TrsmCfg<COLS>::CPT;
    // a thread's 4 rows are 8 apart and its columns 16 apart, so the lanes of a warp read
    // consecutive words of the slab (contiguous tiles put them 32 bytes apart: 4-way conflicts)
    const int tr = tid >> 4, tc = tid & 15;  // 8 row groups x 16 column groups
    const int r0 = CB * 32 + tr, c0 = tc;
    double acc[4][CPT];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < CPT; ++j) acc[i][j] = 0.0;
#pragma unroll 4
    for (int k = 0; k < CB * 32; ++k) {
        double u[4], y[CPT];
#pragma unroll
        for (int i = 0; i < 4; ++i) u[i] = Uc[(r0 + 8 * i) * LDS + k];
#pragma unroll
        for (int j = 0; j < CPT; ++j) y[j] = T[k * TLD + c0 + 16 * j];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < CPT; ++j) acc[i][j] += u[i] * y[j];
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < CPT; ++j) T[(r0 + 8 * i) * TLD + c0 + 16 * j] -= acc[i][j];
}

// --------------------------------------------------------------------------------
// row panel of the blocked factorization: A12 <- U11^-T A12 (U11 128 x 128 upper).
// A CTA stages U11 and a 128 x COLS slab of A12 in shared memory and alternates, per
// 32-row chunk, a register-tiled update with the rows already solved and a forward
// substitution with the 32 x 32 diagonal block (one thread per column, chunk in
// registers).  Substitution only - no explicit inverse (LAPACK dtrsm semantics).
// --------------------------------------------------------------------------------
// stage U11 (nb x nb upper, identity padded to 128) and a 128 x COLS slab of the right-hand sides
template <int COLS>
__device__ __forceinline__ void trsm_stage(const double* __restrict__ U11, int64_t ldu, int nb,
                                           const double* __restrict__ B, int64_t ld, int64_t col0, int ncol, double* Uc,
                                           double* T, double* dinv, int tid) {
    constexpr int TLD = TrsmCfg<COLS>::TLD;
    constexpr int CW = COLS < 32 ? COLS : 32;
    // batches of 32 loads per thread in flight before the stores (see potf2_kernel)
    static_assert(TRSM_THREADS == NB, "staging below assumes one thread per row");
#pragma unroll 1
    for (int i0 = 0; i0 < NB; i0 += 32) {
        double v[32];
#pragma unroll
        for (int t = 0; t < 32; ++t) {
            const int k = tid, i = i0 + t;
            v[t] = (k == i) ? 1.0 : 0.0;
            if (k <= i && i < nb) v[t] = __ldg(U11 + k + (int64_t)i * ldu);
        }
#pragma unroll
        for (int t = 0; t < 32; ++t) Uc[(i0 + t) * LDS + tid] = v[t];
    }
#pragma unroll 1
    for (int c0 = 0; c0 < COLS; c0 += CW) {
        double v[CW];
#pragma unroll
        for (int t = 0; t < CW; ++t) {
            const int r = tid, c = c0 + t;
            v[t] = (c < ncol && r < nb) ? B[r + (col0 + c) * ld] : 0.0;
        }
#pragma unroll
        for (int t = 0; t < CW; ++t) T[tid * TLD + c0 + t] = v[t];
    }
    __syncthreads();
    if (tid < NB) dinv[tid] = 1.0 / Uc[tid * LDS + tid];
    __syncthreads();
}

#ifdef HYB_PROFILE_POTF2
#define TRSM_TS(i) do { if (tid == 0 && blockIdx.x == 0) g_trsm_ts[i] = clock64(); } while (0)
#else
#define TRSM_TS(i)
#endif

template <int COLS>
__global__ void __launch_bounds__(TRSM_THREADS, 1)
trsm_upper_T_kernel(const double* __restrict__ U11, int64_t ldu, int nb, double* __restrict__ A12, int64_t ld,
                    int64_t n2) {
    constexpr int TLD = TrsmCfg<COLS>::TLD;
    extern __shared__ double sm[];
    double* Uc = sm;                     // Uc[i * LDS + k] = U[k][i]  (column i of U11, contiguous in k)
    double* T = sm + NB * LDS;           // T[r * TLD + c]
    double* dinv = T + NB * TLD;         // 1 / U[k][k]
    const int tid = threadIdx.x;
    const int64_t col0 = (int64_t)blockIdx.x * COLS;
    const int ncol = (int)((n2 - col0 < COLS) ? (n2 - col0) : COLS);
    TRSM_TS(0);
    trsm_stage<COLS>(U11, ldu, nb, A12, ld, col0, ncol, Uc, T, dinv, tid);
    TRSM_TS(1);
    const bool solver_thread = tid < COLS;
    if (solver_thread) trsm_solve_chunk<0, COLS>(Uc, dinv, T, tid);
    __syncthreads();
    TRSM_TS(2);
    trsm_update_chunk<1, COLS>(Uc, T, tid);
    __syncthreads();
    TRSM_TS(3);
    if (solver_thread) trsm_solve_chunk<1, COLS>(Uc, dinv, T, tid);
    __syncthreads();
    TRSM_TS(4);
    trsm_update_chunk<2, COLS>(Uc, T, tid);
    __syncthreads();
    TRSM_TS(5);
    if (solver_thread) trsm_solve_chunk<2, COLS>(Uc, dinv, T, tid);
    __syncthreads();
    TRSM_TS(6);
    trsm_update_chunk<3, COLS>(Uc, T, tid);
    __syncthreads();
    TRSM_TS(7);
    if (solver_thread) trsm_solve_chunk<3, COLS>(Uc, dinv, T, tid);
    __syncthreads();
    TRSM_TS(8);
    for (int e = tid; e < NB * COLS; e += TRSM_THREADS) {
        const int r = e & (NB - 1), c = e >> 7;
        if (c < ncol && r < nb) A12[r + (col0 + c) * ld] = T[r * TLD + c];
    }
    TRSM_TS(9);
}

// ---- backward variant: slab <- U11^-1 slab (U11 upper, back substitution) ----------------
constexpr int TLD = TrsmCfg<TRSM_COLS>::TLD;
template <int CB>
__device__ __forceinline__ void trsm_bsolve_chunk(const double* Uc, const double* dinv, double* T, int c) {
    double v[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = T[(CB * 32 + i) * TLD + c];
#pragma unroll
    for (int k = 31; k >= 0; --k) {
        v[k] *= dinv[CB * 32 + k];
#pragma unroll
        for (int i = 0; i < k; ++i) v[i] -= Uc[(CB * 32 + k) * LDS + CB * 32 + i] * v[k];  // U[i][k]
    }
#pragma unroll
    for (int i = 0; i < 32; ++i) T[(CB * 32 + i) * TLD + c] = v[i];
}

// rows of chunk CB -= U[chunk rows, later rows] * X[later rows, :]   (later = (CB+1)*32 .. 127)
template <int CB>
__device__ __forceinline__ void trsm_bupdate_chunk(const double* Uc, double* T, int tid) {
    if (CB == 3) return;
    const int tr = tid >> 4, tc = tid & 15;
    const int r0 = CB * 32 + tr, c0 = tc;  // interleaved like trsm_update_chunk
    double acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.0;
#pragma unroll 4
    for (int k = (CB + 1) * 32; k < NB; ++k) {
        double u[4], y[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) u[i] = Uc[k * LDS + r0 + 8 * i];  // U[r][k]
#pragma unroll
        for (int j = 0; j < 4; ++j) y[j] = T[k * TLD + c0 + 16 * j];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] += u[i] * y[j];
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) T[(r0 + 8 * i) * TLD + c0 + 16 * j] -= acc[i][j];
}

__global__ void __launch_bounds__(TRSM_THREADS, 1)
trsm_upper_N_kernel(const double* __restrict__ U11, int64_t ldu, int nb, double* __restrict__ B, int64_t ld,
                    int64_t n2) {
    extern __shared__ double sm[];
    double* Uc = sm;
    double* T = sm + NB * LDS;
    double* dinv = T + NB * TLD;
    const int tid = threadIdx.x;
    const int64_t col0 = (int64_t)blockIdx.x * TRSM_COLS;
    const int ncol = (int)((n2 - col0 < TRSM_COLS) ? (n2 - col0) : TRSM_COLS);
    trsm_stage<TRSM_COLS>(U11, ldu, nb, B, ld, col0, ncol, Uc, T, dinv, tid);
    const bool solver_thread = tid < TRSM_COLS;
    if (solver_thread) trsm_bsolve_chunk<3>(Uc, dinv, T, tid);
    __syncthreads();
    trsm_bupdate_chunk<2>(Uc, T, tid);
    __syncthreads();
    if (solver_thread) trsm_bsolve_chunk<2>(Uc, dinv, T, tid);
    __syncthreads();
    trsm_bupdate_chunk<1>(Uc, T, tid);
    __syncthreads();
    if (solver_thread) trsm_bsolve_chunk<1>(Uc, dinv, T, tid);
    __syncthreads();
    trsm_bupdate_chunk<0>(Uc, T, tid);
    __syncthreads();
    if (solver_thread) trsm_bsolve_chunk<0>(Uc, dinv, T, tid);
    __syncthreads();
    for (int e = tid; e < NB * TRSM_COLS; e += TRSM_THREADS) {
        const int r = e & (NB - 1), c = e >> 7;
        if (c < ncol && r < nb) B[r + (col0 + c) * ld] = T[r * TLD + c];
    }
}

// dst (cols x rows) = src' ; upper_only: read src's upper triangle only (zeros elsewhere)
__global__ void transpose_kernel(const double* __restrict__ src, int64_t lds_, int64_t rows, int64_t cols,
                                 double* __restrict__ dst, int64_t ldd, int upper_only) {
    __shared__ double tile[32][33];
    const int64_t r0 = (int64_t)blockIdx.x * 32, c0 = (int64_t)blockIdx.y * 32;
    for (int j = threadIdx.y; j < 32; j += 8) {
        const int64_t r = r0 + threadIdx.x, c = c0 + j;
        double v = 0.0;
        if (r < rows && c < cols && (!upper_only || r <= c)) v = src[r + c * lds_];
        tile[j][threadIdx.x] = v;
    }
    __syncthreads();
    for (int j = threadIdx.y; j < 32; j += 8) {
        const int64_t c = c0 + threadIdx.x, r = r0 + j;  // dst(c, r) = src(r, c)
        if (r < rows && c < cols) dst[c + r * ldd] = tile[threadIdx.x][j];
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

// Wavefront solve kernels: 512 threads, one CTA per 128-row block.
//  * tiles of U that a block needs are double-buffered in registers, so the loads of tile
//    jb+1 are in flight while the CTA waits for / consumes x-block jb;
//  * the diagonal block is staged in shared memory pre-scaled by the reciprocal pivots, which
//    shortens the serial chain of the substitution to one shuffle + one FMA per row.
constexpr int TRSV_THREADS = 512;

// Both kernels are templates on R, the number of right-hand sides solved together (columns of x,
// leading dimension ldx): the tiles of U -- all of the traffic -- are loaded once for the R
// vectors, and the R serial substitutions of a block run on R different warps at the same time,
// so R systems cost about the latency of one.

// forward: solve U' y = b in place (x holds b on entry, y on exit)
template <int R>
__global__ void __launch_bounds__(TRSV_THREADS, 1)
trsv_fwd_kernel(const double* __restrict__ U, int64_t ld, double* x, int64_t ldx, int N, int nblk, int* flags, int epoch) {
    extern __shared__ double sm[];
    double* Us = sm;               // Us[k * LDS + i] = U[k][i] / U[i][i]  (k < i)
    double* dinv = sm + NB * LDS;  // 1 / U[i][i]
    double* ys = dinv + NB;        // [R][NB]
    double* rs = ys + R * NB;      // [R][NB]
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int ib = blockIdx.x; ib < nblk; ib += gridDim.x) {
        __syncthreads();  // previous block's substitution is done with Us
        const int nbr = min(NB, N - ib * NB);
        const double* dblk = U + (int64_t)ib * NB + (int64_t)ib * NB * ld;
        if (tid < NB) dinv[tid] = (tid < nbr) ? 1.0 / __ldg(dblk + tid + (int64_t)tid * ld) : 1.0;
        __syncthreads();
        for (int e = tid; e < NB * NB; e += TRSV_THREADS) {
            const int r = e & (NB - 1), c = e >> 7;
            Us[r * LDS + c] = (r < c && c < nbr) ? __ldg(dblk + r + (int64_t)c * ld) * dinv[c] : 0.0;
        }
        // column dots with the tiles above the diagonal: warp w owns columns 8w..8w+7
        double acc[R][8];
#pragma unroll
        for (int rr = 0; rr < R; ++rr)
#pragma unroll
            for (int c = 0; c < 8; ++c) acc[rr][c] = 0.0;
        const int col0 = ib * NB + warp * 8;
        double2 nx0[8], nx1[8];
        auto load_tile = [&](int jb) {
            const double* base = U + (int64_t)jb * NB + 2 * lane;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const int col = col0 + c;
                if (col < N) {
                    nx0[c] = __ldg((const double2*)(base + (int64_t)col * ld));
                    nx1[c] = __ldg((const double2*)(base + (int64_t)col * ld + 64));
                } else {
                    nx0[c] = nx1[c] = make_double2(0.0, 0.0);
                }
            }
        };
        if (ib > 0) load_tile(0);
        for (int jb = 0; jb < ib; ++jb) {
            // tile jb is already in registers (its loads were issued before this wait)
            if (tid == 0) {
                while (ld_acquire(&flags[jb]) != epoch) {
                }
            }
            __syncthreads();
            if (tid < R * NB) ys[tid] = __ldcg(&x[jb * NB + (tid & (NB - 1)) + (int64_t)(tid >> 7) * ldx]);
            __syncthreads();
#pragma unroll
            for (int rr = 0; rr < R; ++rr) {
                const double* yr = ys + rr * NB;
                const double y0 = yr[2 * lane], y1 = yr[2 * lane + 1], y2 = yr[64 + 2 * lane], y3 = yr[65 + 2 * lane];
#pragma unroll
                for (int c = 0; c < 8; ++c)
                    acc[rr][c] += (nx0[c].x * y0 + nx0[c].y * y1) + (nx1[c].x * y2 + nx1[c].y * y3);
            }
            if (jb + 1 < ib) load_tile(jb + 1);
        }
#pragma unroll
        for (int rr = 0; rr < R; ++rr)
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                double v = acc[rr][c];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                acc[rr][c] = v;
            }
        __syncthreads();  // everyone is done with ys; Us fully staged
        if (lane == 0) {
#pragma unroll
            for (int rr = 0; rr < R; ++rr)
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const int col = col0 + c;
                    rs[rr * NB + warp * 8 + c] = (col < N) ? (x[col + (int64_t)rr * ldx] - acc[rr][c]) * dinv[warp * 8 + c] : 0.0;
                }
        }
        __syncthreads();
        // y = U11^-T r by forward substitution on the pre-scaled block: with r~_i = r_i / U_ii,
        // y_k = r~_k and r~_i -= (U[k][i] / U_ii) y_k.  One warp per right-hand side, 4 rows per
        // lane in registers.
        if (warp < R) {
            double rv[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) rv[j] = rs[warp * NB + lane + 32 * j];
#pragma unroll
            for (int slot = 0; slot < 4; ++slot) {
#pragma unroll 8
                for (int kk = 0; kk < 32; ++kk) {
                    const int k = slot * 32 + kk;
                    const double yk = __shfl_sync(0xffffffffu, rv[slot], kk);
                    const double* urow = Us + k * LDS + lane;
#pragma unroll
                    for (int j = 0; j < 4; ++j) rv[j] -= urow[32 * j] * yk;  // entries with i <= k are stored as 0
                }
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int row = ib * NB + lane + 32 * j;
                if (row < N) x[row + (int64_t)warp * ldx] = rv[j];
            }
        }
        __syncthreads();
        if (tid == 0) st_release(&flags[ib], epoch);
    }
}

// backward: solve U x = y in place
template <int R>
__global__ void __launch_bounds__(TRSV_THREADS, 1)
trsv_bwd_kernel(const double* __restrict__ U, int64_t ld, double* x, int64_t ldx, int N, int nblk, int* flags, int epoch) {
    extern __shared__ double sm[];
    double* Us = sm;               // Us[i * LDS + k] = U[i][k] / U[i][i]  (i < k)
    double* dinv = sm + NB * LDS;
    double* xs = dinv + NB;        // [R][NB]
    double* part = xs + R * NB;    // [R][4][NB]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, r = tid & (NB - 1), quarter = tid >> 7;
    for (int ib = nblk - 1 - (int)blockIdx.x; ib >= 0; ib -= gridDim.x) {
        __syncthreads();
        const int nbr = min(NB, N - ib * NB);
        const double* dblk = U + (int64_t)ib * NB + (int64_t)ib * NB * ld;
        if (tid < NB) dinv[tid] = (tid < nbr) ? 1.0 / __ldg(dblk + tid + (int64_t)tid * ld) : 1.0;
        __syncthreads();
        for (int e = tid; e < NB * NB; e += TRSV_THREADS) {
            const int rr = e & (NB - 1), c = e >> 7;
            Us[rr * LDS + c] = (rr < c && c < nbr) ? __ldg(dblk + rr + (int64_t)c * ld) * dinv[rr] : 0.0;
        }
        const int row = ib * NB + r;
        const bool row_ok = row < N;
        double acc[R];
#pragma unroll
        for (int v = 0; v < R; ++v) acc[v] = 0.0;
        double nx[32];
        auto load_tile = [&](int jb) {
            const int ncol = min(NB, N - jb * NB);
            const double* p = U + row + (int64_t)(jb * NB + quarter * 32) * ld;
            const int cend = ncol - quarter * 32;
#pragma unroll
            for (int c = 0; c < 32; ++c) nx[c] = (row_ok && c < cend) ? __ldg(p + (int64_t)c * ld) : 0.0;
        };
        if (ib < nblk - 1) load_tile(nblk - 1);
        for (int jb = nblk - 1; jb > ib; --jb) {
            // tile jb is already in registers (its loads were issued before this wait)
            if (tid == 0) {
                while (ld_acquire(&flags[jb]) != epoch) {
                }
            }
            __syncthreads();
            if (tid < R * NB) {
                const int i = tid & (NB - 1);
                xs[tid] = (jb * NB + i < N) ? __ldcg(&x[jb * NB + i + (int64_t)(tid >> 7) * ldx]) : 0.0;
            }
            __syncthreads();
#pragma unroll
            for (int v = 0; v < R; ++v) {
                const double* xv = xs + v * NB + quarter * 32;
                double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll
                for (int c = 0; c < 32; c += 4) {
                    a0 += nx[c] * xv[c];
                    a1 += nx[c + 1] * xv[c + 1];
                    a2 += nx[c + 2] * xv[c + 2];
                    a3 += nx[c + 3] * xv[c + 3];
                }
                acc[v] += (a0 + a1) + (a2 + a3);
            }
            if (jb - 1 > ib) load_tile(jb - 1);
        }
        __syncthreads();
#pragma unroll
        for (int v = 0; v < R; ++v) part[(v * 4 + quarter) * NB + r] = acc[v];
        __syncthreads();
        // x = U11^-1 (y - acc) by backward substitution on the row-scaled block, one warp per
        // right-hand side
        if (warp < R) {
            const double* pv = part + warp * 4 * NB;
            double rv[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int i = lane + 32 * j;
                const int rw = ib * NB + i;
                const double s = (pv[i] + pv[NB + i]) + (pv[2 * NB + i] + pv[3 * NB + i]);
                rv[j] = (rw < N) ? (x[rw + (int64_t)warp * ldx] - s) * dinv[i] : 0.0;
            }
#pragma unroll
            for (int slot = 3; slot >= 0; --slot) {
#pragma unroll 8
                for (int kk = 31; kk >= 0; --kk) {
                    const int k = slot * 32 + kk;
                    const double xk = __shfl_sync(0xffffffffu, rv[slot], kk);
                    const double* ucol = Us + lane * LDS + k;
#pragma unroll
                    for (int j = 0; j < 4; ++j) rv[j] -= ucol[32 * j * LDS] * xk;  // entries with i >= k are 0
                }
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int rw = ib * NB + lane + 32 * j;
                if (rw < N) x[rw + (int64_t)warp * ldx] = rv[j];
            }
        }
        __syncthreads();
        if (tid == 0) st_release(&flags[ib], epoch);
    }
}

}  // namespace

int hyb_chol_alloc(HybChol& ch, int64_t N) {
    ch.N = N;
    ch.ld = hyb_round_up(N > 0 ? N : 1, 16);
    ch.nblk = (N + NB - 1) / NB;
    HYB_CUDA(cudaMalloc(&ch.F, sizeof(double) * ch.ld * (N > 0 ? N : 1)));
    HYB_CUDA(cudaMalloc(&ch.flags, sizeof(int) * (ch.nblk > 0 ? ch.nblk : 1)));
    HYB_CUDA(cudaMemset(ch.flags, 0, sizeof(int) * (ch.nblk > 0 ? ch.nblk : 1)));
    HYB_CUDA(cudaMalloc(&ch.info, sizeof(int)));
    int lo = 0, hi = 0;
    HYB_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    HYB_CUDA(cudaStreamCreateWithPriority(&ch.panel_stream, cudaStreamNonBlocking, hi));
    HYB_CUDA(cudaEventCreateWithFlags(&ch.ev_head, cudaEventDisableTiming));
    HYB_CUDA(cudaEventCreateWithFlags(&ch.ev_panel, cudaEventDisableTiming));
    HYB_CUDA(cudaStreamCreateWithPriority(&ch.upd_stream, cudaStreamNonBlocking, hi));
    HYB_CUDA(cudaEventCreateWithFlags(&ch.ev_trsm, cudaEventDisableTiming));
    HYB_CUDA(cudaEventCreateWithFlags(&ch.ev_upd, cudaEventDisableTiming));
    for (auto& e : ch.ev_far) HYB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    HYB_CUDA(cudaMalloc(&ch.work, sizeof(double) * (N > 0 ? N : 1)));
    ch.epoch = 0;
    return 0;
}

void hyb_chol_free(HybChol& ch) {
    if (ch.owns_F) cudaFree(ch.F);
    cudaFree(ch.flags);
    cudaFree(ch.info);
    if (ch.panel_stream) cudaStreamDestroy(ch.panel_stream);
    if (ch.ev_head) cudaEventDestroy(ch.ev_head);
    if (ch.ev_panel) cudaEventDestroy(ch.ev_panel);
    if (ch.upd_stream) cudaStreamDestroy(ch.upd_stream);
    if (ch.ev_trsm) cudaEventDestroy(ch.ev_trsm);
    if (ch.ev_upd) cudaEventDestroy(ch.ev_upd);
    for (auto& e : ch.ev_far)
        if (e) cudaEventDestroy(e);
    cudaFree(ch.work);
    cudaFree(ch.ipiv);
    cudaFree(ch.bk_ws);
    cudaFree(ch.bk_state);
    ch = HybChol();
}

constexpr int PANEL_SMEM = (NB * PLD + IB + NB * PTLD) * (int)sizeof(double);
constexpr int TRSM_SMEM = TrsmCfg<TRSM_COLS>::SMEM;
constexpr int TRSV_SMEM = (NB * LDS + NB + 5 * HYB_MAX_RHS * NB) * (int)sizeof(double);
hyb_counter g_potrf_launches{0};

// Right-looking factorization, two-level blocking, look-ahead of two outer blocks.
//  * Outer blocks of OB columns (default 512, HYB_POTRF_OB): everything further than one outer block below a
//    block is updated once per outer block with a rank-OB contraction, so those entries are read and written
//    N/OB times instead of N/128 times (a K = 128 update runs at half the DMMA rate of a K = 512 one: its
//    128 x 128 tile costs as much in launch, pipeline fill and read-modify-write of C as in DMMA work).
//  * Inside an outer block the 128-column panels are factored left to right: panel kernel (diagonal block,
//    including the rank-128 update the previous panel owes it) -> row solve -> rank-128 "near" update of the
//    rest of the outer block and of the whole next one.  See the choreography comment in hyb_potrf.
// The strictly lower triangle of ch.F is workspace (the rectangular updates write there).
static int potrf_outer_block() {
    static int ob = 0;
    if (ob == 0) {
        ob = 512;
        const char* e = getenv("HYB_POTRF_OB");
        if (e) {
            const int v = atoi(e);
            if (v == 128 || v == 256 || v == 512 || v == 1024) ob = v;
        }
    }
    return ob;
}

// per-device attributes of the panel / slab kernels (dynamic shared memory above 48 KB)
static int ensure_panel_attrs() {
    static HybDevOnce once;
    if (HybDevOnce::Guard g_(once); g_.first) {
        HYB_CUDA(cudaFuncSetAttribute(potf2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, PANEL_SMEM));
        HYB_CUDA(cudaFuncSetAttribute(potf2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, PANEL_SMEM));
        HYB_CUDA(cudaFuncSetAttribute(trsm_upper_T_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, TrsmCfg<64>::SMEM));
        HYB_CUDA(cudaFuncSetAttribute(trsm_upper_T_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, TrsmCfg<32>::SMEM));
        HYB_CUDA(cudaFuncSetAttribute(trsm_upper_T_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, TrsmCfg<16>::SMEM));
        HYB_CUDA(cudaFuncSetAttribute(trsm_upper_N_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TRSM_SMEM));
    }
    return 0;
}

// pivot arithmetic of the panel kernel: IEEE sqrt + divide (default) or the polished rsqrt form
// (HYB_POTF2_FAST=1), see potf2_diag_block
static bool potf2_fast() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("HYB_POTF2_FAST");
        v = (e && atoi(e) != 0) ? 1 : 0;
    }
    return v == 1;
}

// A12 (nb x n2) <- U11^-T A12 with the narrowest slabs that still fit one wave of CTAs
static int launch_row_trsm(cudaStream_t s, const double* U11, int64_t ldu, int nb, double* A12, int64_t ld, int64_t n2) {
    const int sms = hyb_num_sms();
    if (n2 <= (int64_t)16 * sms)
        trsm_upper_T_kernel<16><<<(unsigned)((n2 + 15) / 16), TRSM_THREADS, TrsmCfg<16>::SMEM, s>>>(U11, ldu, nb, A12, ld, n2);
    else if (n2 <= (int64_t)32 * sms)
        trsm_upper_T_kernel<32><<<(unsigned)((n2 + 31) / 32), TRSM_THREADS, TrsmCfg<32>::SMEM, s>>>(U11, ldu, nb, A12, ld, n2);
    else
        trsm_upper_T_kernel<64><<<(unsigned)((n2 + 63) / 64), TRSM_THREADS, TrsmCfg<64>::SMEM, s>>>(U11, ldu, nb, A12, ld, n2);
    ++g_potrf_launches;
    HYB_CUDA(cudaGetLastError());
    return 0;
}

int hyb_potrf(cudaStream_t st, HybChol& ch, int* info_out) {
    const int64_t N = ch.N, ld = ch.ld;
    HYB_CHECK(ensure_panel_attrs());
    const int64_t OB = potrf_outer_block();
    const bool fast = potf2_fast();
    ch.kind = 0;
    cudaStream_t sp = ch.panel_stream;
    HYB_CUDA(cudaMemsetAsync(ch.info, 0, sizeof(int), st));
    cudaStream_t su = ch.upd_stream;
    auto factor_block = [&](int64_t k0, int pending) -> int {  // diagonal block, on the panel stream
        const int nb = (int)((N - k0 < NB) ? (N - k0) : NB);
        double* A11 = ch.F + k0 + k0 * ld;
        if (fast)
            potf2_kernel<true><<<1, 512, PANEL_SMEM, sp>>>(A11, ld, nb, ch.info, (int)k0, pending);
        else
            potf2_kernel<false><<<1, 512, PANEL_SMEM, sp>>>(A11, ld, nb, ch.info, (int)k0, pending);
        ++g_potrf_launches;
        HYB_CUDA(cudaGetLastError());
        return 0;
    };
    auto solve_row = [&](int64_t k0) -> int {  // the block row right of it, on the panel stream
        const int nb = (int)((N - k0 < NB) ? (N - k0) : NB);
        const int64_t n2 = N - k0 - nb;
        if (n2 > 0) HYB_CHECK(launch_row_trsm(sp, ch.F + k0 + k0 * ld, ld, nb, ch.F + k0 + (k0 + nb) * ld, ld, n2));
        return 0;
    };
    // HYB_POTRF_TRACE=1: time stamps (CUDA events) of every outer block's phases on both streams, printed
    // as a table at the end (a measurement aid; adds event records only)
    static const bool trace = getenv("HYB_POTRF_TRACE") != nullptr;
    std::vector<cudaEvent_t> tev;
    auto stamp = [&](cudaStream_t s) {
        if (!trace) return;
        cudaEvent_t e;
        cudaEventCreate(&e);
        cudaEventRecord(e, s);
        tev.push_back(e);
    };
    stamp(st);
    HYB_CUDA(cudaEventRecord(ch.ev_head, st));
    // Outer block width by remaining size (measurement knobs; both off by default: with the near / far split
    // below the serial chain is the same whatever the width, and 512 everywhere measured best at N = 465 ... 16384):
    // HYB_POTRF_LATE = n: blocks of 128 once n or fewer columns remain.
    static const int64_t late = [] {
        const char* e = getenv("HYB_POTRF_LATE");
        return (int64_t)(e ? atoi(e) : 0);
    }();
    // HYB_POTRF_EARLY = n: blocks of 2 OB while more than n columns remain.
    static const int64_t early = [] {
        const char* e = getenv("HYB_POTRF_EARLY");
        return (int64_t)(e ? atoi(e) : 0);
    }();
    auto block_width = [&](int64_t remaining) -> int64_t {
        if (remaining <= late) return NB;
        if (early > 0 && remaining > early) return 2 * OB;
        return OB;
    };
    // Choreography.  Three streams: `sp` (high priority) carries the serial chain  panel kernel -> row solve
    // per 128 columns; `su` (high priority) the rank-128 "near" updates; `st` the rank-ob updates of everything
    // further away.  Outer blocks b = [K0, K1), next block [K1, K2), the one after [K2, K3):
    //  * near(j0): right after the row solve of panel j0, its 128 rows update rows [j0 + 128, K2) -- the rest of
    //    its own outer block and ALL rows of the next one -- in columns >= j0 + 256.  The tile column it leaves
    //    out holds the next diagonal block, which the next panel kernel updates itself (potf2_prologue), so that
    //    kernel starts at once and runs beside near(j0); only the next row solve waits for near(j0).
    //  * far(b), after the last row solve of block b: rank-ob update of rows [K2, K3) first (`ev_far[b % 3]`:
    //    near updates from block b + 1 reach into those rows and must follow it), then of the triangle below K3.
    // The chain never waits for a far update younger than two outer blocks, so the latency-bound chain and the
    // DMMA-bound far updates overlap instead of alternating, and the chain is the same whatever `ob` is.
    {
        std::vector<int64_t> edge{0};  // outer block boundaries
        while (edge.back() < N) {
            const int64_t w = block_width(N - edge.back());
            edge.push_back(edge.back() + w < N ? edge.back() + w : N);
        }
        const int64_t nblocks = (int64_t)edge.size() - 1;
        auto edge_at = [&](int64_t i) { return i < (int64_t)edge.size() ? edge[i] : N; };
        bool upd_pending = false;  // a near update the next row solve has to wait for
        HYB_CUDA(cudaStreamWaitEvent(sp, ch.ev_head, 0));  // the matrix and the cleared info word
        for (int64_t b = 0; b < nblocks; ++b) {
            const int64_t K0 = edge[b], K1 = edge[b + 1], K2 = edge_at(b + 2), K3 = edge_at(b + 3), ob = K1 - K0;
            if (b >= 2) HYB_CUDA(cudaStreamWaitEvent(sp, ch.ev_far[(b - 2) % 3], 0));  // far(b - 2) reached these rows
            stamp(sp);  // [1 + 5b] panels start
            bool su_has_far = (b == 0);  // `su` has waited for far(b - 1), which reached the next block's rows
            for (int64_t j0 = K0; j0 < K1; j0 += NB) {
                HYB_CHECK(factor_block(j0, j0 > 0 ? 1 : 0));
                if (upd_pending) HYB_CUDA(cudaStreamWaitEvent(sp, ch.ev_upd, 0));
                upd_pending = false;
                HYB_CHECK(solve_row(j0));
                const int64_t r1 = j0 + NB, nc = N - r1 - NB, M = K2 - r1;
                if (M > 0 && nc > 0) {
                    HYB_CUDA(cudaEventRecord(ch.ev_trsm, sp));
                    HYB_CUDA(cudaStreamWaitEvent(su, ch.ev_trsm, 0));
                    if (!su_has_far) HYB_CUDA(cudaStreamWaitEvent(su, ch.ev_far[(b - 1) % 3], 0));
                    su_has_far = true;
                    const double* U = ch.F + j0 + r1 * ld;
                    HYB_CHECK(hyb_gemm_tn(su, M, nc, NB, -1.0, U, ld, U + NB * ld, ld, 1.0, ch.F + r1 + (r1 + NB) * ld, ld, 0));
                    HYB_CUDA(cudaEventRecord(ch.ev_upd, su));
                    upd_pending = true;
                }
            }
            stamp(sp);  // [2 + 5b] panels end
            HYB_CUDA(cudaEventRecord(ch.ev_panel, sp));
            HYB_CUDA(cudaStreamWaitEvent(st, ch.ev_panel, 0));
            stamp(st);  // [3 + 5b] main stream has the panels (and its own previous far update)
            if (K2 < N) {
                const double* Ub = ch.F + K0 + K2 * ld;
                HYB_CHECK(hyb_gemm_tn(st, K3 - K2, N - K2, ob, -1.0, Ub, ld, Ub, ld, 1.0, ch.F + K2 + K2 * ld, ld, 0));
            }
            HYB_CUDA(cudaEventRecord(ch.ev_far[b % 3], st));
            stamp(st);  // [4 + 5b] rows of block b + 2 done
            if (K3 < N) {
                const double* Ub = ch.F + K0 + K3 * ld;
                HYB_CHECK(hyb_gemm_tn(st, N - K3, N - K3, ob, -1.0, Ub, ld, Ub, ld, 1.0, ch.F + K3 + K3 * ld, ld, HYB_GEMM_UPPER));
            }
            stamp(st);  // [5 + 5b] far update done
        }
    }
    if (trace) {
        cudaStreamSynchronize(sp);
        cudaStreamSynchronize(st);
        auto at = [&](size_t i) {
            float ms = 0;
            if (i < tev.size()) cudaEventElapsedTime(&ms, tev[0], tev[i]);
            return ms * 1e3f;
        };
        fprintf(stderr, "potrf trace N=%lld OB=%lld (us since start): block | panels start..end | main gets panels | head done | bulk done\n",
                (long long)N, (long long)OB);
        for (size_t b = 0; 1 + 5 * b + 2 < tev.size(); ++b)
            fprintf(stderr, "  %3zu | %8.0f .. %8.0f | %8.0f | %8.0f | %8.0f\n", b, at(1 + 5 * b), at(2 + 5 * b), at(3 + 5 * b),
                    at(4 + 5 * b), at(5 + 5 * b));
        for (cudaEvent_t e : tev) cudaEventDestroy(e);
    }
    if (info_out) {
        HYB_CUDA(cudaMemcpyAsync(info_out, ch.info, sizeof(int), cudaMemcpyDeviceToHost, st));
        HYB_CUDA(cudaStreamSynchronize(st));
    }
#ifdef HYB_PROFILE_POTF2
    hyb_potf2_dump_ts();
#endif
    return 0;
}

template <int R>
static int potrs_launch(cudaStream_t st, HybChol& ch, double* x, int64_t ldx) {
    static HybDevOnce attr_once;
    if (HybDevOnce::Guard g_(attr_once); g_.first) {
        HYB_CUDA(cudaFuncSetAttribute(trsv_fwd_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, TRSV_SMEM));
        HYB_CUDA(cudaFuncSetAttribute(trsv_bwd_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, TRSV_SMEM));
    }
    int N = (int)ch.N, nblk = (int)ch.nblk;
    const int num_sms = hyb_num_sms();
    int grid = nblk < num_sms ? nblk : num_sms;
    const double* U = ch.F;
    int64_t ld = ch.ld;
    int* flags = ch.flags;
    int epoch = ++ch.epoch;
    void* args[] = {(void*)&U, (void*)&ld, (void*)&x, (void*)&ldx, (void*)&N, (void*)&nblk, (void*)&flags, (void*)&epoch};
    HYB_CUDA(cudaLaunchCooperativeKernel((void*)trsv_fwd_kernel<R>, dim3(grid), dim3(TRSV_THREADS), args, TRSV_SMEM, st));
    int epoch2 = ++ch.epoch;
    void* args2[] = {(void*)&U, (void*)&ld, (void*)&x, (void*)&ldx, (void*)&N, (void*)&nblk, (void*)&flags, (void*)&epoch2};
    HYB_CUDA(cudaLaunchCooperativeKernel((void*)trsv_bwd_kernel<R>, dim3(grid), dim3(TRSV_THREADS), args2, TRSV_SMEM, st));
    g_potrf_launches += 2;
    return 0;
}

int hyb_potrs_multi(cudaStream_t st, HybChol& ch, double* X, int64_t ldx, int R) {
    if (ch.N == 0 || R <= 0) return 0;
    if (ch.kind == 1) return hyb_sytrs_rook(st, ch, X, ldx, R);
    switch (R) {
        case 1: return potrs_launch<1>(st, ch, X, ldx);
        case 2: return potrs_launch<2>(st, ch, X, ldx);
        case 3: return potrs_launch<3>(st, ch, X, ldx);
        case 4: return potrs_launch<4>(st, ch, X, ldx);
    }
    hyb_set_error("hyb_potrs_multi: %d right-hand sides (max %d)", R, HYB_MAX_RHS);
    return -3;
}

int hyb_potrs(cudaStream_t st, HybChol& ch, double* x) { return hyb_potrs_multi(st, ch, x, ch.N, 1); }

int64_t hyb_potrf_launch_count() { return g_potrf_launches; }

// copy the upper triangle of an N x N matrix (posdef_fact_copy!'s copyto!, dense.jl:232)
int hyb_copy_upper(cudaStream_t st, const double* src, int64_t lds_, double* dst, int64_t ldd, int64_t N) {
    if (N == 0) return 0;
    dim3 grid((unsigned)((N + 255) / 256), (unsigned)N);
    copy_upper_kernel<<<grid, 256, 0, st>>>(src, lds_, dst, ldd, N);
    HYB_CUDA(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------
// dense toolkit used by the large-PSD and WSOS cone engines
// ---------------------------------------------------------------------------------------
// B (N x nrhs) <- U^-T B : blocked forward substitution, DMMA updates (LAPACK dtrsm 'L','U','T')
int hyb_trsm_fwd(cudaStream_t st, const double* U, int64_t ld, int64_t N, double* B, int64_t ldb, int64_t nrhs) {
    if (N <= 0 || nrhs <= 0) return 0;
    HYB_CHECK(ensure_panel_attrs());
    const unsigned grid = (unsigned)((nrhs + TRSM_COLS - 1) / TRSM_COLS);
    for (int64_t k0 = 0; k0 < N; k0 += NB) {
        const int nb = (int)((N - k0 < NB) ? (N - k0) : NB);
        trsm_upper_T_kernel<64><<<grid, TRSM_THREADS, TRSM_SMEM, st>>>(U + k0 + k0 * ld, ld, nb, B + k0, ldb, nrhs);
        ++g_potrf_launches;
        HYB_CUDA(cudaGetLastError());
        const int64_t rest = N - k0 - nb;
        if (rest > 0)
            HYB_CHECK(hyb_gemm_tn(st, rest, nrhs, nb, -1.0, U + k0 + (k0 + nb) * ld, ld, B + k0, ldb, 1.0, B + k0 + nb, ldb, 0));
    }
    return 0;
}

// B (N x nrhs) <- U^-1 B : blocked back substitution; Lc = U' (lower, explicit zeros above the diagonal)
int hyb_trsm_bwd(cudaStream_t st, const double* U, const double* Lc, int64_t ld, int64_t N, double* B, int64_t ldb,
                 int64_t nrhs) {
    if (N <= 0 || nrhs <= 0) return 0;
    HYB_CHECK(ensure_panel_attrs());
    const unsigned grid = (unsigned)((nrhs + TRSM_COLS - 1) / TRSM_COLS);
    const int64_t last = ((N - 1) / NB) * NB;
    for (int64_t k0 = last; k0 >= 0; k0 -= NB) {
        const int nb = (int)((N - k0 < NB) ? (N - k0) : NB);
        trsm_upper_N_kernel<<<grid, TRSM_THREADS, TRSM_SMEM, st>>>(U + k0 + k0 * ld, ld, nb, B + k0, ldb, nrhs);
        ++g_potrf_launches;
        HYB_CUDA(cudaGetLastError());
        if (k0 > 0)  // B[0:k0] -= U[0:k0, k0:k0+nb] X_k = (Lc[k0:k0+nb, 0:k0])' X_k
            HYB_CHECK(hyb_gemm_tn(st, k0, nrhs, nb, -1.0, Lc + k0, ld, B + k0, ldb, 1.0, B, ldb, 0));
    }
    return 0;
}

int hyb_transpose(cudaStream_t st, const double* src, int64_t lds_, int64_t rows, int64_t cols, double* dst,
                  int64_t ldd, int upper_only) {
    if (rows <= 0 || cols <= 0) return 0;
    dim3 grid((unsigned)((rows + 31) / 32), (unsigned)((cols + 31) / 32));
    transpose_kernel<<<grid, dim3(32, 8), 0, st>>>(src, lds_, rows, cols, dst, ldd, upper_only);
    ++g_potrf_launches;
    HYB_CUDA(cudaGetLastError());
    return 0;
}

// Cholesky workspace around caller-owned storage (F is not freed by hyb_chol_free)
int hyb_chol_attach(HybChol& ch, double* F, int64_t ld, int64_t N) {
    HYB_CHECK(hyb_chol_alloc(ch, 0));
    cudaFree(ch.F);
    ch.F = F;
    ch.owns_F = false;
    ch.N = N;
    ch.ld = ld;
    ch.nblk = (N + NB - 1) / NB;
    cudaFree(ch.flags);
    HYB_CUDA(cudaMalloc(&ch.flags, sizeof(int) * (ch.nblk > 0 ? ch.nblk : 1)));
    HYB_CUDA(cudaMemset(ch.flags, 0, sizeof(int) * (ch.nblk > 0 ? ch.nblk : 1)));
    cudaFree(ch.work);
    HYB_CUDA(cudaMalloc(&ch.work, sizeof(double) * (N > 0 ? N : 1)));
    return 0;
}
