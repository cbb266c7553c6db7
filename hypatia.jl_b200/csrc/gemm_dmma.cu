// FP64 tensor-core contraction engine for sm_100a:  C = alpha * A' * B + beta * C.
//
// Replaces the dsyrk / dgemm call sites of the reference's hot path (SURVEY K1, K2,
// K5 trailing updates, K13-K15): src/linearalgebra/dense.jl:72-79 (outer_prod! ->
// BLAS.syrk!('U','T')), src/Solvers/systemsolvers/qrchol.jl:242,253.
//
// Design (B200): neither tcgen05 nor wgmma has an f64 kind; FP64 tensor work is the
// warp-level DMMA (mma.sync m8n8k4 f64 -> SASS DMMA.8x8x4).  Both operands are
// K-major (contraction index contiguous), so one TMA box {16 k, 128 rows} with the
// 128-byte swizzle feeds fragments with the minimum 2 shared-memory wavefronts per
// LDS.64.  One CTA = one 128x128 output tile, 8 consumer warps (64x32 each, 64
// accumulators per thread) + 1 TMA producer warp, 4-stage mbarrier ring of
// (16 KB A + 16 KB B).  Deterministic: no split-K atomics.
#include <cstdlib>

#include "common.cuh"
#include "tma.cuh"

namespace {

constexpr int BM = 128, BN = 128, BK = 16, STAGES = 4;
constexpr int TILE_DOUBLES = BM * BK;           // 2048
constexpr int TILE_BYTES = TILE_DOUBLES * 8;    // 16 KB
constexpr int NUM_CONSUMER_WARPS = 8;
constexpr int THREADS = (NUM_CONSUMER_WARPS + 4) * 32;  // 2 consumer warpgroups + 1 producer warpgroup
constexpr int SMEM_BYTES = STAGES * 2 * TILE_BYTES + 1024 /*align slack*/ + 128 /*barriers*/;

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

// Tile enumeration of the UPPER (syrk) form.  sb == 0: plain column-major walk of the triangle.  sb > 0: the
// triangle is cut into sb x sb super-blocks (walked column-major, tiles inside one column-major too), so the ~148
// CTAs that run at the same time cover a compact square of tiles and share 2 sb operand panels instead of one
// panel per tile row: with the contraction index in lock-step across CTAs that is what the L2 can actually reuse
// (a K x 128 panel is 16.7 MB at K = 16368; the L2 holds seven).  DRAM traffic per wave ~ (distinct panels) x panel.
__device__ __forceinline__ void upper_tile_of(int idx, int T, int sb, int& ti, int& tj) {
    if (sb <= 0) {
        tj = (int)((sqrt(8.0 * idx + 1.0) - 1.0) * 0.5);
        while (tj * (tj + 1) / 2 > idx) --tj;
        while ((tj + 1) * (tj + 2) / 2 <= idx) ++tj;
        ti = idx - tj * (tj + 1) / 2;
        return;
    }
    const int nsb = (T + sb - 1) / sb;
    for (int J = 0; J < nsb; ++J) {
        const int w = min(sb, T - J * sb);
        for (int I = 0; I <= J; ++I) {
            const int h = min(sb, T - I * sb);
            const int cnt = (I == J) ? w * (w + 1) / 2 : h * w;
            if (idx < cnt) {
                int li, lj;
                if (I == J) {
                    lj = (int)((sqrtf(8.0f * idx + 1.0f) - 1.0f) * 0.5f);
                    while (lj * (lj + 1) / 2 > idx) --lj;
                    while ((lj + 1) * (lj + 2) / 2 <= idx) ++lj;
                    li = idx - lj * (lj + 1) / 2;
                } else {
                    lj = idx / h;
                    li = idx - lj * h;
                }
                ti = I * sb + li;
                tj = J * sb + lj;
                return;
            }
            idx -= cnt;
        }
    }
    ti = tj = 0;  // not reached
}

// 12 warps = 3 per SM sub-partition (16K registers each) => 168 registers per thread at launch;
// the producer warpgroup then gives registers back (setmaxnreg.dec) and the two consumer
// warpgroups take them (setmaxnreg.inc 232), the Hopper/Blackwell warp-specialisation idiom.
__global__ void __launch_bounds__(THREADS, 1)
gemm_tn_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, int M, int N, int K,
               double alpha, double beta, double* __restrict__ C, int64_t ldc, int flags, int tiles_m, int a_step,
               int b_step, int64_t c_step, int n_main, int nsplit, double* __restrict__ ws, double* const* __restrict__ own,
               int nown) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    double* sA = (double*)base;
    double* sB = (double*)(base + STAGES * TILE_BYTES);
    uint64_t* full = (uint64_t*)(base + 2 * STAGES * TILE_BYTES);
    uint64_t* empty = full + STAGES;

    // Tail splitting: tiles [n_main, ...) -- the last, partial wave of CTAs -- are cut along K into
    // nsplit CTAs each, which park raw partial tiles in `ws`; gemm_tail_reduce_kernel then adds the
    // partials of a tile in a fixed order (deterministic) and applies alpha / beta.  Without it a
    // result of 14 waves + 8 tiles costs 15 waves.
    int tile_idx = blockIdx.x, split = 0;
    const bool partial = (int)blockIdx.x >= n_main;
    if (partial) {
        const int t = blockIdx.x - n_main;
        tile_idx = n_main + t / nsplit;
        split = t - (t / nsplit) * nsplit;
    }
    int ti, tj;
    if (flags & HYB_GEMM_UPPER) {
        upper_tile_of(tile_idx, (N + BN - 1) / BN, (flags >> 8) & 0xff, ti, tj);
    } else {
        ti = tile_idx % tiles_m;
        tj = tile_idx / tiles_m;
    }
    const int m0 = ti * BM, n0 = tj * BN;
    // batched use (blockIdx.y = batch): A advances by a_step along its outer (M) coordinate, B by
    // b_step along its inner (K) coordinate, C by c_step doubles
    const int a_off = blockIdx.y * a_step, b_off = blockIdx.y * b_step;
    C += (int64_t)blockIdx.y * c_step;
    // owner-scattered result (multi-GPU reduce-scatter): this tile column lives in owner tj % nown's memory at
    // compact column block tj / nown; shifting the base keeps the epilogue's addressing unchanged
    if (own) C = own[tj % nown] + ((int64_t)(tj / nown) - tj) * BN * ldc;
    const int nk_all = (K + BK - 1) / BK;
    int kt0 = 0, nk = nk_all;  // this CTA's k-steps: [kt0, kt0 + nk)
    if (partial) {
        const int per = (nk_all + nsplit - 1) / nsplit;
        kt0 = split * per;
        nk = nk_all - kt0 < per ? nk_all - kt0 : per;
        if (nk < 0) nk = 0;
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], NUM_CONSUMER_WARPS);
        }
        fence_barrier_init();
    }
    __syncthreads();

    if (warp >= NUM_CONSUMER_WARPS) {
        // ---------------- TMA producer warpgroup (one elected thread works) ----------------
        asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
        if (warp == NUM_CONSUMER_WARPS && lane == 0) {
            for (int kt = 0; kt < nk; ++kt) {
                const int s = kt % STAGES;
                if (kt >= STAGES) mbar_wait(&empty[s], ((kt / STAGES) - 1) & 1);
                mbar_arrive_expect_tx(&full[s], 2 * TILE_BYTES);
                tma_load_2d(sA + s * TILE_DOUBLES, &tmA, (kt0 + kt) * BK, m0 + a_off, &full[s]);
                tma_load_2d(sB + s * TILE_DOUBLES, &tmB, (kt0 + kt) * BK + b_off, n0, &full[s]);
            }
        }
        return;
    }

    // ---------------- DMMA consumers ----------------
    asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
    const int wm = warp >> 2, wn = warp & 3;
    const int g = lane >> 2, t = lane & 3;
    double acc[8][4][2];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    const int arow = (wm * 64 + g) * BK;  // + mt*8*BK
    const int brow = (wn * 32 + g) * BK;  // + nt*8*BK

    // The epilogue reads the C tile (beta != 0): ask the L2 for it now, a whole main loop ahead, so those
    // reads do not pay the DRAM latency once per round after the last DMMA (1024 lines of 128 bytes, 4 per thread).
    if (beta != 0.0 && !partial) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int id = threadIdx.x * 4 + u, col = id >> 3, row = (id & 7) * 16;
            if (n0 + col < N && m0 + row < M) {
                const double* p = C + (m0 + row) + (int64_t)(n0 + col) * ldc;
                asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
            }
        }
    }

    for (int kt = 0; kt < nk; ++kt) {
        const int s = kt % STAGES;
        mbar_wait(&full[s], (kt / STAGES) & 1);
        const double* tA = sA + s * TILE_DOUBLES + arow;
        const double* tB = sB + s * TILE_DOUBLES + brow;
#pragma unroll
        for (int kk = 0; kk < BK / 4; ++kk) {
            // 128B swizzle: 16-byte chunk index (k>>1) is XORed with (row & 7) == g
            const int off = ((((kk * 2) + (t >> 1)) ^ g) << 1) + (t & 1);
            double a[8], b[4];
#pragma unroll
            for (int mt = 0; mt < 8; ++mt) a[mt] = tA[mt * 8 * BK + off];
#pragma unroll
            for (int nt = 0; nt < 4; ++nt) b[nt] = tB[nt * 8 * BK + off];
#pragma unroll
            for (int mt = 0; mt < 8; ++mt)
#pragma unroll
                for (int nt = 0; nt < 4; ++nt) dmma884(acc[mt][nt][0], acc[mt][nt][1], a[mt], b[nt]);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
    }

    // ---------------- epilogue ----------------
    // The accumulators are parked in the (now idle) stage buffers as a 128 x 128 column-major
    // tile, then the 256 consumer threads update C with coalesced 16-byte accesses:
    // C = alpha * acc + beta * C.  Loads of a batch are all issued before its stores (a
    // load -> store -> load chain on possibly aliasing pointers serialises the round trips).
    const bool upper = flags & HYB_GEMM_UPPER;
    const bool use_c = (beta != 0.0);
    double* Cs = sA;  // 128 KB = STAGES * (A + B) buffers
    const int ctid = threadIdx.x;  // 0..255 (consumer threads only)
    asm volatile("bar.sync 1, 256;" ::: "memory");  // every warp is done reading the last stages
#pragma unroll
    for (int mt = 0; mt < 8; ++mt) {
        const int row = wm * 64 + mt * 8 + g;
#pragma unroll
        for (int nt = 0; nt < 4; ++nt)
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                const int col = wn * 32 + nt * 8 + 2 * t + i;
                Cs[col * BM + (row ^ (((col >> 1) & 3) << 3))] = acc[mt][nt][i];
            }
    }
    asm volatile("bar.sync 1, 256;" ::: "memory");
    if (partial) {  // raw partial tile, column-major 128 x 128, to this CTA's workspace slot
        double* wt = ws + (int64_t)(blockIdx.x - n_main) * (BM * BN);
#pragma unroll 4
        for (int it = 0; it < 32; ++it) {
            const int idx = it * 256 + ctid;
            const int col = idx >> 6, row = (idx & 63) * 2;
            *(double2*)(wt + col * BM + row) = *(const double2*)(Cs + col * BM + (row ^ (((col >> 1) & 3) << 3)));
        }
        return;
    }
    const bool vec_ok = (flags & HYB_GEMM_CVEC) != 0;
    // EPI row pairs per thread and round: the accumulators already sit in shared memory, so registers are free
    // to keep 16 loads of C in flight (two exposed round trips per tile instead of four)
    constexpr int EPI = 16;
#pragma unroll 1
    for (int round = 0; round < 32 / EPI; ++round) {
        double2 cv[EPI];
        // item = (column, row pair); consecutive threads take consecutive row pairs of a column
#pragma unroll
        for (int j = 0; j < EPI; ++j) {
            const int idx = (round * EPI + j) * 256 + ctid;
            const int col = idx >> 6, row = (idx & 63) * 2;
            const int grow = m0 + row, gcol = n0 + col;
            cv[j] = make_double2(0.0, 0.0);
            if (use_c && gcol < N) {
                const double* p = C + grow + (int64_t)gcol * ldc;
                const bool ok0 = grow < M && !(upper && gcol < grow);
                const bool ok1 = grow + 1 < M && !(upper && gcol < grow + 1);
                if (ok0 && ok1 && vec_ok) {
                    cv[j] = __ldcg((const double2*)p);
                } else {
                    if (ok0) cv[j].x = __ldcg(p);
                    if (ok1) cv[j].y = __ldcg(p + 1);
                }
            }
        }
#pragma unroll
        for (int j = 0; j < EPI; ++j) {
            const int idx = (round * EPI + j) * 256 + ctid;
            const int col = idx >> 6, row = (idx & 63) * 2;
            const int grow = m0 + row, gcol = n0 + col;
            if (gcol < N) {
                const double2 sv = *(const double2*)(Cs + col * BM + (row ^ (((col >> 1) & 3) << 3)));
                double* p = C + grow + (int64_t)gcol * ldc;
                const bool ok0 = grow < M && !(upper && gcol < grow);
                const bool ok1 = grow + 1 < M && !(upper && gcol < grow + 1);
                const double v0 = alpha * sv.x + beta * cv[j].x, v1 = alpha * sv.y + beta * cv[j].y;
                if (ok0 && ok1 && vec_ok) {
                    *(double2*)p = make_double2(v0, v1);
                } else {
                    if (ok0) p[0] = v0;
                    if (ok1) p[1] = v1;
                }
            }
        }
    }
}

// C tile = alpha * (sum of the nsplit partial tiles, in split order) + beta * C for the split tail tiles
__global__ void __launch_bounds__(256)
gemm_tail_reduce_kernel(int M, int N, double alpha, double beta, double* __restrict__ C, int64_t ldc, int flags,
                        int tiles_m, int n_main, int nsplit, const double* __restrict__ ws, double* const* __restrict__ own,
                        int nown) {
    const int tile_idx = n_main + blockIdx.x;
    int ti, tj;
    if (flags & HYB_GEMM_UPPER) {
        upper_tile_of(tile_idx, (N + BN - 1) / BN, (flags >> 8) & 0xff, ti, tj);
    } else {
        ti = tile_idx % tiles_m;
        tj = tile_idx / tiles_m;
    }
    const int m0 = ti * BM, n0 = tj * BN;
    const bool upper = flags & HYB_GEMM_UPPER;
    if (own) C = own[tj % nown] + ((int64_t)(tj / nown) - tj) * BN * ldc;
    const double* w0 = ws + (int64_t)blockIdx.x * nsplit * (BM * BN);
    for (int e = threadIdx.x; e < BM * BN; e += 256) {
        const int row = e & (BM - 1), col = e >> 7;
        const int grow = m0 + row, gcol = n0 + col;
        if (grow >= M || gcol >= N || (upper && gcol < grow)) continue;
        double s = 0.0;
        for (int k = 0; k < nsplit; ++k) s += w0[(int64_t)k * (BM * BN) + e];
        double* p = C + grow + (int64_t)gcol * ldc;
        *p = (beta != 0.0) ? alpha * s + beta * *p : alpha * s;
    }
}

hyb_counter g_gemm_launches{0};
struct TailWs { int dev; cudaStream_t st; double* p; };
std::vector<TailWs> g_tail_ws;
std::mutex g_tail_mutex;  // one workspace per (device, stream) that ever needed one (never shrinks)

}  // namespace

int64_t hyb_gemm_launch_count() { return g_gemm_launches; }

int hyb_gemm_tn(cudaStream_t st, int64_t M, int64_t N, int64_t K, double alpha, const double* A, int64_t lda,
                const double* B, int64_t ldb, double beta, double* C, int64_t ldc, int flags, const HybTileOwners* own) {
    if (M <= 0 || N <= 0) return 0;
    if (K <= 0) {
        hyb_set_error("hyb_gemm_tn: K must be positive");
        return -1;
    }
    if ((lda & 1) || (ldb & 1) || ((uintptr_t)A & 15) || ((uintptr_t)B & 15)) {
        hyb_set_error("hyb_gemm_tn: operands must be 16-byte aligned with even leading dimensions");
        return -1;
    }
    static HybDevOnce attr_once;
    if (HybDevOnce::Guard g_(attr_once); g_.first) {
        HYB_CUDA(cudaFuncSetAttribute(gemm_tn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    }
    CUtensorMap tmA, tmB;
    HYB_CHECK(hyb_make_tmap_2d_f64(&tmA, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda, BK, BM));
    HYB_CHECK(hyb_make_tmap_2d_f64(&tmB, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, BK, BN));
    const int tiles_m = (int)((M + BM - 1) / BM), tiles_n = (int)((N + BN - 1) / BN);
    int64_t ntiles;
    if (flags & HYB_GEMM_UPPER) {
        if (M != N) {
            hyb_set_error("hyb_gemm_tn: UPPER needs a square result");
            return -1;
        }
        ntiles = (int64_t)tiles_n * (tiles_n + 1) / 2;
    } else {
        ntiles = (int64_t)tiles_m * tiles_n;
    }
    double* const* own_bases = own ? own->bases : nullptr;
    const int nown = own ? own->n : 0;
    if (own_bases) {  // owner bases are cudaMalloc'd (256-byte aligned); the compact blocks keep ldc
        if ((ldc & 1) == 0) flags |= HYB_GEMM_CVEC;
    } else if (((uintptr_t)C & 15) == 0 && (ldc & 1) == 0) flags |= HYB_GEMM_CVEC;
    if (flags & HYB_GEMM_UPPER) {  // super-block rasterisation of the triangle (HYB_GEMM_SB, default 12; 0 = plain walk)
        static const int sb = [] {
            const char* e = getenv("HYB_GEMM_SB");
            const int v = e ? atoi(e) : 12;
            return (v < 0 || v > 255) ? 12 : v;
        }();
        flags |= sb << 8;
    }
    // split the tiles of a small last wave along K (long contractions only: the Schur syrk)
    int dev = 0;
    HYB_CUDA(cudaGetDevice(&dev));
    const int P = hyb_num_sms();
    int64_t n_main = ntiles;
    int nsplit = 1;
    double* ws = nullptr;
    const int64_t nk_all = (K + BK - 1) / BK;
    const int64_t r = ntiles % P;
    static const bool no_split = getenv("HYB_GEMM_NO_SPLIT") != nullptr;
    // ... and when the whole result has fewer tiles than half the SMs but the contraction is long (a small Schur
    // matrix under very many cone rows: cfg4-like models), every tile is split the same way
    const bool few_tiles = 2 * ntiles <= P && nk_all >= 64;
    if (!no_split && ((ntiles > P && r > 0 && 2 * r <= P && nk_all >= 128) || few_tiles)) {
        const int64_t rr = few_tiles ? ntiles : r;
        nsplit = (int)(P / rr);
        if (nsplit > nk_all / 32) nsplit = (int)(nk_all / 32);
        if (nsplit >= 2) {
            n_main = ntiles - rr;
            std::lock_guard<std::mutex> lock(g_tail_mutex);
            for (auto& w : g_tail_ws)
                if (w.dev == dev && w.st == st) ws = w.p;
            if (!ws) {
                HYB_CUDA(cudaMalloc(&ws, sizeof(double) * BM * BN * (size_t)P));
                g_tail_ws.push_back({dev, st, ws});
            }
        } else {
            nsplit = 1;
        }
    }
    const int64_t grid = n_main + (ntiles - n_main) * nsplit;
    gemm_tn_kernel<<<(unsigned)grid, THREADS, SMEM_BYTES, st>>>(tmA, tmB, (int)M, (int)N, (int)K, alpha, beta, C, ldc, flags,
                                                              tiles_m, 0, 0, 0, (int)n_main, nsplit, ws, own_bases, nown);
    ++g_gemm_launches;
    HYB_CUDA(cudaGetLastError());
    if (n_main < ntiles) {
        gemm_tail_reduce_kernel<<<(unsigned)(ntiles - n_main), 256, 0, st>>>((int)M, (int)N, alpha, beta, C, ldc, flags, tiles_m,
                                                                            (int)n_main, nsplit, ws, own_bases, nown);
        ++g_gemm_launches;
        HYB_CUDA(cudaGetLastError());
    }
    return 0;
}

// Batched form for block-diagonal left factors: for b = 0..batch-1
//   C[b*step : b*step + M, :] = alpha * A_b' * B[b*step : b*step + K, :] + beta * C[...]
// with A_b = columns b*step .. b*step + M of the K x (batch*step) K-major array A (lda) and B, C tall
// matrices of batch*step rows (ldb, ldc).  Used with M = K = step: every diagonal block maps its
// own row range (the symmetric-Kronecker form of the PSD cone products, cones.cu).  Rows of B
// past the block's K rows are multiplied by zero-filled rows of A (A's K extent is exact), so B
// must be finite there.
int hyb_gemm_tn_blockdiag(cudaStream_t st, int64_t M, int64_t N, int64_t K, int64_t step, int64_t batch, double alpha,
                          const double* A, int64_t lda, const double* B, int64_t ldb, double beta, double* C,
                          int64_t ldc) {
    if (M <= 0 || N <= 0 || batch <= 0) return 0;
    if (K <= 0 || batch > 65535) {
        hyb_set_error("hyb_gemm_tn_blockdiag: bad K or batch");
        return -1;
    }
    if ((lda & 1) || (ldb & 1) || (step & 1) || ((uintptr_t)A & 15) || ((uintptr_t)B & 15)) {
        hyb_set_error("hyb_gemm_tn_blockdiag: operands must be 16-byte aligned with even leading dimensions and step");
        return -1;
    }
    static HybDevOnce attr_once;
    if (HybDevOnce::Guard g_(attr_once); g_.first) {
        HYB_CUDA(cudaFuncSetAttribute(gemm_tn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    }
    CUtensorMap tmA, tmB;
    HYB_CHECK(hyb_make_tmap_2d_f64(&tmA, A, (uint64_t)K, (uint64_t)(step * (batch - 1) + M), (uint64_t)lda, BK, BM));
    HYB_CHECK(hyb_make_tmap_2d_f64(&tmB, B, (uint64_t)(step * (batch - 1) + K), (uint64_t)N, (uint64_t)ldb, BK, BN));
    const int tiles_m = (int)((M + BM - 1) / BM), tiles_n = (int)((N + BN - 1) / BN);
    int flags = 0;
    if (((uintptr_t)C & 15) == 0 && (ldc & 1) == 0) flags |= HYB_GEMM_CVEC;
    dim3 grid((unsigned)(tiles_m * tiles_n), (unsigned)batch);
    gemm_tn_kernel<<<grid, THREADS, SMEM_BYTES, st>>>(tmA, tmB, (int)M, (int)N, (int)K, alpha, beta, C, ldc, flags, tiles_m,
                                                      (int)step, (int)step, step, (int)(tiles_m * tiles_n), 1, nullptr, nullptr, 0);
    ++g_gemm_launches;
    HYB_CUDA(cudaGetLastError());
    return 0;
}
