// Shared declarations for the hypatia_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <atomic>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

// launch counters are bumped from several host threads when one process drives several GPUs (multi.cu)
typedef std::atomic<int64_t> hyb_counter;

#define HYB_EPS 2.220446049250313e-16

// Error plumbing: every internal routine returns cudaError_t-like int (0 = ok) and
// records a message in a thread-local buffer that hyb_last_error() exposes.
void hyb_set_error(const char* fmt, ...);
const char* hyb_get_error();

#define HYB_CUDA(call)                                                                     \
    do {                                                                                   \
        cudaError_t _e = (call);                                                           \
        if (_e != cudaSuccess) {                                                           \
            hyb_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(_e)); \
            return -1;                                                                     \
        }                                                                                  \
    } while (0)

#define HYB_CHECK(call)                 \
    do {                                \
        int _r = (call);                \
        if (_r != 0) return _r;         \
    } while (0)

static inline int64_t hyb_round_up(int64_t x, int64_t m) { return (x + m - 1) / m * m; }

// Per-device one-time set-up.  cudaFuncSetAttribute and the SM count are properties of the CURRENT
// device, so a process that opens contexts on several GPUs must repeat them on each one:
//   static HybDevOnce once;  if (HybDevOnce::Guard g(once); g.first) { cudaFuncSetAttribute(...); }
#define HYB_MAX_DEVICES 64
struct HybDevOnce {
    std::mutex mu;
    bool done[HYB_MAX_DEVICES] = {};
    // `if (HybDevOnce::Guard g(once); g.first) { ... }`: the lock is held until the body has run, so a second
    // host thread (multi.cu issues per-device work from several) cannot launch before the attributes are set
    struct Guard {
        std::unique_lock<std::mutex> lock;
        bool first = true;
        explicit Guard(HybDevOnce& o) : lock(o.mu) {
            int d = 0;
            if (cudaGetDevice(&d) == cudaSuccess && d >= 0 && d < HYB_MAX_DEVICES) {
                first = !o.done[d];
                o.done[d] = true;
            }
        }
    };
};
// multiprocessor count of the current device (cached per device)
static inline int hyb_num_sms() {
    static int sms[HYB_MAX_DEVICES] = {};
    int d = 0;
    if (cudaGetDevice(&d) != cudaSuccess || d < 0 || d >= HYB_MAX_DEVICES) return 148;
    if (sms[d] == 0) cudaDeviceGetAttribute(&sms[d], cudaDevAttrMultiProcessorCount, d);
    return sms[d] > 0 ? sms[d] : 148;
}

// ---- dense engine (gemm_dmma.cu) ------------------------------------------------
// C(MxN) = alpha * A' * B + beta * C.   A: K x M, B: K x N, all column-major with
// even leading dimensions and 16-byte aligned bases (TMA requirement).
// flags: HYB_GEMM_UPPER -> only tiles/entries with col >= row are computed/stored
// (requires the A and B tile grids to coincide, i.e. SYRK-like use).
enum { HYB_GEMM_UPPER = 1, HYB_GEMM_CVEC = 2 /* internal: C is 16-byte aligned with even ldc */ };
// own (optional): the result is scattered by tile column over `n` owners (the reduce-scatter of the multi-GPU
// Schur assembly, multi.cu): tile column tj goes to bases[tj % n] at column block tj / n (compact storage,
// leading dimension ldc); `bases` is a DEVICE array of n device pointers (local or peer memory); C is ignored.
struct HybTileOwners {
    double* const* bases = nullptr;
    int n = 0;
};
int hyb_gemm_tn(cudaStream_t st, int64_t M, int64_t N, int64_t K, double alpha, const double* A, int64_t lda,
                const double* B, int64_t ldb, double beta, double* C, int64_t ldc, int flags,
                const HybTileOwners* own = nullptr);
int hyb_gemm_tn_blockdiag(cudaStream_t st, int64_t M, int64_t N, int64_t K, int64_t step, int64_t batch, double alpha,
                          const double* A, int64_t lda, const double* B, int64_t ldb, double beta, double* C,
                          int64_t ldc);
int64_t hyb_gemm_launch_count();
int64_t hyb_potrf_launch_count();

// ---- factorization / solves (potrf.cu) -------------------------------------------
struct HybChol {
    int64_t N = 0, ld = 0, nblk = 0;
    double* F = nullptr;      // N x N (ld), upper factor after hyb_potrf
    bool owns_F = true;
    int* flags = nullptr;     // nblk ready flags for the wavefront solves
    int* info = nullptr;      // device: index (1-based) of first non-positive pivot, 0 = success
    int epoch = 0;
    double* work = nullptr;   // N doubles
    cudaStream_t panel_stream = nullptr;  // look-ahead panels (high priority)
    cudaEvent_t ev_head = nullptr, ev_panel = nullptr;
    cudaStream_t upd_stream = nullptr;    // rank-128 updates of the rows a panel leaves behind (high priority)
    cudaEvent_t ev_trsm = nullptr, ev_upd = nullptr;
    cudaEvent_t ev_far[3] = {nullptr, nullptr, nullptr};  // far update of outer block b reached block b + 2's rows
    // symmetric-indefinite fallback (sytrf.cu): kind 0 = F holds the Cholesky factor, 1 = F holds the
    // rook-pivoted Bunch-Kaufman factorization U D U' with pivot vector ipiv (LAPACK dsytrf_rook layout)
    int kind = 0;
    int* ipiv = nullptr;       // device, N
    double* bk_ws = nullptr;   // device, 4 N
    void* bk_state = nullptr;  // device step state
};
int hyb_chol_alloc(HybChol& ch, int64_t N);
void hyb_chol_free(HybChol& ch);
// in-place upper Cholesky of ch.F (upper triangle referenced); returns LAPACK-style info via *info_out
int hyb_potrf(cudaStream_t st, HybChol& ch, int* info_out);
int hyb_chol_attach(HybChol& ch, double* F, int64_t ld, int64_t N);
// multi-RHS triangular solves with the upper factor U (ld): B <- U^-T B ; B <- U^-1 B (needs Lc = U')
int hyb_trsm_fwd(cudaStream_t st, const double* U, int64_t ld, int64_t N, double* B, int64_t ldb, int64_t nrhs);
int hyb_trsm_bwd(cudaStream_t st, const double* U, const double* Lc, int64_t ld, int64_t N, double* B, int64_t ldb,
                 int64_t nrhs);
// dst (cols x rows) = src' ; upper_only != 0 reads only src's upper triangle (zeros elsewhere)
int hyb_transpose(cudaStream_t st, const double* src, int64_t lds_, int64_t rows, int64_t cols, double* dst,
                  int64_t ldd, int upper_only);
// x <- (U'U)^-1 x for one device vector of length N
int hyb_potrs(cudaStream_t st, HybChol& ch, double* x);
// symmetric-indefinite fallback of posdef_fact_copy! (dense.jl:199-203,228-249): in-place rook-pivoted
// Bunch-Kaufman of ch.F (LAPACK dsytrf_rook semantics) and the matching solves (dsytrs_rook);
// hyb_potrs / hyb_potrs_multi dispatch on ch.kind, so callers of the solves need not care
int hyb_sytrf_rook(cudaStream_t st, HybChol& ch, int* info_out);
int hyb_sytrs_rook(cudaStream_t st, HybChol& ch, double* X, int64_t ldx, int R);
int64_t hyb_sytrf_launch_count();

// ---- gemv (gemv.cu) -----------------------------------------------------------------
// y = G x (q) ; scratch must hold HYB_GEMV_SPLITS * q doubles
#define HYB_GEMV_SPLITS 8
int hyb_gemv_n(cudaStream_t st, int64_t q, int64_t N, const double* G, int64_t ld, const double* x, double* y,
               double* scratch);
// the same products for M <= HYB_MAX_RHS vectors in one pass over G
#define HYB_MAX_RHS 4
int hyb_gemm_n_small(cudaStream_t st, int64_t q, int64_t N, const double* G, int64_t ld, const double* X, int64_t ldx,
                     int M, double* Y, int64_t ldy, double* scratch);
int hyb_gemm_t_small(cudaStream_t st, int64_t q, int64_t N, const double* G, int64_t ld, const double* Z, int64_t ldz,
                     int M, double* Y, int64_t ldy);
// x <- (U'U)^-1 x for R <= HYB_MAX_RHS device vectors (columns of X, leading dimension ldx)
int hyb_potrs_multi(cudaStream_t st, HybChol& ch, double* X, int64_t ldx, int R);
// y = G' z (N)
int hyb_gemv_t(cudaStream_t st, int64_t q, int64_t N, const double* G, int64_t ld, const double* z, double* y);
