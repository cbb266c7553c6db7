// The library context (one per GPU) shared by api.cu and dirs.cu.
#pragma once
#include <cmath>
#include "../../include/hypatia_b200.h"
#include "common.cuh"
#include "cones.cuh"

extern hyb_counter g_misc_launches;
struct HybDirs;  // device-resident direction engine (dirs.cu)

// aligned staging copy of a run of rows of GQ2 / HGQ2 that starts on an odd row (api.cu: staged_rows)
struct HybStageRun {
    int64_t off = 0, rows = 0, ld = 0;
    bool constant = false;
    double* buf = nullptr;
};

// ----------------------------------------------------------------------------- context
struct HybSymIndef;  // symindef.cu

struct hyb_ctx {
    int device = 0;
    cudaStream_t st = nullptr;
    bool loaded = false;
    int64_t q = 0, N = 0, ldq = 0, ldS = 0;
    double* GQ2 = nullptr;
    double* HGQ2 = nullptr;
    double* S = nullptr;
    HybChol chol;
    HybCones cones;
    double* in_primal = nullptr;
    double* in_dual = nullptr;
    double *vq1 = nullptr, *vq2 = nullptr, *vq3 = nullptr, *vn1 = nullptr, *vn2 = nullptr;
    double* gemv_scratch = nullptr;
    double loaded_scal = 1.0;
    double updated_irtmu = NAN;
    bool point_loaded = false;
    bool fact_ok = false;
    // Cones.jl:193-199: a generic-Hessian cone whose Hessian factorization has not been asked for since the point
    // was loaded (true after hyb_cones_load_point, false once a neighbourhood check or an inverse-Hessian product
    // ran) takes the gemm branch of the Schur assembly when N < dim.  The engines here always have the factor;
    // this flag reproduces the reference's choice (iteration 0 of a solve).
    bool hess_fact_lazy = false;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev2 = nullptr, ev3 = nullptr, evA = nullptr, evB = nullptr;
    // statistics (hyb_stats): device milliseconds between the last H2D and the first D2H of every
    // compute call, and the bytes moved across PCIe by the calls
    double dev_ms = 0.0;
    int64_t h2d_bytes = 0, d2h_bytes = 0;
    double last_t[4] = {0, 0, 0, 0};  // apply-W, syrk, factor, (unused)
    bool in_region = false;
    // peer-memory Schur reduce (multi-GPU): root owns `peer_world` slots of ldS x N doubles; a rank's
    // syrk epilogue stores into `peer_slot` (its slot in the ROOT's memory, mapped through CUDA IPC)
    double* peer_slots = nullptr;   // root only: base of all slots
    double* peer_slot = nullptr;    // this rank's slot (local pointer on root, IPC-mapped elsewhere)
    void* peer_ipc_base = nullptr;  // non-root: pointer returned by cudaIpcOpenMemHandle
    int peer_world = 0, peer_rank = 0;
    // reduce-scatter form (multi.cu): tile column tj of this context's partial Schur matrix goes to owner tj % n,
    // into this source's compact slot there (own_bases: device array of the n slot bases; host copy for memsets)
    double** own_bases_dev = nullptr;
    std::vector<double*> own_bases_host;
    int64_t own_slot_elems = 0;
    HybDirs* dirs = nullptr;  // direction engine state (allocated by hyb_dirs_setup)
    HybSymIndef* symindef = nullptr;  // dense symmetric-indefinite system solver state (hyb_symindef_load)
    std::vector<HybStageRun> stage;
    // contexts that only assemble partial Schur matrices into another context's memory (multi.cu, rootworker.py)
    bool want_assemble_only = false;  // request: skip the Schur matrix and the Cholesky workspace at load
    bool assemble_only = false;
    bool borrowed = false;            // GQ2 / HGQ2 point into another context's arrays (not freed here)
};

static inline cudaError_t h2d(hyb_ctx* c, void* dst, const void* src, size_t bytes) {
    c->h2d_bytes += (int64_t)bytes;
    return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, c->st);
}
static inline cudaError_t d2h(hyb_ctx* c, void* dst, const void* src, size_t bytes) {
    c->d2h_bytes += (int64_t)bytes;
    return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, c->st);
}
// device-time region: KBEGIN after the uploads, KEND before the downloads, KSYNC after the final sync
#define KBEGIN(c) cudaEventRecord((c)->evA, (c)->st)
#define KEND(c) cudaEventRecord((c)->evB, (c)->st)
static inline void ksync(hyb_ctx* c) {
    float ms = 0;
    if (cudaEventElapsedTime(&ms, c->evA, c->evB) == cudaSuccess) c->dev_ms += ms;
}


int hyb_ensure_updated(hyb_ctx* c, double irtmu);
int hyb_model_load_impl(hyb_ctx* c, int64_t q, int64_t N, const double* GQ2, int64_t ldg, int64_t ncones,
                        const hyb_cone_desc* descs, hyb_ctx* borrow, int64_t borrow_row);
int hyb_schur_assemble_into(hyb_ctx* c, double* Sdst, double* t_assemble_ms);
void hyb_dirs_free(hyb_ctx* c);
void hyb_symindef_free(hyb_ctx* c);
