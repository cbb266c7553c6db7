// GPU-resident cone batch: device state and launchers (cones.cu).
#pragma once
#include "common.cuh"

enum HybConeType {
    HYB_NONNEGATIVE = 0,
    HYB_EPINORMEUCL = 1,
    HYB_EPIPERSQUARE = 2,
    HYB_POSSEMIDEFTRI = 3,
    HYB_WSOSINTERPNONNEGATIVE = 4,
    HYB_WSOSINTERPPOSSEMIDEFTRI = 5,
    HYB_EPINORMINF = 6,
    HYB_DOUBLYNONNEGATIVETRI = 7,
    HYB_POSSEMIDEFTRISPARSE = 8,
    HYB_WSOSINTERPEPINORMEUCL = 9,
    HYB_WSOSINTERPEPINORMONE = 10,
};

enum HybProdOp { HYB_OP_HESS = 0, HYB_OP_INV_HESS = 1, HYB_OP_SQRT_HESS = 2, HYB_OP_INV_SQRT_HESS = 3 };

// status bits per item
enum { HYB_ST_FEAS = 1, HYB_ST_DUAL_FEAS = 2, HYB_ST_NUMERICS = 4, HYB_ST_ALL = 7 };

// A block = run of consecutive cones of identical (type, dim, barrier side).
struct HybConeBlock {
    int type = 0;
    int dim = 0;
    int side = 0;          // PSD side
    int use_dual = 0;      // use_dual_barrier of the cones in the block
    int64_t count = 0;     // number of cones
    int64_t offset = 0;    // first row
    int64_t cone0 = 0;     // index of the first cone (into aux)
    int64_t mat_off = 0;   // PSD: offset (doubles) of this block's matrices in `mats`
    void* dense = nullptr; // dense engine object (PosSemidefTri with side > 32, WSOS cones)
    // WSOS parameters (host pointers, valid during hyb_model_load only)
    int64_t U = 0;
    int nPs = 0;
    const double* const* Ps = nullptr;
    const int64_t* Ls = nullptr;
};

struct HybCones {
    int64_t q = 0, K = 0;
    std::vector<HybConeBlock> blocks;
    double* point = nullptr;   // q : loaded primal point (already scaled)
    double* dual = nullptr;    // q : loaded dual point
    double* grad = nullptr;    // q
    double* aux = nullptr;     // K : dist (SOC / RSOC)
    double* mats = nullptr;    // PSD: per cone [X | U | Uinv | Xinv], each side*side column-major
    int* item_flag = nullptr;  // q : status bits per item (cone -> its first row; nonnegative -> every row)
    double* item_prox = nullptr;  // q
    int* red_flag = nullptr;      // 1 : AND over items
    double* red_prox = nullptr;   // 1 : max / sum over items
    double* tmp = nullptr;        // q scratch vector
    double* skron = nullptr;      // PSD blocks, wide products: per cone the dim x dim symmetric-Kronecker matrix
    int64_t skron_cap = 0;        // doubles allocated in skron
    std::vector<char> block_mask; // if non-empty: hybi_cones_prod touches only blocks with mask != 0
    std::vector<char> block_sqrt; // selector 7: per block 1 = (inv_)sqrt product, 0 = block Hessian product
};

int hybi_cones_alloc(HybCones& cs);
void hybi_cones_free(HybCones& cs);

// point <- scal * primal, dual <- dual_in (device vectors), then feas / dual feas / grad /
// check_numerics / proxsqr per item.
// ctl (device, optional): control block of a speculative line-search sequence (dirs.cu): ctl[0] != 0 -> every
// kernel returns at once; else ctl[1] replaces scal and irtmu.  Only for models without dense-engine cones.
int hybi_cones_update(cudaStream_t st, HybCones& cs, const double* primal, const double* dual_in, double scal,
                     double irtmu, int use_max_prox, const double* ctl = nullptr);
bool hybi_cones_all_batched(const HybCones& cs);
int hybi_cones_reduce_dev(cudaStream_t st, HybCones& cs, int use_max_prox, const double* ctl);
// reduce the per-item results: *all_flags = AND of status bits, *agg = max / sum of proxsqr (host outputs, synchronous)
int hybi_cones_reduce(cudaStream_t st, HybCones& cs, int use_max_prox, int* all_flags, double* agg);
// prod = op(H) * arr for every cone; arr/prod are q x m column-major device arrays (ld >= q).
// which: 0 hess, 1 inv_hess, 2 sqrt_hess, 3 inv_sqrt_hess, 4 block (inv_hess for dual-barrier cones, else hess),
//        5 block sqrt (inv_sqrt for dual-barrier cones, else sqrt), 6 hess_prod_slow,
//        7 Schur assembly: per block 5 or 4 as cs.block_sqrt says (qrchol.jl:228-252),
//        8 the inverse of 4 (hess for dual-barrier cones, else inv_hess): symindef.jl:255-259
int hybi_cones_prod(cudaStream_t st, HybCones& cs, int which, const double* arr, int64_t lda, double* prod, int64_t ldp,
                   int64_t m);
int hybi_cones_dder3(cudaStream_t st, HybCones& cs, const double* dir, double* out);
int64_t hybi_cones_launch_count();
