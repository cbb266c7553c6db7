// Dense cone engines: cones whose oracles are sequences of large dense operations
// (Cholesky, multi-RHS triangular solves, FP64 DMMA contractions) rather than one kernel.
#pragma once
#include "common.cuh"

struct HybDensePsd;   // PosSemidefTri with side > 32 (possemideftri.jl)
int hybd_psd_create(HybDensePsd** out, int side);
void hybd_psd_destroy(HybDensePsd* c);
// loads point = scal * primal, dual; feasibility, dual feasibility, grad, check_numerics, proxsqr.
// primal/dual/point/dual_out/grad are this cone's dim-vectors; flag/prox its item slots.
int hybd_psd_update(cudaStream_t st, HybDensePsd* c, const double* primal, const double* dual_in, double scal, double irtmu,
                    double* point, double* dual_out, double* grad, int* flag, double* prox);
// prod columns = op(H) arr columns; arr/prod point at this cone's first row, column strides lda/ldp
int hybd_psd_prod(cudaStream_t st, HybDensePsd* c, int op, const double* arr, int64_t lda, double* prod, int64_t ldp,
                  int64_t m);
int hybd_psd_dder3(cudaStream_t st, HybDensePsd* c, const double* dir, double* out);
int64_t hybd_launch_count();

struct HybDenseWsos;  // WSOSInterpNonnegative (wsosinterpnonnegative.jl), real case
int hybd_wsos_create(HybDenseWsos** out, int64_t U, int nPs, const double* const* Ps_host, const int64_t* Ls);
void hybd_wsos_destroy(HybDenseWsos* c);
int hybd_wsos_update(cudaStream_t st, HybDenseWsos* c, const double* primal, const double* dual_in, double scal,
                     double irtmu, double* point, double* dual_out, double* grad, int* flag, double* prox);
// op: HYB_OP_* or 6 = hess_prod_slow
int hybd_wsos_prod(cudaStream_t st, HybDenseWsos* c, int op, const double* arr, int64_t lda, double* prod, int64_t ldp,
                   int64_t m);
int hybd_wsos_dder3(cudaStream_t st, HybDenseWsos* c, const double* dir, double* out);
// 1 if chol(H) of the loaded point succeeded (sqrt oracles usable, Cones.jl:193-199); synchronizes
int hybd_wsos_sqrt_ok(cudaStream_t st, HybDenseWsos* c, int* ok);

struct HybDenseWsosPsd;  // WSOSInterpPosSemidefTri (wsosinterppossemideftri.jl)
int hybd_wsospsd_create(HybDenseWsosPsd** out, int R, int64_t U, int nPs, const double* const* Ps_host, const int64_t* Ls);
void hybd_wsospsd_destroy(HybDenseWsosPsd* c);
int hybd_wsospsd_update(cudaStream_t st, HybDenseWsosPsd* c, const double* primal, const double* dual_in, double scal,
                        double irtmu, double* point, double* dual_out, double* grad, int* flag, double* prox);
int hybd_wsospsd_prod(cudaStream_t st, HybDenseWsosPsd* c, int op, const double* arr, int64_t lda, double* prod,
                      int64_t ldp, int64_t m);
int hybd_wsospsd_dder3(cudaStream_t st, HybDenseWsosPsd* c, const double* dir, double* out);
int hybd_wsospsd_sqrt_ok(cudaStream_t st, HybDenseWsosPsd* c, int* ok);

struct HybDenseDnn;  // DoublyNonnegativeTri (doublynonnegativetri.jl) and PosSemidefTriSparse (dense implementation)
int hybd_dnn_create(HybDenseDnn** out, int side);
int hybd_spsd_create(HybDenseDnn** out, int side, int64_t dim, const int64_t* rows_cols);
void hybd_dnn_destroy(HybDenseDnn* c);
int hybd_dnn_update(cudaStream_t st, HybDenseDnn* c, const double* primal, const double* dual_in, double scal, double irtmu,
                    double* point, double* dual_out, double* grad, int* flag, double* prox);
int hybd_dnn_prod(cudaStream_t st, HybDenseDnn* c, int op, const double* arr, int64_t lda, double* prod, int64_t ldp,
                  int64_t m);
int hybd_dnn_dder3(cudaStream_t st, HybDenseDnn* c, const double* dir, double* out);
int hybd_dnn_sqrt_ok(cudaStream_t st, HybDenseDnn* c, int* ok);

struct HybDenseWsosEuc;  // WSOSInterpEpiNormEucl (wsosinterpepinormeucl.jl)
int hybd_wsoseuc_create(HybDenseWsosEuc** out, int R, int64_t U, int nPs, const double* const* Ps_host, const int64_t* Ls);
void hybd_wsoseuc_destroy(HybDenseWsosEuc* c);
int hybd_wsoseuc_update(cudaStream_t st, HybDenseWsosEuc* c, const double* primal, const double* dual_in, double scal,
                        double irtmu, double* point, double* dual_out, double* grad, int* flag, double* prox);
int hybd_wsoseuc_prod(cudaStream_t st, HybDenseWsosEuc* c, int op, const double* arr, int64_t lda, double* prod, int64_t ldp,
                      int64_t m);
int hybd_wsoseuc_dder3(cudaStream_t st, HybDenseWsosEuc* c, const double* dir, double* out);
int hybd_wsoseuc_sqrt_ok(cudaStream_t st, HybDenseWsosEuc* c, int* ok);

struct HybDenseWsosOne;  // WSOSInterpEpiNormOne (wsosinterpepinormone.jl)
int hybd_wsosone_create(HybDenseWsosOne** out, int R, int64_t U, int nPs, const double* const* Ps_host, const int64_t* Ls);
void hybd_wsosone_destroy(HybDenseWsosOne* c);
int hybd_wsosone_update(cudaStream_t st, HybDenseWsosOne* c, const double* primal, const double* dual_in, double scal,
                        double irtmu, double* point, double* dual_out, double* grad, int* flag, double* prox);
int hybd_wsosone_prod(cudaStream_t st, HybDenseWsosOne* c, int op, const double* arr, int64_t lda, double* prod, int64_t ldp,
                      int64_t m);
int hybd_wsosone_dder3(cudaStream_t st, HybDenseWsosOne* c, const double* dir, double* out);
int hybd_wsosone_sqrt_ok(cudaStream_t st, HybDenseWsosOne* c, int* ok);
