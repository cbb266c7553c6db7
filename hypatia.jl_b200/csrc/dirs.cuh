// Device-resident direction engine / line search / stepper state shared by dirs.cu and step.cu.
#pragma once
#include "ctx.cuh"

struct HybVecs {  // one set of m <= HYB_MAX_RHS "points" on the device
    double *x = nullptr, *z = nullptr, *s = nullptr, *tk = nullptr;  // x: ldn x M, z/s: ldq x M, tk: 2 x M (tau, kap)
};

struct HybDirs {
    int64_t N = 0, q = 0, ldn = 0, ldq = 0;
    double *c = nullptr, *h = nullptr;
    unsigned char* dual_row = nullptr;  // 1 where the row belongs to a dual-barrier cone
    HybVecs rhs, dir, res, tmp, rrhs, rdir;  // rrhs / rdir: single-column scratch of the refinement rounds
    double *A = nullptr, *B = nullptr, *Vx = nullptr, *Vz = nullptr, *Gx = nullptr, *T = nullptr;  // work, M columns
    double *xc = nullptr, *zc = nullptr;  // sol_const (common.jl:191-209)
    double* scal = nullptr;     // device scalars: [0] = dot_obj(sol_const), [1 + j] = dot_obj(sol_sub_j), [8 + j] = norms
    double* gemm_scratch = nullptr;
    bool have_const = false;
    // device-resident line search (search.jl:72-162): the current point, the four directions of the
    // iteration (slots: 0 cent, 1 centadj, 2 pred, 3 predadj), per-cone data
    double *pz = nullptr, *ps = nullptr;       // point.z, point.s
    double* px = nullptr;                      // point.x (device-resident stepper, step.cu)
    double* kx = nullptr;                      // kept directions, x part: ldn x 4
    double *resx = nullptr, *resz = nullptr;   // x_residual / z_residual of calc_convergence_params (Solvers.jl:631-693)
    double* sout = nullptr;                    // 8 device scalars of the stepper kernels
    double *kz = nullptr, *ks = nullptr;       // kept directions: ldq x 4 each
    double* szk = nullptr;                     // K : <s_k, z_k> of the candidate
    double* nus = nullptr;                     // K : barrier parameters
    int64_t *cone_off = nullptr;               // K : first row of every cone
    int* cone_dim = nullptr;                   // K
    double* sred = nullptr;                    // [0] min szk, [1] sum szk ; int flag in sflag
    int* sflag = nullptr;
    double* sctl = nullptr;                    // 8 : control block of the speculative line search (hyb_search_alpha)
    int64_t K = 0;
    bool search_ready = false;
};


// core of hyb_get_directions on right-hand sides already in d.rhs (device); directions are left in d.dir
int hyb_directions_core(hyb_ctx* c, int m, bool refresh_const, double mu_tt, double res_norm_cutoff, int max_ref_steps,
                        double min_impr_tol, int* num_ref, double* res_norm);
