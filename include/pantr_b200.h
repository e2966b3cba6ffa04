/* pantr_b200 — C ABI of the B200-native batched PANTR+ALM solver.
 *
 * Drop-in boundary for the reference's solver call on the MPC hot path:
 *   ALMSolver<PANTRSolver<NewtonTRDirection<DefaultConfig>>>::operator()(problem, x, y)
 *   (alpaqa/src/alpaqa/include/alpaqa/outer/alm.hpp:157, called per MPC step from
 *   drivers/mpc-driver.cpp:158,199,216-222), batched over B independent instances
 *   (initial state x0, warm start u, multipliers y).  One CUDA warp-block owns one instance; the
 *   problem functions the reference loads from CasADi-generated shared objects
 *   (CasADiProblem.tpp:208-368) are device code inside the kernels.
 *
 * Plain pointers and sizes only.  All functions return 0 on success, a negative pantr_b200_error
 * otherwise (the reference throws C++ exceptions; pantr_b200_last_error() gives the message).
 * A handle is not thread-safe: one handle per host thread / GPU (the reference's solver and
 * problem objects are not thread-safe either: alpaqa/python/test/test_alm_thread.py:37-62).
 */
#ifndef PANTR_B200_H
#define PANTR_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum pantr_b200_error {
    PANTR_B200_OK = 0,
    PANTR_B200_INVALID_ARGUMENT = -1, /* std::invalid_argument in the reference */
    PANTR_B200_CUDA_ERROR = -2,
    PANTR_B200_NO_DEVICE = -3,        /* there is no CPU fallback */
    PANTR_B200_OUT_OF_RANGE = -4,     /* std::out_of_range (e.g. unsupported stop criterion) */
} pantr_b200_error;

/* problem ids: the three models of python/alpaqa_mpc_benchmarks/problems */
enum { PANTR_B200_HANGING_CHAIN = 0, PANTR_B200_QUADCOPTER = 1, PANTR_B200_REALISTIC_QUADCOPTER = 2 };

/* SolverStatus, alpaqa/inner/internal/solverstatus.hpp:11-20 (same integer values) */
enum {
    PANTR_B200_BUSY = 0, PANTR_B200_CONVERGED, PANTR_B200_MAX_TIME, PANTR_B200_MAX_ITER,
    PANTR_B200_NOT_FINITE, PANTR_B200_NO_PROGRESS, PANTR_B200_INTERRUPTED, PANTR_B200_EXCEPTION
};
/* PANOCStopCrit, alpaqa/inner/internal/panoc-stop-crit.hpp:7-102 (same integer values) */
enum {
    PANTR_B200_APPROX_KKT = 0, PANTR_B200_APPROX_KKT2, PANTR_B200_PROJ_GRAD_NORM,
    PANTR_B200_PROJ_GRAD_NORM2, PANTR_B200_PROJ_GRAD_UNIT_NORM, PANTR_B200_PROJ_GRAD_UNIT_NORM2,
    PANTR_B200_FPR_NORM, PANTR_B200_FPR_NORM2, PANTR_B200_IPOPT, PANTR_B200_LBFGSBPP
};

/* ALMParams, alpaqa/outer/alm.hpp:19-90 (printing has no device analogue).  max_time: seconds of DEVICE wall
 * clock (%globaltimer) an instance may spend in its solve before it returns PANTR_B200_MAX_TIME (alm.tpp:120-124,
 * 196-210); inf = no limit. */
typedef struct pantr_b200_alm_params {
    double tolerance, dual_tolerance, penalty_update_factor, penalty_update_factor_lower,
        min_penalty_update_factor, initial_penalty, initial_penalty_factor, initial_penalty_lower,
        initial_tolerance, initial_tolerance_increase, tolerance_update_factor, rho_increase, rho_max,
        rel_penalty_increase_threshold, max_multiplier, max_penalty, min_penalty;
    uint32_t max_iter, max_num_initial_retries, max_num_retries, max_total_num_retries;
    int32_t single_penalty_factor;
    double max_time;
} pantr_b200_alm_params;

/* LipschitzEstimateParams, alpaqa/inner/internal/lipschitz.hpp:8-20 */
typedef struct pantr_b200_lipschitz_params {
    double L_0, epsilon, delta, Lgamma_factor;
} pantr_b200_lipschitz_params;

/* PANTRParams, alpaqa/inner/pantr.hpp:22-91 */
typedef struct pantr_b200_pantr_params {
    pantr_b200_lipschitz_params Lipschitz;
    uint32_t max_iter;
    int32_t stop_crit;
    uint32_t max_no_progress;
    double L_min, L_max;
    double quadratic_upperbound_tolerance_factor, TR_tolerance_factor;
    double ratio_threshold_acceptable, ratio_threshold_good;
    double radius_factor_rejected, radius_factor_acceptable, radius_factor_good;
    double initial_radius, min_radius;
    int32_t compute_ratio_using_new_stepsize, update_direction_on_prox_step,
        recompute_last_prox_step_after_direction_reset, disable_acceleration,
        ratio_approx_fbe_quadratic_model;
    double max_time; /* seconds of device wall clock per inner solve (panoc-helpers.tpp:233-236); inf = no limit */
} pantr_b200_pantr_params;

/* SteihaugCGParams, alpaqa/accelerators/steihaugcg.hpp:9-15 */
typedef struct pantr_b200_steihaug_params {
    double tol_scale, tol_scale_root, tol_max, max_iter_factor;
} pantr_b200_steihaug_params;

/* NewtonTRDirectionParams, alpaqa/inner/directions/pantr/newton-tr.hpp:16-23 */
typedef struct pantr_b200_newton_tr_params {
    int32_t rescale_on_step_size_changes;
    double hessian_vec_factor;
    int32_t finite_diff; /* must be 0: the exact Hessian-vector product is the device path */
    double finite_diff_stepsize;
} pantr_b200_newton_tr_params;

/* Per-instance result: ALMSolver::Stats (alpaqa/outer/alm.hpp:104-149) with the accumulated
 * PANTRStats (alpaqa/inner/pantr.hpp:93-110, :242-258) and the EvalCounter entries that the path
 * touches (alpaqa/problem/problem-counters.hpp:10-58). */
typedef struct pantr_b200_stats {
    int32_t status;
    uint32_t outer_iterations, inner_iterations, inner_convergence_failures;
    uint32_t accelerated_step_rejected, direction_failures, stepsize_backtracks,
        direction_update_rejected;
    double eps, delta, norm_penalty, final_gamma, final_psi, final_h, final_fbe;
    uint32_t ev_proj_multipliers, ev_prox_grad_step, ev_inactive_indices_res_lna, ev_grad_L,
        ev_hess_psi_prod, ev_psi, ev_grad_psi, ev_psi_grad_psi, ev_f, ev_g;
    uint32_t cg_iterations, elapsed_us; /* elapsed_us: wall time (globaltimer) this instance occupied its warp / thread */
    uint64_t cycles; /* SM clock cycles over the same interval: cycles / elapsed_us = the SM clock in MHz under load */
} pantr_b200_stats;

typedef struct pantr_b200_solver pantr_b200_solver;

/* Defaults = the reference struct defaults (files cited above). */
void pantr_b200_default_params(pantr_b200_alm_params *alm, pantr_b200_pantr_params *pantr,
                               pantr_b200_steihaug_params *cg, pantr_b200_newton_tr_params *dir);

/* Replaces: solver construction in drivers/mpc-driver.cpp:47-73,104-115 + problem loading in
 * drivers/casadi-dll-wrapper.cpp:25-67 (problem "<name>_ss2p_<horizon>").  device = CUDA ordinal. */
int pantr_b200_create(int problem_id, int horizon, const pantr_b200_alm_params *alm,
                      const pantr_b200_pantr_params *pantr, const pantr_b200_steihaug_params *cg,
                      const pantr_b200_newton_tr_params *dir, int device, pantr_b200_solver **out);
int pantr_b200_destroy(pantr_b200_solver *s);

/* dims[0..4] = n, m, nx, nu, nc  (TypeErasedProblem::get_n/get_m; extra functions get_nx/nu/nc,
 * drivers/casadi-dll-wrapper.cpp:100-126) */
int pantr_b200_get_dims(const pantr_b200_solver *s, int64_t dims[5]);
/* "ALMSolver<PANTRSolver<NewtonTRDirection<EigenConfigd>>>" (alpaqa/outer/alm.hpp:163-165) */
const char *pantr_b200_get_name(const pantr_b200_solver *s);
/* Box C of the NLP: TypeErasedProblem::get_box_C; lb, ub of length n */
int pantr_b200_get_box_C(const pantr_b200_solver *s, double *lb, double *ub);
/* Box D of the NLP: TypeErasedProblem::get_box_D; lb, ub of length m */
int pantr_b200_get_box_D(const pantr_b200_solver *s, double *lb, double *ub);
/* Initial state and initial guess of the model (extra functions get_initial_state_ptr,
 * initial_guess; problems/problems/hanging_chain.cpp:152-165).  x0: nx, u0: n */
int pantr_b200_get_initial(const pantr_b200_solver *s, double *x0, double *u0);

/* Replaces B calls of ALMSolver::operator()(problem, x, y) (alm.tpp:19-226), HOST buffers:
 * x0 [B*nx] in, u [B*n] in/out, y [B*m] in/out (may be NULL when m == 0), stats [B] out.
 * Host<->device copies happen inside. */
int pantr_b200_solve_batch(pantr_b200_solver *s, int64_t B, const double *x0, double *u, double *y,
                           pantr_b200_stats *stats);
/* Same with DEVICE buffers, asynchronous on `stream` (a cudaStream_t).  NULL = the handle's own
 * non-blocking stream (note: the legacy default stream's handle is also 0; pass cudaStreamLegacy or
 * a created stream to order the solve after other work). */
int pantr_b200_solve_batch_device(pantr_b200_solver *s, int64_t B, const double *d_x0, double *d_u,
                                  double *d_y, pantr_b200_stats *d_stats, void *stream);

/* Batched problem-function evaluation (TypeErasedProblem::eval_psi_grad_psi, eval_psi and
 * eval_hess_psi_prod, type-erased-problem.hpp:31-95), HOST buffers; any output may be NULL.
 * u [B*n], y, Sigma [B*m] (NULL if m == 0), v [B*n] or NULL -> psi [B], grad [B*n], yhat [B*m],
 * Hv [B*n]. */
int pantr_b200_eval_batch(pantr_b200_solver *s, int64_t B, const double *x0, const double *u,
                          const double *y, const double *Sigma, const double *v, double *psi,
                          double *grad, double *yhat, double *Hv);

/* ---- Problem plugin face: the members of alpaqa_problem_functions_t
 * (alpaqa/src/interop/dl-api/include/alpaqa/dl/dl-problem.h:13-153) evaluated for B instances per call,
 * HOST buffers, instance-major ([B*n], [B*m]).  x0 [B*nx] is the NLP parameter (the initial state,
 * CasADiProblem::param); the model comes from the handle.  Semantics: CasADiProblem::eval_*
 * (alpaqa/src/interop/casadi/include/alpaqa/implementation/casadi/CasADiProblem.tpp:208-368) for the
 * smooth functions, BoxConstrProblem (alpaqa/problem/box-constr-problem.hpp:59-185) for the box ones.
 * With m == 0 the m-sized arguments may be NULL and eval_g / eval_proj_* are no-ops. ---- */
/* dl-problem.h:18-20 eval_f */
int pantr_b200_eval_f_batch(pantr_b200_solver *s, int64_t B, const double *x0, const double *x, double *f);
/* dl-problem.h:21-24 eval_grad_f */
int pantr_b200_eval_grad_f_batch(pantr_b200_solver *s, int64_t B, const double *x0, const double *x,
                                 double *grad_fx);
/* dl-problem.h:87-90 eval_f_grad_f */
int pantr_b200_eval_f_grad_f_batch(pantr_b200_solver *s, int64_t B, const double *x0, const double *x,
                                   double *f, double *grad_fx);
/* dl-problem.h:25-28 eval_g : gx [B*m] */
int pantr_b200_eval_g_batch(pantr_b200_solver *s, int64_t B, const double *x0, const double *x, double *gx);
/* dl-problem.h:29-33 eval_grad_g_prod : J_g(x)' y */
int pantr_b200_eval_grad_g_prod_batch(pantr_b200_solver *s, int64_t B, const double *x0, const double *x,
                                      const double *y, double *grad_gxy);
/* dl-problem.h:101-106 eval_grad_L : grad f + J_g' y */
int pantr_b200_eval_grad_L_batch(pantr_b200_solver *s, int64_t B, const double *x0, const double *x,
                                 const double *y, double *grad_L);
/* dl-problem.h:47-53 eval_hess_L_prod : scale * hess_xx L(x, y) v */
int pantr_b200_eval_hess_L_prod_batch(pantr_b200_solver *s, int64_t B, const double *x0, const double *x,
                                      const double *y, double scale, const double *v, double *Hv);
/* dl-problem.h:64-73 eval_hess_psi_prod : scale * hess_xx psi(x; y, Sigma) v  (D = the model's box) */
int pantr_b200_eval_hess_psi_prod_batch(pantr_b200_solver *s, int64_t B, const double *x0, const double *x,
                                        const double *y, const double *Sigma, double scale, const double *v,
                                        double *Hv);
/* dl-problem.h:107-114 eval_psi : psi [B], yhat [B*m] (may be NULL) */
int pantr_b200_eval_psi_batch(pantr_b200_solver *s, int64_t B, const double *x0, const double *x,
                              const double *y, const double *Sigma, double *yhat, double *psi);
/* dl-problem.h:115-124 eval_grad_psi */
int pantr_b200_eval_grad_psi_batch(pantr_b200_solver *s, int64_t B, const double *x0, const double *x,
                                   const double *y, const double *Sigma, double *grad_psi);
/* dl-problem.h:125-134 eval_psi_grad_psi */
int pantr_b200_eval_psi_grad_psi_batch(pantr_b200_solver *s, int64_t B, const double *x0, const double *x,
                                       const double *y, const double *Sigma, double *psi, double *grad_psi);
/* dl-problem.h:135-141, box-constr-problem.hpp:59-64,107-114 eval_prox_grad_step : p = clamp(-gamma grad_psi, C - x), x_hat = x + p
 * (h = 0: the benchmark problems have no l1 term) */
int pantr_b200_eval_prox_grad_step_batch(pantr_b200_solver *s, int64_t B, double gamma, const double *x,
                                         const double *grad_psi, double *x_hat, double *p);
/* box-constr-problem.hpp:117 eval_proj_diff_g : e = z - Pi_D(z) */
int pantr_b200_eval_proj_diff_g_batch(pantr_b200_solver *s, int64_t B, const double *z, double *e);
/* box-constr-problem.hpp:119-143 eval_proj_multipliers : y clamped in place to [-M or 0, +M or 0] */
int pantr_b200_eval_proj_multipliers_batch(pantr_b200_solver *s, int64_t B, double *y, double M);
/* box-constr-problem.hpp:151-185 eval_inactive_indices_res_lna : J [B*n] (first nJ[b] entries of row b,
 * ascending), nJ [B] */
int pantr_b200_eval_inactive_indices_res_lna_batch(pantr_b200_solver *s, int64_t B, double gamma,
                                                   const double *x, const double *grad_psi, int32_t *J,
                                                   int32_t *nJ);

/* Plant step x+ = f_d(x, u) for B instances ("simulate_dynamics" extra function,
 * problems/problems/formulation.hpp:258-261), HOST buffers: x [B*nx], u [B*nu] -> xn [B*nx]. */
int pantr_b200_simulate_dynamics_batch(pantr_b200_solver *s, int64_t B, const double *x,
                                       const double *u, double *xn);
int pantr_b200_simulate_dynamics_batch_device(pantr_b200_solver *s, int64_t B, const double *d_x,
                                              const double *d_u, int64_t u_stride, double *d_xn,
                                              void *stream);
/* Warm-start shift of drivers/mpc-driver.cpp:168-175 on DEVICE buffers: u[0:n-nu] = u[nu:n]
 * (last block duplicated), y[0:m-nc] = y[nc:m]. */
int pantr_b200_shift_warm_start_device(pantr_b200_solver *s, int64_t B, double *d_u, double *d_y,
                                       void *stream);

/* Scheduling hint for the next solves (warp-per-instance path): d_order is a DEVICE array of `length` int32 holding a
 * permutation of 0..length-1; resident blocks take instances in that order instead of by index.  In closed-loop MPC
 * the previous step's per-instance cost (stats[b].cycles) predicts the next one, and starting the expensive
 * instances first shortens the tail of the launch.  Results do not depend on the order.  NULL / 0 = by index.
 * The array is caller-owned and must stay valid while it is set.  A solve whose batch size differs from `length`
 * fails with PANTR_B200_INVALID_ARGUMENT; entries outside 0..B-1 are skipped by the kernel (that instance is then not
 * solved: its stats keep status Busy), so a non-permutation cannot read or write out of bounds.  The
 * thread-per-instance path ignores the hint. */
int pantr_b200_set_order_device(pantr_b200_solver *s, const int32_t *d_order, int64_t length);

/* How instances map onto the GPU for the next solves: 0 = automatic (default), 1 = one warp per instance
 * (hanging chain: lane = coordinate; quadcopters: condensed formulation, lanes = steps / matrix rows),
 * 2 = one thread per instance, forward-over-adjoint sweeps (quadcopter models only; round-1 path kept for A/B).
 * Automatic = 1.  A mode that does not apply to the model falls back to 1. */
int pantr_b200_set_exec_mode(pantr_b200_solver *s, int mode);

/* Own memory checks (the analogue of the reference's Debug-mode NaN poisoning of stale iterates, pantr.tpp:460-467, and
 * of its sanitizer builds, alpaqa/cmake/Sanitizers.cmake:1-15).  flags: 1 = every double of the per-block workspace is
 * set to NaN before each launch (a read of stale or uninitialised workspace then poisons the result: results must stay
 * bit-identical), 2 = the 32-double guard band behind every block's workspace is filled with a canary before each
 * launch; pantr_b200_check_guards counts the canaries that were overwritten since.  grid_limit > 0 caps the number of
 * resident blocks of the solve kernel (results must not depend on it: blocks share nothing but the instance queue). */
int pantr_b200_set_debug(pantr_b200_solver *s, int flags, int grid_limit);
int pantr_b200_check_guards(pantr_b200_solver *s, int64_t *violations);

/* Optional per-iteration trace of one instance of the next solve_batch call (debug / parity):
 * rows of 11 doubles {k, psi, fbe, gamma, eps, radius, rho, q_model, nJ, cg_iters, accepted}. */
int pantr_b200_set_trace(pantr_b200_solver *s, int64_t instance, int max_rows);
/* rows: room for `capacity_rows` rows; at most that many are copied.  n_rows: rows the solve produced. */
int pantr_b200_get_trace(pantr_b200_solver *s, double *rows, int capacity_rows, int *n_rows);

/* Number of kernels this handle has launched so far. */
int64_t pantr_b200_kernel_launches(const pantr_b200_solver *s);
/* Resident blocks the solve kernel uses (grid size), for reporting. */
int pantr_b200_grid_size(const pantr_b200_solver *s);
/* Measured FP64 FMA throughput of the device in TFLOP/s (independent DFMA chains, best of `reps`
 * launches): the denominator of the FP64 roofline bench.py reports. */
int pantr_b200_measure_fp64_peak(int device, int reps, double *tflops);
/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * psi [B] and grad psi [B*n] of B instances, like pantr_b200_eval_psi_grad_psi_batch, but ALLOCATION-FREE: no solver
 * handle, every buffer is a device pointer of the caller (x0 [B*nx], x [B*n], y and Sigma [B*m], NULL when m = 0), and
 * the kernel's per-block model workspace is a static device array of PANTR_B200_NOALLOC_SCRATCH_BYTES inside the
 * library (held by every GPU that loads it, until exit).  Calls cudaMalloc / cudaFree / no allocating routine.
 * Runs on `stream` (a cudaStream_t; NULL = the default stream) of the CURRENT device and returns without
 * synchronising.  Limits: 1 <= horizon <= PANTR_B200_NOALLOC_MAX_HORIZON, 0 <= B <= PANTR_B200_NOALLOC_MAX_BATCH,
 * else PANTR_B200_OUT_OF_RANGE.  All calls on a device share the array: calls that overlap on the GPU race --
 * serialise them (one stream).  The array keeps the previous call's contents; nothing in it needs clearing. */
#define PANTR_B200_NOALLOC_SCRATCH_BYTES (16u << 20)
#define PANTR_B200_NOALLOC_MAX_HORIZON 120
#define PANTR_B200_NOALLOC_MAX_BATCH (1 << 22)
int pantr_b200_eval_psi_grad_psi_noalloc_device(int problem_id, int horizon, int64_t B, const double *d_x0,
                                                const double *d_x, const double *d_y, const double *d_Sigma,
                                                double *d_psi, double *d_grad_psi, void *stream);
const char *pantr_b200_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* PANTR_B200_H */
