// Batched problem-function face: one launch evaluates one member of the reference's problem plugin
// table (alpaqa_problem_functions_t, alpaqa/src/interop/dl-api/include/alpaqa/dl/dl-problem.h:13-153;
// semantics = TypeErasedProblem / BoxConstrProblem / CasADiProblem) for B instances, one warp each.
#pragma once
#include "pantr_device.cuh"

namespace pantr {

enum ProblemFn : int {
    FN_F = 0, FN_GRAD_F, FN_F_GRAD_F, FN_G, FN_GRAD_G_PROD, FN_GRAD_L, FN_HESS_L_PROD, FN_HESS_PSI_PROD,
    FN_PROX_GRAD_STEP, FN_PROJ_DIFF_G, FN_PROJ_MULTIPLIERS, FN_INACTIVE_INDICES, FN_COUNT
};

struct ProblemArgs {
    int fn, N, n, m, nx;
    long long B;
    const double *x0, *x, *y, *Sigma, *v; // v doubles as grad_psi for the box functions, y as z
    double scalar;                        // scale (Hessian products) / gamma / M
    double *out_s, *out_n, *out_n2, *out_m;
    int *out_i, *out_cnt;
    double *ws;
    size_t ws_stride;
};

template <class Model>
__global__ void __launch_bounds__(32) problem_kernel(const __grid_constant__ ProblemArgs A) {
    double *ws = A.ws + size_t(blockIdx.x) * A.ws_stride;
    const int lane = lane_id(), n = A.n, m = A.m;
    for (long long b = blockIdx.x; b < A.B; b += gridDim.x) {
        Model mdl;
        mdl.init(A.N, A.x0 ? A.x0 + b * A.nx : nullptr, ws);
        const double *x = A.x ? A.x + b * n : nullptr, *y = A.y ? A.y + b * m : nullptr,
                     *S = A.Sigma ? A.Sigma + b * m : nullptr, *v = A.v ? A.v + b * n : nullptr;
        double *on = A.out_n ? A.out_n + b * n : nullptr, *on2 = A.out_n2 ? A.out_n2 + b * n : nullptr,
               *om = A.out_m ? A.out_m + b * m : nullptr;
        double f = 0;
        switch (A.fn) {
            case FN_F: f = mdl.eval_f(x); break;
            case FN_GRAD_F:
            case FN_F_GRAD_F: f = mdl.eval_grad_L(x, nullptr, on); break;
            case FN_G: mdl.eval_g(x, om); break;
            case FN_GRAD_G_PROD: mdl.eval_grad_g_prod(x, y, on); break;
            case FN_GRAD_L: f = mdl.eval_grad_L(x, y, on); break;
            // scale multiplies the cost only: s L = s f + y' g, s psi = s f + 1/2 y_hat' d
            // (casadi_generator/__init__.py:65,69); for m = 0 that is the whole product
            case FN_HESS_L_PROD:
                if constexpr (Model::NC == 0) {
                    mdl.hess_L_prod(x, y, v, on);
                    if (A.scalar != 1) for (int i = lane; i < n; i += 32) on[i] *= A.scalar;
                } else mdl.hess_L_prod(x, y, v, on, A.scalar);
                break;
            case FN_HESS_PSI_PROD:
                if constexpr (Model::NC == 0) {
                    mdl.hess_psi_prod(x, y, S, v, on);
                    if (A.scalar != 1) for (int i = lane; i < n; i += 32) on[i] *= A.scalar;
                } else mdl.hess_psi_prod(x, y, S, v, on, A.scalar);
                break;
            case FN_PROX_GRAD_STEP: // box-constr-problem.hpp:59-64,107-114 ; out_n = x_hat, out_n2 = p
                for (int i = lane; i < n; i += 32) {
                    double xi = x[i], s = -A.scalar * v[i];
                    double lo = Model::C_lb(i) - xi, hi = Model::C_ub(i) - xi;
                    double t = s < lo ? lo : s, pi = hi < t ? hi : t;
                    on2[i] = pi; on[i] = xi + pi;
                }
                break;
            case FN_PROJ_DIFF_G: // box-constr-problem.hpp:117, box.hpp:34-53 ; y = z
                for (int i = lane; i < m; i += 32)
                    om[i] = y[i] - fmin(fmax(y[i], Model::D_lb(i, A.N)), Model::D_ub(i, A.N));
                break;
            case FN_PROJ_MULTIPLIERS: // box-constr-problem.hpp:119-143 (penalty_alm_split = 0)
                for (int i = lane; i < m; i += 32) {
                    double lb = Model::D_lb(i, A.N) == -INFINITY ? 0.0 : -A.scalar;
                    double ub = Model::D_ub(i, A.N) == +INFINITY ? 0.0 : +A.scalar;
                    double yi = y[i];
                    yi = yi < lb ? lb : yi; // std::max(y, lb) then std::min(., ub)
                    om[i] = ub < yi ? ub : yi;
                }
                break;
            case FN_INACTIVE_INDICES: { // box-constr-problem.hpp:151-185 (no l1 term)
                const unsigned lt = (1u << lane) - 1u;
                int nJ = 0;
                for (int i0 = 0; i0 < n; i0 += 32) {
                    int i = i0 + lane;
                    bool in = false;
                    if (i < n) {
                        double x_fw = x[i] - A.scalar * v[i];
                        in = Model::C_lb(i) < x_fw && x_fw < Model::C_ub(i);
                    }
                    unsigned bal = __ballot_sync(FULL, in);
                    if (in) A.out_i[b * n + nJ + __popc(bal & lt)] = i;
                    nJ += __popc(bal);
                }
                if (lane == 0) A.out_cnt[b] = nJ;
                break;
            }
            default: break;
        }
        if (A.out_s && lane == 0) A.out_s[b] = f;
        __syncwarp();
    }
}

} // namespace pantr
