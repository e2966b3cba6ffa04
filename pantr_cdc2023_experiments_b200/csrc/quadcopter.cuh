// The two quadcopter MPC models as seen by the device solver (what CasADiProblem::eval_* computes
// for quadcopter_ss2p_<N> / realistic_quadcopter_ss2p_<N>; m = 4(N+1) general constraints, so
// these exercise the ALM outer loop on the device).  The sweeps themselves are the shared
// host/device code of quadcopter_sweeps.h; with nx = 9 there is no intra-instance parallelism
// worth a shuffle, so the 32 lanes of the owning warp execute them redundantly (uniform loads and
// stores) and only the O(n) solver vector work is lane-strided.
#pragma once
#include "quadcopter_sweeps.h"
#include "exec_policy.cuh"

namespace pantr {

template <class G, class E = WarpExec>
struct QuadcopterModel {
    using Vec = typename E::Vec;
    using CVec = typename E::CVec;
    static constexpr int NX = quad::NX, NU = quad::NU, NC = quad::NC;
    static constexpr bool HAS_ROLLOUT_TAG = false;
    static constexpr bool BATCHED_MAPS = false;
    static constexpr int MIN_BLOCKS = 8; // resident warps per SM the solve kernel is compiled for
    using Gen = G;
    __host__ __device__ static int n(int N) { return NU * N; }
    __host__ __device__ static int m(int N) { return NC * (N + 1); }
    __host__ __device__ static int traj_doubles(int N) { return quad::QuadWS::doubles(N); }
    __host__ __device__ static double C_lb(int i) { return quad::C_lb(i); } // quadcopter.py:56-59, ss.py:28-29
    __host__ __device__ static double C_ub(int i) { return quad::C_ub(i); }
    __host__ __device__ static double D_lb(int i, int) { return quad::D_lb(i); } // quadcopter.py:70-73, ss.py:41-42
    __host__ __device__ static double D_ub(int i, int) { return quad::D_ub(i); }
    static void initial_state(double *x) { // quadcopter.py:78-83
        for (int i = 0; i < NX; ++i) x[i] = 0;
        x[0] = -0.2; x[1] = -0.25; x[2] = 0.5;
    }
    static void initial_guess(double *u, int N) { // quadcopter.py:112
        for (int k = 0; k < N; ++k) { u[4 * k] = 9.81; u[4 * k + 1] = u[4 * k + 2] = u[4 * k + 3] = 0; }
    }

    int N;
    const double *x0;
    quad::QuadWST<Vec> w;
    bool hvp_ready;

    __device__ void init(int N_, const double *x0_, Vec ws) { N = N_; x0 = x0_; w.bind(ws, N_); hvp_ready = false; }
    __device__ __noinline__ double eval_f(CVec u) {
        E::sync();
        hvp_ready = false;
        double f = quad::rollout<G>(N, x0, u, w);
        E::sync();
        return f;
    }
    __device__ __noinline__ void eval_g(CVec u, Vec g) {
        (void)eval_f(u);
        quad::constraints<G>(N, w, g);
        E::sync();
    }
    __device__ __noinline__ double eval_psi(CVec x, CVec y, CVec S, Vec yhat) {
        double f   = eval_f(x);
        double psi = f + quad::al_terms<G>(N, w, y, S, yhat, bool(yhat), Vec{}, false);
        E::sync();
        return psi;
    }
    __device__ __noinline__ double eval_psi_grad_psi(CVec x, CVec y, CVec S, Vec grad) {
        double f   = eval_f(x);
        double psi = f + quad::al_terms<G>(N, w, y, S, w.yh, true, w.wact, true);
        quad::adjoint<G>(N, x, w, grad);
        hvp_ready = true;
        E::sync();
        return psi;
    }
    __device__ __noinline__ void hess_psi_prod(CVec x, CVec y, CVec S, CVec v, Vec Hv) {
        E::sync();
        if (!hvp_ready) (void)eval_psi_grad_psi(x, y, S, Hv);
        quad::hvp<G>(N, x, v, w, Hv);
        E::sync();
    }
    // ---- problem plugin face (alpaqa_problem_functions_t members that PANTR itself does not call) ----
    /// y_hat := y (zero if null), no active quadratic penalty: the adjoint / hvp sweeps then give grad L / hess L
    __device__ void set_multipliers(CVec y) {
        const int mm = m(N);
        for (int i = 0; i < mm; ++i) { w.yh[i] = y ? y[i] : 0.0; w.wact[i] = 0.0; }
    }
    /// grad L(x, y) = grad f + J_g' y (CasADiProblem::eval_grad_L, CasADiProblem.tpp:268-276); returns f
    __device__ __noinline__ double eval_grad_L(CVec x, CVec y, Vec grad) {
        double f = eval_f(x);
        set_multipliers(y);
        quad::adjoint<G>(N, x, w, grad);
        hvp_ready = false; // w.yh no longer holds the y_hat of a psi_grad_psi
        E::sync();
        return f;
    }
    /// J_g(x)' y (CasADiProblem::eval_grad_g_prod, CasADiProblem.tpp:226-233)
    __device__ __noinline__ void eval_grad_g_prod(CVec x, CVec y, Vec out) {
        (void)eval_f(x);
        set_multipliers(y);
        quad::adjoint<G, false>(N, x, w, out);
        E::sync();
    }
    /// hess_xx L(x, y) v (CasADiProblem::eval_hess_L_prod, CasADiProblem.tpp:300-313); Hv is also scratch
    __device__ __noinline__ void hess_L_prod(CVec x, CVec y, const CVec v, Vec Hv) {
        (void)eval_grad_L(x, y, Hv);
        quad::hvp<G>(N, x, v, w, Hv);
        hvp_ready = false;
        E::sync();
    }
    /// Plant step x+ = f_d(x, u) (problems/problems/formulation.hpp:258-261)
    __device__ static void plant_step(const double *x, const double *u, double *xn) {
        double xx[NX], uu[NU], o[NX];
        for (int i = 0; i < NX; ++i) xx[i] = x[i];
        for (int j = 0; j < NU; ++j) uu[j] = u[j];
        quad::rk4<G>(xx, uu, o, nullptr);
        for (int i = 0; i < NX; ++i) xn[i] = o[i];
    }
};

using Quadcopter          = QuadcopterModel<quad::SimpleGen>;
using RealisticQuadcopter = QuadcopterModel<quad::RealisticGen>;
/// thread-per-instance variants (32 instances per warp) for large batches
using QuadcopterT          = QuadcopterModel<quad::SimpleGen, ThreadExec>;
using RealisticQuadcopterT = QuadcopterModel<quad::RealisticGen, ThreadExec>;

} // namespace pantr
