// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_core.hpp header).
// The two quadcopter MPC problems (single shooting, m = 4(N+1) general constraints) on top of the
// sequential sweeps of pantr_cdc2023_experiments_b200/csrc/quadcopter_sweeps.h — the same source the
// CUDA library compiles, so both sides follow the same floating point trajectory.
//
// Math spec (reference, paths relative to /root/reference):
//   python/alpaqa_mpc_benchmarks/problems/quadcopter.py:5-121, realistic_quadcopter.py:5-127
//   python/alpaqa_mpc_benchmarks/formulations/ss.py:20-55
//   alpaqa/python/alpaqa/casadi_generator/__init__.py:64-77,181-189 (psi, hess_psi_prod)
//   alpaqa/src/interop/casadi/include/alpaqa/implementation/casadi/CasADiProblem.tpp:208-368 (call sites)
// The reference's derivative code comes from CasADi's AD (absent here): model arithmetic parity
// unpinned; tests check the sweeps against finite differences and an independent numpy model.
#pragma once
#include "oracle_core.hpp"
#include "../pantr_cdc2023_experiments_b200/csrc/quadcopter_sweeps.h"

namespace oracle {

template <class G>
struct Quadcopter : Problem {
    static constexpr int NX = 9, NU = 4, NC = 4;
    int N;
    vec x0;
    mutable vec store;
    mutable pantr::quad::QuadWS w;
    mutable bool hvp_ready = false;

    explicit Quadcopter(int N_) : N(N_), x0(NX, 0.0), store(pantr::quad::QuadWS::doubles(N_)) {
        namespace q = pantr::quad;
        n = NU * N; m = NC * (N + 1);
        w.bind(store.data(), N);
        C.lowerbound.resize(n); C.upperbound.resize(n);
        for (index_t i = 0; i < n; ++i) { C.lowerbound[i] = q::C_lb((int)i); C.upperbound[i] = q::C_ub((int)i); }
        D.lowerbound.resize(m); D.upperbound.resize(m);
        for (index_t i = 0; i < m; ++i) { D.lowerbound[i] = q::D_lb((int)i); D.upperbound[i] = q::D_ub((int)i); }
        x0[0] = -0.2; x0[1] = -0.25; x0[2] = 0.5; // quadcopter.py:78-83
    }
    static void initial_guess(real_t *u, int N) { // quadcopter.py:112
        for (int k = 0; k < N; ++k) { u[4 * k] = 9.81; u[4 * k + 1] = u[4 * k + 2] = u[4 * k + 3] = 0; }
    }
    void plant_step(const real_t *x, const real_t *u, real_t *xn) const { pantr::quad::rk4<G>(x, u, xn, nullptr); }

    real_t eval_f_impl(const real_t *x) const override {
        hvp_ready = false;
        return pantr::quad::rollout<G>(N, x0.data(), x, w);
    }
    void eval_g_impl(const real_t *x, real_t *g) const override {
        hvp_ready = false;
        (void)pantr::quad::rollout<G>(N, x0.data(), x, w);
        pantr::quad::constraints<G>(N, w, g);
    }
    real_t eval_psi_impl(const real_t *x, const real_t *y, const real_t *S, real_t *yhat) const override {
        hvp_ready = false;
        real_t f = pantr::quad::rollout<G>(N, x0.data(), x, w);
        return f + pantr::quad::al_terms<G>(N, w, y, S, yhat, yhat != nullptr, static_cast<real_t *>(nullptr), false);
    }
    real_t eval_psi_grad_psi_impl(const real_t *x, const real_t *y, const real_t *S, real_t *grad) const override {
        real_t f   = pantr::quad::rollout<G>(N, x0.data(), x, w);
        real_t psi = f + pantr::quad::al_terms<G>(N, w, y, S, w.yh, true, w.wact, true);
        pantr::quad::adjoint<G>(N, x, w, grad);
        hvp_ready = true;
        return psi;
    }
    // numerically dead on the PANTR path (pantr.tpp:306-314): counted by the caller, not evaluated
    void eval_grad_L_impl(const real_t *, const real_t *, real_t *) const override {}
    void eval_hess_psi_prod_impl(const real_t *x, const real_t *y, const real_t *S, real_t, const real_t *v,
                                 real_t *Hv) const override {
        if (!hvp_ready) (void)eval_psi_grad_psi_impl(x, y, S, Hv);
        pantr::quad::hvp<G>(N, x, v, w, Hv);
    }
    // ---- problem plugin face ----
    void set_multipliers(const real_t *y) const {
        for (index_t i = 0; i < m; ++i) { w.yh[i] = y ? y[i] : 0.0; w.wact[i] = 0.0; }
    }
    void grad_L_live(const real_t *x, const real_t *y, real_t *g) const override { (void)face_grad_L(x, y, g); }
    real_t face_grad_L(const real_t *x, const real_t *y, real_t *grad) const override {
        real_t f = eval_f_impl(x);
        set_multipliers(y);
        pantr::quad::adjoint<G>(N, x, w, grad);
        return f;
    }
    void face_grad_g_prod(const real_t *x, const real_t *y, real_t *out) const override {
        (void)eval_f_impl(x);
        set_multipliers(y);
        pantr::quad::adjoint<G, false>(N, x, w, out);
    }
    void face_hess_L_prod(const real_t *x, const real_t *y, real_t scale, const real_t *v, real_t *Hv) const override {
        (void)face_grad_L(x, y, Hv);
        pantr::quad::hvp<G>(N, x, v, w, Hv);
        if (scale != 1) for (index_t i = 0; i < n; ++i) Hv[i] *= scale;
    }
};

using SimpleQuadcopter    = Quadcopter<pantr::quad::SimpleGen>;
using RealisticQuadcopter = Quadcopter<pantr::quad::RealisticGen>;

} // namespace oracle
