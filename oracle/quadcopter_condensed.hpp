// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_core.hpp header).
// The two quadcopter MPC problems in the CONDENSED formulation the CUDA library uses
// (csrc/quadcopter_condensed.cuh): same exact psi, grad psi, hess psi * v as the reference's CasADi
// functions (CasADiProblem.tpp:208-368), but organised for a GPU warp:
//   * per-step Jacobians AB_k = d f_d / d(x_k, u_k) of the RK4 step, built from 13 unit tangents,
//   * the adjoint recurrences (gradient, second-order adjoint) and the tangent recurrence as
//     matrix-vector recurrences over AB_k (the only sequential parts besides the rollout),
//   * everything that is local to a step (AB_k, node terms, the second-order source of the product)
//     evaluated independently per step.
// This file is a plain sequential restatement written independently of the device source: it shares
// only the sympy-generated scalar model formulas (csrc/generated/*.h) and the sin/cos restatement
// (csrc/pantr_hd.h) with the product.  It follows the SAME operation order as the device code, lane
// for lane, so results are compared bit for bit.  The reference formulation (forward-over-adjoint
// sweeps, what CasADi's AD generates) is oracle/quadcopter.hpp; tests compare the two to rounding.
//
// Math spec (reference, paths relative to /root/reference):
//   python/alpaqa_mpc_benchmarks/problems/quadcopter.py:5-121, realistic_quadcopter.py:5-127
//   python/alpaqa_mpc_benchmarks/formulations/ss.py:20-55  (rollout, cost sum, g = [c(x_0..x_N)])
//   alpaqa/python/alpaqa/casadi_generator/__init__.py:64-77,181-189 (psi, s psi, hess_psi_prod)
// Model arithmetic: parity unpinned (the reference holds no vectors for these functions).
#pragma once
#include "oracle_core.hpp"
#include "../pantr_cdc2023_experiments_b200/csrc/generated/quadcopter_gen.h"
#include "../pantr_cdc2023_experiments_b200/csrc/generated/realistic_quadcopter_gen.h"

namespace oracle {
namespace qcd {

constexpr int NX = 9, NU = 4, NC = 4, NZ = 13;
constexpr real_t Ts = 0.1, h = Ts, hh = Ts / 2, h6 = Ts / 6, h3 = 2 * (Ts / 6); // quadcopter.py:6,49-55
// quadcopter.py:85-111
constexpr real_t w_v = 1, w_p = 10, w_th = 1, w_u = 1e-4, w_om = 10;
constexpr real_t wN_v = 25 * w_v, wN_p = 25 * w_p, wN_th = 10 * w_th;
constexpr real_t p_f[3] = {0.25, 0.25, 0.5};

struct SimpleFns {
    static constexpr unsigned CONST_COLS = pantr::quadcopter_gen::FC_CONST_COLS;
    template <class... A> static void fc(A... a) { pantr::quadcopter_gen::fc(a...); }
    template <class... A> static void fc_jvp(A... a) { pantr::quadcopter_gen::fc_jvp(a...); }
    template <class... A> static void fc_vjp(A... a) { pantr::quadcopter_gen::fc_vjp(a...); }
    template <class... A> static void fc_hvjp(A... a) { pantr::quadcopter_gen::fc_hvjp(a...); }
    template <class... A> static void con(A... a) { pantr::quadcopter_gen::con(a...); }
    template <class... A> static void con_jvp(A... a) { pantr::quadcopter_gen::con_jvp(a...); }
    template <class... A> static void con_vjp(A... a) { pantr::quadcopter_gen::con_vjp(a...); }
    template <class... A> static void con_hvjp(A... a) { pantr::quadcopter_gen::con_hvjp(a...); }
};
struct RealisticFns {
    static constexpr unsigned CONST_COLS = pantr::realistic_quadcopter_gen::FC_CONST_COLS;
    template <class... A> static void fc(A... a) { pantr::realistic_quadcopter_gen::fc(a...); }
    template <class... A> static void fc_jvp(A... a) { pantr::realistic_quadcopter_gen::fc_jvp(a...); }
    template <class... A> static void fc_vjp(A... a) { pantr::realistic_quadcopter_gen::fc_vjp(a...); }
    template <class... A> static void fc_hvjp(A... a) { pantr::realistic_quadcopter_gen::fc_hvjp(a...); }
    template <class... A> static void con(A... a) { pantr::realistic_quadcopter_gen::con(a...); }
    template <class... A> static void con_jvp(A... a) { pantr::realistic_quadcopter_gen::con_jvp(a...); }
    template <class... A> static void con_vjp(A... a) { pantr::realistic_quadcopter_gen::con_vjp(a...); }
    template <class... A> static void con_hvjp(A... a) { pantr::realistic_quadcopter_gen::con_hvjp(a...); }
};

inline real_t sum_sq3(real_t a, real_t b, real_t c) { return (a * a + b * b) + c * c; }
/// half the diagonal of the cost Hessian wrt state entry i / input entry j
inline real_t wx(int i, bool term) { return i < 3 ? (term ? wN_p : w_p) : i < 6 ? (term ? wN_v : w_v) : (term ? wN_th : w_th); }
inline real_t wu(int j) { return j == 0 ? w_u : w_u + w_om; }
inline real_t D_lo(int i) { return i < 2 ? -1.5707963267948966 : i == 2 ? 0.85252452205950568 : -inf; } // quadcopter.py:70-73
inline real_t D_hi(int i) { return i < 2 ? +1.5707963267948966 : i == 2 ? inf : 0.0; }

} // namespace qcd

template <class F>
struct QuadcopterCondensed : Problem {
    static constexpr int NX = qcd::NX, NU = qcd::NU, NC = qcd::NC, NZ = qcd::NZ;
    int N;
    vec x0;
    // caches of the last evaluation point
    mutable vec traj, sc, AB, lam, yh, wact, dtraj, src;
    mutable bool hvp_ready = false;

    explicit QuadcopterCondensed(int N_)
        : N(N_), x0(NX, 0.0), traj((N_ + 1) * NX), sc((4 * N_ + 1) * 6), AB(size_t(N_) * NZ * NX), lam((N_ + 1) * NX),
          yh(NC * (N_ + 1)), wact(NC * (N_ + 1)), dtraj((N_ + 1) * NX), src(size_t(N_) * NZ) {
        n = NU * N; m = NC * (N + 1);
        C.lowerbound.resize(n); C.upperbound.resize(n);
        for (index_t i = 0; i < n; ++i) { // quadcopter.py:11-14,56-59 ; ss.py:28-29
            C.lowerbound[i] = (i & 3) == 0 ? 0.0 : -0.1;
            C.upperbound[i] = (i & 3) == 0 ? 9.81 * 5 : 0.1;
        }
        D.lowerbound.resize(m); D.upperbound.resize(m);
        for (index_t i = 0; i < m; ++i) { D.lowerbound[i] = qcd::D_lo(int(i & 3)); D.upperbound[i] = qcd::D_hi(int(i & 3)); }
        x0[0] = -0.2; x0[1] = -0.25; x0[2] = 0.5; // quadcopter.py:78-83
    }

    static void angles_sincos(const real_t *x, real_t *s6) {
        for (int a = 0; a < 3; ++a) pantr::pantr_sincos_fma(x[6 + a], &s6[2 * a], &s6[2 * a + 1]);
    }

    // ---- rollout (ss.py:24): states and the sin/cos of every RK4 stage point ----
    void rollout(const real_t *u) const {
        using namespace qcd;
        real_t x[NX], t[NX], S[NX], kk[NX];
        for (int i = 0; i < NX; ++i) traj[i] = x[i] = x0[i];
        for (int k = 0; k < N; ++k) {
            const real_t *uk = u + NU * k;
            real_t *s = &sc[24 * k];
            // the four stages as a running weighted sum (w = 1, 2, 2, 1) and the next stage point x + c k (c = h/2, h/2, h)
            for (int i = 0; i < NX; ++i) { t[i] = x[i]; S[i] = 0.0; }
            for (int st = 0; st < 4; ++st) {
                angles_sincos(t, s + 6 * st);
                F::fc(t, uk, s + 6 * st, kk);
                const real_t w = (st == 0 || st == 3) ? 1.0 : 2.0, c = st == 2 ? h : hh;
                for (int i = 0; i < NX; ++i) { S[i] = std::fma(w, kk[i], S[i]); t[i] = std::fma(c, kk[i], x[i]); }
            }
            for (int i = 0; i < NX; ++i) traj[(k + 1) * NX + i] = x[i] = std::fma(h6, S[i], x[i]);
        }
        angles_sincos(x, &sc[24 * N]);
    }
    // ---- cost (ss.py:32-36): node k < N contributes l(x_k, u_k), node N contributes l_N(x_N); the N+1 node
    // values are summed in warp order (node k belongs to partial sum k mod 32) ----
    real_t cost(const real_t *u) const {
        using namespace qcd;
        real_t p[32] = {};
        for (int k = 0; k <= N; ++k) {
            const real_t *x = &traj[k * NX];
            const bool term = k == N;
            real_t lx = ((term ? wN_v : w_v) * sum_sq3(x[3], x[4], x[5]) +
                         (term ? wN_p : w_p) * sum_sq3(x[0] - p_f[0], x[1] - p_f[1], x[2] - p_f[2])) +
                        (term ? wN_th : w_th) * sum_sq3(x[6], x[7], x[8]);
            real_t node = lx;
            if (!term) {
                const real_t *uk = u + NU * k;
                real_t lu = w_u * ((uk[0] * uk[0] + uk[1] * uk[1]) + (uk[2] * uk[2] + uk[3] * uk[3])) + w_om * sum_sq3(uk[1], uk[2], uk[3]);
                node = lu + lx;
            }
            p[k & 31] = p[k & 31] + node;
        }
        return butterfly_sum(p);
    }
    void constraints(real_t *g) const {
        for (int k = 0; k <= N; ++k) F::con(&traj[k * NX], &sc[24 * k], g + NC * k);
    }
    // ---- augmented-Lagrangian terms (casadi_generator/__init__.py:64-72), node-parallel ----
    real_t al_terms(const real_t *y, const real_t *S, real_t *yhat, real_t *wa) const {
        using namespace qcd;
        real_t p[32] = {};
        for (int k = 0; k <= N; ++k) {
            real_t c[NC], a = 0;
            F::con(&traj[k * NX], &sc[24 * k], c);
            for (int i = 0; i < NC; ++i) {
                const real_t Si = S[NC * k + i];
                const real_t zeta = c[i] + y[NC * k + i] / Si;
                const real_t zh = std::fmax(D_lo(i), std::fmin(zeta, D_hi(i)));
                const real_t d = zeta - zh, yhi = Si * d;
                a = std::fma(yhi, d, a);
                if (yhat) yhat[NC * k + i] = yhi;
                if (wa) wa[NC * k + i] = d != 0 ? Si : 0.0;
            }
            p[k & 31] = p[k & 31] + a;
        }
        return 0.5 * butterfly_sum(p);
    }
    // ---- step Jacobians: column c of AB_k = RK4 tangent of the unit vector e_c of (x, u) ----
    void step_tangent(int k, const real_t *u, const real_t *dx, const real_t *du, real_t *out) const {
        using namespace qcd;
        const real_t *x = &traj[k * NX], *uk = u + NU * k, *s = &sc[24 * k];
        real_t d[NX], dk1[NX], dk2[NX], dk3[NX], dk4[NX];
        F::fc_jvp(x, uk, s, dx, du, dk1);
        for (int i = 0; i < NX; ++i) d[i] = std::fma(hh, dk1[i], dx[i]);
        F::fc_jvp(x, uk, s + 6, d, du, dk2);
        for (int i = 0; i < NX; ++i) d[i] = std::fma(hh, dk2[i], dx[i]);
        F::fc_jvp(x, uk, s + 12, d, du, dk3);
        for (int i = 0; i < NX; ++i) d[i] = std::fma(h, dk3[i], dx[i]);
        F::fc_jvp(x, uk, s + 18, d, du, dk4);
        for (int i = 0; i < NX; ++i) out[i] = std::fma(h6, std::fma(2.0, dk3[i], std::fma(2.0, dk2[i], dk1[i])) + dk4[i], dx[i]);
    }
    void build_AB(const real_t *u) const {
        for (int k = 0; k < N; ++k)
            for (int c = 0; c < NZ; ++c) {
                real_t *col = &AB[(size_t(k) * NZ + c) * NX];
                if (k > 0 && ((F::CONST_COLS >> c) & 1)) { // same constant column at every step
                    for (int i = 0; i < NX; ++i) col[i] = AB[size_t(c) * NX + i];
                    continue;
                }
                real_t e[NZ] = {};
                e[c] = 1.0;
                step_tangent(k, u, e, e + NX, col);
            }
    }
    /// out[c] = sum_r AB_k[r][c] w[r]  (c < 13), one fma chain per column, r ascending
    void ABt_times(int k, const real_t *w, real_t *out) const {
        for (int c = 0; c < NZ; ++c) {
            const real_t *col = &AB[(size_t(k) * NZ + c) * NX];
            real_t a = col[0] * w[0];
            for (int r = 1; r < NX; ++r) a = std::fma(col[r], w[r], a);
            out[c] = a;
        }
    }
    /// out[r] = sum_c AB_k[r][c] z[c]  (r < 9), one fma chain per row, c ascending
    void AB_times(int k, const real_t *z, real_t *out) const {
        for (int r = 0; r < NX; ++r) {
            real_t a = AB[(size_t(k) * NZ) * NX + r] * z[0];
            for (int c = 1; c < NZ; ++c) a = std::fma(AB[(size_t(k) * NZ + c) * NX + r], z[c], a);
            out[r] = a;
        }
    }
    // ---- gradient of  s f + yh' g  wrt u by the adjoint recurrence over AB; stores lambda_1..N ----
    void adjoint(const real_t *u, real_t cost_scale, real_t *grad) const {
        using namespace qcd;
        real_t ax[NX], out[NZ];
        for (int k = N; k >= 1; --k) { // node sources, independent per node: q_k = s dl/dx + J_c' yh_k
            const real_t *x = &traj[k * NX];
            F::con_vjp(x, &sc[24 * k], &yh[NC * k], ax);
            for (int i = 0; i < NX; ++i) {
                const real_t dl = (2 * wx(i, k == N)) * (i < 3 ? x[i] - p_f[i] : x[i]);
                lam[k * NX + i] = std::fma(cost_scale, dl, ax[i]);
            }
        }
        for (int k = N - 1; k >= 0; --k) { // lambda_k = AB_k' lambda_{k+1} + q_k
            ABt_times(k, &lam[(k + 1) * NX], out);
            if (k > 0) for (int i = 0; i < NX; ++i) lam[k * NX + i] = out[i] + lam[k * NX + i];
            for (int j = 0; j < NU; ++j) grad[NU * k + j] = std::fma(cost_scale * (2 * wu(j)), u[NU * k + j], out[NX + j]);
        }
    }
    // ---- second-order source of step k: (d/dz [J_step' lambda_{k+1}]) (dx_k, v_k), z = (x_k, u_k) ----
    void step_second_order(int k, const real_t *u, const real_t *v, real_t *out) const {
        using namespace qcd;
        const real_t *x = &traj[k * NX], *uk = u + NU * k, *vk = v + NU * k, *s = &sc[24 * k];
        const real_t *dx = &dtraj[k * NX], *l = &lam[(k + 1) * NX];
        real_t d2[NX], d3[NX], d4[NX], dk[NX];
        F::fc_jvp(x, uk, s, dx, vk, dk);
        for (int i = 0; i < NX; ++i) d2[i] = std::fma(hh, dk[i], dx[i]);
        F::fc_jvp(x, uk, s + 6, d2, vk, dk);
        for (int i = 0; i < NX; ++i) d3[i] = std::fma(hh, dk[i], dx[i]);
        F::fc_jvp(x, uk, s + 12, d3, vk, dk);
        for (int i = 0; i < NX; ++i) d4[i] = std::fma(h, dk[i], dx[i]);
        // first-order adjoint stage weights of lambda_{k+1}
        real_t w4[NX], w3[NX], w2[NX], w1[NX], ax[NX], au[NU];
        for (int i = 0; i < NX; ++i) w4[i] = h6 * l[i];
        F::fc_vjp(x, uk, s + 18, w4, ax, au);
        for (int i = 0; i < NX; ++i) w3[i] = std::fma(h, ax[i], h3 * l[i]);
        F::fc_vjp(x, uk, s + 12, w3, ax, au);
        for (int i = 0; i < NX; ++i) w2[i] = std::fma(hh, ax[i], h3 * l[i]);
        F::fc_vjp(x, uk, s + 6, w2, ax, au);
        for (int i = 0; i < NX; ++i) w1[i] = std::fma(hh, ax[i], h6 * l[i]);
        // second-order chain, stage 4 -> 1 (zero incoming second-order adjoint: that part is AB' mu)
        real_t bx[NX], bu[NU], cx[NX], cu[NU], sx[NX], su[NU], dw[NX];
        F::fc_hvjp(x, uk, s + 18, w4, d4, vk, cx, cu);
        for (int i = 0; i < NX; ++i) { sx[i] = cx[i]; dw[i] = h * cx[i]; }
        for (int j = 0; j < NU; ++j) su[j] = cu[j];
        F::fc_vjp(x, uk, s + 12, dw, bx, bu);
        F::fc_hvjp(x, uk, s + 12, w3, d3, vk, cx, cu);
        for (int i = 0; i < NX; ++i) { bx[i] = bx[i] + cx[i]; sx[i] = sx[i] + bx[i]; dw[i] = hh * bx[i]; }
        for (int j = 0; j < NU; ++j) su[j] = su[j] + (bu[j] + cu[j]);
        F::fc_vjp(x, uk, s + 6, dw, bx, bu);
        F::fc_hvjp(x, uk, s + 6, w2, d2, vk, cx, cu);
        for (int i = 0; i < NX; ++i) { bx[i] = bx[i] + cx[i]; sx[i] = sx[i] + bx[i]; dw[i] = hh * bx[i]; }
        for (int j = 0; j < NU; ++j) su[j] = su[j] + (bu[j] + cu[j]);
        F::fc_vjp(x, uk, s, dw, bx, bu);
        F::fc_hvjp(x, uk, s, w1, dx, vk, cx, cu);
        for (int i = 0; i < NX; ++i) out[i] = sx[i] + (bx[i] + cx[i]);
        for (int j = 0; j < NU; ++j) out[NX + j] = su[j] + (bu[j] + cu[j]);
    }
    /// node term of the second-order adjoint: (s hess l + sum_i yh_i hess c_i + J_c' W J_c) dx at node k
    void node_second_order(int k, real_t cost_scale, real_t *out) const {
        using namespace qcd;
        const real_t *x = &traj[k * NX], *s = &sc[24 * k], *dx = &dtraj[k * NX];
        real_t bx[NX], dc[NC], t[NC], ax[NX];
        F::con_hvjp(x, s, &yh[NC * k], dx, bx);
        F::con_jvp(x, s, dx, dc);
        for (int i = 0; i < NC; ++i) t[i] = wact[NC * k + i] * dc[i];
        F::con_vjp(x, s, t, ax);
        for (int i = 0; i < NX; ++i) out[i] = std::fma(cost_scale * (2 * wx(i, k == N)), dx[i], bx[i]) + ax[i];
    }
    void hvp(const real_t *u, const real_t *v, real_t cost_scale, real_t *Hv) const {
        using namespace qcd;
        real_t z[NZ], out[NZ], mu[NX], nd[NX];
        for (int i = 0; i < NX; ++i) dtraj[i] = 0.0;
        for (int k = 0; k < N; ++k) { // tangent recurrence dx_{k+1} = AB_k (dx_k, v_k)
            for (int i = 0; i < NX; ++i) z[i] = dtraj[k * NX + i];
            for (int j = 0; j < NU; ++j) z[NX + j] = v[NU * k + j];
            AB_times(k, z, &dtraj[(k + 1) * NX]);
        }
        for (int k = 0; k < N; ++k) { // per-step sources, independent per step
            real_t *s = &src[size_t(k) * NZ];
            step_second_order(k, u, v, s);
            if (k > 0) {
                node_second_order(k, cost_scale, nd);
                for (int i = 0; i < NX; ++i) s[i] = s[i] + nd[i];
            }
            for (int j = 0; j < NU; ++j) s[NX + j] = std::fma(cost_scale * (2 * wu(j)), v[NU * k + j], s[NX + j]);
        }
        node_second_order(N, cost_scale, mu);
        for (int k = N - 1; k >= 0; --k) { // mu_k = AB_k' mu_{k+1} + src_k
            ABt_times(k, mu, out);
            for (int i = 0; i < NX; ++i) mu[i] = out[i] + src[size_t(k) * NZ + i];
            for (int j = 0; j < NU; ++j) Hv[NU * k + j] = out[NX + j] + src[size_t(k) * NZ + NX + j];
        }
    }

    // ---- Problem interface ----
    static void initial_guess(real_t *u, int N) { // quadcopter.py:112
        for (int k = 0; k < N; ++k) { u[4 * k] = 9.81; u[4 * k + 1] = u[4 * k + 2] = u[4 * k + 3] = 0; }
    }
    void plant_step(const real_t *x, const real_t *u, real_t *xn) const {
        QuadcopterCondensed one(1);
        std::copy(x, x + NX, one.x0.begin());
        one.rollout(u);
        std::copy(one.traj.begin() + NX, one.traj.begin() + 2 * NX, xn);
    }
    real_t eval_f_impl(const real_t *x) const override { hvp_ready = false; rollout(x); return cost(x); }
    void eval_g_impl(const real_t *x, real_t *g) const override { hvp_ready = false; rollout(x); constraints(g); }
    real_t eval_psi_impl(const real_t *x, const real_t *y, const real_t *S, real_t *yhat) const override {
        hvp_ready = false;
        rollout(x);
        real_t f = cost(x);
        return f + al_terms(y, S, yhat, nullptr);
    }
    real_t eval_psi_grad_psi_impl(const real_t *x, const real_t *y, const real_t *S, real_t *grad) const override {
        rollout(x);
        real_t f = cost(x);
        real_t psi = f + al_terms(y, S, yh.data(), wact.data());
        build_AB(x);
        adjoint(x, 1.0, grad);
        hvp_ready = true;
        return psi;
    }
    void eval_grad_L_impl(const real_t *, const real_t *, real_t *) const override {} // dead on the PANTR path
    void eval_hess_psi_prod_impl(const real_t *x, const real_t *y, const real_t *S, real_t scale, const real_t *v,
                                 real_t *Hv) const override {
        if (!hvp_ready || scale != 1) { // s psi = s f + 1/2 yh' d : the multipliers of the cost part change
            rollout(x);
            (void)al_terms(y, S, yh.data(), wact.data());
            build_AB(x);
            adjoint(x, scale, Hv);
            hvp_ready = scale == 1;
        }
        hvp(x, v, scale, Hv);
    }
    // ---- problem plugin face ----
    void set_multipliers(const real_t *y) const {
        for (index_t i = 0; i < m; ++i) { yh[i] = y ? y[i] : 0.0; wact[i] = 0.0; }
    }
    void grad_L_live(const real_t *x, const real_t *y, real_t *g) const override { (void)face_grad_L(x, y, g); }
    real_t face_grad_L(const real_t *x, const real_t *y, real_t *grad) const override {
        hvp_ready = false;
        rollout(x);
        real_t f = cost(x);
        set_multipliers(y);
        build_AB(x);
        adjoint(x, 1.0, grad);
        return f;
    }
    void face_grad_g_prod(const real_t *x, const real_t *y, real_t *out) const override {
        hvp_ready = false;
        rollout(x);
        set_multipliers(y);
        build_AB(x);
        adjoint(x, 0.0, out);
    }
    void face_hess_L_prod(const real_t *x, const real_t *y, real_t scale, const real_t *v, real_t *Hv) const override {
        hvp_ready = false;
        rollout(x);
        set_multipliers(y);
        build_AB(x);
        adjoint(x, scale, Hv);
        hvp(x, v, scale, Hv);
    }
};

using SimpleQuadcopterCondensed    = QuadcopterCondensed<qcd::SimpleFns>;
using RealisticQuadcopterCondensed = QuadcopterCondensed<qcd::RealisticFns>;

} // namespace oracle
