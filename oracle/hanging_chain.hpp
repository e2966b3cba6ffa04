// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_core.hpp header).
// Hanging-chain MPC model, single-shooting transcription, exact first and second
// order derivatives by discrete adjoint / forward-over-adjoint.
//
// Math spec (reference, paths relative to /root/reference):
//   python/alpaqa_mpc_benchmarks/problems/hanging_chain.py:7-98   (dynamics, RK4, cost, box, x0)
//   python/alpaqa_mpc_benchmarks/formulations/ss.py:20-55          (rollout, cost sum, C = tile(U,N))
//   alpaqa/python/alpaqa/casadi_generator/__init__.py:64-77,181-189 (psi, hess_psi_prod definitions)
// With nc = nc_N = 0 the problem has m = 0: psi = f, grad_L = grad f, hess_psi_prod = hess f * v.
//
// The arithmetic is written per "lane" l = 3*i + d (link/ball i = 0..9, dimension d = 0..2), the
// same work decomposition the CUDA path uses, so both produce the same bits.  The reference's
// derivative code comes from CasADi's AD (not available here): parity for the model arithmetic is
// unpinned; tests check these sweeps against finite differences and a sympy derivation instead.
#pragma once
#include "oracle_core.hpp"
#include "../pantr_cdc2023_experiments_b200/csrc/hc_formulas.h" // scalar formulas with explicit FMAs, shared with the CUDA path

namespace oracle {
namespace hcf = pantr::hcf;

struct HangingChain : Problem {
    static constexpr int NL = 30, NVL = 27, NX = 57, NU = 3;
    // hanging_chain.py:9-18
    static constexpr real_t alpha = 25, beta = 1, gamma_u = 0.01, mass = 0.03, Dspring = 0.1,
                            Lspring = 0.033, v_max = 1, g_grav = 9.81, Ts = 0.1;
    const real_t inv_m = 1.0 / mass, DL = Dspring * Lspring;
    const real_t h = Ts, hh = Ts / 2, h6 = Ts / 6, h3 = 2 * (Ts / 6);
    const real_t two_alpha = 2 * alpha, two_beta = 2 * beta, two_gamma = 2 * gamma_u;
    int N;
    vec x0; // NLP parameter (initial state), mutable view like CasADiProblem::param
    mutable vec traj, dtraj; // (N+1) x 57 state / tangent trajectories

    explicit HangingChain(int N_) : N(N_), x0(NX, 0.0), traj((N_ + 1) * NX), dtraj((N_ + 1) * NX) {
        n = NU * N; m = 0;
        C.lowerbound.assign(n, -v_max); C.upperbound.assign(n, +v_max); // hanging_chain.py:82-84
        for (int i = 0; i < 10; ++i) x0[3 * i] = real_t(i + 1) / 10;   // hanging_chain.py:92-96
    }
    static real_t ref(int d) { return d == 0 ? 1.0 : 0.0; }  // x_end = e_1
    static real_t grav(int d) { return d == 2 ? -g_grav : 0.0; }

    struct Link { real_t dist[NL], c[NL], e[NL], iw2[NL]; };

    // continuous dynamics f_c: hanging_chain.py:47-56
    void fc(const real_t *X, const real_t *V, const real_t *u, real_t *kX, real_t *kV, Link &ld) const {
        real_t sq[NL], F[NL];
        for (int l = 0; l < NL; ++l) {
            ld.dist[l] = l >= 3 ? X[l] - X[l - 3] : X[l];
            sq[l]      = ld.dist[l] * ld.dist[l];
        }
        for (int i = 0; i < 10; ++i) {
            real_t n2 = (sq[3 * i] + sq[3 * i + 1]) + sq[3 * i + 2];
            real_t w = std::sqrt(n2), iw = 1.0 / w;
            real_t c = hcf::spring_c(iw), iw2 = iw * iw, e = hcf::spring_e(iw, iw2);
            for (int d = 0; d < 3; ++d) { ld.c[3 * i + d] = c; ld.e[3 * i + d] = e; ld.iw2[3 * i + d] = iw2; }
        }
        for (int l = 0; l < NL; ++l) F[l] = ld.dist[l] * ld.c[l];
        for (int l = 0; l < NL; ++l) kX[l] = l < NVL ? V[l] : u[l - NVL];
        for (int l = 0; l < NVL; ++l) kV[l] = hcf::accel(F[l + 3], F[l], grav(l % 3));
    }
    static real_t dot3(const real_t *p, int l) { int b = l - l % 3; return (p[b] + p[b + 1]) + p[b + 2]; }

    // (dfc/d(s,u))^T w
    void vjp(const Link &ld, const real_t *wX, const real_t *wV, real_t *aX, real_t *aV, real_t *au,
             real_t *Fbar_out = nullptr, real_t *dot_out = nullptr) const {
        real_t Fbar[NL], pr[NL], db[NL];
        for (int l = 0; l < NL; ++l) {
            Fbar[l] = ((l >= 3 ? wV[l - 3] : 0.0) - (l < NVL ? wV[l] : 0.0)) * inv_m;
            pr[l]   = ld.dist[l] * Fbar[l];
        }
        for (int l = 0; l < NL; ++l) {
            real_t dt = dot3(pr, l);
            db[l]     = hcf::force_lin(ld.c[l], Fbar[l], ld.e[l], dt, ld.dist[l]);
            if (dot_out) dot_out[l] = dt;
        }
        for (int l = 0; l < NL; ++l) aX[l] = l < NVL ? db[l] - db[l + 3] : db[l];
        for (int l = 0; l < NVL; ++l) aV[l] = wX[l];
        for (int d = 0; d < 3; ++d) au[d] = wX[NVL + d];
        if (Fbar_out) for (int l = 0; l < NL; ++l) Fbar_out[l] = Fbar[l];
    }
    // (dfc/d(s,u)) (ds,du); also returns ddist and dd = dist . ddist
    void jvp(const Link &ld, const real_t *dX, const real_t *dV, const real_t *du, real_t *dkX,
             real_t *dkV, real_t *ddist, real_t *dd) const {
        real_t pr[NL], dF[NL];
        for (int l = 0; l < NL; ++l) {
            ddist[l] = l >= 3 ? dX[l] - dX[l - 3] : dX[l];
            pr[l]    = ld.dist[l] * ddist[l];
        }
        for (int l = 0; l < NL; ++l) {
            dd[l] = dot3(pr, l);
            dF[l] = hcf::force_lin(ld.c[l], ddist[l], ld.e[l], dd[l], ld.dist[l]);
        }
        for (int l = 0; l < NL; ++l) dkX[l] = l < NVL ? dV[l] : du[l - NVL];
        for (int l = 0; l < NVL; ++l) dkV[l] = (dF[l + 3] - dF[l]) * inv_m;
    }
    // d/deps [ vjp(s + eps ds, w + eps dw) ] : second-order adjoint of one stage
    void hvjp(const Link &ld, const real_t *Fbar, const real_t *dotv, const real_t *ddist,
              const real_t *dd, const real_t *dwX, const real_t *dwV, real_t *bX, real_t *bV,
              real_t *bu) const {
        real_t dFbar[NL], p1[NL], p2[NL], ddb[NL];
        for (int l = 0; l < NL; ++l) {
            dFbar[l] = ((l >= 3 ? dwV[l - 3] : 0.0) - (l < NVL ? dwV[l] : 0.0)) * inv_m;
            p1[l]    = ddist[l] * Fbar[l];
            p2[l]    = ld.dist[l] * dFbar[l];
        }
        for (int l = 0; l < NL; ++l) {
            real_t df = dot3(p1, l), dot2 = dot3(p2, l);
            real_t dc = ld.e[l] * dd[l];
            real_t de = ((-3.0 * ld.e[l]) * ld.iw2[l]) * dd[l];
            real_t lin = hcf::force_lin(ld.c[l], dFbar[l], ld.e[l], dot2, ld.dist[l]);
            real_t sec = hcf::force_sec(dc, Fbar[l], de, dotv[l], ld.e[l], df, ld.dist[l], ddist[l]);
            ddb[l] = lin + sec;
        }
        for (int l = 0; l < NL; ++l) bX[l] = l < NVL ? ddb[l] - ddb[l + 3] : ddb[l];
        for (int l = 0; l < NVL; ++l) bV[l] = dwX[l];
        for (int d = 0; d < 3; ++d) bu[d] = dwX[NVL + d];
    }

    // s2 = s + a*k (lane-wise)
    static void axpy_state(real_t a, const real_t *kX, const real_t *kV, const real_t *X,
                           const real_t *V, real_t *oX, real_t *oV) {
        for (int l = 0; l < NL; ++l) oX[l] = hcf::axpy(a, kX[l], X[l]);
        for (int l = 0; l < NVL; ++l) oV[l] = hcf::axpy(a, kV[l], V[l]);
    }

    // One RK4 step of the plant / model: hanging_chain.py:58-66 ; also the driver's
    // simulate_dynamics (problems/problems/formulation.hpp:258-261)
    void rk4(const real_t *s, const real_t *u, real_t *sn) const {
        const real_t *X = s, *V = s + NL;
        real_t k1X[NL], k1V[NVL], k2X[NL], k2V[NVL], k3X[NL], k3V[NVL], k4X[NL], k4V[NVL];
        real_t tX[NL], tV[NVL];
        Link ld;
        fc(X, V, u, k1X, k1V, ld);
        axpy_state(hh, k1X, k1V, X, V, tX, tV); fc(tX, tV, u, k2X, k2V, ld);
        axpy_state(hh, k2X, k2V, X, V, tX, tV); fc(tX, tV, u, k3X, k3V, ld);
        axpy_state(h, k3X, k3V, X, V, tX, tV);  fc(tX, tV, u, k4X, k4V, ld);
        for (int l = 0; l < NL; ++l)
            sn[l] = hcf::rk4c(X[l], k1X[l], k2X[l], k3X[l], k4X[l]);
        for (int l = 0; l < NVL; ++l)
            sn[NL + l] = hcf::rk4c(V[l], k1V[l], k2V[l], k3V[l], k4V[l]);
    }

    // rollout + cost: ss.py:24-37
    real_t rollout(const real_t *u) const {
        real_t acc[32] = {};
        std::copy(x0.begin(), x0.end(), traj.begin());
        for (int k = 0; k < N; ++k) {
            const real_t *s = &traj[k * NX], *uk = u + NU * k;
            for (int l = 0; l < NVL; ++l) acc[l] = std::fma(beta, s[NL + l] * s[NL + l], acc[l]);
            for (int d = 0; d < 3; ++d) {
                real_t dx = s[NVL + d] - ref(d);
                acc[NVL + d] = acc[NVL + d] + std::fma(alpha, dx * dx, gamma_u * (uk[d] * uk[d]));
            }
            rk4(s, uk, &traj[(k + 1) * NX]);
        }
        const real_t *s = &traj[N * NX];
        for (int l = 0; l < NVL; ++l) acc[l] = std::fma(beta, s[NL + l] * s[NL + l], acc[l]);
        for (int d = 0; d < 3; ++d) {
            real_t dx = s[NVL + d] - ref(d);
            acc[NVL + d] = std::fma(alpha, dx * dx, acc[NVL + d]);
        }
        return butterfly_sum(acc);
    }

    // Base data of one RK4 step needed by the reverse sweeps
    struct StepData { Link ld[4]; };
    void stages(const real_t *s, const real_t *u, StepData &sd) const {
        const real_t *X = s, *V = s + NL;
        real_t kX[NL], kV[NVL], tX[NL], tV[NVL];
        fc(X, V, u, kX, kV, sd.ld[0]);
        axpy_state(hh, kX, kV, X, V, tX, tV); fc(tX, tV, u, kX, kV, sd.ld[1]);
        axpy_state(hh, kX, kV, X, V, tX, tV); fc(tX, tV, u, kX, kV, sd.ld[2]);
        axpy_state(h, kX, kV, X, V, tX, tV);  fc(tX, tV, u, kX, kV, sd.ld[3]);
    }
#include "hanging_chain_sweeps.inc"
};

} // namespace oracle
