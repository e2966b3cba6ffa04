// Single-shooting sweeps of the two quadcopter MPC problems: rollout + cost, augmented-Lagrangian
// terms, discrete adjoint (grad psi) and forward-over-adjoint (hess psi * v), written as plain
// sequential host/device code around the sympy-generated f_c / c derivatives (generated/*.h).
// Compiled into the CUDA library and into the CPU oracle: identical operation order on both sides.
//
// Replaces the CasADi-generated f, g, psi, psi_grad_psi, hess_psi_prod that CasADiProblem::eval_*
// calls (alpaqa/src/interop/casadi/include/alpaqa/implementation/casadi/CasADiProblem.tpp:208-368).
// Math spec (paths relative to /root/reference):
//   python/alpaqa_mpc_benchmarks/problems/quadcopter.py:5-121            f_c, RK4, cost, boxes, x0
//   python/alpaqa_mpc_benchmarks/problems/realistic_quadcopter.py:5-127  same with ARB / Omega^-1
//   python/alpaqa_mpc_benchmarks/formulations/ss.py:20-55                rollout, g = [c(x_0..x_N)]
//   alpaqa/python/alpaqa/casadi_generator/__init__.py:64-77,181-189      psi, hess_psi_prod
//
// Caches in the per-instance workspace (QuadWS): the state trajectory, sin/cos of the angles at
// every RK4 stage point (the derivative sweeps need the stage points only through them), the
// multipliers y_hat and the active-set weights of the last psi_grad_psi, the first-order adjoint
// stage weights (they depend on x only, so the Hessian-vector products of one PANTR iteration,
// ~30 per gradient, reuse them) and the tangent stage points of the running product.
#pragma once
#include "pantr_hd.h"
#include "generated/quadcopter_gen.h"
#include "generated/realistic_quadcopter_gen.h"

namespace pantr {
namespace quad {

constexpr int NX = 9, NU = 4, NC = 4;
constexpr double Ts = 0.1, h = Ts, hh = Ts / 2, h6 = Ts / 6, h3 = 2 * (Ts / 6);
// quadcopter.py:85-111 : l = 1e-4 u'u + 10 |omega|^2 + 1 |v|^2 + 10 |p - pf|^2 + 1 |theta|^2 ;
// l_N = 25 |v|^2 + 250 |p - pf|^2 + 10 |theta|^2
constexpr double r_u = 1e-4, r_om = 10, q_v = 1, q_p = 10, q_th = 1;
constexpr double qN_v = 25 * q_v, qN_p = 25 * q_p, qN_th = 10 * q_th;
// quadcopter.py:11-14,56-59,70-73
constexpr double at_max = 9.81 * 5, dtilt_max = 0.1;
constexpr double half_pi = 1.5707963267948966, cos_tilt_max = 0.85252452205950568; // cos(1.1 / 2)

PANTR_HD double pf(int i) { return i == 2 ? 0.5 : 0.25; }
PANTR_HD double C_lb(int i) { return (i & 3) == 0 ? 0.0 : -dtilt_max; }
PANTR_HD double C_ub(int i) { return (i & 3) == 0 ? at_max : dtilt_max; }
PANTR_HD double D_lb(int i) { return (i & 3) < 2 ? -half_pi : ((i & 3) == 2 ? cos_tilt_max : -INFINITY); }
PANTR_HD double D_ub(int i) { return (i & 3) < 2 ? half_pi : ((i & 3) == 2 ? INFINITY : 0.0); }

#define PANTR_QUAD_GEN(Name, ns)                                                                             \
    struct Name {                                                                                            \
        PANTR_HD static void fc(const double *x, const double *u, const double *sc, double *k) { ns::fc(x, u, sc, k); } \
        PANTR_HD static void fc_jvp(const double *x, const double *u, const double *sc, const double *dx,    \
                                    const double *du, double *dk) { ns::fc_jvp(x, u, sc, dx, du, dk); }      \
        PANTR_HD static void fc_vjp(const double *x, const double *u, const double *sc, const double *w,     \
                                    double *ax, double *au) { ns::fc_vjp(x, u, sc, w, ax, au); }             \
        PANTR_HD static void fc_hvjp(const double *x, const double *u, const double *sc, const double *w,    \
                                     const double *dx, const double *du, double *bx, double *bu) {           \
            ns::fc_hvjp(x, u, sc, w, dx, du, bx, bu);                                                        \
        }                                                                                                    \
        PANTR_HD static void con(const double *x, const double *sc, double *k) { ns::con(x, sc, k); }        \
        PANTR_HD static void con_jvp(const double *x, const double *sc, const double *dx, double *dk) {      \
            ns::con_jvp(x, sc, dx, dk);                                                                      \
        }                                                                                                    \
        PANTR_HD static void con_vjp(const double *x, const double *sc, const double *yh, double *ax) {      \
            ns::con_vjp(x, sc, yh, ax);                                                                      \
        }                                                                                                    \
        PANTR_HD static void con_hvjp(const double *x, const double *sc, const double *yh, const double *dx, \
                                      double *bx) { ns::con_hvjp(x, sc, yh, dx, bx); }                       \
    }
PANTR_QUAD_GEN(SimpleGen, quadcopter_gen);
PANTR_QUAD_GEN(RealisticGen, realistic_quadcopter_gen);
#undef PANTR_QUAD_GEN

/// Per-instance workspace (doubles).  Everything is written by every lane of the owning warp with
/// the same value (the sweeps are executed redundantly by the 32 lanes), so no lane ever reads a
/// location it has not written itself.
struct QuadWS {
    double *traj;  // (N+1) x 9   states of the last rollout
    double *sc;    // (4N+1) x 6  sin/cos of the angles at stage point j of step k (4k+j); node N last
    double *wst;   // 4N x 9      first-order adjoint weight of stage j of step k
    double *dtraj; // (N+1) x 9   tangent trajectory of the running Hessian-vector product
    double *dst;   // 4N x 9      tangent stage points
    double *yh;    // 4(N+1)      y_hat of the last psi_grad_psi
    double *wact;  // 4(N+1)      Sigma_i where zeta_i is outside D, else 0
    PANTR_HD static int doubles(int N) { return 2 * (N + 1) * NX + (4 * N + 1) * 6 + 2 * 4 * N * NX + 2 * NC * (N + 1); }
    PANTR_HD void bind(double *p, int N) {
        traj = p; p += (N + 1) * NX; sc = p; p += (4 * N + 1) * 6; wst = p; p += 4 * N * NX;
        dtraj = p; p += (N + 1) * NX; dst = p; p += 4 * N * NX; yh = p; p += NC * (N + 1); wact = p;
    }
};

PANTR_HD void sincos3(const double *x, double *sc) {
    pantr_sincos(x[6], &sc[0], &sc[1]);
    pantr_sincos(x[7], &sc[2], &sc[3]);
    pantr_sincos(x[8], &sc[4], &sc[5]);
}
PANTR_HD double sq3(double a, double b, double c) { return (a * a + b * b) + c * c; }
PANTR_HD double stage_cost(const double *x, const double *u) {
    double lu = r_u * ((u[0] * u[0] + u[1] * u[1]) + (u[2] * u[2] + u[3] * u[3])) + r_om * sq3(u[1], u[2], u[3]);
    double lx = (q_v * sq3(x[3], x[4], x[5]) + q_p * sq3(x[0] - pf(0), x[1] - pf(1), x[2] - pf(2))) +
                q_th * sq3(x[6], x[7], x[8]);
    return lu + lx;
}
PANTR_HD double term_cost(const double *x) {
    return (qN_v * sq3(x[3], x[4], x[5]) + qN_p * sq3(x[0] - pf(0), x[1] - pf(1), x[2] - pf(2))) +
           qN_th * sq3(x[6], x[7], x[8]);
}
/// weight of state entry i in the (diagonal) cost Hessian / 2
PANTR_HD double qx(int i, bool term) {
    return i < 3 ? (term ? qN_p : q_p) : (i < 6 ? (term ? qN_v : q_v) : (term ? qN_th : q_th));
}
PANTR_HD double ru(int j) { return j == 0 ? r_u : r_u + r_om; }
PANTR_HD double dcost_x(const double *x, int i, bool term) {
    return (2 * qx(i, term)) * (i < 3 ? x[i] - pf(i) : x[i]);
}

/// One RK4 step (quadcopter.py:49-55); sc_out (4 x 6) receives the stage sin/cos if non-null.
template <class G>
PANTR_HD void rk4(const double *x, const double *u, double *xn, double *sc_out) {
    double sc[6], k1[NX], k2[NX], k3[NX], k4[NX], t[NX];
    sincos3(x, sc);
    G::fc(x, u, sc, k1);
    if (sc_out) { PANTR_UNROLL for (int i = 0; i < 6; ++i) sc_out[i] = sc[i]; }
    PANTR_UNROLL for (int i = 0; i < NX; ++i) t[i] = x[i] + hh * k1[i];
    sincos3(t, sc);
    G::fc(t, u, sc, k2);
    if (sc_out) { PANTR_UNROLL for (int i = 0; i < 6; ++i) sc_out[6 + i] = sc[i]; }
    PANTR_UNROLL for (int i = 0; i < NX; ++i) t[i] = x[i] + hh * k2[i];
    sincos3(t, sc);
    G::fc(t, u, sc, k3);
    if (sc_out) { PANTR_UNROLL for (int i = 0; i < 6; ++i) sc_out[12 + i] = sc[i]; }
    PANTR_UNROLL for (int i = 0; i < NX; ++i) t[i] = x[i] + h * k3[i];
    sincos3(t, sc);
    G::fc(t, u, sc, k4);
    if (sc_out) { PANTR_UNROLL for (int i = 0; i < 6; ++i) sc_out[18 + i] = sc[i]; }
    PANTR_UNROLL for (int i = 0; i < NX; ++i) xn[i] = x[i] + h6 * (((k1[i] + 2.0 * k2[i]) + 2.0 * k3[i]) + k4[i]);
}

/// Rollout + cost (ss.py:24-37): fills traj and sc, returns f.
template <class G>
PANTR_HD double rollout(int N, const double *x0, const double *u, const QuadWS &w) {
    double x[NX], xn[NX], uk[NU], f = 0;
    PANTR_UNROLL for (int i = 0; i < NX; ++i) { x[i] = x0[i]; w.traj[i] = x[i]; }
    PANTR_NOUNROLL
    for (int k = 0; k < N; ++k) {
        PANTR_UNROLL for (int j = 0; j < NU; ++j) uk[j] = u[NU * k + j];
        f = f + stage_cost(x, uk);
        rk4<G>(x, uk, xn, w.sc + 24 * k);
        PANTR_UNROLL for (int i = 0; i < NX; ++i) { x[i] = xn[i]; w.traj[(k + 1) * NX + i] = x[i]; }
    }
    double sc[6];
    sincos3(x, sc);
    PANTR_UNROLL for (int i = 0; i < 6; ++i) w.sc[24 * N + i] = sc[i];
    return f + term_cost(x);
}

/// g = [c(x_0); ...; c(x_N)] of the last rollout (ss.py:39-40)
template <class G>
PANTR_HD void constraints(int N, const QuadWS &w, double *g) {
    PANTR_NOUNROLL
    for (int k = 0; k <= N; ++k) {
        double c[NC];
        G::con(w.traj + NX * k, w.sc + 24 * k, c);
        PANTR_UNROLL for (int i = 0; i < NC; ++i) g[NC * k + i] = c[i];
    }
}

/// Augmented-Lagrangian terms of the last rollout (casadi_generator/__init__.py:64-72):
/// zeta = g + y / Sigma, z_hat = fmax(zl, fmin(zeta, zu)), d = zeta - z_hat, y_hat = Sigma d;
/// returns 1/2 y_hat' d.  yhat / wact (length m) are optional outputs.
template <class G>
PANTR_HD double al_terms(int N, const QuadWS &w, const double *y, const double *S, double *yhat, double *wact) {
    double dty = 0;
    PANTR_NOUNROLL
    for (int k = 0; k <= N; ++k) {
        double c[NC];
        G::con(w.traj + NX * k, w.sc + 24 * k, c);
        PANTR_UNROLL for (int i = 0; i < NC; ++i) {
            const double Si = S[NC * k + i];
            const double zeta = c[i] + y[NC * k + i] / Si;
            const double zh = fmax(D_lb(i), fmin(zeta, D_ub(i)));
            const double d = zeta - zh, yh = Si * d;
            dty = dty + yh * d;
            if (yhat) yhat[NC * k + i] = yh;
            if (wact) wact[NC * k + i] = d != 0 ? Si : 0.0;
        }
    }
    return 0.5 * dty;
}

/// grad psi = grad f + J_g' y_hat by discrete adjoint; needs the rollout of u and w.yh.  Stores
/// the adjoint stage weights in w.wst.
/// COST = false drops the cost terms: grad = J_g' y_hat (eval_grad_g_prod of the problem plugin face).
template <class G, bool COST = true>
PANTR_HD void adjoint(int N, const double *u, const QuadWS &w, double *grad) {
    double lam[NX], ax[NX], au[NU], s[NX], su[NU], wt[NX], uk[NU];
    {
        const double *x = w.traj + NX * N;
        G::con_vjp(x, w.sc + 24 * N, w.yh + NC * N, ax);
        PANTR_UNROLL for (int i = 0; i < NX; ++i) lam[i] = COST ? dcost_x(x, i, true) + ax[i] : ax[i];
    }
    PANTR_NOUNROLL
    for (int k = N - 1; k >= 0; --k) {
        const double *x = w.traj + NX * k, *sc = w.sc + 24 * k;
        double *ws = w.wst + 4 * NX * k;
        PANTR_UNROLL for (int j = 0; j < NU; ++j) uk[j] = u[NU * k + j];
        // stage 4
        PANTR_UNROLL for (int i = 0; i < NX; ++i) { wt[i] = h6 * lam[i]; ws[3 * NX + i] = wt[i]; }
        G::fc_vjp(x, uk, sc + 18, wt, ax, au);
        PANTR_UNROLL for (int i = 0; i < NX; ++i) s[i] = lam[i] + ax[i];
        PANTR_UNROLL for (int j = 0; j < NU; ++j) su[j] = au[j];
        // stage 3
        PANTR_UNROLL for (int i = 0; i < NX; ++i) { wt[i] = h3 * lam[i] + h * ax[i]; ws[2 * NX + i] = wt[i]; }
        G::fc_vjp(x, uk, sc + 12, wt, ax, au);
        PANTR_UNROLL for (int i = 0; i < NX; ++i) s[i] = s[i] + ax[i];
        PANTR_UNROLL for (int j = 0; j < NU; ++j) su[j] = su[j] + au[j];
        // stage 2
        PANTR_UNROLL for (int i = 0; i < NX; ++i) { wt[i] = h3 * lam[i] + hh * ax[i]; ws[NX + i] = wt[i]; }
        G::fc_vjp(x, uk, sc + 6, wt, ax, au);
        PANTR_UNROLL for (int i = 0; i < NX; ++i) s[i] = s[i] + ax[i];
        PANTR_UNROLL for (int j = 0; j < NU; ++j) su[j] = su[j] + au[j];
        // stage 1
        PANTR_UNROLL for (int i = 0; i < NX; ++i) { wt[i] = h6 * lam[i] + hh * ax[i]; ws[i] = wt[i]; }
        G::fc_vjp(x, uk, sc, wt, ax, au);
        PANTR_UNROLL for (int i = 0; i < NX; ++i) s[i] = s[i] + ax[i];
        PANTR_UNROLL for (int j = 0; j < NU; ++j) su[j] = su[j] + au[j];
        PANTR_UNROLL for (int j = 0; j < NU; ++j) grad[NU * k + j] = COST ? su[j] + (2 * ru(j)) * uk[j] : su[j];
        if (k > 0) { // lambda_0 is not needed: c(x_0) and l_x(x_0) do not depend on u
            G::con_vjp(x, sc, w.yh + NC * k, ax);
            PANTR_UNROLL for (int i = 0; i < NX; ++i) lam[i] = COST ? (s[i] + dcost_x(x, i, false)) + ax[i] : s[i] + ax[i];
        }
    }
}

/// node contribution to the second-order adjoint: (hess l + sum_i y_hat_i hess c_i + J_c' W J_c) dx
template <class G>
PANTR_HD void node_hess(const double *x, const double *sc, const double *yh, const double *wact, const double *dx,
                        bool term, double *out) {
    double bx[NX], dc[NC], t[NC], ax[NX];
    G::con_hvjp(x, sc, yh, dx, bx);
    G::con_jvp(x, sc, dx, dc);
    PANTR_UNROLL for (int i = 0; i < NC; ++i) t[i] = wact[i] * dc[i];
    G::con_vjp(x, sc, t, ax);
    PANTR_UNROLL for (int i = 0; i < NX; ++i) out[i] = ((2 * qx(i, term)) * dx[i] + bx[i]) + ax[i];
}

/// Hv = hess psi(u) v (casadi_generator/__init__.py:181-189, scale s = 1) by forward-over-adjoint.
/// Needs the rollout of u and the caches of adjoint() (w.yh, w.wact, w.wst).
template <class G>
PANTR_HD void hvp(int N, const double *u, const double *v, const QuadWS &w, double *Hv) {
    double dx[NX], dk1[NX], dk2[NX], dk3[NX], dk4[NX], ds[NX], uk[NU], vk[NU];
    // ---- forward tangent sweep ----
    PANTR_UNROLL for (int i = 0; i < NX; ++i) { dx[i] = 0.0; w.dtraj[i] = 0.0; }
    PANTR_NOUNROLL
    for (int k = 0; k < N; ++k) {
        const double *x = w.traj + NX * k, *sc = w.sc + 24 * k;
        double *dst = w.dst + 4 * NX * k;
        PANTR_UNROLL for (int j = 0; j < NU; ++j) { uk[j] = u[NU * k + j]; vk[j] = v[NU * k + j]; }
        PANTR_UNROLL for (int i = 0; i < NX; ++i) dst[i] = dx[i];
        G::fc_jvp(x, uk, sc, dx, vk, dk1);
        PANTR_UNROLL for (int i = 0; i < NX; ++i) { ds[i] = dx[i] + hh * dk1[i]; dst[NX + i] = ds[i]; }
        G::fc_jvp(x, uk, sc + 6, ds, vk, dk2);
        PANTR_UNROLL for (int i = 0; i < NX; ++i) { ds[i] = dx[i] + hh * dk2[i]; dst[2 * NX + i] = ds[i]; }
        G::fc_jvp(x, uk, sc + 12, ds, vk, dk3);
        PANTR_UNROLL for (int i = 0; i < NX; ++i) { ds[i] = dx[i] + h * dk3[i]; dst[3 * NX + i] = ds[i]; }
        G::fc_jvp(x, uk, sc + 18, ds, vk, dk4);
        PANTR_UNROLL for (int i = 0; i < NX; ++i) {
            dx[i] = dx[i] + h6 * (((dk1[i] + 2.0 * dk2[i]) + 2.0 * dk3[i]) + dk4[i]);
            w.dtraj[(k + 1) * NX + i] = dx[i];
        }
    }
    // ---- reverse sweep for mu = d lambda ----
    double mu[NX], bx[NX], bu[NU], cx[NX], cu[NU], t[NX], tu[NU], dw[NX];
    node_hess<G>(w.traj + NX * N, w.sc + 24 * N, w.yh + NC * N, w.wact + NC * N, dx, true, mu);
    PANTR_NOUNROLL
    for (int k = N - 1; k >= 0; --k) {
        const double *x = w.traj + NX * k, *sc = w.sc + 24 * k;
        const double *ws = w.wst + 4 * NX * k, *dst = w.dst + 4 * NX * k;
        PANTR_UNROLL for (int j = 0; j < NU; ++j) { uk[j] = u[NU * k + j]; vk[j] = v[NU * k + j]; }
        // stage 4
        PANTR_UNROLL for (int i = 0; i < NX; ++i) dw[i] = h6 * mu[i];
        G::fc_vjp(x, uk, sc + 18, dw, bx, bu);
        G::fc_hvjp(x, uk, sc + 18, ws + 3 * NX, dst + 3 * NX, vk, cx, cu);
        PANTR_UNROLL for (int i = 0; i < NX; ++i) { bx[i] = bx[i] + cx[i]; t[i] = mu[i] + bx[i]; }
        PANTR_UNROLL for (int j = 0; j < NU; ++j) tu[j] = bu[j] + cu[j];
        // stage 3
        PANTR_UNROLL for (int i = 0; i < NX; ++i) dw[i] = h3 * mu[i] + h * bx[i];
        G::fc_vjp(x, uk, sc + 12, dw, bx, bu);
        G::fc_hvjp(x, uk, sc + 12, ws + 2 * NX, dst + 2 * NX, vk, cx, cu);
        PANTR_UNROLL for (int i = 0; i < NX; ++i) { bx[i] = bx[i] + cx[i]; t[i] = t[i] + bx[i]; }
        PANTR_UNROLL for (int j = 0; j < NU; ++j) tu[j] = tu[j] + (bu[j] + cu[j]);
        // stage 2
        PANTR_UNROLL for (int i = 0; i < NX; ++i) dw[i] = h3 * mu[i] + hh * bx[i];
        G::fc_vjp(x, uk, sc + 6, dw, bx, bu);
        G::fc_hvjp(x, uk, sc + 6, ws + NX, dst + NX, vk, cx, cu);
        PANTR_UNROLL for (int i = 0; i < NX; ++i) { bx[i] = bx[i] + cx[i]; t[i] = t[i] + bx[i]; }
        PANTR_UNROLL for (int j = 0; j < NU; ++j) tu[j] = tu[j] + (bu[j] + cu[j]);
        // stage 1
        PANTR_UNROLL for (int i = 0; i < NX; ++i) dw[i] = h6 * mu[i] + hh * bx[i];
        G::fc_vjp(x, uk, sc, dw, bx, bu);
        G::fc_hvjp(x, uk, sc, ws, dst, vk, cx, cu);
        PANTR_UNROLL for (int i = 0; i < NX; ++i) { bx[i] = bx[i] + cx[i]; t[i] = t[i] + bx[i]; }
        PANTR_UNROLL for (int j = 0; j < NU; ++j) tu[j] = tu[j] + (bu[j] + cu[j]);
        PANTR_UNROLL for (int j = 0; j < NU; ++j) Hv[NU * k + j] = tu[j] + (2 * ru(j)) * vk[j];
        if (k > 0) {
            node_hess<G>(x, sc, w.yh + NC * k, w.wact + NC * k, w.dtraj + NX * k, false, cx);
            PANTR_UNROLL for (int i = 0; i < NX; ++i) mu[i] = t[i] + cx[i];
        }
    }
}

} // namespace quad
} // namespace pantr
