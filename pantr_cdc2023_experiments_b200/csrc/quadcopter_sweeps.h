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

/// Per-instance workspace (doubles), addressed through the pointer-like type P: `double *` for the
/// warp-per-instance path and the CPU oracle, the stride-32 view SVecT<double> for the
/// thread-per-instance path (exec_policy.cuh).  The sweeps copy the few values a stage needs into
/// local arrays before calling the generated model functions, so the layout never leaks into them.
template <class P>
struct QuadWST {
    P traj;  // (N+1) x 9   states of the last rollout
    P sc;    // (4N+1) x 6  sin/cos of the angles at stage point j of step k (4k+j); node N last
    P wst;   // 4N x 9      first-order adjoint weight of stage j of step k
    P dtraj; // (N+1) x 9   tangent trajectory of the running Hessian-vector product
    P dst;   // 4N x 9      tangent stage points
    P yh;    // 4(N+1)      y_hat of the last psi_grad_psi
    P wact;  // 4(N+1)      Sigma_i where zeta_i is outside D, else 0
    PANTR_HD static int doubles(int N) { return 2 * (N + 1) * NX + (4 * N + 1) * 6 + 2 * 4 * N * NX + 2 * NC * (N + 1); }
    PANTR_HD void bind(P p, int N) {
        traj = p; p = p + (N + 1) * NX; sc = p; p = p + (4 * N + 1) * 6; wst = p; p = p + 4 * N * NX;
        dtraj = p; p = p + (N + 1) * NX; dst = p; p = p + 4 * N * NX; yh = p; p = p + NC * (N + 1); wact = p;
    }
};
using QuadWS = QuadWST<double *>;

/// local copy of `count` entries starting at v[off]
template <int count, class V>
PANTR_HD void ld(V v, int off, double *out) {
    PANTR_UNROLL for (int i = 0; i < count; ++i) out[i] = v[off + i];
}

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
template <class G, class SC>
PANTR_HD void rk4(const double *x, const double *u, double *xn, SC sc_out, int so, bool store_sc) {
    double sc[6], k1[NX], k2[NX], k3[NX], k4[NX], t[NX];
    sincos3(x, sc);
    G::fc(x, u, sc, k1);
    if (store_sc) { PANTR_UNROLL for (int i = 0; i < 6; ++i) sc_out[so + i] = sc[i]; }
    PANTR_UNROLL for (int i = 0; i < NX; ++i) t[i] = x[i] + hh * k1[i];
    sincos3(t, sc);
    G::fc(t, u, sc, k2);
    if (store_sc) { PANTR_UNROLL for (int i = 0; i < 6; ++i) sc_out[so + 6 + i] = sc[i]; }
    PANTR_UNROLL for (int i = 0; i < NX; ++i) t[i] = x[i] + hh * k2[i];
    sincos3(t, sc);
    G::fc(t, u, sc, k3);
    if (store_sc) { PANTR_UNROLL for (int i = 0; i < 6; ++i) sc_out[so + 12 + i] = sc[i]; }
    PANTR_UNROLL for (int i = 0; i < NX; ++i) t[i] = x[i] + h * k3[i];
    sincos3(t, sc);
    G::fc(t, u, sc, k4);
    if (store_sc) { PANTR_UNROLL for (int i = 0; i < 6; ++i) sc_out[so + 18 + i] = sc[i]; }
    PANTR_UNROLL for (int i = 0; i < NX; ++i) xn[i] = x[i] + h6 * (((k1[i] + 2.0 * k2[i]) + 2.0 * k3[i]) + k4[i]);
}
/// plant step without caches
template <class G>
PANTR_HD void rk4(const double *x, const double *u, double *xn, double *) {
    rk4<G>(x, u, xn, static_cast<double *>(nullptr), 0, false);
}

/// Rollout + cost (ss.py:24-37): fills traj and sc, returns f.
template <class G, class U, class W>
PANTR_HD double rollout(int N, const double *x0, U u, const W &w) {
    double x[NX], xn[NX], uk[NU], f = 0;
    PANTR_UNROLL for (int i = 0; i < NX; ++i) { x[i] = x0[i]; w.traj[i] = x[i]; }
    PANTR_NOUNROLL
    for (int k = 0; k < N; ++k) {
        PANTR_UNROLL for (int j = 0; j < NU; ++j) uk[j] = u[NU * k + j];
        f = f + stage_cost(x, uk);
        rk4<G>(x, uk, xn, w.sc, 24 * k, true);
        PANTR_UNROLL for (int i = 0; i < NX; ++i) { x[i] = xn[i]; w.traj[(k + 1) * NX + i] = x[i]; }
    }
    double sc[6];
    sincos3(x, sc);
    PANTR_UNROLL for (int i = 0; i < 6; ++i) w.sc[24 * N + i] = sc[i];
    return f + term_cost(x);
}

/// g = [c(x_0); ...; c(x_N)] of the last rollout (ss.py:39-40)
template <class G, class W, class GV>
PANTR_HD void constraints(int N, const W &w, GV g) {
    PANTR_NOUNROLL
    for (int k = 0; k <= N; ++k) {
        double c[NC], x[NX], sc[6];
        ld<NX>(w.traj, NX * k, x); ld<6>(w.sc, 24 * k, sc);
        G::con(x, sc, c);
        PANTR_UNROLL for (int i = 0; i < NC; ++i) g[NC * k + i] = c[i];
    }
}

/// Augmented-Lagrangian terms of the last rollout (casadi_generator/__init__.py:64-72):
/// zeta = g + y / Sigma, z_hat = fmax(zl, fmin(zeta, zu)), d = zeta - z_hat, y_hat = Sigma d;
/// returns 1/2 y_hat' d.  yhat / wact (length m) are optional outputs.
template <class G, class W, class Y, class SV, class YH, class WA>
PANTR_HD double al_terms(int N, const W &w, Y y, SV S, YH yhat, bool want_yhat, WA wact, bool want_wact) {
    double dty = 0;
    PANTR_NOUNROLL
    for (int k = 0; k <= N; ++k) {
        double c[NC], x[NX], sc[6];
        ld<NX>(w.traj, NX * k, x); ld<6>(w.sc, 24 * k, sc);
        G::con(x, sc, c);
        PANTR_UNROLL for (int i = 0; i < NC; ++i) {
            const double Si = S[NC * k + i];
            const double zeta = c[i] + y[NC * k + i] / Si;
            const double zh = fmax(D_lb(i), fmin(zeta, D_ub(i)));
            const double d = zeta - zh, yh = Si * d;
            dty = dty + yh * d;
            if (want_yhat) yhat[NC * k + i] = yh;
            if (want_wact) wact[NC * k + i] = d != 0 ? Si : 0.0;
        }
    }
    return 0.5 * dty;
}

/// grad psi = grad f + J_g' y_hat by discrete adjoint; needs the rollout of u and w.yh.  Stores
/// the adjoint stage weights in w.wst.
/// COST = false drops the cost terms: grad = J_g' y_hat (eval_grad_g_prod of the problem plugin face).
template <class G, bool COST = true, class U, class W, class GV>
PANTR_HD void adjoint(int N, U u, const W &w, GV grad) {
    double lam[NX], ax[NX], au[NU], s[NX], su[NU], wt[NX], uk[NU], x[NX], sc[6], yh[NC];
    {
        ld<NX>(w.traj, NX * N, x); ld<6>(w.sc, 24 * N, sc); ld<NC>(w.yh, NC * N, yh);
        G::con_vjp(x, sc, yh, ax);
        PANTR_UNROLL for (int i = 0; i < NX; ++i) lam[i] = COST ? dcost_x(x, i, true) + ax[i] : ax[i];
    }
    PANTR_NOUNROLL
    for (int k = N - 1; k >= 0; --k) {
        const int so = 24 * k, wo = 4 * NX * k;
        ld<NX>(w.traj, NX * k, x);
        PANTR_UNROLL for (int j = 0; j < NU; ++j) uk[j] = u[NU * k + j];
        // stage 4
        PANTR_UNROLL for (int i = 0; i < NX; ++i) { wt[i] = h6 * lam[i]; w.wst[wo + 3 * NX + i] = wt[i]; }
        ld<6>(w.sc, so + 18, sc);
        G::fc_vjp(x, uk, sc, wt, ax, au);
        PANTR_UNROLL for (int i = 0; i < NX; ++i) s[i] = lam[i] + ax[i];
        PANTR_UNROLL for (int j = 0; j < NU; ++j) su[j] = au[j];
        // stage 3
        PANTR_UNROLL for (int i = 0; i < NX; ++i) { wt[i] = h3 * lam[i] + h * ax[i]; w.wst[wo + 2 * NX + i] = wt[i]; }
        ld<6>(w.sc, so + 12, sc);
        G::fc_vjp(x, uk, sc, wt, ax, au);
        PANTR_UNROLL for (int i = 0; i < NX; ++i) s[i] = s[i] + ax[i];
        PANTR_UNROLL for (int j = 0; j < NU; ++j) su[j] = su[j] + au[j];
        // stage 2
        PANTR_UNROLL for (int i = 0; i < NX; ++i) { wt[i] = h3 * lam[i] + hh * ax[i]; w.wst[wo + NX + i] = wt[i]; }
        ld<6>(w.sc, so + 6, sc);
        G::fc_vjp(x, uk, sc, wt, ax, au);
        PANTR_UNROLL for (int i = 0; i < NX; ++i) s[i] = s[i] + ax[i];
        PANTR_UNROLL for (int j = 0; j < NU; ++j) su[j] = su[j] + au[j];
        // stage 1
        PANTR_UNROLL for (int i = 0; i < NX; ++i) { wt[i] = h6 * lam[i] + hh * ax[i]; w.wst[wo + i] = wt[i]; }
        ld<6>(w.sc, so, sc);
        G::fc_vjp(x, uk, sc, wt, ax, au);
        PANTR_UNROLL for (int i = 0; i < NX; ++i) s[i] = s[i] + ax[i];
        PANTR_UNROLL for (int j = 0; j < NU; ++j) su[j] = su[j] + au[j];
        PANTR_UNROLL for (int j = 0; j < NU; ++j) grad[NU * k + j] = COST ? su[j] + (2 * ru(j)) * uk[j] : su[j];
        if (k > 0) { // lambda_0 is not needed: c(x_0) and l_x(x_0) do not depend on u
            ld<NC>(w.yh, NC * k, yh);
            G::con_vjp(x, sc, yh, ax);
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
template <class G, class U, class V, class W, class HV>
PANTR_HD void hvp(int N, U u, V v, const W &w, HV Hv) {
    double dx[NX], dk1[NX], dk2[NX], dk3[NX], dk4[NX], ds[NX], uk[NU], vk[NU], x[NX], sc[6];
    // ---- forward tangent sweep ----
    PANTR_UNROLL for (int i = 0; i < NX; ++i) { dx[i] = 0.0; w.dtraj[i] = 0.0; }
    PANTR_NOUNROLL
    for (int k = 0; k < N; ++k) {
        const int so = 24 * k, dsto = 4 * NX * k;
        ld<NX>(w.traj, NX * k, x);
        PANTR_UNROLL for (int j = 0; j < NU; ++j) { uk[j] = u[NU * k + j]; vk[j] = v[NU * k + j]; }
        PANTR_UNROLL for (int i = 0; i < NX; ++i) w.dst[dsto + i] = dx[i];
        ld<6>(w.sc, so, sc);
        G::fc_jvp(x, uk, sc, dx, vk, dk1);
        PANTR_UNROLL for (int i = 0; i < NX; ++i) { ds[i] = dx[i] + hh * dk1[i]; w.dst[dsto + NX + i] = ds[i]; }
        ld<6>(w.sc, so + 6, sc);
        G::fc_jvp(x, uk, sc, ds, vk, dk2);
        PANTR_UNROLL for (int i = 0; i < NX; ++i) { ds[i] = dx[i] + hh * dk2[i]; w.dst[dsto + 2 * NX + i] = ds[i]; }
        ld<6>(w.sc, so + 12, sc);
        G::fc_jvp(x, uk, sc, ds, vk, dk3);
        PANTR_UNROLL for (int i = 0; i < NX; ++i) { ds[i] = dx[i] + h * dk3[i]; w.dst[dsto + 3 * NX + i] = ds[i]; }
        ld<6>(w.sc, so + 18, sc);
        G::fc_jvp(x, uk, sc, ds, vk, dk4);
        PANTR_UNROLL for (int i = 0; i < NX; ++i) {
            dx[i] = dx[i] + h6 * (((dk1[i] + 2.0 * dk2[i]) + 2.0 * dk3[i]) + dk4[i]);
            w.dtraj[(k + 1) * NX + i] = dx[i];
        }
    }
    // ---- reverse sweep for mu = d lambda ----
    double mu[NX], bx[NX], bu[NU], cx[NX], cu[NU], t[NX], tu[NU], dw[NX], ws[NX], dst[NX], yh[NC], wa[NC];
    ld<NX>(w.traj, NX * N, x); ld<6>(w.sc, 24 * N, sc); ld<NC>(w.yh, NC * N, yh); ld<NC>(w.wact, NC * N, wa);
    node_hess<G>(x, sc, yh, wa, dx, true, mu);
    PANTR_NOUNROLL
    for (int k = N - 1; k >= 0; --k) {
        const int so = 24 * k, wo = 4 * NX * k;
        ld<NX>(w.traj, NX * k, x);
        PANTR_UNROLL for (int j = 0; j < NU; ++j) { uk[j] = u[NU * k + j]; vk[j] = v[NU * k + j]; }
        // stage 4
        PANTR_UNROLL for (int i = 0; i < NX; ++i) dw[i] = h6 * mu[i];
        ld<6>(w.sc, so + 18, sc); ld<NX>(w.wst, wo + 3 * NX, ws); ld<NX>(w.dst, wo + 3 * NX, dst);
        G::fc_vjp(x, uk, sc, dw, bx, bu);
        G::fc_hvjp(x, uk, sc, ws, dst, vk, cx, cu);
        PANTR_UNROLL for (int i = 0; i < NX; ++i) { bx[i] = bx[i] + cx[i]; t[i] = mu[i] + bx[i]; }
        PANTR_UNROLL for (int j = 0; j < NU; ++j) tu[j] = bu[j] + cu[j];
        // stage 3
        PANTR_UNROLL for (int i = 0; i < NX; ++i) dw[i] = h3 * mu[i] + h * bx[i];
        ld<6>(w.sc, so + 12, sc); ld<NX>(w.wst, wo + 2 * NX, ws); ld<NX>(w.dst, wo + 2 * NX, dst);
        G::fc_vjp(x, uk, sc, dw, bx, bu);
        G::fc_hvjp(x, uk, sc, ws, dst, vk, cx, cu);
        PANTR_UNROLL for (int i = 0; i < NX; ++i) { bx[i] = bx[i] + cx[i]; t[i] = t[i] + bx[i]; }
        PANTR_UNROLL for (int j = 0; j < NU; ++j) tu[j] = tu[j] + (bu[j] + cu[j]);
        // stage 2
        PANTR_UNROLL for (int i = 0; i < NX; ++i) dw[i] = h3 * mu[i] + hh * bx[i];
        ld<6>(w.sc, so + 6, sc); ld<NX>(w.wst, wo + NX, ws); ld<NX>(w.dst, wo + NX, dst);
        G::fc_vjp(x, uk, sc, dw, bx, bu);
        G::fc_hvjp(x, uk, sc, ws, dst, vk, cx, cu);
        PANTR_UNROLL for (int i = 0; i < NX; ++i) { bx[i] = bx[i] + cx[i]; t[i] = t[i] + bx[i]; }
        PANTR_UNROLL for (int j = 0; j < NU; ++j) tu[j] = tu[j] + (bu[j] + cu[j]);
        // stage 1
        PANTR_UNROLL for (int i = 0; i < NX; ++i) dw[i] = h6 * mu[i] + hh * bx[i];
        ld<6>(w.sc, so, sc); ld<NX>(w.wst, wo, ws); ld<NX>(w.dst, wo, dst);
        G::fc_vjp(x, uk, sc, dw, bx, bu);
        G::fc_hvjp(x, uk, sc, ws, dst, vk, cx, cu);
        PANTR_UNROLL for (int i = 0; i < NX; ++i) { bx[i] = bx[i] + cx[i]; t[i] = t[i] + bx[i]; }
        PANTR_UNROLL for (int j = 0; j < NU; ++j) tu[j] = tu[j] + (bu[j] + cu[j]);
        PANTR_UNROLL for (int j = 0; j < NU; ++j) Hv[NU * k + j] = tu[j] + (2 * ru(j)) * vk[j];
        if (k > 0) {
            ld<NC>(w.yh, NC * k, yh); ld<NC>(w.wact, NC * k, wa); ld<NX>(w.dtraj, NX * k, dst);
            node_hess<G>(x, sc, yh, wa, dst, false, cx);
            PANTR_UNROLL for (int i = 0; i < NX; ++i) mu[i] = t[i] + cx[i];
        }
    }
}

} // namespace quad
} // namespace pantr
