// The two quadcopter MPC models (quadcopter_ss2p_<N> / realistic_quadcopter_ss2p_<N>) as
// warp-cooperative device code: one warp = one MPC instance, CONDENSED formulation.
//
// Replaces the CasADi-generated f, g, psi, psi_grad_psi, grad_L, hess_psi_prod / hess_L_prod that
// CasADiProblem::eval_* calls (alpaqa/src/interop/casadi/include/alpaqa/implementation/casadi/
// CasADiProblem.tpp:208-368).  Math spec (paths relative to /root/reference):
//   python/alpaqa_mpc_benchmarks/problems/quadcopter.py:5-121            f_c, RK4, cost, boxes, x0
//   python/alpaqa_mpc_benchmarks/problems/realistic_quadcopter.py:5-127  same with ARB / Omega^-1
//   python/alpaqa_mpc_benchmarks/formulations/ss.py:20-55                rollout, g = [c(x_0..x_N)]
//   alpaqa/python/alpaqa/casadi_generator/__init__.py:64-77,181-189      psi, s psi, hess_psi_prod
//
// Why condensed.  With nx = 9 a rollout / adjoint / forward-over-adjoint sweep is one sequential
// chain with nothing for 32 lanes to share, and a GPU thread runs such a chain ~40x slower than a
// host core.  So the work is reorganised around the per-step Jacobians
//        AB_k = d f_d / d(x_k, u_k)        (9 x 13, RK4 step of the generated f_c),
// built from 13 unit tangents per step -- 13 N independent tasks, lane-parallel.  With them
//   grad psi :  lambda_k = AB_k' lambda_{k+1} + q_k                       (q_k: node terms, lane = node)
//   H v      :  dx_{k+1} = AB_k (dx_k, v_k)                               (tangent recurrence)
//               s_k = d/dz[J_step' lambda_{k+1}] (dx_k, v_k) + node terms  (lane = step: 3 jvp + 6 vjp + 4 hvjp)
//               mu_k = AB_k' mu_{k+1} + s_k                               (second-order adjoint recurrence)
// i.e. the only sequential parts left are the state rollout and three 9 x 13 matrix-vector recurrences
// (one fma chain per lane and step, operands exchanged through shared memory); everything else is
// evaluated independently per step by the 32 lanes.  Same exact derivatives as forward-over-adjoint
// AD, different association of the sums: oracle/quadcopter_condensed.hpp restates this file's
// operation order on the CPU (bit-identical results), oracle/quadcopter.hpp is the reference
// formulation (agreement to rounding, tests/test_oracle_quadcopter.py).
//
// Per-instance model workspace (global memory; L2-resident for the resident warps):
//   traj (N+1) x 9 | sc (4N+1) x 6 sin/cos of every stage point | AB N x 13 x 9 (column-major) |
//   lam (N+1) x 9 adjoint states | yh, wact 4(N+1) | dtraj (N+1) x 9 tangent states |
//   mus (N+1) x 9 and usrc N x 4: per-step sources (then states) of the second-order adjoint recurrence
#pragma once
#include "pantr_hd.h"
#include "generated/quadcopter_gen.h"
#include "generated/realistic_quadcopter_gen.h"
#include "exec_policy.cuh"

namespace pantr {
namespace qc {

constexpr int NX = 9, NU = 4, NC = 4, NZ = 13;
constexpr double Ts = 0.1, h = Ts, hh = Ts / 2, h6 = Ts / 6, h3 = 2 * (Ts / 6); // quadcopter.py:6,49-55
// quadcopter.py:85-111 : l = 1e-4 u'u + 10 |omega|^2 + |v|^2 + 10 |p - pf|^2 + |theta|^2 ;
// l_N = 25 |v|^2 + 250 |p - pf|^2 + 10 |theta|^2
constexpr double w_v = 1, w_p = 10, w_th = 1, w_u = 1e-4, w_om = 10;
constexpr double wN_v = 25 * w_v, wN_p = 25 * w_p, wN_th = 10 * w_th;
// quadcopter.py:11-14,56-59,70-73
constexpr double at_max = 9.81 * 5, dtilt_max = 0.1;
constexpr double half_pi = 1.5707963267948966, cos_tilt_max = 0.85252452205950568; // cos(1.1 / 2)

__host__ __device__ __forceinline__ double pf(int i) { return i == 2 ? 0.5 : 0.25; }
__host__ __device__ __forceinline__ double wx(int i, bool term) {
    return i < 3 ? (term ? wN_p : w_p) : (i < 6 ? (term ? wN_v : w_v) : (term ? wN_th : w_th));
}
__host__ __device__ __forceinline__ double wu(int j) { return j == 0 ? w_u : w_u + w_om; }
__host__ __device__ __forceinline__ double sum_sq3(double a, double b, double c) { return (a * a + b * b) + c * c; }

#define PANTR_QC_FNS(Name, ns)                                                                               \
    struct Name {                                                                                            \
        static constexpr unsigned CONST_COLS = ns::FC_CONST_COLS;                                            \
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
PANTR_QC_FNS(SimpleFns, quadcopter_gen);
PANTR_QC_FNS(RealisticFns, realistic_quadcopter_gen);
#undef PANTR_QC_FNS

template <int count>
__device__ __forceinline__ void ldv(const double *p, double *out) {
#pragma unroll
    for (int i = 0; i < count; ++i) out[i] = p[i];
}

} // namespace qc

#ifndef PANTR_QC_MIN_BLOCKS
#define PANTR_QC_MIN_BLOCKS 16 // resident warps per SM the solve kernel is compiled for (128 registers)
#endif

template <class F>
struct QuadCondensedModel {
    static constexpr int NX = qc::NX, NU = qc::NU, NC = qc::NC, NZ = qc::NZ;
    static constexpr int MIN_BLOCKS = PANTR_QC_MIN_BLOCKS;
    static constexpr unsigned VAR_COLS = ~F::CONST_COLS & 0x1fffu; // columns of AB that differ from step to step
    static constexpr bool HAS_SCALE = true;
    __host__ __device__ static int n(int N) { return NU * N; }
    __host__ __device__ static int m(int N) { return NC * (N + 1); }
    __host__ __device__ static int traj_doubles(int N) {
        const int d = 4 * (N + 1) * NX + (4 * N + 1) * 6 + N * NZ * NX + 2 * NC * (N + 1) + N * NU;
        return (d + 1) / 2 * 2; // the solver vectors behind it stay 16-byte aligned
    }
    __host__ __device__ static double C_lb(int i) { return (i & 3) == 0 ? 0.0 : -qc::dtilt_max; } // quadcopter.py:56-59, ss.py:28-29
    __host__ __device__ static double C_ub(int i) { return (i & 3) == 0 ? qc::at_max : qc::dtilt_max; }
    __host__ __device__ static double D_lb(int i, int) { // quadcopter.py:70-73, ss.py:41-42
        return (i & 3) < 2 ? -qc::half_pi : ((i & 3) == 2 ? qc::cos_tilt_max : -INFINITY);
    }
    __host__ __device__ static double D_ub(int i, int) { return (i & 3) < 2 ? qc::half_pi : ((i & 3) == 2 ? INFINITY : 0.0); }
    static void initial_state(double *x) { // quadcopter.py:78-83
        for (int i = 0; i < NX; ++i) x[i] = 0;
        x[0] = -0.2; x[1] = -0.25; x[2] = 0.5;
    }
    static void initial_guess(double *u, int N) { // quadcopter.py:112
        for (int k = 0; k < N; ++k) { u[4 * k] = 9.81; u[4 * k + 1] = u[4 * k + 2] = u[4 * k + 3] = 0; }
    }

    int N;
    const double *x0;
    double *traj, *sc, *AB, *lam, *yh, *wact, *dtraj, *mus, *usrc;
    bool hvp_ready;

    __device__ void init(int N_, const double *x0_, double *ws) {
        N = N_; x0 = x0_; hvp_ready = false;
        double *p = ws;
        traj = p; p += (N + 1) * NX; lam = p; p += (N + 1) * NX; dtraj = p; p += (N + 1) * NX; mus = p; p += (N + 1) * NX;
        sc = p; p += (4 * N + 1) * 6; AB = p; p += N * NZ * NX; yh = p; p += NC * (N + 1); wact = p; p += NC * (N + 1);
        usrc = p;
    }

    // ---- one RK4 step (quadcopter.py:49-55), executed by every lane; the sin/cos of the three angles of a
    // stage point are computed by lanes 0..2 (one angle each) and broadcast.  sc_out: 4 x 6 or null. ----
    __device__ __forceinline__ static void stage_sincos(const double *x, double *s6, double *sc_out) {
        const int lane = lane_id();
        const double ang = lane == 0 ? x[6] : (lane == 1 ? x[7] : x[8]);
        double s, c;
        pantr_sincos(ang, &s, &c);
        if (sc_out && lane < 3) { sc_out[2 * lane] = s; sc_out[2 * lane + 1] = c; }
        s6[0] = shfl(s, 0); s6[1] = shfl(c, 0); s6[2] = shfl(s, 1); s6[3] = shfl(c, 1); s6[4] = shfl(s, 2); s6[5] = shfl(c, 2);
    }
    __device__ __forceinline__ static void rk4_step(double *x, const double *uk, double *sc_out) {
        using namespace qc;
        double s6[6], t[NX], k1[NX], k2[NX], k3[NX], k4[NX];
        stage_sincos(x, s6, sc_out);
        F::fc(x, uk, s6, k1);
#pragma unroll
        for (int i = 0; i < NX; ++i) t[i] = x[i] + hh * k1[i];
        stage_sincos(t, s6, sc_out ? sc_out + 6 : nullptr);
        F::fc(t, uk, s6, k2);
#pragma unroll
        for (int i = 0; i < NX; ++i) t[i] = x[i] + hh * k2[i];
        stage_sincos(t, s6, sc_out ? sc_out + 12 : nullptr);
        F::fc(t, uk, s6, k3);
#pragma unroll
        for (int i = 0; i < NX; ++i) t[i] = x[i] + h * k3[i];
        stage_sincos(t, s6, sc_out ? sc_out + 18 : nullptr);
        F::fc(t, uk, s6, k4);
#pragma unroll
        for (int i = 0; i < NX; ++i) x[i] = x[i] + h6 * (((k1[i] + 2.0 * k2[i]) + 2.0 * k3[i]) + k4[i]);
    }
    /// Rollout (ss.py:24): fills traj and sc.
    __device__ __noinline__ void rollout(const double *u) {
        double x[NX], uk[NU];
#pragma unroll
        for (int i = 0; i < NX; ++i) { x[i] = x0[i]; traj[i] = x[i]; }
#pragma unroll 1
        for (int k = 0; k < N; ++k) {
            qc::ldv<NU>(u + NU * k, uk);
            rk4_step(x, uk, sc + 24 * k);
#pragma unroll
            for (int i = 0; i < NX; ++i) traj[(k + 1) * NX + i] = x[i];
        }
        double s6[6];
        stage_sincos(x, s6, sc + 24 * N);
        __syncwarp();
    }
    /// Cost (ss.py:32-36): lane = node, node k < N contributes l(x_k, u_k), node N contributes l_N(x_N).
    __device__ __forceinline__ double cost(const double *u) const {
        using namespace qc;
        double p = 0;
        for (int k = lane_id(); k <= N; k += 32) {
            double x[NX];
            ldv<NX>(traj + k * NX, x);
            const bool term = k == N;
            const double lx = ((term ? wN_v : w_v) * sum_sq3(x[3], x[4], x[5]) +
                               (term ? wN_p : w_p) * sum_sq3(x[0] - pf(0), x[1] - pf(1), x[2] - pf(2))) +
                              (term ? wN_th : w_th) * sum_sq3(x[6], x[7], x[8]);
            double node = lx;
            if (!term) {
                double uk[NU];
                ldv<NU>(u + NU * k, uk);
                const double lu = w_u * ((uk[0] * uk[0] + uk[1] * uk[1]) + (uk[2] * uk[2] + uk[3] * uk[3])) + w_om * sum_sq3(uk[1], uk[2], uk[3]);
                node = lu + lx;
            }
            p = p + node;
        }
        return warp_sum(p);
    }
    __device__ __forceinline__ void constraints(double *g) const {
        for (int k = lane_id(); k <= N; k += 32) {
            double x[NX], s6[6], c[NC];
            qc::ldv<NX>(traj + k * NX, x); qc::ldv<6>(sc + 24 * k, s6);
            F::con(x, s6, c);
#pragma unroll
            for (int i = 0; i < NC; ++i) g[NC * k + i] = c[i];
        }
    }
    /// Augmented-Lagrangian terms (casadi_generator/__init__.py:64-72), lane = node: zeta = c + y / Sigma,
    /// z_hat = fmax(zl, fmin(zeta, zu)), d = zeta - z_hat, y_hat = Sigma d; returns 1/2 y_hat' d.
    __device__ __noinline__ double al_terms(const double *y, const double *S, double *yhat, double *wa) const {
        double p = 0;
        for (int k = lane_id(); k <= N; k += 32) {
            double x[NX], s6[6], c[NC], a = 0;
            qc::ldv<NX>(traj + k * NX, x); qc::ldv<6>(sc + 24 * k, s6);
            F::con(x, s6, c);
#pragma unroll
            for (int i = 0; i < NC; ++i) {
                const double Si = S[NC * k + i];
                const double zeta = c[i] + y[NC * k + i] / Si;
                const double zh = fmax(D_lb(i, 0), fmin(zeta, D_ub(i, 0)));
                const double d = zeta - zh, yhi = Si * d;
                a = fma(yhi, d, a);
                if (yhat) yhat[NC * k + i] = yhi;
                if (wa) wa[NC * k + i] = d != 0 ? Si : 0.0;
            }
            p = p + a;
        }
        p = warp_sum(p);
        __syncwarp();
        return 0.5 * p;
    }
    /// RK4 tangent of (dx, du) through step k: one column of AB_k when (dx, du) is a unit vector.
    __device__ __forceinline__ static void step_tangent(const double *xk, const double *uk, const double *s, const double *dx,
                                                        const double *du, double *out) {
        using namespace qc;
        double d[NX], dk1[NX], dk2[NX], dk3[NX], dk4[NX];
        F::fc_jvp(xk, uk, s, dx, du, dk1);
#pragma unroll
        for (int i = 0; i < NX; ++i) d[i] = fma(hh, dk1[i], dx[i]);
        F::fc_jvp(xk, uk, s + 6, d, du, dk2);
#pragma unroll
        for (int i = 0; i < NX; ++i) d[i] = fma(hh, dk2[i], dx[i]);
        F::fc_jvp(xk, uk, s + 12, d, du, dk3);
#pragma unroll
        for (int i = 0; i < NX; ++i) d[i] = fma(h, dk3[i], dx[i]);
        F::fc_jvp(xk, uk, s + 18, d, du, dk4);
#pragma unroll
        for (int i = 0; i < NX; ++i) out[i] = fma(h6, fma(2.0, dk3[i], fma(2.0, dk2[i], dk1[i])) + dk4[i], dx[i]);
    }
    /// j-th set bit of mask (columns of (x, u))
    __device__ __forceinline__ static int nth_col(unsigned mask, int j) {
        int c = 0;
#pragma unroll
        for (int b = 0; b < NZ; ++b) {
            const bool on = (mask >> b) & 1u;
            if (on && j == 0) c = b;
            if (on) --j;
        }
        return c;
    }
    /// Step Jacobians: lane = (step, column) task.  Step 0 gets all 13 columns, the other steps the columns
    /// that vary; the constant columns (FC_CONST_COLS: the same at every step) are then copied from step 0.
    __device__ __noinline__ void build_AB(const double *u) {
        constexpr int NV = __builtin_popcount(VAR_COLS), NCST = NZ - NV;
        const int total = NZ + (N - 1) * NV;
        for (int t = lane_id(); t < total; t += 32) {
            int k = 0, c = t;
            if (t >= NZ) { const int q = t - NZ; k = 1 + q / NV; c = nth_col(VAR_COLS, q % NV); }
            double xk[NX], uk[NU], s[24], e[NZ], out[NX];
            qc::ldv<NX>(traj + k * NX, xk); qc::ldv<NU>(u + NU * k, uk); qc::ldv<24>(sc + 24 * k, s);
#pragma unroll
            for (int i = 0; i < NZ; ++i) e[i] = i == c ? 1.0 : 0.0;
            step_tangent(xk, uk, s, e, e + NX, out);
            double *col = AB + (size_t(k) * NZ + c) * NX;
#pragma unroll
            for (int i = 0; i < NX; ++i) col[i] = out[i];
        }
        __syncwarp();
        if (NCST > 0) {
            const int ncopy = (N - 1) * NCST * NX;
            for (int t = lane_id(); t < ncopy; t += 32) {
                const int k = 1 + t / (NCST * NX), r = t % (NCST * NX), c = nth_col(F::CONST_COLS, r / NX), i = r % NX;
                AB[(size_t(k) * NZ + c) * NX + i] = AB[size_t(c) * NX + i];
            }
        }
        __syncwarp();
    }
    /// Adjoint recurrence w_k = AB_k' w_{k+1} + src_k over k = N-1..0 (lane c < 13 owns entry c of the
    /// product: one fma chain over the 9 rows; the 9 adjoint entries travel through shared memory).
    /// state: [(N+1) x 9], row N = terminal value (in), rows 1..N-1 hold the x-part sources on entry and the
    ///        adjoint states on exit (row 0 is not needed: x_0 does not depend on u).
    /// The u-part of step k, combined with the addend `ufun` supplies, goes to out[4k + j].
    template <class UFun>
    __device__ __forceinline__ void adjoint_recurrence(double *state, double *out, UFun ufun) {
        __shared__ double wbuf[2][16];
        const int lane = lane_id();
        int cur = 0;
        if (lane < NX) wbuf[0][lane] = state[N * NX + lane];
        double col[NX], ncol[NX];
        const int cl = lane < NZ ? lane : NZ - 1;
        qc::ldv<NX>(AB + (size_t(N - 1) * NZ + cl) * NX, col);
        __syncwarp();
#pragma unroll 1
        for (int k = N - 1; k >= 0; --k) {
            double addend = 0;
            if (lane < NX) { if (k > 0) addend = state[k * NX + lane]; }
            else if (lane < NZ) addend = ufun(k, lane - NX);
            if (k > 0) qc::ldv<NX>(AB + (size_t(k - 1) * NZ + cl) * NX, ncol);
            double w[NX];
            qc::ldv<NX>(wbuf[cur], w);
            double a = col[0] * w[0];
#pragma unroll
            for (int r = 1; r < NX; ++r) a = fma(col[r], w[r], a);
            if (lane < NX) {
                if (k > 0) { const double l = a + addend; state[k * NX + lane] = l; wbuf[cur ^ 1][lane] = l; }
            } else if (lane < NZ) out[NU * k + lane - NX] = ufun.finish(lane - NX, a, addend);
#pragma unroll
            for (int r = 0; r < NX; ++r) col[r] = ncol[r];
            __syncwarp();
            cur ^= 1;
        }
    }
    struct GradU { // u-part of the gradient: fma(s * 2 w_u, u_kj, AB' lambda)
        const double *u; double cs;
        __device__ __forceinline__ double operator()(int k, int j) const { return u[NU * k + j]; }
        __device__ __forceinline__ double finish(int j, double a, double ukj) const { return fma(cs * (2 * qc::wu(j)), ukj, a); }
    };
    struct SrcU { // u-part of the Hessian-vector product: AB' mu + usrc
        const double *usrc;
        __device__ __forceinline__ double operator()(int k, int j) const { return usrc[k * NU + j]; }
        __device__ __forceinline__ double finish(int, double a, double s) const { return a + s; }
    };
    /// Gradient of  s f + yh' g  wrt u (s = cost_scale); leaves lambda_1..N in lam.
    __device__ __noinline__ void adjoint(const double *u, double cost_scale, double *grad) {
        using namespace qc;
        for (int k = 1 + lane_id(); k <= N; k += 32) { // node sources q_k = s dl/dx + J_c' yh_k
            double x[NX], s6[6], y4[NC], ax[NX];
            ldv<NX>(traj + k * NX, x); ldv<6>(sc + 24 * k, s6); ldv<NC>(yh + NC * k, y4);
            F::con_vjp(x, s6, y4, ax);
#pragma unroll
            for (int i = 0; i < NX; ++i) {
                const double dl = (2 * wx(i, k == N)) * (i < 3 ? x[i] - pf(i) : x[i]);
                lam[k * NX + i] = fma(cost_scale, dl, ax[i]);
            }
        }
        __syncwarp();
        adjoint_recurrence(lam, grad, GradU{u, cost_scale});
    }
    /// Tangent recurrence dx_{k+1} = AB_k (dx_k, v_k): lane r < 9 owns row r, one fma chain over the 13 columns.
    __device__ __noinline__ void tangent(const double *v) {
        __shared__ double zbuf[2][16];
        const int lane = lane_id();
        const int rl = lane < NX ? lane : NX - 1;
        int cur = 0;
        if (lane < NX) { zbuf[0][lane] = 0.0; dtraj[lane] = 0.0; }
        if (lane < NU) zbuf[0][NX + lane] = v[lane];
        double row[NZ], nrow[NZ];
#pragma unroll
        for (int c = 0; c < NZ; ++c) row[c] = AB[size_t(c) * NX + rl];
        __syncwarp();
#pragma unroll 1
        for (int k = 0; k < N; ++k) {
            double vn = 0;
            if (k + 1 < N) {
                if (lane < NU) vn = v[NU * (k + 1) + lane];
#pragma unroll
                for (int c = 0; c < NZ; ++c) nrow[c] = AB[(size_t(k + 1) * NZ + c) * NX + rl];
            }
            double z[NZ];
            qc::ldv<NZ>(zbuf[cur], z);
            double a = row[0] * z[0];
#pragma unroll
            for (int c = 1; c < NZ; ++c) a = fma(row[c], z[c], a);
            if (lane < NX) { dtraj[(k + 1) * NX + lane] = a; zbuf[cur ^ 1][lane] = a; }
            if (lane < NU) zbuf[cur ^ 1][NX + lane] = vn;
#pragma unroll
            for (int c = 0; c < NZ; ++c) row[c] = nrow[c];
            __syncwarp();
            cur ^= 1;
        }
    }
    /// node term of the second-order adjoint at node k: (s hess l + sum_i yh_i hess c_i + J_c' W J_c) dx_k
    __device__ __forceinline__ void node_second_order(int k, double cost_scale, double *out) const {
        using namespace qc;
        double x[NX], s6[6], y4[NC], wa[NC], dx[NX], bx[NX], dc[NC], t[NC], ax[NX];
        ldv<NX>(traj + k * NX, x); ldv<6>(sc + 24 * k, s6); ldv<NC>(yh + NC * k, y4); ldv<NC>(wact + NC * k, wa);
        ldv<NX>(dtraj + k * NX, dx);
        F::con_hvjp(x, s6, y4, dx, bx);
        F::con_jvp(x, s6, dx, dc);
#pragma unroll
        for (int i = 0; i < NC; ++i) t[i] = wa[i] * dc[i];
        F::con_vjp(x, s6, t, ax);
#pragma unroll
        for (int i = 0; i < NX; ++i) out[i] = fma(cost_scale * (2 * wx(i, k == N)), dx[i], bx[i]) + ax[i];
    }
    /// Per-step sources of the second-order adjoint recurrence, lane = step:
    /// mus_k = d/dx_k[J_step' lambda_{k+1}] (dx_k, v_k) + node term of node k (k >= 1; row N = terminal node term),
    /// usrc_k = d/du_k[...] (dx_k, v_k) + s 2 w_u v_k.
    __device__ __noinline__ void second_order_sources(const double *u, const double *v, double cost_scale) {
        using namespace qc;
        for (int k = lane_id(); k <= N; k += 32) {
            if (k == N) {
                double nd[NX];
                node_second_order(N, cost_scale, nd);
#pragma unroll
                for (int i = 0; i < NX; ++i) mus[N * NX + i] = nd[i];
                continue;
            }
            double xk[NX], uk[NU], vk[NU], dx[NX], l[NX];
            const double *s = sc + 24 * k;
            ldv<NX>(traj + k * NX, xk); ldv<NU>(u + NU * k, uk); ldv<NU>(v + NU * k, vk);
            ldv<NX>(dtraj + k * NX, dx); ldv<NX>(lam + (k + 1) * NX, l);
            double s1[6], s2[6], s3[6], s4[6];
            ldv<6>(s, s1); ldv<6>(s + 6, s2); ldv<6>(s + 12, s3); ldv<6>(s + 18, s4);
            // tangent stage points
            double d2[NX], d3[NX], d4[NX], dk[NX];
            F::fc_jvp(xk, uk, s1, dx, vk, dk);
#pragma unroll
            for (int i = 0; i < NX; ++i) d2[i] = fma(hh, dk[i], dx[i]);
            F::fc_jvp(xk, uk, s2, d2, vk, dk);
#pragma unroll
            for (int i = 0; i < NX; ++i) d3[i] = fma(hh, dk[i], dx[i]);
            F::fc_jvp(xk, uk, s3, d3, vk, dk);
#pragma unroll
            for (int i = 0; i < NX; ++i) d4[i] = fma(h, dk[i], dx[i]);
            // stage 4 .. 1: first-order adjoint stage weight w_j of lambda_{k+1}, then the second-order chain
            // (zero incoming second-order adjoint: that part is AB' mu)
            double w[NX], ax[NX], au[NU], bx[NX], bu[NU], cx[NX], cu[NU], sx[NX], su[NU], dw[NX];
#pragma unroll
            for (int i = 0; i < NX; ++i) w[i] = h6 * l[i];
            F::fc_hvjp(xk, uk, s4, w, d4, vk, cx, cu);
            F::fc_vjp(xk, uk, s4, w, ax, au);
#pragma unroll
            for (int i = 0; i < NX; ++i) { sx[i] = cx[i]; dw[i] = h * cx[i]; w[i] = fma(h, ax[i], h3 * l[i]); }
#pragma unroll
            for (int j = 0; j < NU; ++j) su[j] = cu[j];
            F::fc_vjp(xk, uk, s3, dw, bx, bu);
            F::fc_hvjp(xk, uk, s3, w, d3, vk, cx, cu);
            F::fc_vjp(xk, uk, s3, w, ax, au);
#pragma unroll
            for (int i = 0; i < NX; ++i) {
                bx[i] = bx[i] + cx[i]; sx[i] = sx[i] + bx[i]; dw[i] = hh * bx[i]; w[i] = fma(hh, ax[i], h3 * l[i]);
            }
#pragma unroll
            for (int j = 0; j < NU; ++j) su[j] = su[j] + (bu[j] + cu[j]);
            F::fc_vjp(xk, uk, s2, dw, bx, bu);
            F::fc_hvjp(xk, uk, s2, w, d2, vk, cx, cu);
            F::fc_vjp(xk, uk, s2, w, ax, au);
#pragma unroll
            for (int i = 0; i < NX; ++i) {
                bx[i] = bx[i] + cx[i]; sx[i] = sx[i] + bx[i]; dw[i] = hh * bx[i]; w[i] = fma(hh, ax[i], h6 * l[i]);
            }
#pragma unroll
            for (int j = 0; j < NU; ++j) su[j] = su[j] + (bu[j] + cu[j]);
            F::fc_vjp(xk, uk, s1, dw, bx, bu);
            F::fc_hvjp(xk, uk, s1, w, dx, vk, cx, cu);
#pragma unroll
            for (int i = 0; i < NX; ++i) sx[i] = sx[i] + (bx[i] + cx[i]);
#pragma unroll
            for (int j = 0; j < NU; ++j) su[j] = su[j] + (bu[j] + cu[j]);
            if (k > 0) {
                double nd[NX];
                node_second_order(k, cost_scale, nd);
#pragma unroll
                for (int i = 0; i < NX; ++i) mus[k * NX + i] = sx[i] + nd[i];
            }
#pragma unroll
            for (int j = 0; j < NU; ++j) usrc[k * NU + j] = fma(cost_scale * (2 * wu(j)), vk[j], su[j]);
        }
        __syncwarp();
    }
    __device__ __noinline__ void mu_recurrence(double *Hv) { adjoint_recurrence(mus, Hv, SrcU{usrc}); }
    /// Hv = hess(s f + AL terms)(u) v; needs the caches of the last gradient at u (AB, lam, yh, wact, sc, traj).
    __device__ __forceinline__ void hvp(const double *u, const double *v, double cost_scale, double *Hv) {
        tangent(v);
        second_order_sources(u, v, cost_scale);
        mu_recurrence(Hv);
    }
    /// step Jacobians + adjoint recurrence at the point of the last rollout, cost scaled by s
    __device__ __forceinline__ void gradient_caches(const double *x, double s, double *grad) {
        build_AB(x);
        adjoint(x, s, grad);
    }

    // ---- the problem functions the solver calls (CasADiProblem::eval_*) ----
    __device__ __noinline__ double eval_f(const double *u) {
        __syncwarp();
        hvp_ready = false;
        rollout(u);
        const double f = cost(u);
        __syncwarp();
        return f;
    }
    __device__ __noinline__ void eval_g(const double *u, double *g) {
        __syncwarp();
        hvp_ready = false;
        rollout(u);
        constraints(g);
        __syncwarp();
    }
    __device__ __noinline__ double eval_psi(const double *x, const double *y, const double *S, double *yhat) {
        __syncwarp();
        hvp_ready = false;
        rollout(x);
        const double f = cost(x);
        return f + al_terms(y, S, yhat, nullptr);
    }
    __device__ __noinline__ double eval_psi_grad_psi(const double *x, const double *y, const double *S, double *grad) {
        __syncwarp();
        rollout(x);
        const double f = cost(x);
        const double psi = f + al_terms(y, S, yh, wact);
        gradient_caches(x, 1.0, grad);
        hvp_ready = true;
        __syncwarp();
        return psi;
    }
    /// hess(s f + 1/2 y_hat' d) v (casadi_generator/__init__.py:69,181-189).  Hv doubles as scratch.
    __device__ __noinline__ void hess_psi_prod(const double *x, const double *y, const double *S, const double *v, double *Hv,
                                               double scale = 1.0) {
        __syncwarp();
        if (!hvp_ready || scale != 1.0) { // with s != 1 the multipliers lambda of the cost part change
            rollout(x);
            (void)al_terms(y, S, yh, wact);
            gradient_caches(x, scale, Hv);
            hvp_ready = scale == 1.0;
        }
        hvp(x, v, scale, Hv);
        __syncwarp();
    }
    // ---- problem plugin face (alpaqa_problem_functions_t members that PANTR itself does not call) ----
    /// y_hat := y (zero if null), no active quadratic penalty: the adjoint / hvp sweeps then give grad L / hess L
    __device__ void set_multipliers(const double *y) {
        const int mm = m(N);
        for (int i = lane_id(); i < mm; i += 32) { yh[i] = y ? y[i] : 0.0; wact[i] = 0.0; }
        __syncwarp();
    }
    /// grad L(x, y) = grad f + J_g' y (CasADiProblem::eval_grad_L, CasADiProblem.tpp:268-276); returns f
    __device__ __noinline__ double eval_grad_L(const double *x, const double *y, double *grad) {
        __syncwarp();
        hvp_ready = false;
        rollout(x);
        const double f = cost(x);
        set_multipliers(y);
        gradient_caches(x, 1.0, grad);
        __syncwarp();
        return f;
    }
    /// J_g(x)' y (CasADiProblem::eval_grad_g_prod, CasADiProblem.tpp:226-233)
    __device__ __noinline__ void eval_grad_g_prod(const double *x, const double *y, double *out) {
        __syncwarp();
        hvp_ready = false;
        rollout(x);
        set_multipliers(y);
        gradient_caches(x, 0.0, out);
        __syncwarp();
    }
    /// hess_xx (s f + y' g)(x) v (CasADiProblem::eval_hess_L_prod, CasADiProblem.tpp:300-313; casadi_generator:65)
    __device__ __noinline__ void hess_L_prod(const double *x, const double *y, const double *v, double *Hv, double scale = 1.0) {
        __syncwarp();
        hvp_ready = false;
        rollout(x);
        set_multipliers(y);
        gradient_caches(x, scale, Hv);
        hvp(x, v, scale, Hv);
        __syncwarp();
    }
    /// Plant step x+ = f_d(x, u) (problems/problems/formulation.hpp:258-261)
    __device__ static void plant_step(const double *x, const double *u, double *xn) {
        double xx[NX], uu[NU];
        qc::ldv<NX>(x, xx); qc::ldv<NU>(u, uu);
        rk4_step(xx, uu, nullptr);
#pragma unroll
        for (int i = 0; i < NX; ++i) xn[i] = xx[i];
    }
};

using QuadcopterC          = QuadCondensedModel<qc::SimpleFns>;
using RealisticQuadcopterC = QuadCondensedModel<qc::RealisticFns>;

} // namespace pantr
