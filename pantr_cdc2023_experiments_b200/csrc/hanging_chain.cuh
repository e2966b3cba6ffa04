// Hanging-chain MPC model as warp-cooperative device code (one warp = one instance).
//
// Replaces the CasADi-generated straight-line C (f, f_grad_f, psi_grad_psi, hess_psi_prod) that
// CasADiProblem::eval_* calls (alpaqa/src/interop/casadi/.../CasADiProblem.tpp:208-368) for the
// model of python/alpaqa_mpc_benchmarks/problems/hanging_chain.py:7-98 in the single-shooting
// transcription of python/alpaqa_mpc_benchmarks/formulations/ss.py:20-55.
//
// Lane map: l = 3*i + d, link/ball i = 0..9, dimension d = 0..2 (lanes 30,31 idle).
//   X = position of ball i (state entry l), V = velocity of ball i (state entry 30+l, l < 27).
//   Lanes 27..29 (the free end, whose velocity is the input) carry u[d] in the "V slot".
// The state trajectory (N+1) x 57 and its tangent live in the instance workspace (global memory,
// L2 resident); stage values are recomputed in the reverse sweeps (RK4 checkpointing per step).
#pragma once
#include "warp_utils.cuh"

namespace pantr {

struct HangingChain {
    static constexpr int NX = 57, NU = 3, NL = 30, NVL = 27, NC = 0;
    static constexpr double alpha = 25, beta = 1, gamma_u = 0.01, mass = 0.03, Dspring = 0.1,
                            Lspring = 0.033, v_max = 1, g_grav = 9.81, Ts = 0.1;
    __host__ __device__ static int n(int N) { return NU * N; }
    __host__ __device__ static int m(int) { return 0; }
    __host__ __device__ static int traj_doubles(int N) { return 2 * (N + 1) * NX; }
    __host__ __device__ static double C_lb(int) { return -v_max; }
    __host__ __device__ static double C_ub(int) { return +v_max; }
    __host__ __device__ static double D_lb(int, int) { return 0; }
    __host__ __device__ static double D_ub(int, int) { return 0; }
    __device__ void eval_f_g_sqnorm(const double *u, double &f, double &g2) { f = rollout(u); g2 = 0; }
    static void initial_state(double *x) { // hanging_chain.py:92-96
        for (int i = 0; i < NX; ++i) x[i] = 0;
        for (int i = 0; i < 10; ++i) x[3 * i] = double(i + 1) / 10;
    }
    static void initial_guess(double *u, int N) { for (int i = 0; i < NU * N; ++i) u[i] = 0; }

    struct Link { double dist, c, e, iw2; };

    // per-lane constants
    int l, d, base;
    bool hasV, isEnd, active;
    double grav, ref;
    double inv_m, DL, h, hh, h6, h3, two_alpha, two_beta, two_gamma;
    int N;
    const double *x0;
    double *traj, *dtraj;

    __device__ void init(int N_, const double *x0_, double *ws_traj) {
        l = lane_id(); d = l % 3; base = l - d;
        hasV = l < NVL; active = l < NL; isEnd = active && !hasV;
        grav = d == 2 ? -g_grav : 0.0; ref = d == 0 ? 1.0 : 0.0;
        inv_m = 1.0 / mass; DL = Dspring * Lspring;
        h = Ts; hh = Ts / 2; h6 = Ts / 6; h3 = 2 * (Ts / 6);
        two_alpha = 2 * alpha; two_beta = 2 * beta; two_gamma = 2 * gamma_u;
        N = N_; x0 = x0_; traj = ws_traj; dtraj = ws_traj + (N_ + 1) * NX;
    }

    __device__ __forceinline__ double dot3(double p) const {
        double a = shfl(p, base), b = shfl(p, base + 1), c = shfl(p, base + 2);
        return (a + b) + c;
    }
    // continuous dynamics (hanging_chain.py:47-56). Vu = V for l<27, u[d] for lanes 27..29.
    __device__ __forceinline__ void fc(double X, double Vu, double &kX, double &kV, Link &ld) const {
        double Xm = shfl(X, l - 3);
        ld.dist   = l >= 3 ? X - Xm : X;
        double sq = ld.dist * ld.dist;
        double n2 = dot3(sq);
        if (!active) n2 = 1.0;
        double w = sqrt(n2), iw = 1.0 / w;
        ld.c   = Dspring * (1.0 - Lspring * iw);
        ld.iw2 = iw * iw;
        ld.e   = (DL * iw) * ld.iw2;
        double F  = ld.dist * ld.c;
        double Fu = shfl(F, l + 3);
        kX = Vu;
        kV = (Fu - F) * inv_m + grav;
    }
    // (dfc/d(s,u))^T w ; aVu = adjoint of V (l<27) / of u (lanes 27..29)
    __device__ __forceinline__ void vjp(const Link &ld, double wX, double wV, double &aX, double &aVu,
                                        double &Fbar, double &dotv) const {
        double wVm = shfl(wV, l - 3);
        Fbar       = ((l >= 3 ? wVm : 0.0) - (hasV ? wV : 0.0)) * inv_m;
        dotv       = dot3(ld.dist * Fbar);
        double db  = ld.c * Fbar + (ld.e * dotv) * ld.dist;
        double dbu = shfl(db, l + 3);
        aX  = hasV ? db - dbu : db;
        aVu = wX;
    }
    // (dfc/d(s,u)) (ds,du)
    __device__ __forceinline__ void jvp(const Link &ld, double dX, double dVu, double &dkX, double &dkV,
                                        double &ddist, double &dd) const {
        double dXm = shfl(dX, l - 3);
        ddist      = l >= 3 ? dX - dXm : dX;
        dd         = dot3(ld.dist * ddist);
        double dF  = ld.c * ddist + (ld.e * dd) * ld.dist;
        double dFu = shfl(dF, l + 3);
        dkX = dVu;
        dkV = (dFu - dF) * inv_m;
    }
    // second-order adjoint of one stage
    __device__ __forceinline__ void hvjp(const Link &ld, double Fbar, double dotv, double ddist, double dd,
                                         double dwX, double dwV, double &bX, double &bVu) const {
        double dwVm  = shfl(dwV, l - 3);
        double dFbar = ((l >= 3 ? dwVm : 0.0) - (hasV ? dwV : 0.0)) * inv_m;
        double df    = dot3(ddist * Fbar);
        double dot2  = dot3(ld.dist * dFbar);
        double dc    = ld.e * dd;
        double de    = ((-3.0 * ld.e) * ld.iw2) * dd;
        double lin   = ld.c * dFbar + (ld.e * dot2) * ld.dist;
        double sec   = (dc * Fbar + (de * dotv + ld.e * df) * ld.dist) + (ld.e * dotv) * ddist;
        double ddb   = lin + sec;
        double ddbu  = shfl(ddb, l + 3);
        bX  = hasV ? ddb - ddbu : ddb;
        bVu = dwX;
    }

    __device__ __forceinline__ void load_state(const double *t, int k, double &X, double &V) const {
        const double *s = t + k * NX;
        X = active ? s[l] : 0.0;
        V = hasV ? s[NL + l] : 0.0;
    }
    __device__ __forceinline__ void store_state(double *t, int k, double X, double V) const {
        double *s = t + k * NX;
        if (active) s[l] = X;
        if (hasV) s[NL + l] = V;
    }
    __device__ __forceinline__ double load_u(const double *u, int k) const {
        return isEnd ? u[NU * k + (l - NVL)] : 0.0;
    }

    // base stage link data of one RK4 step (X,V = s_k ; ud = u_k on lanes 27..29)
    __device__ __forceinline__ void stages(double X, double V, double ud, Link (&ld)[4]) const {
        double Vu = hasV ? V : ud, kX, kV;
        fc(X, Vu, kX, kV, ld[0]);
        double tX = X + hh * kX, tV = V + hh * kV;
        fc(tX, hasV ? tV : ud, kX, kV, ld[1]);
        tX = X + hh * kX; tV = V + hh * kV;
        fc(tX, hasV ? tV : ud, kX, kV, ld[2]);
        tX = X + h * kX; tV = V + h * kV;
        fc(tX, hasV ? tV : ud, kX, kV, ld[3]);
    }

    // One RK4 step (hanging_chain.py:58-66)
    __device__ __forceinline__ void rk4(double X, double V, double ud, double &Xn, double &Vn) const {
        Link ld;
        double k1X, k1V, k2X, k2V, k3X, k3V, k4X, k4V;
        fc(X, hasV ? V : ud, k1X, k1V, ld);
        double tX = X + hh * k1X, tV = V + hh * k1V;
        fc(tX, hasV ? tV : ud, k2X, k2V, ld);
        tX = X + hh * k2X; tV = V + hh * k2V;
        fc(tX, hasV ? tV : ud, k3X, k3V, ld);
        tX = X + h * k3X; tV = V + h * k3V;
        fc(tX, hasV ? tV : ud, k4X, k4V, ld);
        Xn = X + h6 * (((k1X + 2.0 * k2X) + 2.0 * k3X) + k4X);
        Vn = V + h6 * (((k1V + 2.0 * k2V) + 2.0 * k3V) + k4V);
    }

    // Rollout + cost (ss.py:24-37); fills traj.  Returns f on all lanes.
    __device__ double rollout(const double *u) {
        double acc = 0, X, V;
        X = active ? x0[l] : 0.0;
        V = hasV ? x0[NL + l] : 0.0;
        store_state(traj, 0, X, V);
        double ud = load_u(u, 0);
        for (int k = 0; k < N; ++k) {
            double un = k + 1 < N ? load_u(u, k + 1) : 0.0;
            if (hasV) acc = acc + beta * (V * V);
            else if (isEnd) { double dx = X - ref; acc = acc + (alpha * (dx * dx) + gamma_u * (ud * ud)); }
            double Xn, Vn;
            rk4(X, V, ud, Xn, Vn);
            X = Xn; V = Vn; ud = un;
            store_state(traj, k + 1, X, V);
        }
        if (hasV) acc = acc + beta * (V * V);
        else if (isEnd) { double dx = X - ref; acc = acc + alpha * (dx * dx); }
        return warp_sum(acc);
    }
#include "hanging_chain_sweeps.cuh"
};

} // namespace pantr
