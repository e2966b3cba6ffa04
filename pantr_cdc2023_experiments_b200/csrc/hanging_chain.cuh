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
//
// Per-instance model workspace (global memory, L1/L2 resident):
//   traj  (N+1) x 57   state trajectory of the last rollout
//   dtraj (N+1) x 57   tangent trajectory of the last Hessian-vector product
//   iwc   N x 4 x 10   1/|dist| of every link at every RK4 stage point of the last rollout: the
//                      adjoint and Hessian-vector sweeps rebuild the stage points from it without any
//                      sqrt / division (same bits as recomputing: identical operations on identical
//                      inputs), which is what keeps them FP64-add/mul bound.
//   adjc  N x 4 x 40   first-order adjoint data of every stage point (Fbar per lane, dist . Fbar per
//                      link), written by the adjoint sweep: they depend on u only, and PANTR takes
//                      ~30 Hessian-vector products per gradient, so the products' reverse sweep reads
//                      them instead of carrying the lambda recurrence (no vjp, 25 % fewer instructions).
//   dic   N x 160      link vectors `dist` of the 4 stage points (4 x 30) + the step's iwc block (4 x 10),
//                      written by the adjoint sweep: the Hessian-vector product reads its stage points from
//                      here (PANTR_HC_DIST_CACHE, default) and stages dic / adjc / dtraj rows through
//                      shared memory with cp.async; dtraj rows are then padded to 58 doubles and dtraj
//                      comes first in the workspace (16-byte alignment of every staged block).
// The three sweeps are __noinline__ so that the solver contains exactly one copy of each: the
// solve kernel's code then fits the SM instruction caches (the fully inlined version was 1.1 MB of
// SASS and stalled 14 of 19 cycles per issue on instruction fetch: profiles/r01_v0_*).
#pragma once
#include "warp_utils.cuh"
#include "hc_formulas.h"

namespace pantr {
namespace hc {

constexpr int NX = 57, NU = 3, NL = 30, NVL = 27, NLINK = 10;
// hanging_chain.py:9-18
constexpr double alpha = 25, beta = 1, gamma_u = 0.01, mass = 0.03, Dspring = 0.1, Lspring = 0.033,
                 v_max = 1, g_grav = 9.81, Ts = 0.1;
constexpr double inv_m = 1.0 / mass, DL = Dspring * Lspring;
constexpr double h = Ts, hh = Ts / 2, h6 = Ts / 6, h3 = 2 * (Ts / 6);
constexpr double two_alpha = 2 * alpha, two_beta = 2 * beta, two_gamma = 2 * gamma_u;

struct Lane {
    int l, base, link;
    bool hasV, isEnd, active;
    double grav, ref;
    __device__ __forceinline__ Lane() {
        l = lane_id();
        int d = l % 3;
        base = l - d;
        link = l / 3 < NLINK ? l / 3 : NLINK - 1;
        hasV = l < NVL; active = l < NL; isEnd = active && !hasV;
        grav = d == 2 ? -g_grav : 0.0;
        ref  = d == 0 ? 1.0 : 0.0;
    }
    __device__ __forceinline__ double dot3(double p) const {
        double a = shfl(p, base), b = shfl(p, base + 1), c = shfl(p, base + 2);
        return (a + b) + c;
    }
};

struct Link { double dist, c, e; };

// continuous dynamics (hanging_chain.py:47-56). Vu = V for l<27, u[d] for lanes 27..29.
// Returns iw = 1/|dist| of the lane's link.
__device__ __forceinline__ double fc(const Lane &L, double X, double Vu, double &kX, double &kV) {
    double Xm   = shfl(X, L.l - 3);
    double dist = L.l >= 3 ? X - Xm : X;
    double n2   = L.dot3(dist * dist);
    if (!L.active) n2 = 1.0;
    double w = sqrt(n2), iw = 1.0 / w;
    double c  = hcf::spring_c(iw);
    double F  = dist * c;
    double Fu = shfl(F, L.l + 3);
    kX = Vu;
    kV = hcf::accel(Fu, F, L.grav);
    return iw;
}
// the same stage evaluation from the cached iw (no sqrt, no division)
__device__ __forceinline__ void fc_cached(const Lane &L, double X, double Vu, double iw, double &kX,
                                          double &kV, Link &ld, double &iw2) {
    double Xm = shfl(X, L.l - 3);
    ld.dist   = L.l >= 3 ? X - Xm : X;
    ld.c      = hcf::spring_c(iw);
    iw2       = iw * iw;
    ld.e      = hcf::spring_e(iw, iw2);
    double F  = ld.dist * ld.c;
    double Fu = shfl(F, L.l + 3);
    kX = Vu;
    kV = hcf::accel(Fu, F, L.grav);
}
// (dfc/d(s,u))^T w ; aVu = adjoint of V (l<27) / of u (lanes 27..29)
__device__ __forceinline__ void vjp(const Lane &L, const Link &ld, double wX, double wV, double &aX,
                                    double &aVu, double &Fbar, double &dotv) {
    double wVm = shfl(wV, L.l - 3);
    Fbar       = ((L.l >= 3 ? wVm : 0.0) - (L.hasV ? wV : 0.0)) * inv_m;
    dotv       = L.dot3(ld.dist * Fbar);
    double db  = hcf::force_lin(ld.c, Fbar, ld.e, dotv, ld.dist);
    double dbu = shfl(db, L.l + 3);
    aX  = L.hasV ? db - dbu : db;
    aVu = wX;
}
// (dfc/d(s,u)) (ds,du)
__device__ __forceinline__ void jvp(const Lane &L, const Link &ld, double dX, double dVu, double &dkX,
                                    double &dkV, double &ddist, double &dd) {
    double dXm = shfl(dX, L.l - 3);
    ddist      = L.l >= 3 ? dX - dXm : dX;
    dd         = L.dot3(ld.dist * ddist);
    double dF  = hcf::force_lin(ld.c, ddist, ld.e, dd, ld.dist);
    double dFu = shfl(dF, L.l + 3);
    dkX = dVu;
    dkV = (dFu - dF) * inv_m;
}
// second-order adjoint of one stage
__device__ __forceinline__ void hvjp(const Lane &L, const Link &ld, double iw2, double Fbar, double dotv,
                                     double ddist, double dd, double dwX, double dwV, double &bX,
                                     double &bVu) {
    double dwVm  = shfl(dwV, L.l - 3);
    double dFbar = ((L.l >= 3 ? dwVm : 0.0) - (L.hasV ? dwV : 0.0)) * inv_m;
    double df    = L.dot3(ddist * Fbar);
    double dot2  = L.dot3(ld.dist * dFbar);
    double dc    = ld.e * dd;
    double de    = ((-3.0 * ld.e) * iw2) * dd;
    double lin   = hcf::force_lin(ld.c, dFbar, ld.e, dot2, ld.dist);
    double sec   = hcf::force_sec(dc, Fbar, de, dotv, ld.e, df, ld.dist, ddist);
    double ddb   = lin + sec;
    double ddbu  = shfl(ddb, L.l + 3);
    bX  = L.hasV ? ddb - ddbu : ddb;
    bVu = dwX;
}

__device__ __forceinline__ void load_state(const Lane &L, const double *t, int k, double &X, double &V) {
    const double *s = t + k * NX;
    X = L.active ? __ldcg(s + L.l) : 0.0; // streamed once per sweep: L2 only, L1 stays with the solver vectors
    V = L.hasV ? __ldcg(s + NL + L.l) : 0.0;
}
__device__ __forceinline__ void store_state(const Lane &L, double *t, int k, double X, double V) {
    double *s = t + k * NX;
    if (L.active) __stcg(s + L.l, X);
    if (L.hasV) __stcg(s + NL + L.l, V);
}
__device__ __forceinline__ double load_u(const Lane &L, const double *u, int k) {
    return L.isEnd ? u[NU * k + (L.l - NVL)] : 0.0;
}

// One RK4 step (hanging_chain.py:58-66); iw_out (optional): the 4 stage values of 1/|dist|
__device__ __forceinline__ void rk4(const Lane &L, double X, double V, double ud, double &Xn, double &Vn,
                                    double *iw_out) {
    double k1X, k1V, k2X, k2V, k3X, k3V, k4X, k4V;
    const bool wr = iw_out != nullptr && L.active && L.l == L.base;
    double iw = fc(L, X, L.hasV ? V : ud, k1X, k1V);
    if (wr) __stcg(iw_out + L.link, iw);
    double tX = hcf::axpy(hh, k1X, X), tV = hcf::axpy(hh, k1V, V);
    iw = fc(L, tX, L.hasV ? tV : ud, k2X, k2V);
    if (wr) __stcg(iw_out + NLINK + L.link, iw);
    tX = hcf::axpy(hh, k2X, X); tV = hcf::axpy(hh, k2V, V);
    iw = fc(L, tX, L.hasV ? tV : ud, k3X, k3V);
    if (wr) __stcg(iw_out + 2 * NLINK + L.link, iw);
    tX = hcf::axpy(h, k3X, X); tV = hcf::axpy(h, k3V, V);
    iw = fc(L, tX, L.hasV ? tV : ud, k4X, k4V);
    if (wr) __stcg(iw_out + 3 * NLINK + L.link, iw);
    Xn = hcf::rk4c(X, k1X, k2X, k3X, k4X);
    Vn = hcf::rk4c(V, k1V, k2V, k3V, k4V);
}

// base stage link data of one RK4 step from the cached iw (X,V = s_k ; ud = u_k on lanes 27..29)
struct IW { double i0, i1, i2, i3; };
__device__ __forceinline__ IW load_iw(const Lane &L, const double *iwc, int k) {
    const double *p = iwc + k * (4 * NLINK) + L.link;
    return IW{__ldcg(p), __ldcg(p + NLINK), __ldcg(p + 2 * NLINK), __ldcg(p + 3 * NLINK)};
}
// Per-stage caches, 40 doubles per stage point (4 per RK4 step): entries 0..29 are lane values,
// 30..39 the per-link scalars (written by the link's base lane).
constexpr int SC = NL + NLINK, SCSTEP = 4 * SC;
__device__ __forceinline__ void store_sc(const Lane &L, double *c, double lanev, double linkv) {
    if (L.active) __stcg(c + L.l, lanev);
    if (L.active && L.l == L.base) __stcg(c + NL + L.link, linkv);
}
__device__ __forceinline__ void load_sc(const Lane &L, const double *c, double &lanev, double &linkv) {
    lanev = L.active ? c[L.l] : 0.0;
    linkv = c[NL + L.link];
}
// ---- shared-memory staging of the reverse sweep's per-step cache block (adjc: 160 doubles, iwc: 40) ----
// cp.async (LDGSTS, L2 -> shared, bypassing L1) fetches step k-1 while step k computes; the
// L1-prefetch version still stalled ~1300 cycles per step on the first use of iw / Fbar
// (profiles/r01_v7_*: 10 % of all warp samples on one DFMA).
constexpr int STAGE_D = SCSTEP + 4 * NLINK; // 200 doubles
__device__ __forceinline__ void cp_async16(double *smem, const double *g) {
    unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(g) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void stage_fetch(const Lane &L, double *buf, const double *ac, const double *iw) {
    const int l = lane_id();
    cp_async16(buf + 2 * l, ac + 2 * l);
    cp_async16(buf + 2 * (l + 32), ac + 2 * (l + 32));
    if (l < 16) cp_async16(buf + 2 * (l + 64), ac + 2 * (l + 64));
    if (l < 20) cp_async16(buf + SCSTEP + 2 * l, iw + 2 * l);
    (void)L;
}
// ---- PANTR_HC_DIST_CACHE: the Hessian-vector product reads the stage points' link vectors from a cache ----
// dic  N x 160: per RK4 step the 4 x 30 lane values of `dist` at the stage points followed by the 4 x 10
// values of 1/|dist| (a copy of the step's iwc block, so that one contiguous block feeds both sweeps of
// the product).  Written by the adjoint sweep (depends on u only, like adjc); with it the product does
// not rebuild the stage points (no state trajectory reads, no k_j chain, 8 shuffles fewer per stage).
#ifndef PANTR_HC_TMA
#define PANTR_HC_TMA 0 // 1: the per-step blocks of the product arrive by cp.async.bulk + mbarrier (one elected lane) instead of cp.async
#endif
#ifndef PANTR_HC_DIST_CACHE
#define PANTR_HC_DIST_CACHE 1 // 0: the product rebuilds the stage points from traj + iwc (the r01_v9 kernel, kept for A/B runs)
#endif
constexpr int DSTEP = 4 * NL + 4 * NLINK; // 160 doubles
__device__ __forceinline__ void link_cached(double dist, double iw, Link &ld, double &iw2) {
    ld.dist = dist;
    ld.c    = hcf::spring_c(iw);
    iw2     = iw * iw;
    ld.e    = hcf::spring_e(iw, iw2);
}
/// 160 doubles global -> shared with cp.async (80 chunks of 16 bytes)
__device__ __forceinline__ void fetch160(double *buf, const double *g) {
    const int l = lane_id();
    cp_async16(buf + 2 * l, g + 2 * l);
    cp_async16(buf + 2 * (l + 32), g + 2 * (l + 32));
    if (l < 16) cp_async16(buf + 2 * (l + 64), g + 2 * (l + 64));
}
__device__ __forceinline__ void cp_async_wait_1() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }
__device__ __forceinline__ void stages_cached(const Lane &L, double X, double V, double ud, const IW &iw,
                                              Link (&ld)[4], double (&iw2)[4]) {
    double Vu = L.hasV ? V : ud, kX, kV;
    const double i0 = iw.i0, i1 = iw.i1, i2 = iw.i2, i3 = iw.i3;
    fc_cached(L, X, Vu, i0, kX, kV, ld[0], iw2[0]);
    double tX = hcf::axpy(hh, kX, X), tV = hcf::axpy(hh, kV, V);
    fc_cached(L, tX, L.hasV ? tV : ud, i1, kX, kV, ld[1], iw2[1]);
    tX = hcf::axpy(hh, kX, X); tV = hcf::axpy(hh, kV, V);
    fc_cached(L, tX, L.hasV ? tV : ud, i2, kX, kV, ld[2], iw2[2]);
    tX = hcf::axpy(h, kX, X); tV = hcf::axpy(h, kV, V);
    fc_cached(L, tX, L.hasV ? tV : ud, i3, kX, kV, ld[3], iw2[3]);
}

// Rollout + cost (ss.py:24-37); fills traj and iwc.  Returns f on all lanes.
__device__ __noinline__ double rollout(int N, const double *x0, const double *u, double *traj, double *iwc) {
    __builtin_assume(__isGlobal(x0)); __builtin_assume(__isGlobal(u)); __builtin_assume(__isGlobal(traj));
    __builtin_assume(__isGlobal(iwc));
    const Lane L;
    double acc = 0, X, V;
    X = L.active ? x0[L.l] : 0.0;
    V = L.hasV ? x0[NL + L.l] : 0.0;
    store_state(L, traj, 0, X, V);
    double ud = load_u(L, u, 0);
#pragma unroll 1
    for (int k = 0; k < N; ++k) {
        double un = k + 1 < N ? load_u(L, u, k + 1) : 0.0;
        if (L.hasV) acc = fma(beta, V * V, acc);
        else if (L.isEnd) { double dx = X - L.ref; acc = acc + fma(alpha, dx * dx, gamma_u * (ud * ud)); }
        double Xn, Vn;
        rk4(L, X, V, ud, Xn, Vn, iwc + k * (4 * NLINK));
        X = Xn; V = Vn; ud = un;
        store_state(L, traj, k + 1, X, V);
    }
    if (L.hasV) acc = fma(beta, V * V, acc);
    else if (L.isEnd) { double dx = X - L.ref; acc = fma(alpha, dx * dx, acc); }
    __syncwarp();
    return warp_sum(acc);
}

// grad f by discrete adjoint; traj / iwc hold the rollout of `u`
// Also fills adjc (N x 4 x 40): Fbar (lane) and dist . Fbar (link) of every stage, which depend on
// u only and are reused by every Hessian-vector product at this u (~30 per gradient).
__device__ __noinline__ void adjoint(int N, const double *u, const double *traj, const double *iwc,
                                     double *grad, double *adjc, double *dic) {
    __builtin_assume(__isGlobal(dic));
    __builtin_assume(__isGlobal(u)); __builtin_assume(__isGlobal(traj)); __builtin_assume(__isGlobal(iwc));
    __builtin_assume(__isGlobal(adjc)); // grad may be shared memory
    const Lane L;
    double X, V, lX, lV;
    load_state(L, traj, N, X, V);
    lX = L.isEnd ? two_alpha * (X - L.ref) : 0.0;
    lV = L.hasV ? two_beta * V : 0.0;
    load_state(L, traj, N - 1, X, V);
    double ud = load_u(L, u, N - 1);
    IW iw = load_iw(L, iwc, N - 1);
#pragma unroll 1
    for (int k = N - 1; k >= 0; --k) {
        // prefetch the next (earlier) step while this one computes
        double Xp = 0, Vp = 0, up = 0;
        IW iwp = iw;
        if (k > 0) { load_state(L, traj, k - 1, Xp, Vp); up = load_u(L, u, k - 1); iwp = load_iw(L, iwc, k - 1); }
        Link ld[4]; double iw2[4];
        stages_cached(L, X, V, ud, iw, ld, iw2);
#if PANTR_HC_DIST_CACHE
        {
            double *dc = dic + k * DSTEP;
            if (L.active) {
#pragma unroll
                for (int j = 0; j < 4; ++j) __stcg(dc + j * NL + L.l, ld[j].dist);
                if (L.l == L.base) {
                    __stcg(dc + 4 * NL + L.link, iw.i0); __stcg(dc + 4 * NL + NLINK + L.link, iw.i1);
                    __stcg(dc + 4 * NL + 2 * NLINK + L.link, iw.i2); __stcg(dc + 4 * NL + 3 * NLINK + L.link, iw.i3);
                }
            }
        }
#endif
        double wX, wV, aX, aV, sX, sV, Fb, dt;
        wX = h6 * lX; wV = h6 * lV;
        double *ac = adjc + k * SCSTEP;
        vjp(L, ld[3], wX, wV, aX, aV, Fb, dt);
        store_sc(L, ac + 3 * SC, Fb, dt);
        sX = lX + aX; sV = lV + aV;
        wX = hcf::lin2(h3, lX, h, aX); wV = hcf::lin2(h3, lV, h, aV);
        vjp(L, ld[2], wX, wV, aX, aV, Fb, dt);
        store_sc(L, ac + 2 * SC, Fb, dt);
        sX = sX + aX; sV = sV + aV;
        wX = hcf::lin2(h3, lX, hh, aX); wV = hcf::lin2(h3, lV, hh, aV);
        vjp(L, ld[1], wX, wV, aX, aV, Fb, dt);
        store_sc(L, ac + SC, Fb, dt);
        sX = sX + aX; sV = sV + aV;
        wX = hcf::lin2(h6, lX, hh, aX); wV = hcf::lin2(h6, lV, hh, aV);
        vjp(L, ld[0], wX, wV, aX, aV, Fb, dt);
        store_sc(L, ac, Fb, dt);
        sX = sX + aX; sV = sV + aV;
        lX = L.isEnd ? fma(two_alpha, X - L.ref, sX) : sX;
        if (L.hasV) lV = fma(two_beta, V, sV);
        else {
            if (L.isEnd) grad[NU * k + (L.l - NVL)] = fma(two_gamma, ud, sV);
            lV = 0.0;
        }
        X = Xp; V = Vp; ud = up; iw = iwp;
    }
#if PANTR_HC_TMA
    __threadfence(); // dic / adjc are read through the async proxy (TMA reads L2) by the products that follow
#endif
}

// Hv = hess f(u) v by forward-over-adjoint (the exact Hessian-vector product that
// casadi_generator/__init__.py:181-189 defines as hess_psi_prod; with m = 0, psi = f).
// traj / iwc hold the rollout of `u`; dtraj is scratch for the tangent trajectory.
#if !PANTR_HC_DIST_CACHE
__device__ __noinline__ void hvp(int N, const double *u, const double *v, const double *traj, double *dtraj,
                                 const double *iwc, const double *adjc, const double *, double *Hv) {
    __builtin_assume(__isGlobal(u)); __builtin_assume(__isGlobal(traj)); // v, Hv may be shared memory
    __builtin_assume(__isGlobal(dtraj)); __builtin_assume(__isGlobal(iwc));
    __builtin_assume(__isGlobal(adjc));
    const Lane L;
    double X, V, dX = 0.0, dV = 0.0;
    // ---- forward tangent sweep ----
    store_state(L, dtraj, 0, dX, dV);
    load_state(L, traj, 0, X, V);
    double ud = load_u(L, u, 0), vd = load_u(L, v, 0);
    IW iw = load_iw(L, iwc, 0);
#pragma unroll 1
    for (int k = 0; k < N; ++k) {
        double Xp, Vp, up = 0, vp = 0;
        IW iwp = iw;
        load_state(L, traj, k + 1, Xp, Vp);
        if (k + 1 < N) { up = load_u(L, u, k + 1); vp = load_u(L, v, k + 1); iwp = load_iw(L, iwc, k + 1); }
        Link ld[4]; double iw2[4];
        stages_cached(L, X, V, ud, iw, ld, iw2);
        double k1X, k1V, k2X, k2V, k3X, k3V, k4X, k4V, ddist, dd;
        jvp(L, ld[0], dX, L.hasV ? dV : vd, k1X, k1V, ddist, dd);
        double tX = hcf::axpy(hh, k1X, dX), tV = hcf::axpy(hh, k1V, dV);
        jvp(L, ld[1], tX, L.hasV ? tV : vd, k2X, k2V, ddist, dd);
        tX = hcf::axpy(hh, k2X, dX); tV = hcf::axpy(hh, k2V, dV);
        jvp(L, ld[2], tX, L.hasV ? tV : vd, k3X, k3V, ddist, dd);
        tX = hcf::axpy(h, k3X, dX); tV = hcf::axpy(h, k3V, dV);
        jvp(L, ld[3], tX, L.hasV ? tV : vd, k4X, k4V, ddist, dd);
        dX = hcf::rk4c(dX, k1X, k2X, k3X, k4X);
        dV = hcf::rk4c(dV, k1V, k2V, k3V, k4V);
        store_state(L, dtraj, k + 1, dX, dV);
        X = Xp; V = Vp; ud = up; vd = vp; iw = iwp;
    }
    __syncwarp();
    // ---- reverse sweep for mu = d lambda ----  (X, V = s_N here)
    // The first-order quantities of every stage (Fbar, dist . Fbar) come from the cache the adjoint
    // sweep filled at this u: no lambda recurrence and no vjp here.  (Caching the tangent stage
    // data of the forward sweep as well removes the jvp recomputation but triples the streamed
    // bytes and measured 6 % slower: profiles/r01_v5_*.)
    double mX, mV;
    mX = L.isEnd ? two_alpha * dX : 0.0;
    mV = L.hasV ? two_beta * dV : 0.0;
    load_state(L, traj, N - 1, X, V);
    load_state(L, dtraj, N - 1, dX, dV);
    ud = load_u(L, u, N - 1); vd = load_u(L, v, N - 1);
    __shared__ __align__(16) double stage[2][STAGE_D];
    __syncwarp();
    stage_fetch(L, stage[0], adjc + (N - 1) * SCSTEP, iwc + (N - 1) * (4 * NLINK));
    cp_async_commit();
    int cur = 0;
#pragma unroll 1
    for (int k = N - 1; k >= 0; --k) {
        cp_async_wait_all();
        __syncwarp();
        const double *sb = stage[cur];
        if (k > 0) stage_fetch(L, stage[cur ^ 1], adjc + (k - 1) * SCSTEP, iwc + (k - 1) * (4 * NLINK));
        cp_async_commit();
        cur ^= 1;
        double Fb[4], dt[4], ddist[4], dd[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) { Fb[j] = L.active ? sb[j * SC + L.l] : 0.0; dt[j] = sb[j * SC + NL + L.link]; }
        const IW iw{sb[SCSTEP + L.link], sb[SCSTEP + NLINK + L.link], sb[SCSTEP + 2 * NLINK + L.link],
                    sb[SCSTEP + 3 * NLINK + L.link]};
        double Xp = 0, Vp = 0, dXp = 0, dVp = 0, up = 0, vp = 0;
        if (k > 0) {
            load_state(L, traj, k - 1, Xp, Vp); load_state(L, dtraj, k - 1, dXp, dVp);
            up = load_u(L, u, k - 1); vp = load_u(L, v, k - 1);
        }
        Link ld[4]; double iw2[4];
        stages_cached(L, X, V, ud, iw, ld, iw2);
        {
            double kX, kV;
            jvp(L, ld[0], dX, L.hasV ? dV : vd, kX, kV, ddist[0], dd[0]);
            double sX = hcf::axpy(hh, kX, dX), sV = hcf::axpy(hh, kV, dV);
            jvp(L, ld[1], sX, L.hasV ? sV : vd, kX, kV, ddist[1], dd[1]);
            sX = hcf::axpy(hh, kX, dX); sV = hcf::axpy(hh, kV, dV);
            jvp(L, ld[2], sX, L.hasV ? sV : vd, kX, kV, ddist[2], dd[2]);
            sX = hcf::axpy(h, kX, dX); sV = hcf::axpy(h, kV, dV);
            jvp(L, ld[3], sX, L.hasV ? sV : vd, kX, kV, ddist[3], dd[3]);
        }
        double dwX, dwV, bX, bV, tX, tV;
        // stage 4
        dwX = h6 * mX; dwV = h6 * mV;
        hvjp(L, ld[3], iw2[3], Fb[3], dt[3], ddist[3], dd[3], dwX, dwV, bX, bV);
        tX = mX + bX; tV = mV + bV;
        // stage 3
        dwX = hcf::lin2(h3, mX, h, bX); dwV = hcf::lin2(h3, mV, h, bV);
        hvjp(L, ld[2], iw2[2], Fb[2], dt[2], ddist[2], dd[2], dwX, dwV, bX, bV);
        tX = tX + bX; tV = tV + bV;
        // stage 2
        dwX = hcf::lin2(h3, mX, hh, bX); dwV = hcf::lin2(h3, mV, hh, bV);
        hvjp(L, ld[1], iw2[1], Fb[1], dt[1], ddist[1], dd[1], dwX, dwV, bX, bV);
        tX = tX + bX; tV = tV + bV;
        // stage 1
        dwX = hcf::lin2(h6, mX, hh, bX); dwV = hcf::lin2(h6, mV, hh, bV);
        hvjp(L, ld[0], iw2[0], Fb[0], dt[0], ddist[0], dd[0], dwX, dwV, bX, bV);
        tX = tX + bX; tV = tV + bV;
        // stage cost
        mX = L.isEnd ? fma(two_alpha, dX, tX) : tX;
        if (L.hasV) mV = fma(two_beta, dV, tV);
        else {
            if (L.isEnd) Hv[NU * k + (L.l - NVL)] = fma(two_gamma, vd, tV);
            mV = 0.0;
        }
        X = Xp; V = Vp; dX = dXp; dV = dVp; ud = up; vd = vp;
    }
}
#else
// The stage points' link vectors come from dic, so neither sweep of the product touches the state
// trajectory; same operations on the same operands as the recomputing version above, hence the same bits.
// (Tried on top of this and dropped: software-pipelining the reverse sweep so that the tangent chain of
// step k-1 issues alongside the second-order chain of step k; no gain, profiles/r01_v10_variants_*.log.)
/// v_k on lanes 27..29 by an unpredicated load (the other lanes read v[3k], which nothing consumes: a
/// predicated load followed by a zeroing move exposed the whole load latency once per step,
/// 10 % of all warp samples in the first capture of this variant)
__device__ __forceinline__ double load_v(const Lane &L, const double *v, int k) {
    return v[NU * k + (L.isEnd ? L.l - NVL : 0)];
}
// Tangent trajectory rows are padded to 58 doubles (16-byte aligned) so that the reverse sweep can
// stage them with 16-byte cp.async like the other per-step blocks.
constexpr int DTS = NX + 1;
__device__ __forceinline__ void dt_store(const Lane &L, double *t, int k, double X, double V) {
    double *s = t + k * DTS;
    if (L.active) __stcg(s + L.l, X);
    if (L.hasV) __stcg(s + NL + L.l, V);
}
__device__ __forceinline__ void dt_load(const Lane &L, const double *t, int k, double &X, double &V) {
    const double *s = t + k * DTS;
    X = L.active ? __ldcg(s + L.l) : 0.0;
    V = L.hasV ? __ldcg(s + NL + L.l) : 0.0;
}
__device__ __forceinline__ void fetch_row(double *buf, const double *g) { // 58 doubles = 29 chunks
    const int l = lane_id();
    if (l < DTS / 2) cp_async16(buf + 2 * l, g + 2 * l);
}
// Shared-memory operands are read unpredicated: lanes 30 / 31 pick up a neighbouring entry of the slot,
// which only their own (never consumed) results see; `active ? smem : 0` cost a predicated load, a
// zeroing CS2R and two predicated moves per operand (16 operands per step).
struct Tan { double ddist[4], dd[4]; };
struct Links { Link ld[4]; double iw2[4]; };
/// tangent stage chain of one RK4 step; db = the step's dic block (shared memory)
__device__ __forceinline__ void jvp_chain(const Lane &L, const double *db, double dX, double dV, double vd,
                                          Links &K, Tan &T, double (&kX)[4], double (&kV)[4]) {
#pragma unroll
    for (int j = 0; j < 4; ++j) link_cached(db[j * NL + L.l], db[4 * NL + j * NLINK + L.link], K.ld[j], K.iw2[j]);
    jvp(L, K.ld[0], dX, L.hasV ? dV : vd, kX[0], kV[0], T.ddist[0], T.dd[0]);
    double sX = hcf::axpy(hh, kX[0], dX), sV = hcf::axpy(hh, kV[0], dV);
    jvp(L, K.ld[1], sX, L.hasV ? sV : vd, kX[1], kV[1], T.ddist[1], T.dd[1]);
    sX = hcf::axpy(hh, kX[1], dX); sV = hcf::axpy(hh, kV[1], dV);
    jvp(L, K.ld[2], sX, L.hasV ? sV : vd, kX[2], kV[2], T.ddist[2], T.dd[2]);
    sX = hcf::axpy(h, kX[2], dX); sV = hcf::axpy(h, kV[2], dV);
    jvp(L, K.ld[3], sX, L.hasV ? sV : vd, kX[3], kV[3], T.ddist[3], T.dd[3]);
}
/// second-order adjoint chain of one RK4 step; ab = the step's adjc block (shared memory)
__device__ __forceinline__ void hvjp_chain(const Lane &L, const Links &K, const double *ab, const Tan &T,
                                           double mX, double mV, double &tX, double &tV) {
    double dwX, dwV, bX, bV;
    // stage 4
    dwX = h6 * mX; dwV = h6 * mV;
    hvjp(L, K.ld[3], K.iw2[3], ab[3 * SC + L.l], ab[3 * SC + NL + L.link], T.ddist[3], T.dd[3], dwX, dwV, bX, bV);
    tX = mX + bX; tV = mV + bV;
    // stage 3
    dwX = hcf::lin2(h3, mX, h, bX); dwV = hcf::lin2(h3, mV, h, bV);
    hvjp(L, K.ld[2], K.iw2[2], ab[2 * SC + L.l], ab[2 * SC + NL + L.link], T.ddist[2], T.dd[2], dwX, dwV, bX, bV);
    tX = tX + bX; tV = tV + bV;
    // stage 2
    dwX = hcf::lin2(h3, mX, hh, bX); dwV = hcf::lin2(h3, mV, hh, bV);
    hvjp(L, K.ld[1], K.iw2[1], ab[SC + L.l], ab[SC + NL + L.link], T.ddist[1], T.dd[1], dwX, dwV, bX, bV);
    tX = tX + bX; tV = tV + bV;
    // stage 1
    dwX = hcf::lin2(h6, mX, hh, bX); dwV = hcf::lin2(h6, mV, hh, bV);
    hvjp(L, K.ld[0], K.iw2[0], ab[L.l], ab[NL + L.link], T.ddist[0], T.dd[0], dwX, dwV, bX, bV);
    tX = tX + bX; tV = tV + bV;
}

#if PANTR_HC_TMA
// Same product, same operands, same bits; only the transport of the per-step blocks differs: ONE lane issues one bulk
// copy per block (dic 1280 B; reverse sweep also adjc 1280 B and the 464-byte tangent row) that completes on the slot's
// mbarrier, instead of every lane issuing 2-5 16-byte cp.async plus commit / wait_group.
__device__ __noinline__ void hvp(int N, const double *u, const double *v, const double *traj, double *dtraj,
                                 const double *iwc, const double *adjc, const double *dic, double *Hv) {
    (void)u; (void)traj; (void)iwc; // v, Hv may be shared memory
    __builtin_assume(__isGlobal(dtraj)); __builtin_assume(__isGlobal(adjc)); __builtin_assume(__isGlobal(dic));
    const Lane L;
    constexpr int RSLOT = DSTEP + SCSTEP + DTS + 2; // reverse sweep: dic + adjc + dtraj row of one step (380 doubles)
    __shared__ __align__(16) double ring[2 * RSLOT];
    __shared__ __align__(8) unsigned long long bars[3];
    double (*dring)[DSTEP] = reinterpret_cast<double (*)[DSTEP]>(ring); // forward sweep: 3 dic slots in the same bytes
    static_assert(3 * DSTEP <= 2 * RSLOT, "ring");
    const bool lead = lane_id() == 0;
    double dX = 0.0, dV = 0.0;
    // ---- forward tangent sweep: dic blocks two steps ahead through the 3-slot ring ----
    dt_store(L, dtraj, 0, dX, dV);
    if (lead) {
        mbar_init(bars, 1); mbar_init(bars + 1, 1); mbar_init(bars + 2, 1);
        mbar_init_fence();
        for (int k = 0; k < 2 && k < N; ++k) { mbar_expect_tx(bars + k, DSTEP * 8); bulk_g2s(dring[k], dic + k * DSTEP, DSTEP * 8, bars + k); }
    }
    __syncwarp();
    double vd = load_v(L, v, 0);
    int s0 = 0, s2 = 2;
    unsigned ph = 0; // bit j = parity of slot j's next completion
#pragma unroll 1
    for (int k = 0; k < N; ++k) {
        mbar_wait(bars + s0, (ph >> s0) & 1u);
        ph ^= 1u << s0;
        __syncwarp(); // every lane has finished with slot s2 (read in step k - 1)
        if (lead && k + 2 < N) { mbar_expect_tx(bars + s2, DSTEP * 8); bulk_g2s(dring[s2], dic + (k + 2) * DSTEP, DSTEP * 8, bars + s2); }
        const double vp = load_v(L, v, k + 1 < N ? k + 1 : k);
        Tan T; Links K; double kX[4], kV[4];
        jvp_chain(L, dring[s0], dX, dV, vd, K, T, kX, kV);
        dX = hcf::rk4c(dX, kX[0], kX[1], kX[2], kX[3]);
        dV = hcf::rk4c(dV, kV[0], kV[1], kV[2], kV[3]);
        dt_store(L, dtraj, k + 1, dX, dV);
        vd = vp;
        s0 = s0 == 2 ? 0 : s0 + 1; s2 = s2 == 2 ? 0 : s2 + 1;
    }
    __threadfence(); // the tangent rows are read back through the async proxy: they must have reached L2
    __syncwarp();
    // ---- reverse sweep for mu = d lambda ----
    double mX = L.isEnd ? two_alpha * dX : 0.0;
    double mV = L.hasV ? two_beta * dV : 0.0;
    auto fetch = [&](int slot, int k) { // lead lane only
        double *b = ring + slot * RSLOT;
        mbar_expect_tx(bars + slot, (DSTEP + SCSTEP + DTS) * 8);
        bulk_g2s(b, dic + k * DSTEP, DSTEP * 8, bars + slot);
        bulk_g2s(b + DSTEP, adjc + k * SCSTEP, SCSTEP * 8, bars + slot);
        bulk_g2s(b + DSTEP + SCSTEP, dtraj + k * DTS, DTS * 8, bars + slot);
    };
    if (lead) {
        mbar_init(bars, 1); mbar_init(bars + 1, 1);
        mbar_init_fence();
        fetch(0, N - 1);
    }
    __syncwarp();
    vd = load_v(L, v, N - 1);
    int cur = 0;
    ph = 0;
#pragma unroll 1
    for (int k = N - 1; k >= 0; --k) {
        mbar_wait(bars + cur, (ph >> cur) & 1u);
        ph ^= 1u << cur;
        __syncwarp(); // every lane has finished with the other slot (read in the previous step)
        const double *db = ring + cur * RSLOT, *ab = db + DSTEP, *tb = ab + SCSTEP;
        if (lead && k > 0) fetch(cur ^ 1, k - 1);
        cur ^= 1;
        const double vp = load_v(L, v, k > 0 ? k - 1 : 0);
        dX = tb[L.l];                      // lanes 30, 31 / 27..29: unused garbage (see above)
        dV = tb[L.l < NVL ? NL + L.l : 0];
        Tan T; Links K; double kX[4], kV[4], tX, tV;
        jvp_chain(L, db, dX, dV, vd, K, T, kX, kV);
        hvjp_chain(L, K, ab, T, mX, mV, tX, tV);
        mX = L.isEnd ? fma(two_alpha, dX, tX) : tX;
        if (L.hasV) mV = fma(two_beta, dV, tV);
        else {
            if (L.isEnd) Hv[NU * k + (L.l - NVL)] = fma(two_gamma, vd, tV);
            mV = 0.0;
        }
        vd = vp;
    }
    __syncwarp();
}
#else
__device__ __noinline__ void hvp(int N, const double *u, const double *v, const double *traj, double *dtraj,
                                 const double *iwc, const double *adjc, const double *dic, double *Hv) {
    (void)u; (void)traj; (void)iwc; // v, Hv may be shared memory
    __builtin_assume(__isGlobal(dtraj)); __builtin_assume(__isGlobal(adjc)); __builtin_assume(__isGlobal(dic));
    const Lane L;
    constexpr int RSLOT = DSTEP + SCSTEP + DTS + 2; // reverse sweep: dic + adjc + dtraj row of one step (380 doubles)
    __shared__ __align__(16) double ring[2 * RSLOT];
    double (*dring)[DSTEP] = reinterpret_cast<double (*)[DSTEP]>(ring); // forward sweep: 3 dic slots in the same bytes
    static_assert(3 * DSTEP <= 2 * RSLOT, "ring");
    double dX = 0.0, dV = 0.0;
    // ---- forward tangent sweep: dic blocks staged two steps ahead through the 3-slot ring ----
    dt_store(L, dtraj, 0, dX, dV);
    __syncwarp();
    fetch160(dring[0], dic); cp_async_commit();
    fetch160(dring[1], dic + (N > 1 ? DSTEP : 0)); cp_async_commit();
    double vd = load_v(L, v, 0);
    int s0 = 0, s1 = 1, s2 = 2;
#pragma unroll 1
    for (int k = 0; k < N; ++k) {
        cp_async_wait_1();
        __syncwarp();
        const int kf = k + 2 < N ? k + 2 : N - 1;
        fetch160(dring[s2], dic + kf * DSTEP); cp_async_commit();
        const double vp = load_v(L, v, k + 1 < N ? k + 1 : k);
        Tan T; Links K; double kX[4], kV[4];
        jvp_chain(L, dring[s0], dX, dV, vd, K, T, kX, kV);
        dX = hcf::rk4c(dX, kX[0], kX[1], kX[2], kX[3]);
        dV = hcf::rk4c(dV, kV[0], kV[1], kV[2], kV[3]);
        dt_store(L, dtraj, k + 1, dX, dV);
        vd = vp;
        const int t = s0; s0 = s1; s1 = s2; s2 = t;
    }
    cp_async_wait_all();
    __syncwarp();
    // ---- reverse sweep for mu = d lambda ----
    double mX = L.isEnd ? two_alpha * dX : 0.0;
    double mV = L.hasV ? two_beta * dV : 0.0;
    // Everything a step needs from global memory (dic, adjc, the tangent state) arrives through one
    // cp.async slot fetched a step ahead: no register-destination global loads inside the loop, whose
    // scoreboard the first shared-memory read of a step otherwise ends up waiting on
    // (15 % of all warp samples on one DMUL when the tangent state came by __ldcg: profiles/r01_v10_variant1_ncu_full_summary.csv).
    {
        double *b0 = ring;
        fetch160(b0, dic + (N - 1) * DSTEP); fetch160(b0 + DSTEP, adjc + (N - 1) * SCSTEP);
        fetch_row(b0 + DSTEP + SCSTEP, dtraj + (N - 1) * DTS);
        cp_async_commit();
    }
    vd = load_v(L, v, N - 1);
    int cur = 0;
#pragma unroll 1
    for (int k = N - 1; k >= 0; --k) {
        cp_async_wait_all();
        __syncwarp();
        const double *db = ring + cur * RSLOT, *ab = db + DSTEP, *tb = ab + SCSTEP;
        {
            const int kn = k > 0 ? k - 1 : 0;
            double *nb = ring + (cur ^ 1) * RSLOT;
            fetch160(nb, dic + kn * DSTEP); fetch160(nb + DSTEP, adjc + kn * SCSTEP);
            fetch_row(nb + DSTEP + SCSTEP, dtraj + kn * DTS);
            cp_async_commit();
        }
        cur ^= 1;
        const double vp = load_v(L, v, k > 0 ? k - 1 : 0);
        dX = tb[L.l];                      // lanes 30, 31 / 27..29: unused garbage (see above)
        dV = tb[L.l < NVL ? NL + L.l : 0];
        Tan T; Links K; double kX[4], kV[4], tX, tV;
        jvp_chain(L, db, dX, dV, vd, K, T, kX, kV);
        hvjp_chain(L, K, ab, T, mX, mV, tX, tV);
        mX = L.isEnd ? fma(two_alpha, dX, tX) : tX;
        if (L.hasV) mV = fma(two_beta, dV, tV);
        else {
            if (L.isEnd) Hv[NU * k + (L.l - NVL)] = fma(two_gamma, vd, tV);
            mV = 0.0;
        }
        vd = vp;
    }
    cp_async_wait_all();
    __syncwarp();
}
#endif
#endif

} // namespace hc

// Problem-function surface the solver sees (what CasADiProblem::eval_* would compute, m = 0).
struct HangingChain {
    static constexpr int NX = hc::NX, NU = hc::NU, NC = 0;
#ifndef PANTR_MIN_BLOCKS
#define PANTR_MIN_BLOCKS 12 // resident warps per SM the solve kernel is compiled for (168 registers)
#endif
    static constexpr bool HAS_ROLLOUT_TAG = false;
    static constexpr bool BATCHED_MAPS = true;  // solver vector loops: loads of four elements before the first store (+1.2 %)
    static constexpr int MAX_SMEM_SET = 3; // solver vectors in shared memory: the largest set that keeps the occupancy (Pantr::smem_bytes)
    static constexpr bool OBJECTS_IN_SHARED = false; // solve kernel: model / solver objects stay per-lane locals (measured faster)
    static constexpr int MIN_BLOCKS = PANTR_MIN_BLOCKS;
    __host__ __device__ static int n(int N) { return NU * N; }
    __host__ __device__ static int m(int) { return 0; }
    // traj + dtraj + iwc + adjc
    __host__ __device__ static int traj_doubles(int N) {
#if PANTR_HC_DIST_CACHE
        return (N + 1) * hc::DTS + ((N + 1) * NX + 1) / 2 * 2 + 4 * hc::NLINK * N + hc::SCSTEP * N + hc::DSTEP * N;
#else
        return 2 * (N + 1) * NX + 4 * hc::NLINK * N + hc::SCSTEP * N;
#endif
    }
    __host__ __device__ static double C_lb(int) { return -hc::v_max; } // hanging_chain.py:82-84
    __host__ __device__ static double C_ub(int) { return +hc::v_max; }
    __host__ __device__ static double D_lb(int, int) { return 0; }
    __host__ __device__ static double D_ub(int, int) { return 0; }
    static void initial_state(double *x) { // hanging_chain.py:92-96
        for (int i = 0; i < NX; ++i) x[i] = 0;
        for (int i = 0; i < 10; ++i) x[3 * i] = double(i + 1) / 10;
    }
    static void initial_guess(double *u, int N) { for (int i = 0; i < NU * N; ++i) u[i] = 0; }

    int N;
    const double *x0;
    double *traj, *dtraj, *iwc, *adjc, *dic;
    bool hvp_ready; // adjc holds the first-order stage data of the last rollout

    __device__ void init(int N_, const double *x0_, double *ws) {
        N = N_; x0 = x0_; hvp_ready = false;
#if PANTR_HC_DIST_CACHE
        dtraj = ws; traj = ws + (N_ + 1) * hc::DTS; iwc = traj + ((N_ + 1) * NX + 1) / 2 * 2;
#else
        traj = ws; dtraj = ws + (N_ + 1) * NX; iwc = ws + 2 * (N_ + 1) * NX;
#endif
        adjc = iwc + 4 * hc::NLINK * N_; dic = adjc + hc::SCSTEP * N_;
    }
    __device__ double eval_f(const double *u) { return eval_psi(u, nullptr, nullptr, nullptr); }
    __device__ void eval_g(const double *, double *) {}
    __device__ double eval_psi(const double *x, const double *, const double *, double *) {
        __syncwarp();
        hvp_ready = false;
        double f = hc::rollout(N, x0, x, traj, iwc);
        __syncwarp();
        return f;
    }
    __device__ double eval_psi_grad_psi(const double *x, const double *, const double *, double *grad) {
        __syncwarp();
        double f = hc::rollout(N, x0, x, traj, iwc);
        __syncwarp();
        hc::adjoint(N, x, traj, iwc, grad, adjc, dic);
        hvp_ready = true;
        __syncwarp();
        return f;
    }
    /// Requires that the last rollout was at x (PANTR guarantees this: pantr.tpp:314 then :373).
    __device__ void hess_psi_prod(const double *x, const double *, const double *, const double *v, double *Hv) {
        __syncwarp();
        if (!hvp_ready) (void)eval_psi_grad_psi(x, nullptr, nullptr, Hv); // Hv as scratch for the gradient
        hc::hvp(N, x, v, traj, dtraj, iwc, adjc, dic, Hv);
        __syncwarp();
    }
    // ---- problem plugin face (m = 0: L = f) ----
    __device__ double eval_grad_L(const double *x, const double *, double *grad) {
        return eval_psi_grad_psi(x, nullptr, nullptr, grad);
    }
    __device__ void eval_grad_g_prod(const double *, const double *, double *out) {
        for (int i = lane_id(); i < n(N); i += 32) out[i] = 0.0;
        __syncwarp();
    }
    __device__ void hess_L_prod(const double *x, const double *, const double *v, double *Hv) {
        hvp_ready = false;
        hess_psi_prod(x, nullptr, nullptr, v, Hv);
    }
    /// Plant step x+ = f_d(x, u) (problems/problems/formulation.hpp:258-261)
    __device__ static void plant_step(const double *x, const double *u, double *xn) {
        const hc::Lane L;
        double X = L.active ? x[L.l] : 0.0, V = L.hasV ? x[hc::NL + L.l] : 0.0, ud = L.isEnd ? u[L.l - hc::NVL] : 0.0;
        double Xn, Vn;
        hc::rk4(L, X, V, ud, Xn, Vn, nullptr);
        if (L.active) xn[L.l] = Xn;
        if (L.hasV) xn[hc::NL + L.l] = Vn;
    }
};

} // namespace pantr
