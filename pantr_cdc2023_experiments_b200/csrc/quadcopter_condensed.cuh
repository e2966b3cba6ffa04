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
// Per-instance model workspace (global memory):
//   ABv N x NV x 10: the columns of AB_k that vary from step to step (FC_CONST_COLS marks the others: 7 of 13 vary for
//   both models, 560 bytes per step) | ABc 13 x 10: the constant columns, once per instance (the recurrences keep them
//   in registers) | lam, mus (N+1) x 14 rows
//   [x-part source -> adjoint state (9) | u-part addend (4) | pad] of the two adjoint recurrences |
//   traj, dtraj (N+1) x 9 | sc (4N+1) x 6 sin/cos of every stage point | yh, wact 4(N+1)
// The three recurrences are latency chains: every operand of a step must already sit in shared memory when the
// step starts.  A step's record is therefore fetched PANTR_QC_RING_DEPTH steps ahead into a shared-memory ring:
// the variable columns of AB_k (560 bytes) by ONE elected lane with cp.async.bulk (TMA; SASS: UBLKCP) completing on a per-slot mbarrier the
// warp waits on; the 112-byte source row of the adjoint recurrences by cp.async (LDGSTS) of 7 lanes -- the rows are
// written by the lanes just before the recurrence starts, and cp.async is a generic-proxy read ordered by
// __syncwarp like any load, whereas a bulk copy of freshly written global data needs a gpu-scope fence first (AB gets
// one __threadfence per build).  Results leave with plain stores.
// History (profiles/): v0, register prefetch one step ahead + global loads of the source rows: ~2000 cycles per
// step with the resident warps' working set streaming from HBM (r02_a_quad_condensed_speed_v0.log); v1, everything
// by TMA with a fence.proxy.async before every refill: the fence alone cost as much (35-40 % of the recurrences'
// stall samples on `membar`, r02_b_quadc_ncu_*).
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

// resident warps per SM the solve kernel is compiled for: 16 (128 registers) for the simplified model, 12 (168 registers)
// for the realistic one, whose generated derivative functions are twice as large (profiles/r02_o_variants.log: 16 -> 12
// is -5 % for the simplified and +8 % for the realistic model in a two-wave run; 20 and 24 are slower for both)
#ifndef PANTR_QC_MIN_BLOCKS
#define PANTR_QC_MIN_BLOCKS 16
#endif
#ifndef PANTR_QC_MIN_BLOCKS_REALISTIC
#define PANTR_QC_MIN_BLOCKS_REALISTIC 12
#endif
#define PANTR_QC_FNS(Name, ns, MB)                                                                             \
    struct Name {                                                                                            \
        static constexpr unsigned CONST_COLS = ns::FC_CONST_COLS;                                            \
        static constexpr int MIN_BLOCKS = MB;                                                                \
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
PANTR_QC_FNS(SimpleFns, quadcopter_gen, PANTR_QC_MIN_BLOCKS);
PANTR_QC_FNS(RealisticFns, realistic_quadcopter_gen, PANTR_QC_MIN_BLOCKS_REALISTIC);
#undef PANTR_QC_FNS

template <int count>
__device__ __forceinline__ void ldv(const double *p, double *out) {
#pragma unroll
    for (int i = 0; i < count; ++i) out[i] = p[i];
}

} // namespace qc

// ---- per-step record ring of the matrix-vector recurrences: cp.async.bulk + mbarrier ----
#ifndef PANTR_QC_RING_DEPTH
#define PANTR_QC_RING_DEPTH 4
#endif
namespace qc {
constexpr int COL = 10;        // doubles per column of AB_k (9 + 1: columns start 16-byte aligned, LDS.128)
constexpr int ABC = NZ * COL;  // doubles of the constant-column block (one per instance)
constexpr int ROW = 14;        // doubles per row of lam / mus (112 bytes)
constexpr int RING_D = PANTR_QC_RING_DEPTH;
static_assert((RING_D & (RING_D - 1)) == 0 && RING_D >= 2, "ring depth: a power of two");
struct StepRing {
    double *slot;              // RING_D slots: [the step's variable columns of AB | source row]
    double *xch;               // 2 x 16: the recurrence's state / operand exchange (ping-pong)
    double *cst;               // the constant columns of AB (adjoint recurrences: one column per lane)
    unsigned long long *bar;   // RING_D mbarriers
};
/// One shared-memory ring per block, shared by the three recurrences (they never overlap in time).
template <int SLOT, int CST>
__device__ __forceinline__ StepRing step_ring() {
    __shared__ __align__(16) double storage[RING_D * SLOT + 32 + CST];
    __shared__ __align__(8) unsigned long long bars[RING_D];
    return StepRing{storage, storage + RING_D * SLOT, storage + RING_D * SLOT + 32, bars};
}
} // namespace qc


template <class F>
struct QuadCondensedModel {
    static constexpr int NX = qc::NX, NU = qc::NU, NC = qc::NC, NZ = qc::NZ;
    static constexpr int MIN_BLOCKS = F::MIN_BLOCKS;
    static constexpr unsigned VAR_COLS = ~F::CONST_COLS & 0x1fffu; // columns of AB that differ from step to step
    static constexpr int NV = __builtin_popcount(VAR_COLS), NCST = NZ - NV;
    static constexpr int ABS = NV * qc::COL;            // doubles per step of ABv (a multiple of 2: 16-byte bulk copies)
    static constexpr int RING_SLOT = ABS + qc::ROW;     // doubles per ring slot
    static constexpr bool HAS_SCALE = true;
    static constexpr int MAX_SMEM_SET = 0; // no solver vectors in shared memory: +2 % / +6 % with the L1 that leaves (r02_n_smem_sets.log)
    static constexpr bool BATCHED_MAPS = false; // solver vector loops stay plain (-3 % batched, exec_policy.cuh)
    static constexpr bool OBJECTS_IN_SHARED = true; // solve kernel: model / solver objects in shared memory (pantr_b200.cu)
    __host__ __device__ static int n(int N) { return NU * N; }
    __host__ __device__ static int m(int N) { return NC * (N + 1); }
    __host__ __device__ static int traj_doubles(int N) {
        const int d = N * ABS + qc::ABC + 2 * (N + 1) * qc::ROW + 2 * (N + 1) * NX + (4 * N + 1) * 6 + 2 * NC * (N + 1);
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

    // The model object lives in local memory (its methods are __noinline__ and take `this`), and every store through a
    // double* may alias it: a sweep that reads N or an array pointer inside its loop reloads it from memory in every step
    // (three exposed loads per recurrence step in the first version, profiles/r02_c_*).  Every sweep therefore starts by
    // copying the members it uses into locals of the same name.
#define PANTR_QC_LOCALS                                                                                                  \
    const int N = this->N;                                                                                               \
    const double *const x0 = this->x0;                                                                                   \
    double *const traj = this->traj, *const sc = this->sc, *const AB = this->AB, *const lam = this->lam,                 \
                  *const yh = this->yh, *const wact = this->wact, *const dtraj = this->dtraj, *const mus = this->mus,    \
                  *const ABc = this->ABc;                                                                                \
    (void)N; (void)x0; (void)traj; (void)sc; (void)AB; (void)lam; (void)yh; (void)wact; (void)dtraj; (void)mus; (void)ABc
    int N;
    const double *x0;
    double *traj, *sc, *AB, *ABc, *lam, *yh, *wact, *dtraj, *mus;
    bool hvp_ready;
    /// The vector whose rollout traj / sc currently hold.  PANTR evaluates psi(x_hat) and then, one step later,
    /// psi_grad_psi at a copy of the same point (pantr.tpp:303-316: prox.x = curr.x_hat): the solver moves the tag
    /// along with the copy (retag) and eval_psi_grad_psi skips the second rollout -- same values, a third fewer
    /// sequential sweeps.  Every rollout sets the tag, only eval_psi_grad_psi ever trusts it.
    const double *rolled;
    static constexpr bool HAS_ROLLOUT_TAG = true;
    __device__ __forceinline__ void retag(const double *from, const double *to) { if (rolled == from) rolled = to; }
    __device__ __forceinline__ void untag() { rolled = nullptr; }

    __device__ void init(int N_, const double *x0_, double *ws) {
        N = N_; x0 = x0_; hvp_ready = false; rolled = nullptr;
        double *p = ws;
        // the bulk-copied arrays first: the block's workspace starts 128-byte aligned, AB steps are 944 bytes, rows 112
        AB = p; p += N * ABS; ABc = p; p += qc::ABC; lam = p; p += (N + 1) * qc::ROW; mus = p; p += (N + 1) * qc::ROW;
        traj = p; p += (N + 1) * NX; dtraj = p; p += (N + 1) * NX;
        sc = p; p += (4 * N + 1) * 6; yh = p; p += NC * (N + 1); wact = p;
    }

    // ---- one RK4 step (quadcopter.py:49-55), executed by every lane; the sin/cos of the three angles of a
    // stage point are computed by lanes 0..2 (one angle each, pantr_sincos_fma) and handed to the other lanes through
    // shared memory: one 16-byte store, one __syncwarp and three 16-byte loads instead of twelve shuffles.  `buf`
    // alternates 0 / 1 from stage to stage: the __syncwarp of the stage in between orders the reads of a buffer
    // before its next writes.  sc_out: 4 x 6 or null. ----
    __device__ __forceinline__ static void stage_sincos(const double *x, double *s6, double *sc_out, int buf) {
        __shared__ __align__(16) double xs[2][8];
        const int lane = lane_id();
        const double ang = lane == 0 ? x[6] : (lane == 1 ? x[7] : x[8]);
        double s, c;
        pantr_sincos_fma(ang, &s, &c);
        if (lane < 3) {
            *reinterpret_cast<double2 *>(&xs[buf][2 * lane]) = make_double2(s, c);
            if (sc_out) { sc_out[2 * lane] = s; sc_out[2 * lane + 1] = c; }
        }
        __syncwarp();
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const double2 v = *reinterpret_cast<const double2 *>(&xs[buf][2 * a]);
            s6[2 * a] = v.x; s6[2 * a + 1] = v.y;
        }
    }
    /// The four stages as ONE rolled loop over a running weighted sum S += w_s k_s (w = 1, 2, 2, 1; 0 + 1 k_1 = k_1) and the
    /// next stage point t = x + c_s k_s (c = h/2, h/2, h), both as fma: one stage of straight-line code (2.7 KB) instead of four.
    /// Unrolled, the rollout loop body was 11.5 KB and its warps spent half of their stall samples waiting for
    /// instructions (profiles/r02_k_quadc_ncu_regions.txt: no_inst 0.52); a first rolled version that kept k_1..k_4 in
    /// arrays went through local memory and executed 50 % more instructions (profiles/r02_d_quadc_ncu_*).
    __device__ __forceinline__ static void rk4_step(double *x, const double *uk, double *sc_out) {
        using namespace qc;
        double s6[6], t[NX], S[NX], k[NX];
#pragma unroll
        for (int i = 0; i < NX; ++i) { t[i] = x[i]; S[i] = 0.0; }
#pragma unroll 1
        for (int st = 0; st < 4; ++st) {
            stage_sincos(t, s6, sc_out ? sc_out + 6 * st : nullptr, st & 1);
            F::fc(t, uk, s6, k);
            const double w = (st == 0 || st == 3) ? 1.0 : 2.0, c = st == 2 ? h : hh;
#pragma unroll
            for (int i = 0; i < NX; ++i) { S[i] = fma(w, k[i], S[i]); t[i] = fma(c, k[i], x[i]); } // (fma: +2.5 %, r02_q_variants.log)
        }
#pragma unroll
        for (int i = 0; i < NX; ++i) x[i] = fma(h6, S[i], x[i]);
    }
    /// Rollout (ss.py:24): fills traj and sc.
    __device__ __noinline__ void rollout(const double *u) {
        PANTR_QC_LOCALS;
        double x[NX], uk[NU], un[NU];
#pragma unroll
        for (int i = 0; i < NX; ++i) { x[i] = x0[i]; traj[i] = x[i]; }
        qc::ldv<NU>(u, un);
#pragma unroll 1
        for (int k = 0; k < N; ++k) {
#pragma unroll
            for (int j = 0; j < NU; ++j) uk[j] = un[j];
            qc::ldv<NU>(u + NU * (k + 1 < N ? k + 1 : k), un); // the next step's input travels while this step computes
            rk4_step(x, uk, sc + 24 * k);
#pragma unroll
            for (int i = 0; i < NX; ++i) traj[(k + 1) * NX + i] = x[i];
        }
        double s6[6];
        stage_sincos(x, s6, sc + 24 * N, 0);
        this->rolled = u;
        __syncwarp();
    }
    /// Cost (ss.py:32-36): lane = node, node k < N contributes l(x_k, u_k), node N contributes l_N(x_N).
    __device__ __forceinline__ double cost(const double *u) const {
        PANTR_QC_LOCALS;
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
        PANTR_QC_LOCALS;
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
        PANTR_QC_LOCALS;
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
    /// Step Jacobians: lane = (step, column) task.  Every step gets the columns that vary (ABv); the constant columns
    /// (FC_CONST_COLS: the same at every step of every instance) are computed once, at step 0 (ABc).
    __device__ __noinline__ void build_AB(const double *u) {
        PANTR_QC_LOCALS;
        const int total = N * NV + NCST;
        for (int t = lane_id(); t < total; t += 32) {
            int k = 0, c;
            double *col;
            if (t < N * NV) { k = t / NV; const int j = t % NV; c = nth_col(VAR_COLS, j); col = AB + size_t(k) * ABS + j * qc::COL; }
            else { c = nth_col(F::CONST_COLS, t - N * NV); col = ABc + (t - N * NV) * qc::COL; } // ABc: the constant columns, compacted
            double xk[NX], uk[NU], s[24], e[NZ], out[NX];
            qc::ldv<NX>(traj + k * NX, xk); qc::ldv<NU>(u + NU * k, uk); qc::ldv<24>(sc + 24 * k, s);
#pragma unroll
            for (int i = 0; i < NZ; ++i) e[i] = i == c ? 1.0 : 0.0;
            step_tangent(xk, uk, s, e, e + NX, out);
#pragma unroll
            for (int i = 0; i < NX; ++i) col[i] = out[i];
        }
        __threadfence(); // the recurrences read ABv through the async proxy (TMA reads L2): the stores must have landed
        __syncwarp();
    }
    /// Adjoint recurrence w_k = AB_k' w_{k+1} + src_k over k = N-1..0 (lane c < 13 owns entry c of the product: one
    /// fma chain over the 9 rows; the 9 adjoint entries travel through shared memory).
    /// rows: [(N+1) x ROW]; row N = terminal value (in); row k < N holds the x-part source (entries 0..8) and the
    /// u-part addend (9..12) on entry, the adjoint state lambda_k in 0..8 on exit (row 0 keeps its source: x_0 does
    /// not depend on u).  The u-part of step k goes to out[4k + j]: product + addend (second-order adjoint), or, with
    /// `grad`, fma(cs 2 w_u, addend, product) where the addend is u_kj (gradient).  Both are val = fma(coef, addend,
    /// product) with a per-lane coef (fma(1, b, a) is a + b exactly), so the x- and u-lanes share one code path.
    /// Step i (k = N-1-i) reads its record -- AB_k and row k -- from ring slot i % D, filled D steps earlier; slot,
    /// barrier, parity and exchange-buffer addresses are running registers (32-bit shared-window addresses).
#ifndef PANTR_QC_ROWS_TMA
#define PANTR_QC_ROWS_TMA 0
#endif
#ifndef PANTR_QC_SPLIT_CHAIN
#define PANTR_QC_SPLIT_CHAIN 0
#endif
    __device__ __forceinline__ static void row_commit() { if (!PANTR_QC_ROWS_TMA) cp_async_commit_group(); }
    template <int PENDING>
    __device__ __forceinline__ static void row_wait() { if (!PANTR_QC_ROWS_TMA) cp_async_wait_group<PENDING>(); }
    __device__ __noinline__ void adjoint_recurrence(double *rows, double *out, bool grad, double cs) {
        using namespace qc;
        const int N = this->N;
        const double *const AB = this->AB;
        const StepRing R = step_ring<RING_SLOT, NCST * COL>();
        const unsigned slot0 = smem_addr(R.slot), bar0 = smem_addr(R.bar), xch0 = smem_addr(R.xch);
        constexpr unsigned SLOT_B = RING_SLOT * 8, RING_B = RING_D * SLOT_B;
        const int lane = lane_id();
        const int cl = lane < NZ ? lane : NZ - 1;
        // this lane's column of AB_k: the vj-th variable column of the step's record, or a constant column (shared memory,
        // loaded once per sweep)
        const bool is_var = (VAR_COLS >> cl) & 1u;
        const int vj = __popc(VAR_COLS & ((1u << cl) - 1u));
        const unsigned voff = is_var ? vj * (COL * 8) : 0u;
        const unsigned caddr = smem_addr(R.cst) + (cl - vj) * (COL * 8);
        const double coef = (grad && lane >= NX && lane < NZ) ? cs * (2 * wu(lane - NX)) : 1.0;
        double *dst = lane < NX ? rows + size_t(N - 1) * ROW + lane : out + size_t(N - 1) * NU + (cl - NX);
        const int dstride = lane < NX ? -ROW : -NU;
        auto refill = [&](int k, unsigned rec, unsigned bar) { // record of step k: ABv_k by lane 0 (TMA), the row by lanes 0..6 (cp.async)
#if PANTR_QC_ROWS_TMA // A/B build: the row by TMA as well (needs the gpu-scope fence below, once per sweep); bit-identical, measured -0.4 % / -0.9 % (r02_q_variants.log): off
            if (lane == 0) {
                mbar_expect_tx_a(bar, (ABS + ROW) * 8);
                bulk_g2s_a(rec, AB + size_t(k) * ABS, ABS * 8, bar);
                bulk_g2s_a(rec + ABS * 8, rows + size_t(k) * ROW, ROW * 8, bar);
            }
#else
            if (lane == 0) {
                mbar_expect_tx_a(bar, ABS * 8);
                bulk_g2s_a(rec, AB + size_t(k) * ABS, ABS * 8, bar);
            }
            if (lane < ROW / 2) cp_async_16_a(rec + ABS * 8 + 16 * lane, rows + size_t(k) * ROW + 2 * lane);
#endif
        };
        if (lane == 0) {
#pragma unroll
            for (int sl = 0; sl < RING_D; ++sl) mbar_init(R.bar + sl, 1);
            mbar_init_fence();
        }
        if (lane < NX) R.xch[lane] = rows[N * ROW + lane];
        for (int t = lane; t < NCST * COL; t += 32) R.cst[t] = this->ABc[t];
#if PANTR_QC_ROWS_TMA
        __threadfence();
        proxy_async_fence();
#endif
        __syncwarp(); // the rows the lanes have just written are ordered before the copies below
#pragma unroll
        for (int i = 0; i < RING_D; ++i) {
            if (i < N) refill(N - 1 - i, slot0 + i * SLOT_B, bar0 + i * 8);
            row_commit();
        }
        row_wait<RING_D - 1>(); // step 0's row has landed ...
        __syncwarp();                      // ... in every lane's view
        unsigned rec = slot0, bar = bar0, par = 0, xr = xch0, xw = xch0 + 128;
#pragma unroll 1
        for (int i = 0; i < N; ++i) {
            mbar_wait_a(bar, par);
            double col[NX], w[NX];
            lds_vec<NX>(is_var ? rec + voff : caddr, col);
            const double addend = lds_f64(rec + ABS * 8 + 8 * cl);
            lds_vec<NX>(xr, w);
#if PANTR_QC_SPLIT_CHAIN // A/B build (changes the rounding): three chains of three instead of one of nine; measured +1.4 % / -0.4 % (r02_q_variants.log): off
            const double a0 = fma(col[2], w[2], fma(col[1], w[1], col[0] * w[0]));
            const double a1 = fma(col[5], w[5], fma(col[4], w[4], col[3] * w[3]));
            const double a2 = fma(col[8], w[8], fma(col[7], w[7], col[6] * w[6]));
            const double a = (a0 + a1) + a2;
#else
            double a = col[0] * w[0];
#pragma unroll
            for (int r = 1; r < NX; ++r) a = fma(col[r], w[r], a);
#endif
            const double val = fma(coef, addend, a);
            if (lane < NX ? i + 1 < N : lane < NZ) *dst = val;
            dst += dstride;
            if (lane < NX) sts_f64(xw + 8 * lane, val);
            row_wait<RING_D - 2>(); // the next step's row has landed
            __syncwarp();                      // slot `rec` has been read by every lane: refill it
            const int kn = N - 1 - RING_D - i;
            if (kn >= 0) refill(kn, rec, bar);
            row_commit();
            { const unsigned t = xr; xr = xw; xw = t; }
            rec += SLOT_B; bar += 8;
            if (rec == slot0 + RING_B) { rec = slot0; bar = bar0; par ^= 1u; }
        }
        row_wait<0>();
    }
    /// Gradient of  s f + yh' g  wrt u (s = cost_scale); leaves lambda_1..N in lam.
    __device__ __noinline__ void adjoint(const double *u, double cost_scale, double *grad) {
        PANTR_QC_LOCALS;
        using namespace qc;
        for (int k = lane_id(); k <= N; k += 32) { // rows of the recurrence: node sources q_k = s dl/dx + J_c' yh_k, and u_k
            if (k < N) {
#pragma unroll
                for (int j = 0; j < NU; ++j) lam[k * ROW + NX + j] = u[NU * k + j];
            }
            if (k == 0) continue;
            double x[NX], s6[6], y4[NC], ax[NX];
            ldv<NX>(traj + k * NX, x); ldv<6>(sc + 24 * k, s6); ldv<NC>(yh + NC * k, y4);
            F::con_vjp(x, s6, y4, ax);
#pragma unroll
            for (int i = 0; i < NX; ++i) {
                const double dl = (2 * wx(i, k == N)) * (i < 3 ? x[i] - pf(i) : x[i]);
                lam[k * ROW + i] = fma(cost_scale, dl, ax[i]);
            }
        }
        __syncwarp();
        adjoint_recurrence(this->lam, grad, true, cost_scale);
    }
    /// Tangent recurrence dx_{k+1} = AB_k (dx_k, v_k): lane r < 9 owns row r, one fma chain over the 13 columns;
    /// AB_k arrives through the ring, (dx_k, v_k) through the exchange buffer.
    __device__ __noinline__ void tangent(const double *v) {
        using namespace qc;
        const int N = this->N;
        const double *const AB = this->AB;
        const StepRing R = step_ring<RING_SLOT, NCST * COL>();
        const unsigned slot0 = smem_addr(R.slot), bar0 = smem_addr(R.bar), xch0 = smem_addr(R.xch);
        constexpr unsigned SLOT_B = RING_SLOT * 8, RING_B = RING_D * SLOT_B;
        const int lane = lane_id();
        const int rl = lane < NX ? lane : NX - 1;
        double crow[NZ]; // this lane's row of the constant columns (the entries of the variable columns stay unused)
        {
            int jc = 0; // compile-time after unrolling: index among the constant columns (ABc is compacted)
#pragma unroll
            for (int c = 0; c < NZ; ++c) {
                if ((VAR_COLS >> c) & 1u) crow[c] = 0.0;
                else { crow[c] = this->ABc[jc * COL + rl]; ++jc; }
            }
        }
        if (lane == 0) {
#pragma unroll
            for (int sl = 0; sl < RING_D; ++sl) mbar_init(R.bar + sl, 1);
            mbar_init_fence();
            for (int k = 0; k < RING_D && k < N; ++k) {
                mbar_expect_tx_a(bar0 + k * 8, ABS * 8);
                bulk_g2s_a(slot0 + k * SLOT_B, AB + size_t(k) * ABS, ABS * 8, bar0 + k * 8);
            }
        }
        double *dt = this->dtraj + lane; // lanes 0..8
        if (lane < NX) { R.xch[lane] = 0.0; *dt = 0.0; }
        if (lane < NU) R.xch[NX + lane] = v[lane];
        __syncwarp();
        const double *vnext = v + NU + (lane < NU ? lane : 0);
        unsigned rec = slot0, bar = bar0, par = 0, xr = xch0, xw = xch0 + 128;
#pragma unroll 1
        for (int k = 0; k < N; ++k) {
            double vn = 0;
            if (k + 1 < N && lane < NU) vn = *vnext;
            vnext += NU;
            mbar_wait_a(bar, par);
            double row[NZ], z[NZ];
            {
                int j = 0; // compile-time after unrolling: index among the variable columns
#pragma unroll
                for (int c = 0; c < NZ; ++c) {
                    if ((VAR_COLS >> c) & 1u) { row[c] = lds_f64(rec + (j * COL + rl) * 8); ++j; }
                    else row[c] = crow[c];
                }
            }
            lds_vec<NZ>(xr, z);
#if PANTR_QC_SPLIT_CHAIN
            const double t0 = fma(row[4], z[4], fma(row[3], z[3], fma(row[2], z[2], fma(row[1], z[1], row[0] * z[0]))));
            const double t1 = fma(row[8], z[8], fma(row[7], z[7], fma(row[6], z[6], row[5] * z[5])));
            const double t2 = fma(row[12], z[12], fma(row[11], z[11], fma(row[10], z[10], row[9] * z[9])));
            const double a = (t0 + t1) + t2;
#else
            double a = row[0] * z[0];
#pragma unroll
            for (int c = 1; c < NZ; ++c) a = fma(row[c], z[c], a);
#endif
            dt += NX;
            if (lane < NX) { *dt = a; sts_f64(xw + 8 * lane, a); }
            if (lane < NU) sts_f64(xw + 8 * (NX + lane), vn);
            __syncwarp(); // slot `rec` has been read by every lane: refill it
            if (lane == 0 && k + RING_D < N) {
                mbar_expect_tx_a(bar, ABS * 8);
                bulk_g2s_a(rec, AB + size_t(k + RING_D) * ABS, ABS * 8, bar);
            }
            { const unsigned t = xr; xr = xw; xw = t; }
            rec += SLOT_B; bar += 8;
            if (rec == slot0 + RING_B) { rec = slot0; bar = bar0; par ^= 1u; }
        }
    }
    /// node term of the second-order adjoint at node k: (s hess l + sum_i yh_i hess c_i + J_c' W J_c) dx_k
    __device__ __forceinline__ void node_second_order(int k, double cost_scale, double *out) const {
        PANTR_QC_LOCALS;
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
    /// row k of mus = [ d/dx_k[J_step' lambda_{k+1}] (dx_k, v_k) + node term of node k (k >= 1; row N = terminal node term) |
    ///                  d/du_k[...] (dx_k, v_k) + s 2 w_u v_k ].
    __device__ __noinline__ void second_order_sources(const double *u, const double *v, double cost_scale) {
        PANTR_QC_LOCALS;
        using namespace qc;
        for (int k = lane_id(); k <= N; k += 32) {
            if (k == N) {
                double nd[NX];
                node_second_order(N, cost_scale, nd);
#pragma unroll
                for (int i = 0; i < NX; ++i) mus[N * ROW + i] = nd[i];
                continue;
            }
            double xk[NX], uk[NU], vk[NU], dx[NX], l[NX];
            const double *s = sc + 24 * k;
            ldv<NX>(traj + k * NX, xk); ldv<NU>(u + NU * k, uk); ldv<NU>(v + NU * k, vk);
            ldv<NX>(dtraj + k * NX, dx); ldv<NX>(lam + (k + 1) * ROW, l);
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
                for (int i = 0; i < NX; ++i) mus[k * ROW + i] = sx[i] + nd[i];
            }
#pragma unroll
            for (int j = 0; j < NU; ++j) mus[k * ROW + NX + j] = fma(cost_scale * (2 * wu(j)), vk[j], su[j]);
        }
        __syncwarp();
    }
    __device__ __forceinline__ void mu_recurrence(double *Hv) { adjoint_recurrence(this->mus, Hv, false, 0.0); }
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
        if (rolled != x) rollout(x);
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
