// Device-side PANTR (prox-gradient + Newton trust-region / Steihaug-CG) for one instance per warp.
// Follows, step by step, the reference's
//   PANTRSolver::operator()        alpaqa/implementation/inner/pantr.tpp:25-473
//   NewtonTRDirection::apply       alpaqa/inner/directions/pantr/newton-tr.hpp:91-155
//   SteihaugCG::solve              alpaqa/accelerators/steihaugcg.hpp:39-140
//   BoxConstrProblem::*            alpaqa/problem/box-constr-problem.hpp:59-185
//   PANOCHelpers::*                alpaqa/implementation/inner/panoc-helpers.tpp:62-137,215-249,288-334
// All scalars are warp-uniform; vectors are lane-strided in the instance workspace.
#pragma once
#include "../../include/pantr_b200.h"
#include "exec_policy.cuh"

namespace pantr {

struct SolverParams {
    pantr_b200_alm_params alm;
    pantr_b200_pantr_params pantr;
    pantr_b200_steihaug_params cg;
    pantr_b200_newton_tr_params dir;
};

struct IterScalars { double psi_x, psi_xh, gamma, L, pTp, gTp, h; };

struct Trace { double *rows; int *n_rows; int max_rows; long long instance; };

// Number of n-vectors / m-vectors / ints in the workspace (see Workspace::bind)
constexpr int WS_NVEC = 26, WS_MVEC = 10;

template <class Model, class E = WarpExec>
struct Pantr {
    using Vec = typename E::Vec;
    using CVec = typename E::CVec;
    using IVec = typename E::IVec;
    Model &mdl;
    const SolverParams &P;
    IterScalars *it; // [3] in shared memory
    int n, m;
    // iterate storage (slot-major)
    Vec X, XH, G, PP, YH;
    Vec q, work, work2, rJ, qJ, z, r, d, Bd, we, wn1, wn2, gxh;
    Vec x, y, Sig, Sig_old, e1, e2; // ALM-level vectors
    IVec JK;
    pantr_b200_stats st;
    unsigned cg_last;
    unsigned long long t_alm_start = 0, t_inner_start = 0; // %globaltimer at the start of the ALM / the inner solve
    bool tracing;
    Trace tr;

    /// sm / smem_set: the Steihaug-CG vectors (set >= 1: z, r, d, Bd; set 2: also work, work2 and the
    /// J/K index list; set 3: also the CG iterate qJ) live in the block's shared memory instead of the global workspace.
    __host__ __device__ static size_t smem_bytes(int n, int set) {
        return set == 0 ? 0 : (set == 1 ? size_t(4) * n * 8
                                        : size_t(6) * n * 8 + size_t(n + 2) / 2 * 8 + (set >= 3 ? size_t(n) * 8 : 0));
    }
    __device__ Pantr(Model &mdl_, const SolverParams &P_, IterScalars *it_, int n_, int m_, Vec ws,
                     double *sm = nullptr, int smem_set = 0)
        : mdl(mdl_), P(P_), it(it_), n(n_), m(m_) {
        Vec p = ws;
        X = p; p += 3 * n; XH = p; p += 3 * n; G = p; p += 3 * n; PP = p; p += 3 * n;
        q = p; p += n; work = p; p += n; work2 = p; p += n; rJ = p; p += n; qJ = p; p += n;
        z = p; p += n; r = p; p += n; d = p; p += n; Bd = p; p += n; we = p; p += n;
        wn1 = p; p += n; wn2 = p; p += n; gxh = p; p += n; // 12 + 13 = 25 n-vectors
        YH = p; p += 3 * m;
        x = p; p += n; // ThreadExec only: the iterate, copied in / out of the batch arrays (26th n-vector)
        y = p; p += m; Sig = p; p += m; Sig_old = p; p += m; e1 = p; p += m; e2 = p; p += m;
        p += 2 * m; // spare (10 m-vectors)
        if constexpr (E::THREAD) {
            // 32 n ints of the warp's block fit in its 32 (n / 2 + 2) doubles; the lane offset is in int units
            JK = IVec{reinterpret_cast<int *>(p.p - lane_id()) + lane_id()};
        } else {
            JK = reinterpret_cast<int *>(p);
            if (smem_set >= 1) { z = sm; r = sm + n; d = sm + 2 * n; Bd = sm + 3 * n; }
            if (smem_set >= 2) { work = sm + 4 * n; work2 = sm + 5 * n; JK = reinterpret_cast<int *>(sm + 6 * n); }
            if (smem_set >= 3) qJ = sm + 6 * n + (n + 2) / 2; // the CG iterate s: written, normed and copied every CG iteration
        }
    }
    __host__ __device__ static size_t ws_doubles(int n, int m) {
        return size_t(WS_NVEC) * n + size_t(WS_MVEC) * m + (n + 1) / 2 + 2;
    }

    /// element-wise vector loop of the execution policy, batched where the model asks for it (exec_policy.cuh)
    template <class L, class S>
    __device__ __forceinline__ void emap(int nn, L load, S store) const { E::template map<Model::BATCHED_MAPS>(nn, load, store); }
    // ---- problem function wrappers (+ counters, problem-with-counters.hpp:24-112) ----
    __device__ __noinline__ void eval_psi_grad_psi(int s) {
        ++st.ev_psi_grad_psi;
        double v = mdl.eval_psi_grad_psi(X + s * n, y, Sig, G + s * n);
        it[s].psi_x = v;
    }
    // box-constr-problem.hpp:59-64 ; returns h = 0
    __device__ __noinline__ void prox_grad_step_raw(double gamma, CVec xx, CVec g, Vec xh, Vec p) {
        ++st.ev_prox_grad_step;
        emap(n, [&](int i) {
            double xi = xx[i], s = -gamma * g[i];
            double lo = Model::C_lb(i) - xi, hi = Model::C_ub(i) - xi;
            double t  = s < lo ? lo : s;
            double pi = hi < t ? hi : t;
            return make_double2(pi, xi + pi);
        }, [&](int i, double2 e) { p[i] = e.x; xh[i] = e.y; });
    }
    __device__ __noinline__ void eval_prox_grad_step(int s) { // pantr.tpp:88-92
        prox_grad_step_raw(it[s].gamma, X + s * n, G + s * n, XH + s * n, PP + s * n);
        E::sync();
        double pTp = xsqnorm<E>(PP + s * n, n), gTp = xdot<E>(PP + s * n, G + s * n, n);
        it[s].h = 0; it[s].pTp = pTp; it[s].gTp = gTp;
    }
    __device__ __noinline__ void eval_psi_xh(int s) { // pantr.tpp:93-95
        ++st.ev_psi;
        double v = mdl.eval_psi(XH + s * n, y, Sig, YH + s * m);
        it[s].psi_xh = v;
    }
    __device__ double fbe(int s) const { // pantr.tpp:69
        return it[s].psi_x + it[s].h + it[s].pTp / (2 * it[s].gamma) + it[s].gTp;
    }
    __device__ bool qub_violated(int s) const { // pantr.tpp:102-106
        double margin = (1 + fabs(it[s].psi_x)) * P.pantr.quadratic_upperbound_tolerance_factor;
        return it[s].psi_xh > it[s].psi_x + it[s].gTp + 0.5 * it[s].L * it[s].pTp + margin;
    }
    __device__ __noinline__ void backtrack_qub(int s) { // pantr.tpp:107-115
        while (it[s].L < P.pantr.L_max && qub_violated(s)) {
            E::sync();
            double g2 = it[s].gamma / 2, L2 = it[s].L * 2;
            E::sync();
            it[s].gamma = g2; it[s].L = L2;
            E::sync();
            eval_prox_grad_step(s);
            eval_psi_xh(s);
            E::sync();
        }
    }
#include "pantr_direction.cuh"
#include "pantr_inner.cuh"
    InnerState ist_;
    InnerResult ires_;
};

} // namespace pantr
