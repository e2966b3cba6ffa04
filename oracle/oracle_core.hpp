// ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product path.
//
// CPU restatement (plain C++17, no Eigen) of the reference's PANTR + ALM hot
// path.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may build or call this.  Every function cites the
// reference file:line (paths relative to /root/reference/alpaqa/src/alpaqa/
// include/alpaqa unless stated otherwise).
//
// Parity status: the solver-independent known-answer tests of the reference
// (test-alm.cpp QPs, test-casadi.cpp Rosenbrock through the committed
// CasADi-generated C) pin this restatement; the three MPC models themselves
// have no golden vectors in the reference => "parity unpinned" for the model
// arithmetic (finite-difference + sympy cross-checks are used instead).
//
// Arithmetic conventions (shared with the CUDA path so that both follow the
// same floating point trajectory): compile with -ffp-contract=off; all
// reductions (dot, squared norm) use the "warp order": 32 strided partial
// sums followed by an xor-butterfly (16,8,4,2,1).  Eigen's own reduction order
// is unspecified, so this is as valid a restatement as any other order.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <limits>
#include <vector>

namespace oracle {

using real_t  = double;
using index_t = std::ptrdiff_t;
using vec     = std::vector<real_t>;
constexpr real_t inf = std::numeric_limits<real_t>::infinity();
constexpr real_t NaN = std::numeric_limits<real_t>::quiet_NaN();
constexpr real_t eps = std::numeric_limits<real_t>::epsilon();

// ---- warp-order reductions -------------------------------------------------
inline real_t butterfly_sum(real_t *p) {
    for (int m = 16; m >= 1; m >>= 1) {
        real_t q[32];
        for (int l = 0; l < 32; ++l) q[l] = p[l] + p[l ^ m];
        for (int l = 0; l < 32; ++l) p[l] = q[l];
    }
    return p[0];
}
inline real_t dot(const real_t *a, const real_t *b, index_t n) {
    real_t p[32] = {};
    for (index_t i = 0; i < n; ++i) p[i & 31] = p[i & 31] + a[i] * b[i];
    return butterfly_sum(p);
}
inline real_t sqnorm(const real_t *a, index_t n) { return dot(a, a, n); }
inline real_t norm2(const real_t *a, index_t n) { return std::sqrt(sqnorm(a, n)); }
inline real_t norm1(const real_t *a, index_t n) {
    real_t p[32] = {};
    for (index_t i = 0; i < n; ++i) p[i & 31] = p[i & 31] + std::fabs(a[i]);
    return butterfly_sum(p);
}
/// max |a_i|; NaN if any entry is NaN.
inline real_t norm_inf(const real_t *a, index_t n) {
    real_t m = 0;
    for (index_t i = 0; i < n; ++i) {
        real_t v = std::fabs(a[i]);
        if (v > m || std::isnan(v)) m = std::isnan(m) ? m : v;
    }
    return m;
}

// ---- enums (integer values matter for results) ------------------------------
// inner/internal/solverstatus.hpp:11-20
enum SolverStatus : int {
    Busy = 0, Converged, MaxTime, MaxIter, NotFinite, NoProgress, Interrupted, Exception,
};
// inner/internal/panoc-stop-crit.hpp:7-102 (same order)
enum PANOCStopCrit : int {
    ApproxKKT = 0, ApproxKKT2, ProjGradNorm, ProjGradNorm2, ProjGradUnitNorm,
    ProjGradUnitNorm2, FPRNorm, FPRNorm2, Ipopt, LBFGSBpp,
};

// ---- evaluation counters: problem/problem-counters.hpp:10-58 ----------------
struct EvalCounter {
    unsigned proj_diff_g = 0, proj_multipliers = 0, prox_grad_step = 0,
             inactive_indices_res_lna = 0, f = 0, grad_f = 0, f_grad_f = 0, f_g = 0,
             grad_f_grad_g_prod = 0, g = 0, grad_g_prod = 0, grad_gi = 0, grad_L = 0,
             hess_L_prod = 0, hess_L = 0, hess_psi_prod = 0, hess_psi = 0, psi = 0,
             grad_psi = 0, psi_grad_psi = 0;
};

// ---- Box: problem/box.hpp:9-53 ----------------------------------------------
struct Box {
    vec lowerbound, upperbound;
};

// ---- Problem API surface ------------------------------------------------------
// problem/type-erased-problem.hpp:31-95 (the subset PANTR+ALM touches), with the
// BoxConstrProblem behaviour (problem/box-constr-problem.hpp) built in.
struct Problem {
    index_t n = 0, m = 0;
    Box C, D;
    index_t penalty_alm_split = 0;
    mutable EvalCounter evaluations;
    virtual ~Problem() = default;

    // -- to be supplied by the concrete model (CasADiProblem.tpp:208-368) --
    virtual real_t eval_f_impl(const real_t *x) const                         = 0;
    virtual void eval_g_impl(const real_t *x, real_t *g) const                = 0;
    virtual real_t eval_psi_impl(const real_t *x, const real_t *y, const real_t *S,
                                 real_t *yhat) const                          = 0;
    virtual real_t eval_psi_grad_psi_impl(const real_t *x, const real_t *y,
                                          const real_t *S, real_t *grad) const = 0;
    virtual void eval_grad_L_impl(const real_t *x, const real_t *y,
                                  real_t *grad_L) const                       = 0;
    virtual void eval_hess_psi_prod_impl(const real_t *x, const real_t *y,
                                         const real_t *S, real_t scale,
                                         const real_t *v, real_t *Hv) const   = 0;

    // -- problem plugin face beyond what PANTR calls (CasADiProblem.tpp:226-233,268-276,300-313):
    //    J_g' y, grad L = grad f + J_g' y (returns f), hess_xx L v --
    //    (only the MPC models implement them; the KAT problems do not need them)
    virtual void face_grad_g_prod(const real_t *, const real_t *, real_t *) const {}
    virtual real_t face_grad_L(const real_t *, const real_t *, real_t *) const { return 0; }
    virtual void face_hess_L_prod(const real_t *, const real_t *, real_t, const real_t *, real_t *) const {}

    // grad L where PANTR really needs it (stop criteria on grad psi(x_hat), pantr.tpp:246-249).  The MPC models keep
    // eval_grad_L_impl empty because its other call site is numerically dead (pantr.tpp:306-314); they override this.
    virtual void grad_L_live(const real_t *x, const real_t *y, real_t *g) const { eval_grad_L_impl(x, y, g); }

    // -- counted wrappers: problem/problem-with-counters.hpp:24-112 --
    real_t eval_f(const real_t *x) const { ++evaluations.f; return eval_f_impl(x); }
    void eval_g(const real_t *x, real_t *g) const { ++evaluations.g; eval_g_impl(x, g); }
    real_t eval_psi(const real_t *x, const real_t *y, const real_t *S, real_t *yhat) const {
        ++evaluations.psi; return eval_psi_impl(x, y, S, yhat);
    }
    real_t eval_psi_grad_psi(const real_t *x, const real_t *y, const real_t *S, real_t *g) const {
        ++evaluations.psi_grad_psi; return eval_psi_grad_psi_impl(x, y, S, g);
    }
    // CasADiProblem.tpp:243-258: grad_psi is evaluated through psi_grad_psi
    void eval_grad_psi(const real_t *x, const real_t *y, const real_t *S, real_t *g) const {
        ++evaluations.grad_psi; (void)eval_psi_grad_psi_impl(x, y, S, g);
    }
    void eval_grad_L(const real_t *x, const real_t *y, real_t *g) const {
        ++evaluations.grad_L; eval_grad_L_impl(x, y, g);
    }
    void eval_hess_psi_prod(const real_t *x, const real_t *y, const real_t *S, real_t scale,
                            const real_t *v, real_t *Hv) const {
        ++evaluations.hess_psi_prod; eval_hess_psi_prod_impl(x, y, S, scale, v, Hv);
    }

    // box-constr-problem.hpp:59-64,107-114 (l1_reg empty => h = 0)
    real_t eval_prox_grad_step(real_t gamma, const real_t *x, const real_t *grad_psi,
                               real_t *xhat, real_t *p) const {
        ++evaluations.prox_grad_step;
        for (index_t i = 0; i < n; ++i) {
            real_t s = -gamma * grad_psi[i];
            // Eigen cwiseMax/cwiseMin: max(a,b) = a<b ? b : a ; min(a,b) = b<a ? b : a
            real_t lo = C.lowerbound[i] - x[i], hi = C.upperbound[i] - x[i];
            real_t t = s < lo ? lo : s;
            p[i]    = hi < t ? hi : t;
            xhat[i] = x[i] + p[i];
        }
        return 0;
    }
    // box-constr-problem.hpp:117, box.hpp:34-53
    void eval_proj_diff_g(const real_t *z, real_t *p) const {
        ++evaluations.proj_diff_g;
        for (index_t i = 0; i < m; ++i)
            p[i] = z[i] - std::fmin(std::fmax(z[i], D.lowerbound[i]), D.upperbound[i]);
    }
    // box-constr-problem.hpp:119-143
    void eval_proj_multipliers(real_t *y, real_t M) const {
        ++evaluations.proj_multipliers;
        for (index_t i = 0; i < m; ++i) {
            if (i < penalty_alm_split) { y[i] = 0; continue; }
            real_t y_lb = D.lowerbound[i] == -inf ? 0 : -M;
            real_t y_ub = D.upperbound[i] == +inf ? 0 : +M;
            y[i] = std::min(std::max(y[i], y_lb), y_ub);
        }
    }
    // box-constr-problem.hpp:151-185 (λ = 0 branch)
    index_t eval_inactive_indices_res_lna(real_t gamma, const real_t *x, const real_t *grad_psi,
                                          index_t *J) const {
        ++evaluations.inactive_indices_res_lna;
        index_t nJ = 0;
        for (index_t i = 0; i < n; ++i) {
            real_t x_fw = x[i] - gamma * grad_psi[i];
            if (C.lowerbound[i] < x_fw && x_fw < C.upperbound[i]) J[nJ++] = i;
        }
        return nJ;
    }
};

// util/index-set.hpp:46-50,81-94 — K = {0..n-1} \ J, ascending
inline void compute_complement(const index_t *J, index_t nJ, index_t *K, index_t n) {
    index_t j = 0, k = 0;
    for (index_t i = 0; i < n; ++i) {
        if (j < nJ && J[j] == i) ++j; else K[k++] = i;
    }
}

} // namespace oracle
