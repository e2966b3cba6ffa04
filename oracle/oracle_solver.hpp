// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_core.hpp header).
// SteihaugCG, NewtonTRDirection, PANTRSolver, ALMSolver restated from the reference.
#pragma once
#include "oracle_core.hpp"
#include <utility>

namespace oracle {

// accelerators/steihaugcg.hpp:9-15
struct SteihaugCGParams {
    real_t tol_scale = 1, tol_scale_root = 0.5, tol_max = inf, max_iter_factor = 1;
};
// inner/directions/pantr/newton-tr.hpp:16-23
struct NewtonTRDirectionParams {
    bool rescale_on_step_size_changes = false;
    real_t hessian_vec_factor         = 0.5;
    bool finite_diff                  = false;
    real_t finite_diff_stepsize       = 1.4901161193847656e-08; // sqrt(eps)
};
// inner/internal/lipschitz.hpp:8-20
struct LipschitzEstimateParams {
    real_t L_0 = 0, eps = 1e-6, delta = 1e-12, Lgamma_factor = 0.95;
};
// inner/pantr.hpp:22-91
struct PANTRParams {
    LipschitzEstimateParams Lipschitz;
    unsigned max_iter        = 100;
    real_t L_min = 1e-5, L_max = 1e20;
    int stop_crit            = ApproxKKT;
    unsigned max_no_progress = 10;
    real_t quadratic_upperbound_tolerance_factor = 10 * eps;
    real_t TR_tolerance_factor                   = 10 * eps;
    real_t ratio_threshold_acceptable = 0.2, ratio_threshold_good = 0.8;
    real_t radius_factor_rejected = 0.35, radius_factor_acceptable = 0.999,
           radius_factor_good = 2.5;
    real_t initial_radius = NaN, min_radius = 100 * eps;
    bool compute_ratio_using_new_stepsize               = false;
    bool update_direction_on_prox_step                  = true;
    bool recompute_last_prox_step_after_direction_reset = false;
    bool disable_acceleration                           = false;
    bool ratio_approx_fbe_quadratic_model               = true;
};
// inner/pantr.hpp:93-110
struct PANTRStats {
    int status = Busy;
    real_t eps_ = inf;
    unsigned iterations = 0, accelerated_step_rejected = 0, stepsize_backtracks = 0,
             direction_failures = 0, direction_update_rejected = 0;
    real_t final_gamma = 0, final_psi = 0, final_h = 0, final_fbe = 0;
};
// outer/alm.hpp:19-90
struct ALMParams {
    real_t tolerance = 1e-5, dual_tolerance = 1e-5, penalty_update_factor = 10,
           penalty_update_factor_lower = 0.8, min_penalty_update_factor = 1.1,
           initial_penalty = 1, initial_penalty_factor = 20, initial_penalty_lower = 0.6,
           initial_tolerance = 1, initial_tolerance_increase = 1.1,
           tolerance_update_factor = 1e-1, rho_increase = 2, rho_max = 0.5,
           rel_penalty_increase_threshold = 0.1, max_multiplier = 1e9, max_penalty = 1e9,
           min_penalty = 1e-9;
    unsigned max_iter = 100;
    unsigned max_num_initial_retries = 0, max_num_retries = 0, max_total_num_retries = 0;
    bool single_penalty_factor = false;
};
// outer/alm.hpp:104-149 + inner accumulation inner/pantr.hpp:242-258
struct ALMStats {
    unsigned outer_iterations = 0, initial_penalty_reduced = 0, penalty_reduced = 0,
             inner_convergence_failures = 0;
    real_t eps_ = inf, delta = inf, norm_penalty = 0;
    int status = Busy;
    // accumulated inner stats
    unsigned inner_iterations = 0, accelerated_step_rejected = 0, direction_failures = 0,
             direction_update_rejected = 0, stepsize_backtracks = 0;
    real_t final_gamma = 0, final_psi = 0, final_h = 0, final_fbe = 0;
};
// inner/inner-solve-options.hpp:11-28 (time limits are not modelled)
struct InnerSolveOptions {
    bool always_overwrite_results = true;
    real_t tolerance              = 0;
    unsigned outer_iter           = 0;
};

// ---- SteihaugCG: accelerators/steihaugcg.hpp:39-140 --------------------------
struct SteihaugCG {
    SteihaugCGParams params;
    mutable vec z, r, d, Bd, work_eval;
    mutable unsigned last_iters = 0;
    void resize(index_t n) { z.resize(n); r.resize(n); d.resize(n); Bd.resize(n); work_eval.resize(n); }

    // steihaugcg.hpp:122-140
    static std::pair<real_t, real_t> get_boundaries_intersections(const real_t *z, const real_t *d,
                                                                  index_t n, real_t trust_radius) {
        real_t a = sqnorm(d, n);
        real_t b = 2 * dot(z, d, n);
        real_t c = sqnorm(z, n) - trust_radius * trust_radius;
        real_t sqrt_discriminant = std::sqrt(b * b - 4 * a * c);
        real_t aux = b + std::copysign(sqrt_discriminant, b);
        real_t ta  = -aux / (2 * a);
        real_t tb  = -2 * c / aux;
        return {std::fmin(ta, tb), std::fmax(ta, tb)};
    }

    // steihaugcg.hpp:39-117
    template <class HessFun>
    real_t solve(const real_t *g, index_t n, const HessFun &hess_prod, real_t trust_radius,
                 real_t *s) const {
        real_t *z = this->z.data(), *r = this->r.data(), *d = this->d.data(),
               *Bd = this->Bd.data(), *we = this->work_eval.data();
        for (index_t i = 0; i < n; ++i) { z[i] = 0; r[i] = g[i]; d[i] = -r[i]; }
        real_t r_sq     = sqnorm(r, n);
        real_t grad_mag = norm2(g, n);
        real_t tolerance = std::fmin(params.tol_max,
                                     params.tol_scale * grad_mag *
                                         std::fmin(params.tol_scale_root, std::sqrt(grad_mag)));
        auto eval = [&](const real_t *p) {
            hess_prod(p, we);
            return dot(p, g, n) + real_t(0.5) * dot(p, we, n);
        };
        index_t i           = 0;
        const auto max_iter = static_cast<index_t>(
            std::round(static_cast<real_t>(n) * params.max_iter_factor));
        last_iters = 0;
        while (true) {
            ++last_iters;
            hess_prod(d, Bd);
            real_t dBd = dot(d, Bd, n);
            if (dBd <= 0) {
                auto [ta, tb] = get_boundaries_intersections(z, d, n, trust_radius);
                real_t *pa = r, *pb = d; // storage reuse as in the reference
                for (index_t j = 0; j < n; ++j) {
                    real_t zj = z[j], dj = d[j];
                    pa[j] = zj + ta * dj;
                    pb[j] = zj + tb * dj;
                }
                real_t q_a = eval(pa), q_b = eval(pb);
                real_t q_min = std::fmin(q_a, q_b);
                if (q_a == q_min) { std::copy(pa, pa + n, s); return q_a; }
                else              { std::copy(pb, pb + n, s); return q_b; }
            }
            real_t alpha = r_sq / dBd;
            for (index_t j = 0; j < n; ++j) s[j] = z[j] + alpha * d[j];
            if (norm2(s, n) >= trust_radius) {
                auto [ta, tb] = get_boundaries_intersections(z, d, n, trust_radius);
                (void)ta;
                for (index_t j = 0; j < n; ++j) s[j] = z[j] + tb * d[j];
                return eval(s);
            }
            for (index_t j = 0; j < n; ++j) r[j] = r[j] + alpha * Bd[j];
            real_t r_next_sq = sqnorm(r, n);
            if (std::sqrt(r_next_sq) < tolerance || i > max_iter) return eval(s);
            real_t beta_next = r_next_sq / r_sq;
            r_sq             = r_next_sq;
            for (index_t j = 0; j < n; ++j) { d[j] = beta_next * d[j] - r[j]; z[j] = s[j]; }
            ++i;
        }
    }
};

// ---- NewtonTRDirection: inner/directions/pantr/newton-tr.hpp:47-155 -----------
struct NewtonTRDirection {
    SteihaugCG steihaug;
    NewtonTRDirectionParams direction_params;
    const Problem *problem = nullptr;
    const real_t *y = nullptr, *S = nullptr;
    mutable std::vector<index_t> JK;
    mutable vec rJ, qJ, work, work_2;
    mutable index_t last_nJ = 0;

    void initialize(const Problem &p, const real_t *y_, const real_t *S_) {
        problem = &p; y = y_; S = S_;
        JK.resize(p.n); rJ.resize(p.n); qJ.resize(p.n); work.resize(p.n); work_2.resize(p.n);
        steihaug.resize(p.n);
    }
    bool has_initial_direction() const { return true; }

    // newton-tr.hpp:91-155 (finite_diff = false branch)
    real_t apply(real_t gamma, const real_t *xk, const real_t *pk, const real_t *grad_psi,
                 real_t radius, real_t *qk) const {
        const index_t n = problem->n;
        index_t nJ = problem->eval_inactive_indices_res_lna(gamma, xk, grad_psi, JK.data());
        index_t *J = JK.data(), *K = JK.data() + nJ;
        compute_complement(J, nJ, K, n);
        last_nJ = nJ;
        const index_t nK = n - nJ;
        for (index_t j = 0; j < nJ; ++j) { rJ[j] = (-real_t(1) / gamma) * pk[J[j]]; qk[J[j]] = 0; }
        vec pK(nK);
        for (index_t k = 0; k < nK; ++k) { qk[K[k]] = pk[K[k]]; pK[k] = pk[K[k]]; }
        real_t norm_qK_sq = sqnorm(pK.data(), nK);
        if (direction_params.hessian_vec_factor != 0) {
            problem->eval_hess_psi_prod(xk, y, S, 1, qk, work.data());
            for (index_t j = 0; j < nJ; ++j)
                rJ[j] = rJ[j] + work[J[j]] * direction_params.hessian_vec_factor;
        }
        auto hess_vec_mult = [&](const real_t *p, real_t *Bp) {
            std::fill(work.begin(), work.end(), real_t(0));
            for (index_t j = 0; j < nJ; ++j) work[J[j]] = p[j];
            problem->eval_hess_psi_prod(xk, y, S, 1, work.data(), work_2.data());
            for (index_t j = 0; j < nJ; ++j) Bp[j] = work_2[J[j]];
        };
        real_t qJ_model = steihaug.solve(rJ.data(), nJ, hess_vec_mult, radius, qJ.data());
        for (index_t j = 0; j < nJ; ++j) qk[J[j]] = qJ[j];
        return qJ_model - norm_qK_sq / (2 * gamma);
    }
};

} // namespace oracle
