// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_core.hpp header).
// ALMSolver::operator() restated from implementation/outer/alm.tpp:19-226 and
// implementation/outer/internal/alm-helpers.tpp:16-59.
#pragma once
#include "oracle_pantr.hpp"

namespace oracle {

// alm-helpers.tpp:16-45
inline void update_penalty_weights(const ALMParams &params, real_t Delta, bool first_iter,
                                   const vec &e, const vec &old_e, real_t norm_e, real_t old_norm_e,
                                   const vec &S_old, vec &S, bool monotone) {
    const real_t theta = params.rel_penalty_increase_threshold;
    if (norm_e <= params.dual_tolerance) { S = S_old; return; }
    if (params.single_penalty_factor) {
        if (first_iter || norm_e > theta * old_norm_e) {
            real_t new_S = std::fmin(params.max_penalty, Delta * S_old[0]);
            std::fill(S.begin(), S.end(), new_S);
        } else S = S_old;
    } else {
        for (size_t i = 0; i < e.size(); ++i) {
            if (first_iter || std::abs(e[i]) > theta * std::abs(old_e[i]))
                S[i] = std::fmin(params.max_penalty,
                                 std::fmax(Delta * std::abs(e[i]) / norm_e, real_t(1) * monotone) * S_old[i]);
            else
                S[i] = S_old[i];
        }
    }
}

// alm-helpers.tpp:47-59
inline void initialize_penalty(const Problem &p, const ALMParams &params, const vec &x0, vec &S) {
    real_t f0 = p.eval_f(x0.data());
    vec g0(p.m);
    p.eval_g(x0.data(), g0.data());
    real_t sigma = params.initial_penalty_factor * std::max(real_t(1), std::abs(f0)) /
                   std::max(real_t(1), real_t(0.5) * sqnorm(g0.data(), p.m));
    sigma = std::clamp(sigma, params.min_penalty, params.max_penalty);
    std::fill(S.begin(), S.end(), sigma);
}

struct ALMSolver {
    ALMParams params;
    PANTRSolver inner_solver;

    void accumulate(ALMStats &s, const PANTRStats &ps) const { // inner/pantr.hpp:242-258
        s.inner_iterations += ps.iterations;
        s.accelerated_step_rejected += ps.accelerated_step_rejected;
        s.stepsize_backtracks += ps.stepsize_backtracks;
        s.direction_failures += ps.direction_failures;
        s.direction_update_rejected += ps.direction_update_rejected;
        s.final_gamma = ps.final_gamma; s.final_psi = ps.final_psi;
        s.final_h = ps.final_h; s.final_fbe = ps.final_fbe;
    }

    // alm.tpp:19-226 (wall-clock limits not modelled)
    ALMStats operator()(const Problem &p, vec &x, vec &y) {
        const index_t m = p.m;
        ALMStats s;
        if (m == 0) { // alm.tpp:29-52
            vec S(0), error(0);
            InnerSolveOptions opts{true, params.tolerance, 0};
            auto ps = inner_solver(p, opts, x, y, S, error);
            s.inner_convergence_failures = ps.status != Converged;
            accumulate(s, ps);
            s.eps_ = ps.eps_; s.delta = 0; s.norm_penalty = 0; s.outer_iterations = 1;
            s.status = ps.status;
            return s;
        }
        vec S(m, NaN), S_old(m, NaN), error_1(m, NaN), error_2(m, NaN);
        real_t norm_e_1 = NaN, norm_e_2 = NaN;
        if (params.initial_penalty > 0) std::fill(S.begin(), S.end(), params.initial_penalty);
        else initialize_penalty(p, params, x, S);
        real_t eps_ = params.initial_tolerance, eps_old = NaN;
        real_t Delta = params.penalty_update_factor, rho = params.tolerance_update_factor;
        unsigned num_successful_iters = 0;
        for (unsigned i = 0; i < params.max_iter; ++i) {
            p.eval_proj_multipliers(y.data(), params.max_multiplier);
            bool out_of_penalty_factor_updates =
                (num_successful_iters == 0 ? s.initial_penalty_reduced == params.max_num_initial_retries
                                           : s.penalty_reduced == params.max_num_retries) ||
                (s.initial_penalty_reduced + s.penalty_reduced == params.max_total_num_retries);
            bool out_of_iter       = i + 1 == params.max_iter;
            bool overwrite_results = out_of_iter || out_of_penalty_factor_updates;
            InnerSolveOptions opts{overwrite_results, eps_, i};
            auto ps              = inner_solver(p, opts, x, y, S, error_2);
            bool inner_converged = ps.status == Converged;
            s.inner_convergence_failures += !inner_converged;
            accumulate(s, ps);
            if (ps.status == Exception) { // std::logic_error of the direction (newton-tr.hpp:96-99) leaves operator()
                s.eps_ = ps.eps_; s.delta = 0; s.outer_iterations = i + 1; s.status = Exception;
                return s;
            }
            bool backtrack = !inner_converged && !overwrite_results;
            if (backtrack) { // alm.tpp:155-187
                if (num_successful_iters > 0) {
                    Delta = std::fmax(params.min_penalty_update_factor, Delta * params.penalty_update_factor_lower);
                    update_penalty_weights(params, Delta, false, error_1, error_2, norm_e_1, norm_e_2, S_old, S, true);
                    rho  = std::fmin(params.rho_max, rho * params.rho_increase);
                    eps_ = std::fmax(rho * eps_old, params.tolerance);
                    ++s.penalty_reduced;
                } else {
                    for (auto &v : S) v *= params.initial_penalty_lower;
                    eps_ *= params.initial_tolerance_increase;
                    ++s.initial_penalty_reduced;
                }
            } else { // alm.tpp:190-222
                error_2.swap(error_1);
                norm_e_2 = std::exchange(norm_e_1, norm_inf(error_1.data(), m));
                bool alm_converged = ps.eps_ <= params.tolerance && inner_converged &&
                                     norm_e_1 <= params.dual_tolerance;
                if (alm_converged || out_of_iter) {
                    s.eps_ = ps.eps_; s.delta = norm_e_1; s.norm_penalty = norm2(S.data(), m);
                    s.outer_iterations = i + 1;
                    s.status = alm_converged ? Converged : MaxIter;
                    return s;
                }
                S_old.swap(S);
                update_penalty_weights(params, Delta, num_successful_iters == 0, error_1, error_2,
                                       norm_e_1, norm_e_2, S_old, S, true);
                eps_old = std::exchange(eps_, std::fmax(rho * eps_, params.tolerance));
                ++num_successful_iters;
            }
        }
        s.status = Exception;
        return s;
    }
};

} // namespace oracle
