// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_core.hpp header).
// PANTRSolver::operator() and ALMSolver::operator() restated from the reference.
#pragma once
#include "oracle_solver.hpp"

namespace oracle {

// Per-iteration trace record (for iteration-level parity tests against the GPU path)
struct TraceRec {
    unsigned k; real_t psi, fbe, gamma, eps_, radius, rho, q_model; int nJ; unsigned cg_iters; int accepted;
};

struct PANTRSolver {
    PANTRParams params;
    NewtonTRDirection direction;
    std::vector<TraceRec> *trace = nullptr;

    // implementation/inner/pantr.tpp:53-75
    struct Iterate {
        vec x, xhat, grad_psi, p, yhat;
        real_t psi_x = NaN, psi_xhat = NaN, gamma = NaN, L = NaN, pTp = NaN, grad_psiTp = NaN,
               h_xhat = NaN;
        real_t fbe() const { return psi_x + h_xhat + pTp / (2 * gamma) + grad_psiTp; }
        Iterate(index_t n, index_t m) : x(n), xhat(n), grad_psi(n), p(n), yhat(m) {}
    };

    // implementation/inner/panoc-helpers.tpp:62-137 (branches PANTR configs use)
    // panoc-helpers.tpp:41-59
    static bool stop_crit_requires_grad_psi_xhat(int crit) { return crit == ApproxKKT || crit == ApproxKKT2 || crit == Ipopt; }
    static real_t calc_error_stop_crit(const Problem &problem, int crit, const vec &pk, real_t gamma,
                                       const vec &xk, const vec &xhatk, const vec &yhatk, const vec &grad_psik,
                                       const vec &grad_psi_hatk, vec &work_n1, vec &work_n2) {
        const index_t n = problem.n;
        switch (crit) {
            case ApproxKKT:
            case ApproxKKT2: {
                const real_t ig = 1 / gamma;
                for (index_t i = 0; i < n; ++i) work_n2[i] = ig * pk[i] + (grad_psik[i] - grad_psi_hatk[i]);
                return crit == ApproxKKT ? norm_inf(work_n2.data(), n) : norm2(work_n2.data(), n);
            }
            case ProjGradNorm: return norm_inf(pk.data(), n);
            case ProjGradNorm2: return norm2(pk.data(), n);
            case ProjGradUnitNorm:
                problem.eval_prox_grad_step(1, xk.data(), grad_psik.data(), work_n1.data(), work_n2.data());
                return norm_inf(work_n2.data(), n);
            case ProjGradUnitNorm2:
                problem.eval_prox_grad_step(1, xk.data(), grad_psik.data(), work_n1.data(), work_n2.data());
                return norm2(work_n2.data(), n);
            case FPRNorm: return norm_inf(pk.data(), n) / gamma;
            case FPRNorm2: return norm2(pk.data(), n) / gamma;
            case Ipopt: {
                problem.eval_prox_grad_step(1, xhatk.data(), grad_psi_hatk.data(), work_n1.data(), work_n2.data());
                real_t err = norm_inf(work_n2.data(), n);
                auto nn    = 2 * (index_t(yhatk.size()) + n);
                if (nn == 0) return err;
                for (index_t i = 0; i < n; ++i) work_n2[i] = work_n2[i] - grad_psi_hatk[i];
                real_t C_lagr_mult = norm1(work_n2.data(), n), D_lagr_mult = norm1(yhatk.data(), index_t(yhatk.size()));
                const real_t s_max = 100, s_n = static_cast<real_t>(nn);
                real_t s_d = std::fmax(s_max, (C_lagr_mult + D_lagr_mult) / s_n) / s_max;
                return err / s_d;
            }
            case LBFGSBpp:
                problem.eval_prox_grad_step(1, xk.data(), grad_psik.data(), work_n1.data(), work_n2.data());
                return norm_inf(work_n2.data(), n) / std::fmax(real_t(1), norm2(xk.data(), n));
            default: return NaN; // the reference throws std::out_of_range (panoc-helpers.tpp:136)
        }
    }

    // implementation/inner/panoc-helpers.tpp:215-249 (no wall-clock limit, no stop signal)
    SolverStatus check_all_stop_conditions(const InnerSolveOptions &opts, unsigned iteration,
                                           real_t eps_k, unsigned no_progress) const {
        real_t tolerance     = opts.tolerance > 0 ? opts.tolerance : real_t(1e-8);
        bool out_of_iter     = iteration == params.max_iter;
        bool not_finite      = !std::isfinite(eps_k);
        bool converged       = eps_k <= tolerance;
        bool max_no_progress = no_progress > params.max_no_progress;
        return converged ? Converged : out_of_iter ? MaxIter : not_finite ? NotFinite
               : max_no_progress ? NoProgress : Busy;
    }

    // implementation/inner/panoc-helpers.tpp:288-334
    static real_t initial_lipschitz_estimate(const Problem &problem, const vec &x, const real_t *y,
                                             const real_t *S, real_t e, real_t delta, real_t L_min,
                                             real_t L_max, real_t &psi, vec &grad_psi, vec &work_x,
                                             vec &work_grad_psi) {
        const index_t n = problem.n;
        psi = problem.eval_psi_grad_psi(x.data(), y, S, grad_psi.data());
        vec h(n);
        for (index_t i = 0; i < n; ++i) {
            real_t g = grad_psi[i];
            h[i]      = g > 0 ? std::max(g * e, delta) : std::min(g * e, -delta);
            work_x[i] = x[i] - h[i];
        }
        real_t norm_h = norm2(h.data(), n);
        problem.eval_grad_psi(work_x.data(), y, S, work_grad_psi.data());
        vec diff(n);
        for (index_t i = 0; i < n; ++i) diff[i] = work_grad_psi[i] - grad_psi[i];
        real_t L = norm2(diff.data(), n) / norm_h;
        return std::clamp(L, L_min, L_max);
    }

    // implementation/inner/pantr.tpp:25-473
    PANTRStats operator()(const Problem &problem, const InnerSolveOptions &opts, vec &x, vec &y,
                          const vec &S, vec &err_z) {
        PANTRStats s;
        const index_t n = problem.n, m = problem.m;
        Iterate iterates[3]{{n, m}, {n, m}, {n, m}};
        Iterate *curr = &iterates[0], *prox = &iterates[1], *cand = &iterates[2];
        vec grad_psi_xhat(n), work_n(n), q(n);

        auto eval_psi_grad_psi = [&](Iterate &i) {
            i.psi_x = problem.eval_psi_grad_psi(i.x.data(), y.data(), S.data(), i.grad_psi.data());
        };
        auto eval_prox_grad_step = [&](Iterate &i) {
            i.h_xhat = problem.eval_prox_grad_step(i.gamma, i.x.data(), i.grad_psi.data(),
                                                   i.xhat.data(), i.p.data());
            i.pTp        = sqnorm(i.p.data(), n);
            i.grad_psiTp = dot(i.p.data(), i.grad_psi.data(), n);
        };
        auto eval_psi_xhat = [&](Iterate &i) {
            i.psi_xhat = problem.eval_psi(i.xhat.data(), y.data(), S.data(), i.yhat.data());
        };
        auto eval_grad_psi_xhat = [&](Iterate &i, vec &g) {
            problem.eval_grad_L(i.xhat.data(), i.yhat.data(), g.data());
        };
        auto qub_violated = [&](const Iterate &i) { // pantr.tpp:102-106
            real_t margin = (1 + std::abs(i.psi_x)) * params.quadratic_upperbound_tolerance_factor;
            return i.psi_xhat > i.psi_x + i.grad_psiTp + real_t(0.5) * i.L * i.pTp + margin;
        };
        auto backtrack_qub = [&](Iterate &i) { // pantr.tpp:107-115
            while (i.L < params.L_max && qub_violated(i)) {
                i.gamma /= 2;
                i.L *= 2;
                eval_prox_grad_step(i);
                eval_psi_xhat(i);
            }
        };

        curr->x = x; // pantr.tpp:195
        if (params.Lipschitz.L_0 <= 0) { // pantr.tpp:200-212
            curr->L = initial_lipschitz_estimate(problem, curr->x, y.data(), S.data(),
                                                 params.Lipschitz.eps, params.Lipschitz.delta,
                                                 params.L_min, params.L_max, curr->psi_x,
                                                 curr->grad_psi, curr->xhat, cand->grad_psi);
        } else {
            curr->L = params.Lipschitz.L_0;
            eval_psi_grad_psi(*curr);
        }
        if (!std::isfinite(curr->L)) { s.status = NotFinite; return s; }
        curr->gamma = params.Lipschitz.Lgamma_factor / curr->L;
        eval_prox_grad_step(*curr); // pantr.tpp:221-223
        eval_psi_xhat(*curr);
        backtrack_qub(*curr);

        unsigned k = 0, no_progress = 0;
        bool accept_candidate = false;
        real_t Delta = params.initial_radius; // pantr.tpp:232-235
        if (!std::isfinite(Delta) || Delta == 0) Delta = real_t(0.1) * norm2(curr->grad_psi.data(), n);
        Delta   = std::fmax(Delta, params.min_radius);
        real_t rho = NaN;
        direction.initialize(problem, y.data(), S.data());

        while (true) {
            const bool need_grad = stop_crit_requires_grad_psi_xhat(params.stop_crit); // pantr.tpp:77,246-249
            if (need_grad) {
                ++problem.evaluations.grad_L;
                problem.grad_L_live(curr->xhat.data(), curr->yhat.data(), grad_psi_xhat.data());
            }
            real_t eps_k = calc_error_stop_crit(problem, params.stop_crit, curr->p, curr->gamma, curr->x, curr->xhat,
                                                curr->yhat, curr->grad_psi, grad_psi_xhat, work_n, cand->p);
            auto stop_status = check_all_stop_conditions(opts, k, eps_k, no_progress);
            if (stop_status != Busy) { // pantr.tpp:269-298
                if (stop_status == Converged || stop_status == Interrupted ||
                    opts.always_overwrite_results) {
                    if (!err_z.empty())
                        for (index_t i = 0; i < m; ++i) err_z[i] = (real_t(1) / S[i]) * (curr->yhat[i] - y[i]);
                    x = curr->xhat;
                    y = curr->yhat;
                }
                s.iterations = k; s.eps_ = eps_k; s.status = stop_status;
                s.final_gamma = curr->gamma; s.final_psi = curr->psi_xhat;
                s.final_h = curr->h_xhat; s.final_fbe = curr->fbe();
                return s;
            }
            // compute_FBS_step: pantr.tpp:303-319
            if (!need_grad) eval_grad_psi_xhat(*curr, grad_psi_xhat); // result is dead (overwritten below), but counted
            prox->x     = curr->xhat;
            prox->psi_x = curr->psi_xhat;
            prox->grad_psi.swap(grad_psi_xhat);
            prox->gamma = curr->gamma;
            prox->L     = curr->L;
            eval_psi_grad_psi(*prox);
            eval_prox_grad_step(*prox);

            accept_candidate           = false;
            bool accelerated_iteration = k > 0 || direction.has_initial_direction();
            TraceRec tr{k, curr->psi_x, curr->fbe(), curr->gamma, eps_k, Delta, NaN, NaN, -1, 0, 0};
            if (accelerated_iteration && !params.disable_acceleration) {
                // NewtonTRDirection::apply throws std::logic_error on these radii (newton-tr.hpp:96-99); the
                // exception leaves the reference's solvers: reported as status Exception
                if (!std::isfinite(Delta) || Delta < eps) {
                    s.iterations = k; s.eps_ = eps_k; s.status = Exception;
                    s.final_gamma = curr->gamma; s.final_psi = curr->psi_xhat;
                    s.final_h = curr->h_xhat; s.final_fbe = curr->fbe();
                    return s;
                }
                // compute_trust_region_step: pantr.tpp:371-392
                real_t q_model = direction.apply(prox->gamma, prox->x.data(), prox->p.data(),
                                                 prox->grad_psi.data(), Delta, q.data());
                tr.nJ = (int)direction.last_nJ; tr.cg_iters = direction.steihaug.last_iters;
                bool finite = true;
                for (index_t i = 0; i < n; ++i) finite = finite && std::isfinite(q[i]);
                if (!finite) { ++s.direction_failures; q_model = +inf; }
                else if (q_model >= 0) ++s.direction_failures;
                tr.q_model = q_model;
                if (q_model < 0) {
                    // compute_candidate_fbe: pantr.tpp:329-344
                    for (index_t i = 0; i < n; ++i) cand->x[i] = prox->x[i] + q[i];
                    eval_psi_grad_psi(*cand);
                    cand->gamma = prox->gamma;
                    cand->L     = prox->L;
                    eval_prox_grad_step(*cand);
                    if (params.compute_ratio_using_new_stepsize) { eval_psi_xhat(*cand); backtrack_qub(*cand); }
                    // compute_candidate_ratio: pantr.tpp:347-355
                    real_t fbe      = prox->fbe();
                    real_t fbe_next = cand->fbe();
                    real_t margin   = (1 + std::abs(fbe)) * params.TR_tolerance_factor;
                    rho             = (fbe - fbe_next + margin) / (-q_model);
                    if (params.ratio_approx_fbe_quadratic_model)
                        rho = rho / (1 - params.Lipschitz.Lgamma_factor);
                    accept_candidate = rho >= params.ratio_threshold_acceptable;
                    // compute_updated_radius: pantr.tpp:358-368
                    real_t newD;
                    if (rho >= params.ratio_threshold_good)
                        newD = std::max(params.radius_factor_good * norm2(q.data(), n), Delta);
                    else if (rho >= params.ratio_threshold_acceptable)
                        newD = Delta * params.radius_factor_acceptable;
                    else
                        newD = params.radius_factor_rejected * norm2(q.data(), n);
                    Delta = std::fmax(newD, params.min_radius);
                }
            }
            tr.rho = rho; tr.accepted = accept_candidate;
            if (trace) trace->push_back(tr);
            if (accept_candidate) { // pantr.tpp:411-434
                if (!params.compute_ratio_using_new_stepsize) { eval_psi_xhat(*cand); backtrack_qub(*cand); }
                if (prox->gamma != cand->gamma && params.recompute_last_prox_step_after_direction_reset) {
                    prox->gamma = cand->gamma; prox->L = cand->L;
                    eval_prox_grad_step(*prox);
                }
                std::swap(curr, cand);
            } else { // pantr.tpp:436-458
                if (accelerated_iteration) ++s.accelerated_step_rejected;
                eval_psi_xhat(*prox);
                backtrack_qub(*prox);
                if (prox->gamma != curr->gamma && params.recompute_last_prox_step_after_direction_reset) {
                    curr->gamma = prox->gamma; curr->L = prox->L;
                    eval_prox_grad_step(*curr);
                }
                std::swap(curr, prox);
            }
            ++k;
        }
    }
};

} // namespace oracle
