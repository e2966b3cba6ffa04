// Included inside struct pantr::Pantr<Model> (via pantr_inner.cuh).
// ALMSolver::operator() (alpaqa/implementation/outer/alm.tpp:19-226) and
// ALMHelpers::update_penalty_weights (alpaqa/implementation/outer/internal/alm-helpers.tpp:16-45).

    __device__ void accumulate(const InnerResult &ps) { // inner/pantr.hpp:242-258
        st.inner_iterations += ps.iterations;
        st.inner_convergence_failures += ps.status != PANTR_B200_CONVERGED;
    }

    // BoxConstrProblem::eval_proj_multipliers (box-constr-problem.hpp:119-143), penalty_alm_split = 0
    __device__ void proj_multipliers(double M) {
        ++st.ev_proj_multipliers;
        for (int i = E::first(); i < m; i += E::STEP) {
            double y_lb = Model::D_lb(i, mdl.N) == -INFINITY ? 0.0 : -M;
            double y_ub = Model::D_ub(i, mdl.N) == +INFINITY ? 0.0 : +M;
            double v = y[i];
            v = v < y_lb ? y_lb : v; // std::max(y, y_lb)
            v = y_ub < v ? y_ub : v; // std::min(., y_ub)
            y[i] = v;
        }
        E::sync();
    }

    // alm-helpers.tpp:16-45 (monotone = true)
    __device__ void update_penalty_weights(double Delta, bool first_iter, CVec e, CVec old_e, double norm_e,
                                           double old_norm_e, CVec S_old, Vec S) {
        const auto &a = P.alm;
        const double theta = a.rel_penalty_increase_threshold;
        if (norm_e <= a.dual_tolerance) { xcopy<E, Model::BATCHED_MAPS>(S, S_old, m); E::sync(); return; }
        if (a.single_penalty_factor) {
            if (first_iter || norm_e > theta * old_norm_e) {
                double new_S = fmin(a.max_penalty, Delta * S_old[0]);
                E::sync();
                for (int i = E::first(); i < m; i += E::STEP) S[i] = new_S;
            } else xcopy<E, Model::BATCHED_MAPS>(S, S_old, m);
        } else {
            for (int i = E::first(); i < m; i += E::STEP) {
                if (first_iter || fabs(e[i]) > theta * fabs(old_e[i]))
                    S[i] = fmin(a.max_penalty, fmax(Delta * fabs(e[i]) / norm_e, 1.0) * S_old[i]);
                else
                    S[i] = S_old[i];
            }
        }
        E::sync();
    }

    // alm.tpp:19-226.  x, y are this->x / this->y (in/out).
    __device__ void alm_solve() {
        const auto &a = P.alm;
        st.status = PANTR_B200_BUSY;
        t_alm_start = globaltimer_ns();
        if (Model::NC == 0) { // alm.tpp:29-52 (m == 0  <=>  the model has no general constraints)
            InnerResult ps = inner_solve(a.tolerance, true, Vec{}, false);
            accumulate(ps);
            st.eps = ps.eps; st.delta = 0; st.norm_penalty = 0; st.outer_iterations = 1;
            st.status = ps.status;
            return;
        }
        Vec error_1 = e1, error_2 = e2, S = Sig, S_old = Sig_old;
        double norm_e_1 = nan(""), norm_e_2 = nan("");
        // alm.tpp:54-58: Sigma, Sigma_old, error_1, error_2 start as NaN (a backtrack before the second successful
        // iteration compares against them: NaN compares false and keeps Sigma_old)
        for (int i = E::first(); i < m; i += E::STEP) e1[i] = e2[i] = Sig[i] = Sig_old[i] = nan("");
        E::sync();
        if (a.initial_penalty > 0) {
            for (int i = E::first(); i < m; i += E::STEP) S[i] = a.initial_penalty;
        } else { // alm-helpers.tpp:47-59
            ++st.ev_f; ++st.ev_g;
            double f0 = mdl.eval_f(x);
            mdl.eval_g(x, e1);
            double g0sq = xsqnorm<E>(e1, m);
            double sigma = a.initial_penalty_factor * fmax(1.0, fabs(f0)) / fmax(1.0, 0.5 * g0sq);
            sigma = sigma < a.min_penalty ? a.min_penalty : (a.max_penalty < sigma ? a.max_penalty : sigma);
            for (int i = E::first(); i < m; i += E::STEP) S[i] = sigma;
        }
        E::sync();
        double eps_ = a.initial_tolerance, eps_old = nan("");
        double Delta = a.penalty_update_factor, rho = a.tolerance_update_factor;
        unsigned num_successful_iters = 0, initial_penalty_reduced = 0, penalty_reduced = 0;
        for (unsigned i = 0; i < a.max_iter; ++i) {
            proj_multipliers(a.max_multiplier);
            bool out_of_penalty_factor_updates =
                (num_successful_iters == 0 ? initial_penalty_reduced == a.max_num_initial_retries
                                           : penalty_reduced == a.max_num_retries) ||
                (initial_penalty_reduced + penalty_reduced == a.max_total_num_retries);
            bool out_of_iter       = i + 1 == a.max_iter;
            bool overwrite_results = out_of_iter || out_of_penalty_factor_updates;
            Sig = S; // the inner solver and the model read this->Sig
            InnerResult ps = inner_solve(eps_, overwrite_results, error_2, true);
            bool inner_converged = ps.status == PANTR_B200_CONVERGED;
            accumulate(ps);
            if (ps.status == PANTR_B200_EXCEPTION) { // an exception thrown by the direction leaves ALMSolver::operator() as well
                st.eps = ps.eps; st.outer_iterations = i + 1; st.status = PANTR_B200_EXCEPTION;
                return;
            }
            const bool out_of_time = (double)(globaltimer_ns() - t_alm_start) * 1e-9 > a.max_time; // alm.tpp:123-124
            bool backtrack = !inner_converged && !overwrite_results && !out_of_time;
            if (backtrack) { // alm.tpp:155-187
                if (num_successful_iters > 0) {
                    Delta = fmax(a.min_penalty_update_factor, Delta * a.penalty_update_factor_lower);
                    update_penalty_weights(Delta, false, error_1, error_2, norm_e_1, norm_e_2, S_old, S);
                    rho  = fmin(a.rho_max, rho * a.rho_increase);
                    eps_ = fmax(rho * eps_old, a.tolerance);
                    ++penalty_reduced;
                } else {
                    for (int j = E::first(); j < m; j += E::STEP) S[j] = S[j] * a.initial_penalty_lower;
                    E::sync();
                    eps_ *= a.initial_tolerance_increase;
                    ++initial_penalty_reduced;
                }
            } else { // alm.tpp:190-222
                { Vec t = error_2; error_2 = error_1; error_1 = t; }
                norm_e_2 = norm_e_1;
                norm_e_1 = xnorm_inf<E>(error_1, m);
                bool alm_converged = ps.eps <= a.tolerance && inner_converged && norm_e_1 <= a.dual_tolerance;
                if (alm_converged || out_of_iter || out_of_time) { // alm.tpp:200-212
                    st.eps = ps.eps; st.delta = norm_e_1; st.norm_penalty = xnorm2<E>(S, m);
                    st.outer_iterations = i + 1;
                    st.status = alm_converged ? PANTR_B200_CONVERGED : out_of_time ? PANTR_B200_MAX_TIME : PANTR_B200_MAX_ITER;
                    return;
                }
                { Vec t = S_old; S_old = S; S = t; }
                update_penalty_weights(Delta, num_successful_iters == 0, error_1, error_2, norm_e_1, norm_e_2, S_old, S);
                eps_old = eps_;
                eps_    = fmax(rho * eps_, a.tolerance);
                ++num_successful_iters;
            }
        }
        st.status = PANTR_B200_EXCEPTION;
    }
