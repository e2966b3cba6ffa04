// Included inside struct pantr::Pantr<Model> (pantr_device.cuh).
// PANTRSolver::operator() (pantr.tpp:25-473) and ALMSolver::operator() (alm.tpp:19-226).

    struct InnerResult { int status; double eps; unsigned iterations; };

    // panoc-helpers.tpp:41-59
    __device__ __forceinline__ bool stop_crit_requires_grad_psi_xhat() const {
        const int c = P.pantr.stop_crit;
        return c == PANTR_B200_APPROX_KKT || c == PANTR_B200_APPROX_KKT2 || c == PANTR_B200_IPOPT;
    }
    // panoc-helpers.tpp:62-137 ; gxh = grad psi(x_hat) = grad L(x_hat, y_hat) where the criterion needs it
    __device__ __noinline__ double calc_error_stop_crit(int c) {
        const int crit = P.pantr.stop_crit;
        const CVec pk = PP + c * n, gk = G + c * n;
        switch (crit) {
            case PANTR_B200_APPROX_KKT:
            case PANTR_B200_APPROX_KKT2: {
                const double ig = 1 / it[c].gamma;
                for (int i = E::first(); i < n; i += E::STEP) wn2[i] = ig * pk[i] + (gk[i] - gxh[i]);
                E::sync();
                return crit == PANTR_B200_APPROX_KKT ? xnorm_inf<E>(wn2, n) : xnorm2<E>(wn2, n);
            }
            case PANTR_B200_PROJ_GRAD_NORM: return xnorm_inf<E>(pk, n);
            case PANTR_B200_PROJ_GRAD_NORM2: return xnorm2<E>(pk, n);
            case PANTR_B200_PROJ_GRAD_UNIT_NORM:
                prox_grad_step_raw(1.0, X + c * n, G + c * n, wn1, wn2);
                E::sync();
                return xnorm_inf<E>(wn2, n);
            case PANTR_B200_PROJ_GRAD_UNIT_NORM2:
                prox_grad_step_raw(1.0, X + c * n, G + c * n, wn1, wn2);
                E::sync();
                return xnorm2<E>(wn2, n);
            case PANTR_B200_FPR_NORM: return xnorm_inf<E>(pk, n) / it[c].gamma;
            case PANTR_B200_FPR_NORM2: return xnorm2<E>(pk, n) / it[c].gamma;
            case PANTR_B200_IPOPT: {
                prox_grad_step_raw(1.0, XH + c * n, gxh, wn1, wn2);
                E::sync();
                const double err = xnorm_inf<E>(wn2, n);
                const int nn = 2 * (m + n);
                if (nn == 0) return err;
                for (int i = E::first(); i < n; i += E::STEP) wn2[i] = wn2[i] - gxh[i];
                E::sync();
                const double C_lagr_mult = xnorm1<E>(wn2, n), D_lagr_mult = xnorm1<E>(YH + c * m, m);
                const double s_max = 100, s_n = (double)nn;
                const double s_d = fmax(s_max, (C_lagr_mult + D_lagr_mult) / s_n) / s_max;
                return err / s_d;
            }
            case PANTR_B200_LBFGSBPP:
                prox_grad_step_raw(1.0, X + c * n, G + c * n, wn1, wn2);
                E::sync();
                return xnorm_inf<E>(wn2, n) / fmax(1.0, xnorm2<E>(X + c * n, n));
            default: return nan(""); // rejected by the host (std::out_of_range in the reference)
        }
    }
    __device__ __forceinline__ bool out_of_time() const { // panoc-helpers.tpp:233-236: min(params.max_time, opts.max_time)
        const unsigned long long now = globaltimer_ns();
        return (double)(now - t_inner_start) * 1e-9 > P.pantr.max_time || (double)(now - t_alm_start) * 1e-9 > P.alm.max_time;
    }

    // x (in/out) lives in this->x ; y in this->y ; Sigma in this->Sig ; err_z -> errz (may be null)
    /// State of one inner solve between PANTR iterations (everything inner_iterate needs besides the
    /// iterate slots X/XH/G/PP/YH and it[]): lets the scheduler hand out iterations instead of solves.
    struct InnerState {
        int curr, prox, cand;
        unsigned k, no_progress;
        double Delta, rho, tolerance;
        int always_overwrite, has_errz;
    };

    /// pantr.tpp:195-235: everything before the main loop.  Returns false if the solve ended here.
    __device__ __noinline__ bool inner_begin(InnerState &s, InnerResult &res, double tolerance_opt, bool always_overwrite,
                                             bool has_errz) {
        const auto &pp = P.pantr;
        s.curr = 0; s.prox = 1; s.cand = 2;
        const int curr = 0, cand = 2;
        t_inner_start = globaltimer_ns();
        res = InnerResult{PANTR_B200_BUSY, INFINITY, 0};
        s.always_overwrite = always_overwrite; s.has_errz = has_errz;
        if constexpr (Model::HAS_ROLLOUT_TAG) mdl.untag(); // the iterate slots are about to be overwritten
        xcopy<E, Model::BATCHED_MAPS>(X + curr * n, x, n); // pantr.tpp:195
        E::sync();
        // ---- initial Lipschitz estimate: pantr.tpp:200-217, panoc-helpers.tpp:288-334 ----
        if (pp.Lipschitz.L_0 <= 0) {
            eval_psi_grad_psi(curr);
            Vec wx = XH + curr * n, wg = G + cand * n, hv = wn1;
            for (int i = E::first(); i < n; i += E::STEP) {
                double g = G[curr * n + i];
                double hi = g > 0 ? fmax(g * pp.Lipschitz.epsilon, pp.Lipschitz.delta)
                                  : fmin(g * pp.Lipschitz.epsilon, -pp.Lipschitz.delta);
                hv[i] = hi;
                wx[i] = X[curr * n + i] - hi;
            }
            E::sync();
            double norm_h = xnorm2<E>(hv, n);
            ++st.ev_grad_psi;
            (void)mdl.eval_psi_grad_psi(wx, y, Sig, wg);
            for (int i = E::first(); i < n; i += E::STEP) hv[i] = wg[i] - G[curr * n + i];
            E::sync();
            double L = xnorm2<E>(hv, n) / norm_h;
            // std::clamp(L, L_min, L_max)
            L = L < pp.L_min ? pp.L_min : (pp.L_max < L ? pp.L_max : L);
            it[curr].L = L;
        } else {
            it[curr].L = pp.Lipschitz.L_0;
            eval_psi_grad_psi(curr);
        }
        E::sync();
        if (!isfinite(it[curr].L)) { res.status = PANTR_B200_NOT_FINITE; return false; }
        it[curr].gamma = pp.Lipschitz.Lgamma_factor / it[curr].L;
        E::sync();
        eval_prox_grad_step(curr); // pantr.tpp:221-223
        eval_psi_xh(curr);
        backtrack_qub(curr);

        s.k = 0; s.no_progress = 0;
        double Delta = pp.initial_radius; // pantr.tpp:232-235
        if (!isfinite(Delta) || Delta == 0) Delta = 0.1 * xnorm2<E>(G + curr * n, n);
        s.Delta     = fmax(Delta, pp.min_radius);
        s.rho       = nan("");
        s.tolerance = tolerance_opt > 0 ? tolerance_opt : 1e-8;
        return true;
    }


    /// One pass of the main loop (pantr.tpp:237-462).  Returns true when the solve has ended (res is set).
    __device__ __noinline__ bool inner_iterate(InnerState &s, InnerResult &res, Vec errz) {
        const auto &pp = P.pantr;
        int &curr = s.curr, &prox = s.prox, &cand = s.cand;
        unsigned &k = s.k, &no_progress = s.no_progress;
        double &Delta = s.Delta, &rho = s.rho;
        const double tolerance = s.tolerance;
        const bool always_overwrite = s.always_overwrite, has_errz = s.has_errz;
        E::sync();
        const bool need_grad = stop_crit_requires_grad_psi_xhat();
        if (need_grad) { // pantr.tpp:246-249: grad psi(x_hat) = grad L(x_hat, y_hat)
            ++st.ev_grad_L;
            (void)mdl.eval_grad_L(XH + curr * n, m > 0 ? CVec(YH + curr * m) : CVec{}, gxh);
            E::sync();
        }
        double eps_k = calc_error_stop_crit(curr);
        // check_all_stop_conditions: panoc-helpers.tpp:215-249 (device clock; no stop signal on device)
        int stop = eps_k <= tolerance                    ? PANTR_B200_CONVERGED
                   : out_of_time()                       ? PANTR_B200_MAX_TIME
                   : k == pp.max_iter                    ? PANTR_B200_MAX_ITER
                   : !isfinite(eps_k)                    ? PANTR_B200_NOT_FINITE
                   : no_progress > pp.max_no_progress    ? PANTR_B200_NO_PROGRESS
                                                         : PANTR_B200_BUSY;
        if (stop != PANTR_B200_BUSY) { // pantr.tpp:269-298
            if (stop == PANTR_B200_CONVERGED || always_overwrite) {
                const CVec yh = YH + curr * m;
                for (int i = E::first(); i < m; i += E::STEP) {
                    double yhi = yh[i];
                    if (has_errz) errz[i] = (1.0 / Sig[i]) * (yhi - y[i]);
                    y[i] = yhi;
                }
                xcopy<E, Model::BATCHED_MAPS>(x, XH + curr * n, n);
            }
            st.final_gamma = it[curr].gamma; st.final_psi = it[curr].psi_xh;
            st.final_h = it[curr].h; st.final_fbe = fbe(curr);
            res.status = stop; res.eps = eps_k; res.iterations = k;
            E::sync();
            return true;
        }
        // ---- compute_FBS_step: pantr.tpp:303-319 ----
        // Unless the stop criterion already needed it, the reference evaluates grad_L(x_hat, y_hat) here and
        // immediately overwrites the result with psi_grad_psi(prox.x): numerically dead, so only counted.
        if (!need_grad) ++st.ev_grad_L;
        xcopy<E, Model::BATCHED_MAPS>(X + prox * n, XH + curr * n, n);
        if constexpr (Model::HAS_ROLLOUT_TAG) mdl.retag(XH + curr * n, X + prox * n); // psi(x_hat) rolled this very point out
        double cg_ = it[curr].gamma, cL_ = it[curr].L, cpsi = it[curr].psi_xh;
        E::sync();
        it[prox].gamma = cg_; it[prox].L = cL_; it[prox].psi_x = cpsi;
        E::sync();
        eval_psi_grad_psi(prox);
        eval_prox_grad_step(prox);
        E::sync();

        bool accept_candidate = false;
        const bool accelerated_iteration = true; // k > 0 || has_initial_direction() (newton-tr.hpp:79)
        double q_model = nan(""); int nJ = -1; unsigned cgi = 0;
        const double tr_psi = it[curr].psi_x, tr_fbe = fbe(curr), tr_gamma = it[curr].gamma, tr_Delta = Delta;
        if (accelerated_iteration && !pp.disable_acceleration) {
            // NewtonTRDirection::apply throws std::logic_error on these radii (newton-tr.hpp:96-99); the exception
            // leaves the reference's solver, here the instance ends with status Exception
            if (!isfinite(Delta) || Delta < 2.220446049250313e-16) {
                res.status = PANTR_B200_EXCEPTION; res.eps = eps_k; res.iterations = k;
                st.final_gamma = it[curr].gamma; st.final_psi = it[curr].psi_xh; st.final_h = it[curr].h; st.final_fbe = fbe(curr);
                E::sync();
                return true;
            }
            // compute_trust_region_step: pantr.tpp:371-392
            q_model = direction_apply(it[prox].gamma, X + prox * n, PP + prox * n, G + prox * n, Delta, q, nJ);
            cgi = cg_last; st.cg_iterations += cg_last;
            if (!xall_finite<E>(q, n)) { ++st.direction_failures; q_model = INFINITY; }
            else if (q_model >= 0) ++st.direction_failures;
            if (q_model < 0) {
                // compute_candidate_fbe: pantr.tpp:329-344
                emap(n, [&](int i) { return X[prox * n + i] + q[i]; }, [&](int i, double v) { X[cand * n + i] = v; });
                E::sync();
                eval_psi_grad_psi(cand);
                double pg = it[prox].gamma, pL = it[prox].L;
                E::sync();
                it[cand].gamma = pg; it[cand].L = pL;
                E::sync();
                eval_prox_grad_step(cand);
                if (pp.compute_ratio_using_new_stepsize) { eval_psi_xh(cand); backtrack_qub(cand); }
                E::sync();
                // compute_candidate_ratio: pantr.tpp:347-355
                double phi = fbe(prox), phi_next = fbe(cand);
                double margin = (1 + fabs(phi)) * pp.TR_tolerance_factor;
                rho = (phi - phi_next + margin) / (-q_model);
                if (pp.ratio_approx_fbe_quadratic_model) rho = rho / (1 - pp.Lipschitz.Lgamma_factor);
                accept_candidate = rho >= pp.ratio_threshold_acceptable;
                // compute_updated_radius: pantr.tpp:358-368,402-403
                double newD;
                if (rho >= pp.ratio_threshold_good) {
                    double a = pp.radius_factor_good * xnorm2<E>(q, n);
                    newD = a < Delta ? Delta : a; // std::max(a, old)
                } else if (rho >= pp.ratio_threshold_acceptable)
                    newD = Delta * pp.radius_factor_acceptable;
                else
                    newD = pp.radius_factor_rejected * xnorm2<E>(q, n);
                Delta = fmax(newD, pp.min_radius);
            }
        }
        if (tracing && E::leader()) {
            int row = *tr.n_rows;
            if (row < tr.max_rows) {
                double *o = tr.rows + 11 * row;
                o[0] = k; o[1] = tr_psi; o[2] = tr_fbe; o[3] = tr_gamma; o[4] = eps_k; o[5] = tr_Delta;
                o[6] = rho; o[7] = q_model; o[8] = nJ; o[9] = cgi; o[10] = accept_candidate;
            }
            *tr.n_rows = row + 1;
        }
        if (accept_candidate) { // pantr.tpp:411-434
            if (!pp.compute_ratio_using_new_stepsize) { eval_psi_xh(cand); backtrack_qub(cand); }
            E::sync();
            if (it[prox].gamma != it[cand].gamma && pp.recompute_last_prox_step_after_direction_reset) {
                double g2 = it[cand].gamma, L2 = it[cand].L;
                E::sync();
                it[prox].gamma = g2; it[prox].L = L2;
                E::sync();
                eval_prox_grad_step(prox);
            }
            int t = curr; curr = cand; cand = t;
        } else { // pantr.tpp:436-458
            if (accelerated_iteration) ++st.accelerated_step_rejected;
            eval_psi_xh(prox);
            backtrack_qub(prox);
            E::sync();
            if (it[prox].gamma != it[curr].gamma && pp.recompute_last_prox_step_after_direction_reset) {
                double g2 = it[prox].gamma, L2 = it[prox].L;
                E::sync();
                it[curr].gamma = g2; it[curr].L = L2;
                E::sync();
                eval_prox_grad_step(curr);
            }
            int t = curr; curr = prox; prox = t;
        }
        ++k;
        return false;
    }

    /// The loop state lives in the solver object (ist_, ires_), i.e. in shared memory for the warp kernel: as locals
    /// passed by reference to the __noinline__ steps they were per-lane local-memory copies.
    __device__ InnerResult inner_solve(double tolerance_opt, bool always_overwrite, Vec errz, bool has_errz) {
        if (!inner_begin(ist_, ires_, tolerance_opt, always_overwrite, has_errz)) return ires_;
        while (!inner_iterate(ist_, ires_, errz)) {}
        return ires_;
    }
#include "pantr_alm.cuh"
