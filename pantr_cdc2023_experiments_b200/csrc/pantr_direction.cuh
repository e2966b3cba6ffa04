// Included inside struct pantr::Pantr<Model> (pantr_device.cuh).
// NewtonTRDirection::apply (newton-tr.hpp:91-155) and SteihaugCG::solve (steihaugcg.hpp:39-140).

    // eval_inactive_indices_res_lna (box-constr-problem.hpp:151-185) + IndexSet::compute_complement
    // (util/index-set.hpp:46-50): ascending J in JK[0..nJ), ascending K in JK[nJ..n), by
    // ballot / prefix-popcount compaction.
    __device__ __noinline__ int inactive_indices(double gamma, CVec xx, CVec g) {
        ++st.ev_inactive_indices_res_lna;
        auto inside = [&](int i) {
            double x_fw = xx[i] - gamma * g[i];
            return Model::C_lb(i) < x_fw && x_fw < Model::C_ub(i);
        };
        int nJ = E::compact(n, JK, 0, inside);
        (void)E::compact(n, JK, nJ, [&](int i) { return !inside(i); });
        E::sync();
        return nJ;
    }

    // hess_vec_mult of newton-tr.hpp:134-149: Bp = (H work)(J), work(J) = p, zero elsewhere
    __device__ __noinline__ void hess_vec_mult(CVec xx, CVec p, Vec Bp, int nJ) {
        const IVec J = JK;
        for (int i = E::first(); i < n; i += E::STEP) work[i] = 0.0;
        E::sync();
        struct IdxVal { int i; double v; };
        emap(nJ, [&](int j) { return IdxVal{J[j], p[j]}; }, [&](int, IdxVal e) { work[e.i] = e.v; });
        ++st.ev_hess_psi_prod;
        mdl.hess_psi_prod(xx, y, Sig, work, work2);
        emap(nJ, [&](int j) { return work2[J[j]]; }, [&](int j, double v) { Bp[j] = v; });
        E::sync();
    }
    // eval lambda of steihaugcg.hpp:62-65
    __device__ __noinline__ double cg_eval(CVec xx, CVec p, CVec g, int nJ) {
        hess_vec_mult(xx, p, we, nJ);
        return xdot<E>(p, g, nJ) + 0.5 * xdot<E>(p, we, nJ);
    }
    // steihaugcg.hpp:122-140
    __device__ __noinline__ void boundaries(CVec zz, CVec dd_, int nJ, double radius, double &lo, double &hi) {
        double a = xsqnorm<E>(dd_, nJ);
        double b = 2 * xdot<E>(zz, dd_, nJ);
        double c = xsqnorm<E>(zz, nJ) - radius * radius;
        double sqrt_discriminant = sqrt(b * b - 4 * a * c);
        double aux = b + copysign(sqrt_discriminant, b);
        double ta  = -aux / (2 * a);
        double tb  = -2 * c / aux;
        lo = fmin(ta, tb); hi = fmax(ta, tb);
    }
    // steihaugcg.hpp:39-117 ; g = rJ, step -> s (both compressed, length nJ)
    __device__ __noinline__ double steihaug(CVec xx, CVec g, int nJ, double radius, Vec s) {
        emap(nJ, [&](int j) { return g[j]; }, [&](int j, double gj) { z[j] = 0.0; r[j] = gj; d[j] = -gj; });
        E::sync();
        double r_sq = xsqnorm<E>(r, nJ), grad_mag = xnorm2<E>(g, nJ);
        double tolerance = fmin(P.cg.tol_max,
                                P.cg.tol_scale * grad_mag * fmin(P.cg.tol_scale_root, sqrt(grad_mag)));
        long long i = 0;
        const long long max_iter = (long long)round((double)nJ * P.cg.max_iter_factor);
        cg_last = 0;
        while (true) {
            ++cg_last;
            hess_vec_mult(xx, d, Bd, nJ);
            double dBd = xdot<E>(d, Bd, nJ);
            if (dBd <= 0) {
                double ta, tb;
                boundaries(z, d, nJ, radius, ta, tb);
                Vec pa = r, pb = d; // storage reuse as in the reference
                emap(nJ, [&](int j) { return make_double2(z[j], d[j]); },
                       [&](int j, double2 e) { pa[j] = e.x + ta * e.y; pb[j] = e.x + tb * e.y; }); // pa = r, pb = d: own index only
                E::sync();
                double q_a = cg_eval(xx, pa, g, nJ), q_b = cg_eval(xx, pb, g, nJ);
                double q_min = fmin(q_a, q_b);
                if (q_a == q_min) { xcopy<E, Model::BATCHED_MAPS>(s, pa, nJ); E::sync(); return q_a; }
                else              { xcopy<E, Model::BATCHED_MAPS>(s, pb, nJ); E::sync(); return q_b; }
            }
            double alpha = r_sq / dBd;
            emap(nJ, [&](int j) { return z[j] + alpha * d[j]; }, [&](int j, double v) { s[j] = v; });
            E::sync();
            if (xnorm2<E>(s, nJ) >= radius) {
                double ta, tb;
                boundaries(z, d, nJ, radius, ta, tb);
                emap(nJ, [&](int j) { return z[j] + tb * d[j]; }, [&](int j, double v) { s[j] = v; });
                E::sync();
                return cg_eval(xx, s, g, nJ);
            }
            emap(nJ, [&](int j) { return r[j] + alpha * Bd[j]; }, [&](int j, double v) { r[j] = v; });
            E::sync();
            double r_next_sq = xsqnorm<E>(r, nJ);
            if (sqrt(r_next_sq) < tolerance || i > max_iter) return cg_eval(xx, s, g, nJ);
            double beta_next = r_next_sq / r_sq;
            r_sq             = r_next_sq;
            emap(nJ, [&](int j) { return make_double2(beta_next * d[j] - r[j], s[j]); },
                   [&](int j, double2 e) { d[j] = e.x; z[j] = e.y; });
            E::sync();
            ++i;
        }
    }

    // newton-tr.hpp:91-155 (exact Hessian branch). xx, pk, g: prox iterate. Returns q_model.
    __device__ __noinline__ double direction_apply(double gamma, CVec xx, CVec pk, CVec g, double radius, Vec qk, int &nJ_out) {
        const int nJ = inactive_indices(gamma, xx, g), nK = n - nJ;
        nJ_out = nJ;
        const IVec J = JK, K = JK + nJ;
        const double mig = -1.0 / gamma;
        struct IdxVal { int i; double v; };
        emap(nJ, [&](int j) { const int i = J[j]; return IdxVal{i, mig * pk[i]}; }, [&](int j, IdxVal e) { rJ[j] = e.v; qk[e.i] = 0.0; });
        emap(nK, [&](int k) { const int i = K[k]; return IdxVal{i, pk[i]}; }, [&](int, IdxVal e) { qk[e.i] = e.v; });
        double norm_qK_sq = E::sum(nK, [&](int k) { double pi = pk[K[k]]; return pi * pi; });
        E::sync();
        const double hvf = P.dir.hessian_vec_factor;
        if (hvf != 0) {
            ++st.ev_hess_psi_prod;
            mdl.hess_psi_prod(xx, y, Sig, qk, work);
            emap(nJ, [&](int j) { return rJ[j] + work[J[j]] * hvf; }, [&](int j, double v) { rJ[j] = v; });
            E::sync();
        }
        double qJ_model = steihaug(xx, rJ, nJ, radius, qJ);
        emap(nJ, [&](int j) { return IdxVal{J[j], qJ[j]}; }, [&](int, IdxVal e) { qk[e.i] = e.v; });
        E::sync();
        return qJ_model - norm_qK_sq / (2 * gamma);
    }
