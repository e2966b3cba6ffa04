// Included inside struct pantr::Pantr<Model> (pantr_device.cuh).
// NewtonTRDirection::apply (newton-tr.hpp:91-155) and SteihaugCG::solve (steihaugcg.hpp:39-140).

    // eval_inactive_indices_res_lna (box-constr-problem.hpp:151-185) + IndexSet::compute_complement
    // (util/index-set.hpp:46-50): ascending J in JK[0..nJ), ascending K in JK[nJ..n), by
    // ballot / prefix-popcount compaction.
    __device__ __noinline__ int inactive_indices(double gamma, const double *xx, const double *g) {
        ++st.ev_inactive_indices_res_lna;
        const int lane = lane_id();
        const unsigned lt = (1u << lane) - 1u;
        int nJ = 0;
        for (int i0 = 0; i0 < n; i0 += 32) {
            int i = i0 + lane;
            bool in = false;
            if (i < n) {
                double x_fw = xx[i] - gamma * g[i];
                in = Model::C_lb(i) < x_fw && x_fw < Model::C_ub(i);
            }
            unsigned b = __ballot_sync(FULL, in);
            if (in) JK[nJ + __popc(b & lt)] = i;
            nJ += __popc(b);
        }
        int nK = 0;
        for (int i0 = 0; i0 < n; i0 += 32) {
            int i = i0 + lane;
            bool out = false;
            if (i < n) {
                double x_fw = xx[i] - gamma * g[i];
                out = !(Model::C_lb(i) < x_fw && x_fw < Model::C_ub(i));
            }
            unsigned b = __ballot_sync(FULL, out);
            if (out) JK[nJ + nK + __popc(b & lt)] = i;
            nK += __popc(b);
        }
        __syncwarp();
        return nJ;
    }

    // hess_vec_mult of newton-tr.hpp:134-149: Bp = (H work)(J), work(J) = p, zero elsewhere
    __device__ __noinline__ void hess_vec_mult(const double *xx, const double *p, double *Bp, int nJ) {
        const int *J = JK;
        for (int i = lane_id(); i < n; i += 32) work[i] = 0.0;
        __syncwarp();
        for (int j = lane_id(); j < nJ; j += 32) work[J[j]] = p[j];
        ++st.ev_hess_psi_prod;
        mdl.hess_psi_prod(xx, y, Sig, work, work2);
        for (int j = lane_id(); j < nJ; j += 32) Bp[j] = work2[J[j]];
        __syncwarp();
    }
    // eval lambda of steihaugcg.hpp:62-65
    __device__ __noinline__ double cg_eval(const double *xx, const double *p, const double *g, int nJ) {
        hess_vec_mult(xx, p, we, nJ);
        return vdot(p, g, nJ) + 0.5 * vdot(p, we, nJ);
    }
    // steihaugcg.hpp:122-140
    __device__ __noinline__ void boundaries(const double *zz, const double *dd_, int nJ, double radius, double &lo,
                               double &hi) {
        double a = vsqnorm(dd_, nJ);
        double b = 2 * vdot(zz, dd_, nJ);
        double c = vsqnorm(zz, nJ) - radius * radius;
        double sqrt_discriminant = sqrt(b * b - 4 * a * c);
        double aux = b + copysign(sqrt_discriminant, b);
        double ta  = -aux / (2 * a);
        double tb  = -2 * c / aux;
        lo = fmin(ta, tb); hi = fmax(ta, tb);
    }
    // steihaugcg.hpp:39-117 ; g = rJ, step -> s (both compressed, length nJ)
    __device__ __noinline__ double steihaug(const double *xx, const double *g, int nJ, double radius, double *s) {
        const int lane = lane_id();
        for (int j = lane; j < nJ; j += 32) { double gj = g[j]; z[j] = 0.0; r[j] = gj; d[j] = -gj; }
        __syncwarp();
        double r_sq = vsqnorm(r, nJ), grad_mag = vnorm2(g, nJ);
        double tolerance = fmin(P.cg.tol_max,
                                P.cg.tol_scale * grad_mag * fmin(P.cg.tol_scale_root, sqrt(grad_mag)));
        long long i = 0;
        const long long max_iter = (long long)round((double)nJ * P.cg.max_iter_factor);
        cg_last = 0;
        while (true) {
            ++cg_last;
            hess_vec_mult(xx, d, Bd, nJ);
            double dBd = vdot(d, Bd, nJ);
            if (dBd <= 0) {
                double ta, tb;
                boundaries(z, d, nJ, radius, ta, tb);
                double *pa = r, *pb = d; // storage reuse as in the reference
                for (int j = lane; j < nJ; j += 32) {
                    double zj = z[j], dj = d[j];
                    pa[j] = zj + ta * dj;
                    pb[j] = zj + tb * dj;
                }
                __syncwarp();
                double q_a = cg_eval(xx, pa, g, nJ), q_b = cg_eval(xx, pb, g, nJ);
                double q_min = fmin(q_a, q_b);
                if (q_a == q_min) { vcopy(s, pa, nJ); __syncwarp(); return q_a; }
                else              { vcopy(s, pb, nJ); __syncwarp(); return q_b; }
            }
            double alpha = r_sq / dBd;
            for (int j = lane; j < nJ; j += 32) s[j] = z[j] + alpha * d[j];
            __syncwarp();
            if (vnorm2(s, nJ) >= radius) {
                double ta, tb;
                boundaries(z, d, nJ, radius, ta, tb);
                for (int j = lane; j < nJ; j += 32) s[j] = z[j] + tb * d[j];
                __syncwarp();
                return cg_eval(xx, s, g, nJ);
            }
            for (int j = lane; j < nJ; j += 32) r[j] = r[j] + alpha * Bd[j];
            __syncwarp();
            double r_next_sq = vsqnorm(r, nJ);
            if (sqrt(r_next_sq) < tolerance || i > max_iter) return cg_eval(xx, s, g, nJ);
            double beta_next = r_next_sq / r_sq;
            r_sq             = r_next_sq;
            for (int j = lane; j < nJ; j += 32) { d[j] = beta_next * d[j] - r[j]; z[j] = s[j]; }
            __syncwarp();
            ++i;
        }
    }

    // newton-tr.hpp:91-155 (exact Hessian branch). xx, pk, g: prox iterate. Returns q_model.
    __device__ __noinline__ double direction_apply(double gamma, const double *xx, const double *pk, const double *g,
                                      double radius, double *qk, int &nJ_out) {
        const int lane = lane_id();
        const int nJ = inactive_indices(gamma, xx, g), nK = n - nJ;
        nJ_out = nJ;
        const int *J = JK, *K = JK + nJ;
        const double mig = -1.0 / gamma;
        for (int j = lane; j < nJ; j += 32) { int i = J[j]; rJ[j] = mig * pk[i]; qk[i] = 0.0; }
        double part = 0;
        for (int k = lane; k < nK; k += 32) { int i = K[k]; double pi = pk[i]; qk[i] = pi; part = part + pi * pi; }
        double norm_qK_sq = warp_sum(part);
        __syncwarp();
        const double hvf = P.dir.hessian_vec_factor;
        if (hvf != 0) {
            ++st.ev_hess_psi_prod;
            mdl.hess_psi_prod(xx, y, Sig, qk, work);
            for (int j = lane; j < nJ; j += 32) rJ[j] = rJ[j] + work[J[j]] * hvf;
            __syncwarp();
        }
        double qJ_model = steihaug(xx, rJ, nJ, radius, qJ);
        for (int j = lane; j < nJ; j += 32) qk[J[j]] = qJ[j];
        __syncwarp();
        return qJ_model - norm_qK_sq / (2 * gamma);
    }
