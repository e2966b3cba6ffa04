// Included inside struct pantr::HangingChain (hanging_chain.cuh).
// Reverse (adjoint) sweep for grad f, and forward-over-adjoint sweep for hess f * v
// (the exact Hessian-vector product that casadi_generator/__init__.py:181-189 defines as
// hess_psi_prod; with m = 0, psi = f).  Both assume traj holds the rollout of `u`.

    __device__ void adjoint(const double *u, double *grad) {
        double X, V, lX, lV;
        load_state(traj, N, X, V);
        lX = isEnd ? two_alpha * (X - ref) : 0.0;
        lV = hasV ? two_beta * V : 0.0;
        for (int k = N - 1; k >= 0; --k) {
            load_state(traj, k, X, V);
            double ud = load_u(u, k);
            Link ld[4];
            stages(X, V, ud, ld);
            double wX, wV, aX, aV, sX, sV, Fb, dt;
            wX = h6 * lX; wV = h6 * lV;
            vjp(ld[3], wX, wV, aX, aV, Fb, dt);
            sX = lX + aX; sV = lV + aV;
            wX = h3 * lX + h * aX; wV = h3 * lV + h * aV;
            vjp(ld[2], wX, wV, aX, aV, Fb, dt);
            sX = sX + aX; sV = sV + aV;
            wX = h3 * lX + hh * aX; wV = h3 * lV + hh * aV;
            vjp(ld[1], wX, wV, aX, aV, Fb, dt);
            sX = sX + aX; sV = sV + aV;
            wX = h6 * lX + hh * aX; wV = h6 * lV + hh * aV;
            vjp(ld[0], wX, wV, aX, aV, Fb, dt);
            sX = sX + aX; sV = sV + aV;
            lX = isEnd ? sX + two_alpha * (X - ref) : sX;
            if (hasV) lV = sV + two_beta * V;
            else {
                if (isEnd) grad[NU * k + (l - NVL)] = sV + two_gamma * ud;
                lV = 0.0;
            }
        }
    }

    __device__ void hvp(const double *u, const double *v, double *Hv) {
        double X, V, dX = 0.0, dV = 0.0;
        // ---- forward tangent sweep ----
        store_state(dtraj, 0, dX, dV);
        for (int k = 0; k < N; ++k) {
            load_state(traj, k, X, V);
            double ud = load_u(u, k), vd = load_u(v, k);
            Link ld[4];
            stages(X, V, ud, ld);
            double k1X, k1V, k2X, k2V, k3X, k3V, k4X, k4V, ddist, dd;
            jvp(ld[0], dX, hasV ? dV : vd, k1X, k1V, ddist, dd);
            double tX = dX + hh * k1X, tV = dV + hh * k1V;
            jvp(ld[1], tX, hasV ? tV : vd, k2X, k2V, ddist, dd);
            tX = dX + hh * k2X; tV = dV + hh * k2V;
            jvp(ld[2], tX, hasV ? tV : vd, k3X, k3V, ddist, dd);
            tX = dX + h * k3X; tV = dV + h * k3V;
            jvp(ld[3], tX, hasV ? tV : vd, k4X, k4V, ddist, dd);
            dX = dX + h6 * (((k1X + 2.0 * k2X) + 2.0 * k3X) + k4X);
            dV = dV + h6 * (((k1V + 2.0 * k2V) + 2.0 * k3V) + k4V);
            store_state(dtraj, k + 1, dX, dV);
        }
        __syncwarp();
        // ---- reverse sweep for lambda and mu = d lambda ----
        double lX, lV, mX, mV;
        load_state(traj, N, X, V);
        lX = isEnd ? two_alpha * (X - ref) : 0.0;
        mX = isEnd ? two_alpha * dX : 0.0;
        lV = hasV ? two_beta * V : 0.0;
        mV = hasV ? two_beta * dV : 0.0;
        for (int k = N - 1; k >= 0; --k) {
            load_state(traj, k, X, V);
            load_state(dtraj, k, dX, dV);
            double ud = load_u(u, k), vd = load_u(v, k);
            Link ld[4];
            stages(X, V, ud, ld);
            double ddist[4], dd[4];
            {
                double kX, kV;
                jvp(ld[0], dX, hasV ? dV : vd, kX, kV, ddist[0], dd[0]);
                double tX = dX + hh * kX, tV = dV + hh * kV;
                jvp(ld[1], tX, hasV ? tV : vd, kX, kV, ddist[1], dd[1]);
                tX = dX + hh * kX; tV = dV + hh * kV;
                jvp(ld[2], tX, hasV ? tV : vd, kX, kV, ddist[2], dd[2]);
                tX = dX + h * kX; tV = dV + h * kV;
                jvp(ld[3], tX, hasV ? tV : vd, kX, kV, ddist[3], dd[3]);
            }
            double wX, wV, aX, aV, sX, sV, dwX, dwV, bX, bV, tX, tV, Fb, dt;
            // stage 4
            wX = h6 * lX; wV = h6 * lV; dwX = h6 * mX; dwV = h6 * mV;
            vjp(ld[3], wX, wV, aX, aV, Fb, dt);
            hvjp(ld[3], Fb, dt, ddist[3], dd[3], dwX, dwV, bX, bV);
            sX = lX + aX; sV = lV + aV; tX = mX + bX; tV = mV + bV;
            // stage 3
            wX = h3 * lX + h * aX; wV = h3 * lV + h * aV; dwX = h3 * mX + h * bX; dwV = h3 * mV + h * bV;
            vjp(ld[2], wX, wV, aX, aV, Fb, dt);
            hvjp(ld[2], Fb, dt, ddist[2], dd[2], dwX, dwV, bX, bV);
            sX = sX + aX; sV = sV + aV; tX = tX + bX; tV = tV + bV;
            // stage 2
            wX = h3 * lX + hh * aX; wV = h3 * lV + hh * aV; dwX = h3 * mX + hh * bX; dwV = h3 * mV + hh * bV;
            vjp(ld[1], wX, wV, aX, aV, Fb, dt);
            hvjp(ld[1], Fb, dt, ddist[1], dd[1], dwX, dwV, bX, bV);
            sX = sX + aX; sV = sV + aV; tX = tX + bX; tV = tV + bV;
            // stage 1
            wX = h6 * lX + hh * aX; wV = h6 * lV + hh * aV; dwX = h6 * mX + hh * bX; dwV = h6 * mV + hh * bV;
            vjp(ld[0], wX, wV, aX, aV, Fb, dt);
            hvjp(ld[0], Fb, dt, ddist[0], dd[0], dwX, dwV, bX, bV);
            sX = sX + aX; sV = sV + aV; tX = tX + bX; tV = tV + bV;
            // stage cost
            lX = isEnd ? sX + two_alpha * (X - ref) : sX;
            mX = isEnd ? tX + two_alpha * dX : tX;
            if (hasV) { lV = sV + two_beta * V; mV = tV + two_beta * dV; }
            else {
                if (isEnd) Hv[NU * k + (l - NVL)] = tV + two_gamma * vd;
                lV = 0.0; mV = 0.0;
            }
        }
    }

    // ---- problem evaluators (what CasADiProblem::eval_* would compute, m = 0) ----
    __device__ double eval_psi(const double *x, const double *, const double *, double *) {
        __syncwarp();
        double f = rollout(x);
        __syncwarp();
        return f;
    }
    __device__ double eval_psi_grad_psi(const double *x, const double *, const double *, double *grad) {
        __syncwarp();
        double f = rollout(x);
        __syncwarp();
        adjoint(x, grad);
        __syncwarp();
        return f;
    }
    /// Requires that the last rollout was at x (PANTR guarantees this: pantr.tpp:314 then :373).
    __device__ void hess_psi_prod(const double *x, const double *, const double *, const double *v,
                                  double *Hv) {
        __syncwarp();
        hvp(x, v, Hv);
        __syncwarp();
    }
    /// Plant step x+ = f_d(x, u) (problems/problems/formulation.hpp:258-261)
    __device__ void plant_step(const double *x, const double *u, double *xn) const {
        double X = active ? x[l] : 0.0, V = hasV ? x[NL + l] : 0.0, ud = isEnd ? u[l - NVL] : 0.0;
        double Xn, Vn;
        rk4(X, V, ud, Xn, Vn);
        if (active) xn[l] = Xn;
        if (hasV) xn[NL + l] = Vn;
    }
