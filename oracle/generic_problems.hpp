// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_core.hpp header).
// Two generic problems used to pin the PANTR+ALM restatement with the reference's own
// known-answer tests (which are solver independent):
//   * DenseQP       — the FunctionalProblem QPs of alpaqa/test/outer/test-alm.cpp:7-304, with the
//                     default AL compositions of implementation/problem/type-erased-problem.tpp:8-32,
//                     137-196 and an exact Hessian-vector product of the augmented Lagrangian.
//   * CasadiCProblem — a CasADi-generated shared object driven exactly like
//                     interop/casadi/.../implementation/casadi/CasADiProblem.tpp:208-368 does
//                     (argument order f(x,p), psi_grad_psi(x,p,y,S,zl,zu), psi(...), grad_L(x,p,y),
//                     hess_psi_prod(x,p,y,S,scale,zl,zu,v)); used with the committed
//                     alpaqa/test/outer/rosenbrock_functions_test.c compiled into oracle/_ref/.
#pragma once
#include "oracle_core.hpp"
#include <dlfcn.h>
#include <stdexcept>
#include <string>

namespace oracle {

// Default AL compositions of the reference's problem API for a problem that supplies f, grad f,
// g and grad g * y: implementation/problem/type-erased-problem.tpp:8-32 (calc_yhat_dTyhat),
// :141-150 (grad_L), :155-166 (psi), :184-198 (psi_grad_psi).
struct ComposedProblem : Problem {
    virtual real_t f_impl(const real_t *x) const                               = 0;
    virtual void grad_f_impl(const real_t *x, real_t *g) const                 = 0;
    virtual void g_impl(const real_t *x, real_t *g) const                      = 0;
    virtual void grad_g_prod_impl(const real_t *x, const real_t *y, real_t *o) const = 0;

    real_t eval_f_impl(const real_t *x) const override { return f_impl(x); }
    void eval_g_impl(const real_t *x, real_t *g) const override { g_impl(x, g); }
    real_t calc_yhat_dTyhat(real_t *g_yhat, const real_t *y, const real_t *S) const {
        for (index_t i = 0; i < m; ++i) g_yhat[i] += (1 / S[i]) * y[i];
        eval_proj_diff_g(g_yhat, g_yhat);
        real_t d = 0;
        for (index_t i = 0; i < m; ++i) { d += g_yhat[i] * S[i] * g_yhat[i]; g_yhat[i] = S[i] * g_yhat[i]; }
        return d;
    }
    real_t eval_psi_impl(const real_t *x, const real_t *y, const real_t *S, real_t *yhat) const override {
        if (m == 0) return f_impl(x);
        vec tmp(yhat ? 0 : m);
        real_t *yh = yhat ? yhat : tmp.data();
        real_t f = f_impl(x);
        g_impl(x, yh);
        return f + real_t(0.5) * calc_yhat_dTyhat(yh, y, S);
    }
    void eval_grad_L_impl(const real_t *x, const real_t *y, real_t *gL) const override {
        grad_f_impl(x, gL);
        if (m == 0) return;
        vec w(n);
        grad_g_prod_impl(x, y, w.data());
        for (index_t j = 0; j < n; ++j) gL[j] += w[j];
    }
    real_t eval_psi_grad_psi_impl(const real_t *x, const real_t *y, const real_t *S, real_t *grad) const override {
        if (m == 0) { grad_f_impl(x, grad); return f_impl(x); }
        vec yhat(m);
        real_t psi = eval_psi_impl(x, y, S, yhat.data());
        eval_grad_L_impl(x, yhat.data(), grad);
        return psi;
    }
};

// f(x) = 1/2 x'Hx + c'x + f0 ,  g(x) = Gx + g0   (H: n x n, G: m x n, row major)
struct DenseQP : ComposedProblem {
    vec H, c, G, g0;
    real_t f0 = 0;
    DenseQP(index_t n_, index_t m_) : H(n_ * n_), c(n_), G(m_ * n_), g0(m_) {
        n = n_; m = m_;
        C.lowerbound.assign(n, -inf); C.upperbound.assign(n, +inf);
        D.lowerbound.assign(m, -inf); D.upperbound.assign(m, +inf);
    }
    void Hx(const real_t *x, real_t *o) const {
        for (index_t i = 0; i < n; ++i) { real_t s = 0; for (index_t j = 0; j < n; ++j) s += H[i * n + j] * x[j]; o[i] = s; }
    }
    real_t f_impl(const real_t *x) const override {
        vec t(n); Hx(x, t.data());
        real_t s = f0;
        for (index_t i = 0; i < n; ++i) s += (real_t(0.5) * t[i] + c[i]) * x[i];
        return s;
    }
    void grad_f_impl(const real_t *x, real_t *g) const override { Hx(x, g); for (index_t i = 0; i < n; ++i) g[i] += c[i]; }
    void g_impl(const real_t *x, real_t *g) const override {
        for (index_t i = 0; i < m; ++i) { real_t s = g0[i]; for (index_t j = 0; j < n; ++j) s += G[i * n + j] * x[j]; g[i] = s; }
    }
    void grad_g_prod_impl(const real_t *, const real_t *y, real_t *o) const override {
        for (index_t j = 0; j < n; ++j) o[j] = 0;
        for (index_t i = 0; i < m; ++i) for (index_t j = 0; j < n; ++j) o[j] += G[i * n + j] * y[i];
    }
    // exact Hessian of the AL: scale*H + G' diag(S_i [zeta_i outside D]) G
    // (casadi_generator/__init__.py:181-189 differentiates psi symbolically)
    void eval_hess_psi_prod_impl(const real_t *x, const real_t *y, const real_t *S, real_t scale,
                                 const real_t *v, real_t *Hv) const override {
        Hx(v, Hv);
        for (index_t i = 0; i < n; ++i) Hv[i] *= scale;
        if (m == 0) return;
        vec z(m);
        g_impl(x, z.data());
        for (index_t i = 0; i < m; ++i) {
            real_t zeta = z[i] + y[i] / S[i];
            bool outside = zeta < D.lowerbound[i] || zeta > D.upperbound[i];
            if (!outside) continue;
            real_t Gv = 0;
            for (index_t j = 0; j < n; ++j) Gv += G[i * n + j] * v[j];
            for (index_t j = 0; j < n; ++j) Hv[j] += G[i * n + j] * (S[i] * Gv);
        }
    }
};

// CasADi-generated C functions behind dlopen (CasADiLoader-util.hpp:24-29 uses casadi::external;
// the generated C ABI is int fn(const double **arg, double **res, long long *iw, double *w, int mem)).
struct CasadiCProblem : Problem {
    using fn_t   = int (*)(const double **, double **, long long *, double *, int);
    using work_t = int (*)(long long *, long long *, long long *, long long *);
    struct Fun {
        fn_t fn = nullptr;
        mutable std::vector<const double *> arg;
        mutable std::vector<double *> res;
        mutable std::vector<long long> iw;
        mutable vec w;
        void load(void *h, const std::string &name) {
            fn = reinterpret_cast<fn_t>(dlsym(h, name.c_str()));
            auto wk = reinterpret_cast<work_t>(dlsym(h, (name + "_work").c_str()));
            if (!fn || !wk) throw std::runtime_error("CasadiCProblem: missing function " + name);
            long long a = 0, r = 0, i = 0, ww = 0;
            wk(&a, &r, &i, &ww);
            arg.assign(a + 8, nullptr); res.assign(r + 8, nullptr); iw.assign(i + 1, 0); w.assign(ww + 1, 0);
        }
        void operator()(std::initializer_list<const double *> in, std::initializer_list<double *> out) const {
            size_t k = 0; for (auto p : in) arg[k++] = p;
            k = 0; for (auto p : out) res[k++] = p;
            fn(arg.data(), res.data(), iw.data(), w.data(), 0);
        }
    };
    void *handle = nullptr;
    Fun f, g, psi_grad_psi, psi, grad_L, hess_psi_prod;
    vec param;
    mutable vec work_m;

    CasadiCProblem(const char *so, index_t n_, index_t m_, const real_t *p, index_t np) : param(p, p + np), work_m(m_) {
        n = n_; m = m_;
        handle = dlopen(so, RTLD_NOW | RTLD_LOCAL);
        if (!handle) throw std::runtime_error(std::string("dlopen: ") + dlerror());
        f.load(handle, "f"); g.load(handle, "g"); psi_grad_psi.load(handle, "psi_grad_psi");
        psi.load(handle, "psi"); grad_L.load(handle, "grad_L"); hess_psi_prod.load(handle, "hess_psi_prod");
        C.lowerbound.assign(n, -inf); C.upperbound.assign(n, +inf);
        D.lowerbound.assign(m, -inf); D.upperbound.assign(m, +inf);
    }
    ~CasadiCProblem() override { if (handle) dlclose(handle); }
    real_t eval_f_impl(const real_t *x) const override { real_t v; f({x, param.data()}, {&v}); return v; }
    void eval_g_impl(const real_t *x, real_t *gx) const override { g({x, param.data()}, {gx}); }
    // CasADiProblem.tpp:279-297
    real_t eval_psi_impl(const real_t *x, const real_t *y, const real_t *S, real_t *yhat) const override {
        real_t v;
        psi({x, param.data(), y, S, D.lowerbound.data(), D.upperbound.data()}, {&v, yhat ? yhat : work_m.data()});
        return v;
    }
    // CasADiProblem.tpp:243-266
    real_t eval_psi_grad_psi_impl(const real_t *x, const real_t *y, const real_t *S, real_t *grad) const override {
        real_t v;
        psi_grad_psi({x, param.data(), y, S, D.lowerbound.data(), D.upperbound.data()}, {&v, grad});
        return v;
    }
    // CasADiProblem.tpp:268-277
    void eval_grad_L_impl(const real_t *x, const real_t *y, real_t *gL) const override {
        grad_L({x, param.data(), y}, {gL});
    }
    // CasADiProblem.tpp:342-353
    void eval_hess_psi_prod_impl(const real_t *x, const real_t *y, const real_t *S, real_t scale,
                                 const real_t *v, real_t *Hv) const override {
        hess_psi_prod({x, param.data(), y, S, &scale, D.lowerbound.data(), D.upperbound.data(), v}, {Hv});
    }
};

// The same shared object, but psi / grad psi / yhat composed by the oracle's own AL formulas from
// the generated f_grad_f, g and grad_L (grad g * y = grad_L(x, y) - grad f): cross-checks
// ComposedProblem against CasADi's symbolic psi, psi_grad_psi (casadi_generator/__init__.py:64-77).
struct CasadiComposedProblem : ComposedProblem {
    CasadiCProblem inner;
    CasadiCProblem::Fun f_grad_f;
    CasadiComposedProblem(const char *so, index_t n_, index_t m_, const real_t *p, index_t np)
        : inner(so, n_, m_, p, np) {
        n = n_; m = m_;
        f_grad_f.load(inner.handle, "f_grad_f");
        C = inner.C; D = inner.D;
    }
    real_t f_impl(const real_t *x) const override { return inner.eval_f_impl(x); }
    void grad_f_impl(const real_t *x, real_t *g) const override { real_t v; f_grad_f({x, inner.param.data()}, {&v, g}); }
    void g_impl(const real_t *x, real_t *g) const override { inner.eval_g_impl(x, g); }
    void grad_g_prod_impl(const real_t *x, const real_t *y, real_t *o) const override {
        vec gf(n);
        grad_f_impl(x, gf.data());
        inner.eval_grad_L_impl(x, y, o);
        for (index_t j = 0; j < n; ++j) o[j] -= gf[j];
    }
    void eval_hess_psi_prod_impl(const real_t *x, const real_t *y, const real_t *S, real_t scale,
                                 const real_t *v, real_t *Hv) const override {
        inner.eval_hess_psi_prod_impl(x, y, S, scale, v, Hv);
    }
};

} // namespace oracle
