// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_core.hpp header).
// C entry points for ctypes (tests/, smoke(), bench.py cpu_baseline / --impl reference).
#include "hanging_chain.hpp"
#include "quadcopter.hpp"
#include "quadcopter_condensed.hpp"
#include "generic_problems.hpp"
#include "oracle_alm.hpp"
#include <atomic>
#include <cstring>
#include <memory>
#include <thread>

using namespace oracle;

namespace {

constexpr int NPARAMS = 48, NSTATS = 24;

std::unique_ptr<Problem> make_problem(int problem_id, int N) {
    switch (problem_id) {
        case 0: return std::make_unique<HangingChain>(N);
        case 1: return std::make_unique<SimpleQuadcopter>(N);
        case 2: return std::make_unique<RealisticQuadcopter>(N);
        // 5, 6: the condensed formulation of the same two problems (operation-order mirror of the CUDA path)
        case 5: return std::make_unique<SimpleQuadcopterCondensed>(N);
        case 6: return std::make_unique<RealisticQuadcopterCondensed>(N);
        default: return nullptr;
    }
}
int nx_of(int problem_id) { return problem_id == 0 ? HangingChain::NX : 9; }
void set_x0(Problem &p, int problem_id, const double *x0) {
    if (problem_id == 0) { auto &hc = static_cast<HangingChain &>(p); std::copy(x0, x0 + hc.NX, hc.x0.begin()); }
    if (problem_id == 1) { auto &q = static_cast<SimpleQuadcopter &>(p); std::copy(x0, x0 + 9, q.x0.begin()); }
    if (problem_id == 2) { auto &q = static_cast<RealisticQuadcopter &>(p); std::copy(x0, x0 + 9, q.x0.begin()); }
    if (problem_id == 5) { auto &q = static_cast<SimpleQuadcopterCondensed &>(p); std::copy(x0, x0 + 9, q.x0.begin()); }
    if (problem_id == 6) { auto &q = static_cast<RealisticQuadcopterCondensed &>(p); std::copy(x0, x0 + 9, q.x0.begin()); }
}

ALMSolver make_solver(const double *P) {
    ALMSolver s;
    auto &a = s.params;
    a.tolerance = P[0]; a.dual_tolerance = P[1]; a.penalty_update_factor = P[2];
    a.initial_penalty = P[3]; a.initial_tolerance = P[4]; a.tolerance_update_factor = P[5];
    a.rel_penalty_increase_threshold = P[6]; a.max_multiplier = P[7]; a.max_penalty = P[8];
    a.min_penalty = P[9]; a.max_iter = (unsigned)P[10]; a.initial_penalty_factor = P[11];
    a.single_penalty_factor = P[12] != 0;
    a.max_num_initial_retries = (unsigned)P[13]; a.max_num_retries = (unsigned)P[14]; a.max_total_num_retries = (unsigned)P[15];
    a.penalty_update_factor_lower = P[37]; a.min_penalty_update_factor = P[38]; a.initial_penalty_lower = P[39];
    a.initial_tolerance_increase = P[41]; a.rho_increase = P[42]; a.rho_max = P[43];
    auto &p = s.inner_solver.params;
    p.max_iter = (unsigned)P[16]; p.stop_crit = (int)P[17]; p.L_min = P[18]; p.L_max = P[19];
    p.quadratic_upperbound_tolerance_factor = P[20]; p.TR_tolerance_factor = P[21];
    p.ratio_threshold_acceptable = P[22]; p.ratio_threshold_good = P[23];
    p.radius_factor_rejected = P[24]; p.radius_factor_acceptable = P[25]; p.radius_factor_good = P[26];
    p.initial_radius = P[27]; p.min_radius = P[28];
    p.Lipschitz.L_0 = P[29]; p.Lipschitz.eps = P[30]; p.Lipschitz.delta = P[31];
    p.Lipschitz.Lgamma_factor = P[32];
    p.compute_ratio_using_new_stepsize = P[33] != 0;
    p.recompute_last_prox_step_after_direction_reset = P[34] != 0;
    p.disable_acceleration = P[35] != 0; p.ratio_approx_fbe_quadratic_model = P[36] != 0;
    s.inner_solver.direction.direction_params.hessian_vec_factor = P[40];
    auto &c = s.inner_solver.direction.steihaug.params;
    c.tol_scale = P[44]; c.tol_scale_root = P[45]; c.tol_max = P[46]; c.max_iter_factor = P[47];
    return s;
}

void write_stats(const ALMStats &s, const EvalCounter &e, double *o) {
    o[0] = s.status; o[1] = s.outer_iterations; o[2] = s.inner_iterations; o[3] = s.eps_;
    o[4] = s.delta; o[5] = s.norm_penalty; o[6] = s.final_psi; o[7] = s.final_gamma;
    o[8] = s.final_fbe; o[9] = s.accelerated_step_rejected; o[10] = s.direction_failures;
    o[11] = s.inner_convergence_failures; o[12] = e.psi; o[13] = e.psi_grad_psi; o[14] = e.grad_L;
    o[15] = e.hess_psi_prod; o[16] = e.prox_grad_step; o[17] = e.grad_psi;
    o[18] = e.inactive_indices_res_lna; o[19] = e.proj_multipliers; o[20] = o[21] = o[22] = o[23] = 0;
}

} // namespace

extern "C" {

int oracle_nparams() { return NPARAMS; }
int oracle_nstats() { return NSTATS; }
int oracle_trace_width() { return 11; }

// dims[0..3] = n, m, nx, nu
int oracle_dims(int problem_id, int N, long *dims) {
    auto p = make_problem(problem_id, N);
    if (!p) return -1;
    dims[0] = p->n; dims[1] = p->m; dims[2] = nx_of(problem_id); dims[3] = p->n / N;
    return 0;
}

// f and g = [c(x_0); ...; c(x_N)] at u (TypeErasedProblem::eval_f / eval_g)
int oracle_eval_f_g(int problem_id, int N, const double *x0, const double *u, double *f, double *g) {
    auto p = make_problem(problem_id, N);
    if (!p) return -1;
    set_x0(*p, problem_id, x0);
    *f = p->eval_f(u);
    if (g && p->m) p->eval_g(u, g);
    return 0;
}

// sin/cos restatement shared with the CUDA path (csrc/pantr_hd.h)
void oracle_sincos(long n, const double *x, double *s, double *c) {
    for (long i = 0; i < n; ++i) pantr::pantr_sincos(x[i], s + i, c + i);
}
// the fma form the condensed quadcopter formulation uses (device: stage_sincos, oracle: quadcopter_condensed.hpp)
void oracle_sincos_fma(long n, const double *x, double *s, double *c) {
    for (long i = 0; i < n; ++i) pantr::pantr_sincos_fma(x[i], s + i, c + i);
}

// Evaluate psi, grad psi, yhat and (if v) hess_psi_prod at (u, y, Sigma) for initial state x0.
int oracle_eval(int problem_id, int N, const double *x0, const double *u, const double *y,
                const double *Sigma, const double *v, double *psi, double *grad, double *yhat,
                double *Hv) {
    auto p = make_problem(problem_id, N);
    if (!p) return -1;
    set_x0(*p, problem_id, x0);
    if (grad) *psi = p->eval_psi_grad_psi(u, y, Sigma, grad);
    if (yhat || !grad) *psi = p->eval_psi(u, y, Sigma, yhat);
    if (v && Hv) p->eval_hess_psi_prod(u, y, Sigma, 1, v, Hv);
    return 0;
}

// One member of the problem plugin table (fn = pantr::ProblemFn of csrc/problem_face.cuh) for one instance.
// v doubles as grad_psi for the box functions, y as z; scalar = scale / gamma / M.
int oracle_problem_eval(int problem_id, int N, int fn, const double *x0, const double *x, const double *y,
                        const double *Sigma, const double *v, double scalar, double *out_s, double *out_n,
                        double *out_n2, double *out_m, int *out_i, int *out_cnt) {
    auto p = make_problem(problem_id, N);
    if (!p) return -1;
    if (x0) set_x0(*p, problem_id, x0);
    switch (fn) {
        case 0: *out_s = p->eval_f(x); break;
        case 1: (void)p->face_grad_L(x, nullptr, out_n); break;
        case 2: *out_s = p->face_grad_L(x, nullptr, out_n); break;
        case 3: p->eval_g(x, out_m); break;
        case 4: p->face_grad_g_prod(x, y, out_n); break;
        case 5: (void)p->face_grad_L(x, y, out_n); break;
        case 6: p->face_hess_L_prod(x, y, scalar, v, out_n); break;
        case 7:
            p->eval_hess_psi_prod(x, y, Sigma, scalar, v, out_n);
            if (scalar != 1 && (problem_id == 1 || problem_id == 2)) for (index_t i = 0; i < p->n; ++i) out_n[i] *= scalar;
            break;
        case 8: (void)p->eval_prox_grad_step(scalar, x, v, out_n, out_n2); break;
        case 9: p->eval_proj_diff_g(y, out_m); break;
        case 10:
            for (index_t i = 0; i < p->m; ++i) out_m[i] = y[i];
            p->eval_proj_multipliers(out_m, scalar);
            break;
        case 11: {
            std::vector<index_t> J(p->n);
            index_t nJ = p->eval_inactive_indices_res_lna(scalar, x, v, J.data());
            for (index_t j = 0; j < nJ; ++j) out_i[j] = (int)J[j];
            *out_cnt = (int)nJ;
            break;
        }
        default: return -1;
    }
    return 0;
}

// Plant step x+ = f_d(x, u): problems/problems/formulation.hpp:258-261
int oracle_simulate_dynamics(int problem_id, const double *x, const double *u, double *xn) {
    if (problem_id == 0) { HangingChain hc(1); hc.rk4(x, u, xn); return 0; }
    if (problem_id == 1) { pantr::quad::rk4<pantr::quad::SimpleGen>(x, u, xn, nullptr); return 0; }
    if (problem_id == 2) { pantr::quad::rk4<pantr::quad::RealisticGen>(x, u, xn, nullptr); return 0; }
    if (problem_id == 5) { SimpleQuadcopterCondensed(1).plant_step(x, u, xn); return 0; }
    if (problem_id == 6) { RealisticQuadcopterCondensed(1).plant_step(x, u, xn); return 0; }
    return -1;
}

// One ALM+PANTR solve.  u (n) and y (m) are in/out.  trace (optional): max_trace rows of 11 doubles.
int oracle_solve(int problem_id, int N, const double *params, const double *x0, double *u, double *y,
                 double *stats, double *trace, int max_trace, int *n_trace) {
    auto p = make_problem(problem_id, N);
    if (!p) return -1;
    set_x0(*p, problem_id, x0);
    auto solver = make_solver(params);
    std::vector<TraceRec> tr;
    if (trace) solver.inner_solver.trace = &tr;
    vec x(u, u + p->n), yy(y, y + p->m);
    auto s = solver(*p, x, yy);
    std::copy(x.begin(), x.end(), u);
    std::copy(yy.begin(), yy.end(), y);
    write_stats(s, p->evaluations, stats);
    if (trace) {
        int nt = std::min<int>(max_trace, (int)tr.size());
        for (int i = 0; i < nt; ++i) {
            double *r = trace + 11 * i; const auto &t = tr[i];
            r[0] = t.k; r[1] = t.psi; r[2] = t.fbe; r[3] = t.gamma; r[4] = t.eps_; r[5] = t.radius;
            r[6] = t.rho; r[7] = t.q_model; r[8] = t.nJ; r[9] = t.cg_iters; r[10] = t.accepted;
        }
        if (n_trace) *n_trace = (int)tr.size();
    }
    return 0;
}

// B independent solves, one instance per thread at a time (the reference runs one process per
// experiment: new-benchmarks-paper/run-single-mpc-exp.sh:19, README.md:19).
int oracle_solve_batch(int problem_id, int N, const double *params, long B, const double *x0,
                       double *u, double *y, double *stats, int nthreads) {
    auto probe = make_problem(problem_id, N);
    if (!probe) return -1;
    const long n = probe->n, m = probe->m, nx = nx_of(problem_id);
    std::atomic<long> next{0};
    auto worker = [&] {
        auto p = make_problem(problem_id, N);
        auto solver = make_solver(params);
        for (long b; (b = next.fetch_add(1)) < B;) {
            set_x0(*p, problem_id, x0 + b * nx);
            p->evaluations = {};
            vec x(u + b * n, u + (b + 1) * n), yy(y + b * m, y + (b + 1) * m);
            auto s = solver(*p, x, yy);
            std::copy(x.begin(), x.end(), u + b * n);
            std::copy(yy.begin(), yy.end(), y + b * m);
            write_stats(s, p->evaluations, stats + b * NSTATS);
        }
    };
    if (nthreads <= 1) { worker(); return 0; }
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; ++t) th.emplace_back(worker);
    for (auto &t : th) t.join();
    return 0;
}

// ---- generic problems (known-answer tests of the reference, see generic_problems.hpp) ----
static void set_boxes(Problem &p, const double *Clb, const double *Cub, const double *Dlb, const double *Dub) {
    if (Clb) p.C.lowerbound.assign(Clb, Clb + p.n);
    if (Cub) p.C.upperbound.assign(Cub, Cub + p.n);
    if (Dlb) p.D.lowerbound.assign(Dlb, Dlb + p.m);
    if (Dub) p.D.upperbound.assign(Dub, Dub + p.m);
}
static int solve_generic(Problem &p, const double *params, double *x, double *y, double *stats) {
    auto solver = make_solver(params);
    vec xx(x, x + p.n), yy(y, y + p.m);
    auto s = solver(p, xx, yy);
    std::copy(xx.begin(), xx.end(), x);
    std::copy(yy.begin(), yy.end(), y);
    write_stats(s, p.evaluations, stats);
    return 0;
}

// min 1/2 x'Hx + c'x + f0  s.t. x in C, Gx + g0 in D   (test-alm.cpp QPs)
int oracle_solve_qp(long n, long m, const double *H, const double *c, double f0, const double *G,
                    const double *g0, const double *Clb, const double *Cub, const double *Dlb,
                    const double *Dub, const double *params, double *x, double *y, double *stats) {
    DenseQP p(n, m);
    std::copy(H, H + n * n, p.H.begin()); std::copy(c, c + n, p.c.begin()); p.f0 = f0;
    if (m) { std::copy(G, G + m * n, p.G.begin()); std::copy(g0, g0 + m, p.g0.begin()); }
    set_boxes(p, Clb, Cub, Dlb, Dub);
    return solve_generic(p, params, x, y, stats);
}

// CasADi-generated shared object (test-casadi.cpp); composed != 0 uses the oracle's own AL formulas
int oracle_solve_casadi(const char *so, long n, long m, const double *param, long np, const double *Clb,
                        const double *Cub, const double *Dlb, const double *Dub, int composed,
                        const double *params, double *x, double *y, double *stats) {
    try {
        if (composed) {
            CasadiComposedProblem p(so, n, m, param, np);
            set_boxes(p, Clb, Cub, Dlb, Dub); set_boxes(p.inner, Clb, Cub, Dlb, Dub);
            return solve_generic(p, params, x, y, stats);
        }
        CasadiCProblem p(so, n, m, param, np);
        set_boxes(p, Clb, Cub, Dlb, Dub);
        return solve_generic(p, params, x, y, stats);
    } catch (std::exception &) { return -2; }
}

// psi, grad psi, yhat, Hv of a CasADi-generated problem, either straight from the generated
// psi_grad_psi / psi / hess_psi_prod (composed = 0) or through the oracle's AL compositions.
int oracle_eval_casadi(const char *so, long n, long m, const double *param, long np, const double *Dlb,
                       const double *Dub, int composed, const double *x, const double *y,
                       const double *S, const double *v, double *psi, double *grad, double *yhat,
                       double *Hv, double *psi2) {
    try {
        CasadiComposedProblem pc(so, n, m, param, np);
        set_boxes(pc, nullptr, nullptr, Dlb, Dub); set_boxes(pc.inner, nullptr, nullptr, Dlb, Dub);
        Problem &p = composed ? static_cast<Problem &>(pc) : static_cast<Problem &>(pc.inner);
        *psi  = p.eval_psi_grad_psi(x, y, S, grad);
        *psi2 = p.eval_psi(x, y, S, yhat);
        if (v && Hv) p.eval_hess_psi_prod(x, y, S, 1, v, Hv);
        return 0;
    } catch (std::exception &) { return -2; }
}

} // extern "C"
