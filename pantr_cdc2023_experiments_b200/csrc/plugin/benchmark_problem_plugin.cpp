// The reference's problem plugin ABI on top of the B200 library.
//
// A shared object built from this file exports
//     extern "C" alpaqa_problem_register_t benchmark_problem_register(void *user_data)
// exactly like the reference's own `dl:` problems (problems/problems/hanging_chain.cpp:139-170, quadcopter.cpp,
// realistic_quadcopter.cpp): alpaqa's DLProblem (alpaqa/src/interop/dl/src/dl-problem.cpp:35-66) and the benchmark
// driver (drivers/casadi-dll-wrapper.cpp:69-129) can load it with prefix "benchmark_problem", passing
// std::any{ProblemConfig{size, options}} as user data, and get
//   * an alpaqa_problem_functions_t (alpaqa/src/interop/dl-api/include/alpaqa/dl/dl-problem.h:13-155) whose members
//     forward to the batched C ABI of libpantr_b200.so with B = 1 (every call = H2D copy, one kernel, D2H copy:
//     this is the drop-in face, not the fast path -- the fast path is pantr_b200_solve_batch),
//   * the extra functions the driver looks up by name (casadi-dll-wrapper.cpp:100-126): initial_guess,
//     get_initial_state_ptr, get_initial_state, set_initial_state, get_nx, get_nu, get_nc, get_nc_N, simulate_dynamics,
//     registered through the header's own register_member_function so that the std::function signatures are the
//     ones DLProblem::call_extra_func casts to.
// Members the reference's ss problems provide and this plugin does not: eval_jac_g, eval_hess_L and their
// nonzero counts (full sparse matrices; PANTR never calls them: newton-tr.hpp only needs Hessian-vector products).
// Members provided in addition: eval_hess_L_prod, eval_f_grad_f, eval_grad_psi.
//
// Compiled against the reference headers where they lie (-I /root/reference/alpaqa/src/interop/dl-api/include
// -I /root/reference/problems), never copied.  One object per problem: -DPANTR_PLUGIN_PROBLEM=0|1|2
// (hanging_chain | quadcopter | realistic_quadcopter), written to pantr_cdc2023_experiments_b200/plugins/<name>.so.
#include <alpaqa/dl/dl-problem.h>
#include <problem-config.hpp>

#include "../../../include/pantr_b200.h"

#include <algorithm>
#include <any>
#include <cstring>
#include <span>
#include <stdexcept>
#include <string>
#include <vector>

#ifndef PANTR_PLUGIN_PROBLEM
#define PANTR_PLUGIN_PROBLEM 0
#endif

namespace {

using real_t   = alpaqa_real_t;
using length_t = alpaqa_length_t;

void check(int rc, const char *what) {
    if (rc != PANTR_B200_OK) throw std::runtime_error(std::string(what) + ": " + pantr_b200_last_error());
}

struct Problem {
    pantr_b200_solver *h = nullptr;
    length_t n = 0, m = 0, nx = 0, nu = 0, nc = 0;
    std::vector<real_t> initial_state; // the NLP parameter; the driver writes through get_initial_state_ptr()
    std::vector<real_t> guess;
    alpaqa_problem_functions_t funcs;

    explicit Problem(int horizon) {
        check(pantr_b200_create(PANTR_PLUGIN_PROBLEM, horizon, nullptr, nullptr, nullptr, nullptr, 0, &h), "pantr_b200_create");
        int64_t d[5];
        check(pantr_b200_get_dims(h, d), "pantr_b200_get_dims");
        n = d[0]; m = d[1]; nx = d[2]; nu = d[3]; nc = d[4];
        initial_state.resize(nx); guess.resize(n);
        check(pantr_b200_get_initial(h, initial_state.data(), guess.data()), "pantr_b200_get_initial");
        std::memset(&funcs, 0, sizeof(funcs));
        using alpaqa::member_caller;
        using P                = Problem;
        funcs.n                = n;
        funcs.m                = m;
        funcs.initialize_box_C = member_caller<&P::get_C>();
        funcs.initialize_box_D = member_caller<&P::get_D>();
        funcs.eval_f           = member_caller<&P::eval_f>();
        funcs.eval_grad_f      = member_caller<&P::eval_grad_f>();
        funcs.eval_f_grad_f    = member_caller<&P::eval_f_grad_f>();
        funcs.eval_g           = member_caller<&P::eval_g>();
        funcs.eval_grad_g_prod = member_caller<&P::eval_grad_g_prod>();
        funcs.eval_grad_L      = member_caller<&P::eval_grad_L>();
        funcs.eval_hess_L_prod = member_caller<&P::eval_hess_L_prod>();
        funcs.eval_hess_ψ_prod = member_caller<&P::eval_hess_ψ_prod>();
        funcs.eval_ψ           = member_caller<&P::eval_ψ>();
        funcs.eval_grad_ψ      = member_caller<&P::eval_grad_ψ>();
        funcs.eval_ψ_grad_ψ    = member_caller<&P::eval_ψ_grad_ψ>();
    }
    ~Problem() { pantr_b200_destroy(h); }
    Problem(const Problem &)            = delete;
    Problem &operator=(const Problem &) = delete;

    const real_t *x0() const { return initial_state.data(); }

    // ---- alpaqa_problem_functions_t members (dl-problem.h:18-153), B = 1 ----
    real_t eval_f(const real_t *x) const {
        real_t f;
        check(pantr_b200_eval_f_batch(h, 1, x0(), x, &f), "eval_f");
        return f;
    }
    void eval_grad_f(const real_t *x, real_t *grad_fx) const { check(pantr_b200_eval_grad_f_batch(h, 1, x0(), x, grad_fx), "eval_grad_f"); }
    real_t eval_f_grad_f(const real_t *x, real_t *grad_fx) const {
        real_t f;
        check(pantr_b200_eval_f_grad_f_batch(h, 1, x0(), x, &f, grad_fx), "eval_f_grad_f");
        return f;
    }
    void eval_g(const real_t *x, real_t *gx) const { check(pantr_b200_eval_g_batch(h, 1, x0(), x, gx), "eval_g"); }
    void eval_grad_g_prod(const real_t *x, const real_t *y, real_t *grad_gxy) const {
        if (m == 0) { std::fill_n(grad_gxy, n, real_t(0)); return; } // J_g is empty
        check(pantr_b200_eval_grad_g_prod_batch(h, 1, x0(), x, y, grad_gxy), "eval_grad_g_prod");
    }
    void eval_grad_L(const real_t *x, const real_t *y, real_t *grad_L, real_t *) const {
        check(pantr_b200_eval_grad_L_batch(h, 1, x0(), x, y, grad_L), "eval_grad_L");
    }
    void eval_hess_L_prod(const real_t *x, const real_t *y, real_t scale, const real_t *v, real_t *Hv) const {
        check(pantr_b200_eval_hess_L_prod_batch(h, 1, x0(), x, y, scale, v, Hv), "eval_hess_L_prod");
    }
    // zl, zu: the caller passes the problem's own box D (TypeErasedProblem); the device model has it built in
    void eval_hess_ψ_prod(const real_t *x, const real_t *y, const real_t *Σ, real_t scale, const real_t *, const real_t *,
                          const real_t *v, real_t *Hv) const {
        check(pantr_b200_eval_hess_psi_prod_batch(h, 1, x0(), x, y, Σ, scale, v, Hv), "eval_hess_psi_prod");
    }
    real_t eval_ψ(const real_t *x, const real_t *y, const real_t *Σ, const real_t *, const real_t *, real_t *ŷ) const {
        real_t psi;
        check(pantr_b200_eval_psi_batch(h, 1, x0(), x, y, Σ, ŷ, &psi), "eval_psi");
        return psi;
    }
    void eval_grad_ψ(const real_t *x, const real_t *y, const real_t *Σ, const real_t *, const real_t *, real_t *grad_ψ, real_t *,
                     real_t *) const {
        check(pantr_b200_eval_grad_psi_batch(h, 1, x0(), x, y, Σ, grad_ψ), "eval_grad_psi");
    }
    real_t eval_ψ_grad_ψ(const real_t *x, const real_t *y, const real_t *Σ, const real_t *, const real_t *, real_t *grad_ψ,
                         real_t *, real_t *) const {
        real_t psi;
        check(pantr_b200_eval_psi_grad_psi_batch(h, 1, x0(), x, y, Σ, &psi, grad_ψ), "eval_psi_grad_psi");
        return psi;
    }
    void get_C(real_t *lb, real_t *ub) const { check(pantr_b200_get_box_C(h, lb, ub), "get_box_C"); }
    void get_D(real_t *lb, real_t *ub) const { check(pantr_b200_get_box_D(h, lb, ub), "get_box_D"); }

    // ---- extra functions (problems/problems/formulation.hpp:236-261) ----
    void get_initial_guess(real_t *u0) const { std::copy(guess.begin(), guess.end(), u0); }
    void get_initial_state(real_t *x) const { std::copy(initial_state.begin(), initial_state.end(), x); }
    void set_initial_state(const real_t *x) { std::copy_n(x, nx, initial_state.begin()); }
    [[nodiscard]] length_t get_nx() const { return nx; }
    [[nodiscard]] length_t get_nu() const { return nu; }
    [[nodiscard]] length_t get_nc() const { return nc; }
    [[nodiscard]] length_t get_nc_N() const { return nc; } // stage = terminal constraints in all three models
    std::span<real_t> get_initial_state_ptr() { return {initial_state.data(), initial_state.size()}; }
    void simulate_dynamics(const real_t *x, const real_t *u, real_t *x_next) const {
        check(pantr_b200_simulate_dynamics_batch(h, 1, x, u, x_next), "simulate_dynamics");
    }
};

} // namespace

extern "C" __attribute__((visibility("default"))) alpaqa_problem_register_t benchmark_problem_register(void *user_data) {
    if (!user_data)
        throw std::invalid_argument("Missing user data"); // problems/problems/hanging_chain.cpp:141-142
    const auto *user        = reinterpret_cast<std::any *>(user_data);
    const auto &user_config = std::any_cast<const ProblemConfig &>(*user);
    int horizon = 30; // the reference's default (hanging_chain.cpp:29, quadcopter.cpp:29)
    if (user_config.size > 0)
        horizon = user_config.size;
    auto *problem = new Problem{horizon};
    alpaqa_problem_register_t result;
    std::memset(&result, 0, sizeof(result));
    alpaqa::register_member_function(result, "initial_guess", &Problem::get_initial_guess);
    alpaqa::register_member_function(result, "get_initial_state_ptr", &Problem::get_initial_state_ptr);
    alpaqa::register_member_function(result, "get_initial_state", &Problem::get_initial_state);
    alpaqa::register_member_function(result, "set_initial_state", &Problem::set_initial_state);
    alpaqa::register_member_function(result, "get_nx", &Problem::get_nx);
    alpaqa::register_member_function(result, "get_nu", &Problem::get_nu);
    alpaqa::register_member_function(result, "get_nc", &Problem::get_nc);
    alpaqa::register_member_function(result, "get_nc_N", &Problem::get_nc_N);
    alpaqa::register_member_function(result, "simulate_dynamics", &Problem::simulate_dynamics);
    result.instance  = problem;
    result.cleanup   = [](void *instance) { delete reinterpret_cast<Problem *>(instance); };
    result.functions = &problem->funcs;
    return result;
}
