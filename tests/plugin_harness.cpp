// TEST INFRASTRUCTURE.  Plays the part of the reference's plugin loader for one problem plugin:
// what alpaqa::dl::DLProblem (alpaqa/src/interop/dl/src/dl-problem.cpp:19-66,99-119) and the benchmark driver
// (drivers/casadi-dll-wrapper.cpp:69-129) do -- dlopen, dlsym "<prefix>_register", call it with
// std::any{ProblemConfig{size, options}}, read n / m / the function table, look the extra functions up by name
// with the exact std::function signatures call_extra_func casts to -- behind a flat C API for ctypes
// (tests/test_gpu_plugin.py).  Compiled against the reference headers where they lie, like the plugin itself.
#include <alpaqa/dl/dl-problem.h>
#include <problem-config.hpp>

#include <any>
#include <cstring>
#include <dlfcn.h>
#include <functional>
#include <memory>
#include <span>
#include <stdexcept>
#include <string>

namespace {
using real_t   = alpaqa_real_t;
using length_t = alpaqa_length_t;
using instance_t = void;

struct Loaded {
    void *handle = nullptr;
    alpaqa_problem_register_t r{};
    std::unique_ptr<alpaqa_function_dict_t> extra;
    ~Loaded() {
        if (r.cleanup && r.instance) r.cleanup(r.instance);
        extra.reset();
        if (handle) dlclose(handle);
    }
    // DLProblem::call_extra_func (alpaqa/src/interop/dl/include/alpaqa/dl/dl-problem.hpp): any_cast to std::function<Signature>
    template <class Signature, class... Args>
    decltype(auto) call_extra(const std::string &name, Args &&...args) {
        auto it = extra->dict.find(name);
        if (it == extra->dict.end()) throw std::out_of_range("extra function not found: " + name);
        auto &f = std::any_cast<std::function<Signature> &>(it->second); // throws std::bad_any_cast on a signature mismatch
        return f(r.instance, std::forward<Args>(args)...);
    }
};
thread_local std::string g_err;
} // namespace

extern "C" {

const char *harness_last_error() { return g_err.c_str(); }

// size <= 0 with null_user_data != 0: call the registration function without user data (must throw invalid_argument)
void *harness_open(const char *so, int size, int null_user_data) {
    auto L = std::make_unique<Loaded>();
    try {
        L->handle = dlopen(so, RTLD_LOCAL | RTLD_NOW);
        if (!L->handle) throw std::runtime_error(dlerror());
        auto *reg = reinterpret_cast<alpaqa_problem_register_t (*)(void *)>(dlsym(L->handle, "benchmark_problem_register"));
        if (!reg) throw std::runtime_error("symbol benchmark_problem_register not found");
        std::any user{ProblemConfig{static_cast<int32_t>(size), {}}};
        L->r = reg(null_user_data ? nullptr : &user);
        L->extra.reset(L->r.extra_functions);
        return L.release();
    } catch (std::invalid_argument &e) {
        g_err = std::string("invalid_argument: ") + e.what();
    } catch (std::exception &e) {
        g_err = std::string("exception: ") + e.what();
    }
    return nullptr;
}
void harness_close(void *p) { delete static_cast<Loaded *>(p); }
long harness_n(void *p) { return static_cast<Loaded *>(p)->r.functions->n; }
long harness_m(void *p) { return static_cast<Loaded *>(p)->r.functions->m; }

// bit i set <=> member i of the table below is non-null (dl-problem.cpp:99-119: null = "not provided")
unsigned harness_provided(void *p) {
    auto *f = static_cast<Loaded *>(p)->r.functions;
    const void *members[] = {(void *)f->eval_f, (void *)f->eval_grad_f, (void *)f->eval_g, (void *)f->eval_grad_g_prod,
                             (void *)f->eval_jac_g, (void *)f->get_jac_g_num_nonzeros, (void *)f->eval_grad_gi,
                             (void *)f->eval_hess_L_prod, (void *)f->eval_hess_L, (void *)f->get_hess_L_num_nonzeros,
                             (void *)f->eval_hess_ψ_prod, (void *)f->eval_hess_ψ, (void *)f->get_hess_ψ_num_nonzeros,
                             (void *)f->eval_f_grad_f, (void *)f->eval_f_g, (void *)f->eval_grad_f_grad_g_prod,
                             (void *)f->eval_grad_L, (void *)f->eval_ψ, (void *)f->eval_grad_ψ, (void *)f->eval_ψ_grad_ψ,
                             (void *)f->eval_prox_grad_step, (void *)f->initialize_box_C, (void *)f->initialize_box_D,
                             (void *)f->initialize_l1_reg};
    unsigned mask = 0;
    for (unsigned i = 0; i < sizeof(members) / sizeof(members[0]); ++i)
        if (members[i]) mask |= 1u << i;
    return mask;
}

#define L_ static_cast<Loaded *>(p)
#define F_ L_->r.functions
#define I_ L_->r.instance
int harness_call(void *p, const char *name, const real_t *x, const real_t *y, const real_t *S, const real_t *v, real_t scale,
                 const real_t *zl, const real_t *zu, real_t *out_s, real_t *out_n, real_t *out_m, real_t *work_n, real_t *work_m) {
    try {
        const std::string f = name;
        if (f == "eval_f") *out_s = F_->eval_f(I_, x);
        else if (f == "eval_grad_f") F_->eval_grad_f(I_, x, out_n);
        else if (f == "eval_f_grad_f") *out_s = F_->eval_f_grad_f(I_, x, out_n);
        else if (f == "eval_g") F_->eval_g(I_, x, out_m);
        else if (f == "eval_grad_g_prod") F_->eval_grad_g_prod(I_, x, y, out_n);
        else if (f == "eval_grad_L") F_->eval_grad_L(I_, x, y, out_n, work_n);
        else if (f == "eval_hess_L_prod") F_->eval_hess_L_prod(I_, x, y, scale, v, out_n);
        else if (f == "eval_hess_psi_prod") F_->eval_hess_ψ_prod(I_, x, y, S, scale, zl, zu, v, out_n);
        else if (f == "eval_psi") *out_s = F_->eval_ψ(I_, x, y, S, zl, zu, out_m);
        else if (f == "eval_grad_psi") F_->eval_grad_ψ(I_, x, y, S, zl, zu, out_n, work_n, work_m);
        else if (f == "eval_psi_grad_psi") *out_s = F_->eval_ψ_grad_ψ(I_, x, y, S, zl, zu, out_n, work_n, work_m);
        else if (f == "initialize_box_C") F_->initialize_box_C(I_, out_n, work_n);
        else if (f == "initialize_box_D") F_->initialize_box_D(I_, out_m, work_m);
        else throw std::invalid_argument("unknown member " + f);
        return 0;
    } catch (std::exception &e) { g_err = e.what(); return -1; }
}

// the extra functions with the signatures drivers/casadi-dll-wrapper.cpp:100-126 uses
int harness_extra(void *p, const char *name, const real_t *in1, const real_t *in2, real_t *out, long *out_len) {
    try {
        const std::string f = name;
        if (f == "initial_guess") L_->call_extra<void(const instance_t *, real_t *)>(f, out);
        else if (f == "get_nx" || f == "get_nu" || f == "get_nc" || f == "get_nc_N")
            *out_len = L_->call_extra<length_t(const instance_t *)>(f);
        else if (f == "get_initial_state_ptr") { // mutable view of the NLP parameter: write in1 through it if given, then read it back
            std::span<real_t> s = L_->call_extra<std::span<real_t>(instance_t *)>(f);
            *out_len = (long)s.size();
            if (in1) std::copy(in1, in1 + s.size(), s.begin());
            std::copy(s.begin(), s.end(), out);
        } else if (f == "get_initial_state") L_->call_extra<void(const instance_t *, real_t *)>(f, out);
        else if (f == "set_initial_state") L_->call_extra<void(instance_t *, const real_t *)>(f, in1);
        else if (f == "simulate_dynamics")
            L_->call_extra<void(const instance_t *, const real_t *, const real_t *, real_t *)>(f, in1, in2, out);
        else throw std::invalid_argument("unknown extra function " + f);
        return 0;
    } catch (std::exception &e) { g_err = e.what(); return -1; }
}

} // extern "C"
