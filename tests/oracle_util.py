"""ctypes access to the CPU oracle (oracle/liboracle.so) for the tests.  Test infrastructure."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
EPS = 2.220446049250313e-16
NSTATS = 24
STAT_KEYS = ["status", "outer_iterations", "inner_iterations", "eps", "delta", "norm_penalty",
             "final_psi", "final_gamma", "final_fbe", "accelerated_step_rejected",
             "direction_failures", "inner_convergence_failures", "ev_psi", "ev_psi_grad_psi",
             "ev_grad_L", "ev_hess_psi_prod", "ev_prox_grad_step", "ev_grad_psi",
             "ev_inactive_indices_res_lna", "ev_proj_multipliers"]
_lib = None


def load():
    global _lib
    if _lib is None:
        so = os.path.join(ORACLE_DIR, "liboracle.so")
        if not os.path.exists(so):
            subprocess.check_call(["make", "-C", ORACLE_DIR, "liboracle.so"])
        _lib = C.CDLL(so)
    return _lib


def P(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def pack_params(solver):
    """Flatten the package's param dataclasses into the oracle's params[48] layout."""
    from pantr_cdc2023_experiments_b200.params import PANOCStopCrit
    a, p = solver.params, solver.inner_solver.params
    d, c = solver.inner_solver.direction.direction, solver.inner_solver.direction.accelerator
    par = np.zeros(48)
    par[0:13] = [a.tolerance, a.dual_tolerance, a.penalty_update_factor, a.initial_penalty,
                 a.initial_tolerance, a.tolerance_update_factor, a.rel_penalty_increase_threshold,
                 a.max_multiplier, a.max_penalty, a.min_penalty, a.max_iter, a.initial_penalty_factor,
                 a.single_penalty_factor]
    par[16:37] = [p.max_iter, PANOCStopCrit[p.stop_crit], p.L_min, p.L_max,
                  p.quadratic_upperbound_tolerance_factor, p.TR_tolerance_factor,
                  p.ratio_threshold_acceptable, p.ratio_threshold_good, p.radius_factor_rejected,
                  p.radius_factor_acceptable, p.radius_factor_good, p.initial_radius, p.min_radius,
                  p.Lipschitz.L_0, p.Lipschitz.ε, p.Lipschitz.δ, p.Lipschitz.Lγ_factor,
                  p.compute_ratio_using_new_stepsize, p.recompute_last_prox_step_after_direction_reset,
                  p.disable_acceleration, p.ratio_approx_fbe_quadratic_model]
    par[13:16] = [a.max_num_initial_retries, a.max_num_retries, a.max_total_num_retries]
    par[37:40] = [a.penalty_update_factor_lower, a.min_penalty_update_factor, a.initial_penalty_lower]
    par[41:44] = [a.initial_tolerance_increase, a.ρ_increase, a.ρ_max]
    par[40] = d.hessian_vec_factor
    par[44:48] = [c.tol_scale, c.tol_scale_root, c.tol_max, c.max_iter_factor]
    return par


def dims(pid, N):
    d = (C.c_long * 4)()
    assert load().oracle_dims(pid, N, d) == 0
    return tuple(d)


def eval_(pid, N, x0, u, y=None, Sigma=None, v=None, want_yhat=False):
    n, m, nx, nu = dims(pid, N)
    psi = C.c_double()
    grad = np.zeros(n); Hv = np.zeros(n) if v is not None else None
    yhat = np.zeros(m) if want_yhat else None
    assert load().oracle_eval(pid, N, P(x0), P(u), P(y), P(Sigma), P(v), C.byref(psi), P(grad), P(yhat), P(Hv)) == 0
    return psi.value, grad, yhat, Hv


def simulate(pid, x, u):
    xn = np.zeros_like(x)
    assert load().oracle_simulate_dynamics(pid, P(np.ascontiguousarray(x)), P(np.ascontiguousarray(u)), P(xn)) == 0
    return xn


def solve(pid, N, par, x0, u, y=None, trace_rows=0):
    n, m, nx, nu = dims(pid, N)
    u = np.array(u, dtype=np.float64); y = np.zeros(m) if y is None else np.array(y, dtype=np.float64)
    st = np.zeros(NSTATS)
    tr = np.zeros((max(trace_rows, 1), 11)); nt = C.c_int(0)
    assert load().oracle_solve(pid, N, P(par), P(np.ascontiguousarray(x0)), P(u), P(y), P(st),
                               P(tr) if trace_rows else None, trace_rows, C.byref(nt)) == 0
    stats = dict(zip(STAT_KEYS, st))
    return u, y, stats, tr[:min(nt.value, trace_rows)]


def solve_batch(pid, N, par, x0, u, y=None, nthreads=1):
    n, m, nx, nu = dims(pid, N)
    B = x0.shape[0]
    u = np.array(u, dtype=np.float64); y = np.zeros((B, m)) if y is None else np.array(y, dtype=np.float64)
    st = np.zeros((B, NSTATS))
    assert load().oracle_solve_batch(pid, N, P(par), C.c_long(B), P(np.ascontiguousarray(x0)), P(u), P(y), P(st),
                                     nthreads) == 0
    return u, y, st


FN = {"f": 0, "grad_f": 1, "f_grad_f": 2, "g": 3, "grad_g_prod": 4, "grad_L": 5, "hess_L_prod": 6,
      "hess_psi_prod": 7, "prox_grad_step": 8, "proj_diff_g": 9, "proj_multipliers": 10, "inactive_indices": 11}


def problem_eval(pid, N, fn, x0=None, x=None, y=None, Sigma=None, v=None, scalar=1.0):
    """One member of the problem plugin table for one instance (oracle_problem_eval)."""
    n, m, nx, nu = dims(pid, N)
    c = lambda a: None if a is None else np.ascontiguousarray(a, dtype=np.float64)
    x0, x, y, Sigma, v = c(x0), c(x), c(y), c(Sigma), c(v)
    s = C.c_double(0.0); cnt = C.c_int(0)
    on, on2, om = np.zeros(n), np.zeros(n), np.zeros(max(m, 1))
    oi = np.zeros(n, dtype=np.int32)
    rc = load().oracle_problem_eval(pid, N, FN[fn], P(x0), P(x), P(y), P(Sigma), P(v), C.c_double(scalar),
                                    C.byref(s), P(on), P(on2), P(om), P(oi), C.byref(cnt))
    assert rc == 0
    return {"s": s.value, "n": on, "n2": on2, "m": om[:m], "J": oi[:cnt.value].copy()}
