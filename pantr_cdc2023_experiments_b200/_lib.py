"""ctypes binding of the C ABI in include/pantr_b200.h (libpantr_b200.so, built in-tree).

There is no CPU fallback: if the shared library is missing this module raises, and if no CUDA
device is present `pantr_b200_create` returns PANTR_B200_NO_DEVICE which is raised as an error.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PANTR_B200_LIB", os.path.join(_HERE, "libpantr_b200.so"))  # override: experiments only


class ALMParamsC(C.Structure):
    _fields_ = [(k, C.c_double) for k in (
        "tolerance", "dual_tolerance", "penalty_update_factor", "penalty_update_factor_lower",
        "min_penalty_update_factor", "initial_penalty", "initial_penalty_factor",
        "initial_penalty_lower", "initial_tolerance", "initial_tolerance_increase",
        "tolerance_update_factor", "rho_increase", "rho_max", "rel_penalty_increase_threshold",
        "max_multiplier", "max_penalty", "min_penalty")] + [(k, C.c_uint32) for k in (
        "max_iter", "max_num_initial_retries", "max_num_retries", "max_total_num_retries")] + [
        ("single_penalty_factor", C.c_int32), ("max_time", C.c_double)]


class LipschitzParamsC(C.Structure):
    _fields_ = [(k, C.c_double) for k in ("L_0", "epsilon", "delta", "Lgamma_factor")]


class PANTRParamsC(C.Structure):
    _fields_ = [("Lipschitz", LipschitzParamsC), ("max_iter", C.c_uint32), ("stop_crit", C.c_int32),
                ("max_no_progress", C.c_uint32)] + [(k, C.c_double) for k in (
        "L_min", "L_max", "quadratic_upperbound_tolerance_factor", "TR_tolerance_factor",
        "ratio_threshold_acceptable", "ratio_threshold_good", "radius_factor_rejected",
        "radius_factor_acceptable", "radius_factor_good", "initial_radius", "min_radius")] + [
        (k, C.c_int32) for k in (
        "compute_ratio_using_new_stepsize", "update_direction_on_prox_step",
        "recompute_last_prox_step_after_direction_reset", "disable_acceleration",
        "ratio_approx_fbe_quadratic_model")] + [("max_time", C.c_double)]


class SteihaugParamsC(C.Structure):
    _fields_ = [(k, C.c_double) for k in ("tol_scale", "tol_scale_root", "tol_max", "max_iter_factor")]


class NewtonTRParamsC(C.Structure):
    _fields_ = [("rescale_on_step_size_changes", C.c_int32), ("hessian_vec_factor", C.c_double),
                ("finite_diff", C.c_int32), ("finite_diff_stepsize", C.c_double)]


class StatsC(C.Structure):
    _fields_ = [("status", C.c_int32)] + [(k, C.c_uint32) for k in (
        "outer_iterations", "inner_iterations", "inner_convergence_failures",
        "accelerated_step_rejected", "direction_failures", "stepsize_backtracks",
        "direction_update_rejected")] + [(k, C.c_double) for k in (
        "eps", "delta", "norm_penalty", "final_gamma", "final_psi", "final_h", "final_fbe")] + [
        (k, C.c_uint32) for k in (
        "ev_proj_multipliers", "ev_prox_grad_step", "ev_inactive_indices_res_lna", "ev_grad_L",
        "ev_hess_psi_prod", "ev_psi", "ev_grad_psi", "ev_psi_grad_psi", "ev_f", "ev_g",
        "cg_iterations", "elapsed_us")] + [("cycles", C.c_uint64)]


# every symbol include/pantr_b200.h declares
SYMBOLS = [
    "pantr_b200_default_params", "pantr_b200_create", "pantr_b200_destroy", "pantr_b200_get_dims",
    "pantr_b200_get_name", "pantr_b200_get_box_C", "pantr_b200_get_box_D", "pantr_b200_get_initial",
    "pantr_b200_solve_batch", "pantr_b200_solve_batch_device", "pantr_b200_eval_batch",
    "pantr_b200_simulate_dynamics_batch", "pantr_b200_simulate_dynamics_batch_device",
    "pantr_b200_shift_warm_start_device", "pantr_b200_set_trace", "pantr_b200_get_trace",
    "pantr_b200_kernel_launches", "pantr_b200_grid_size", "pantr_b200_measure_fp64_peak",
    "pantr_b200_last_error", "pantr_b200_set_exec_mode", "pantr_b200_set_order_device", "pantr_b200_set_debug",
    "pantr_b200_check_guards",
    # problem plugin face (alpaqa_problem_functions_t members, batched)
    "pantr_b200_eval_f_batch", "pantr_b200_eval_grad_f_batch", "pantr_b200_eval_f_grad_f_batch",
    "pantr_b200_eval_g_batch", "pantr_b200_eval_grad_g_prod_batch", "pantr_b200_eval_grad_L_batch",
    "pantr_b200_eval_hess_L_prod_batch", "pantr_b200_eval_hess_psi_prod_batch", "pantr_b200_eval_psi_batch",
    "pantr_b200_eval_grad_psi_batch", "pantr_b200_eval_psi_grad_psi_batch",
    "pantr_b200_eval_prox_grad_step_batch", "pantr_b200_eval_proj_diff_g_batch",
    "pantr_b200_eval_proj_multipliers_batch", "pantr_b200_eval_inactive_indices_res_lna_batch",
    # allocation-free variant (no counterpart in the reference)
    "pantr_b200_eval_psi_grad_psi_noalloc_device",
]

_lib = None


def load():
    """Load libpantr_b200.so (raises if it has not been built: no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'`"
            " (pantr_b200 has no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    vp, dp, i64 = C.c_void_p, C.POINTER(C.c_double), C.c_int64
    lib.pantr_b200_last_error.restype = C.c_char_p
    lib.pantr_b200_get_name.restype = C.c_char_p
    lib.pantr_b200_get_name.argtypes = [vp]
    lib.pantr_b200_default_params.restype = None
    lib.pantr_b200_default_params.argtypes = [C.POINTER(ALMParamsC), C.POINTER(PANTRParamsC),
                                              C.POINTER(SteihaugParamsC), C.POINTER(NewtonTRParamsC)]
    lib.pantr_b200_create.argtypes = [C.c_int, C.c_int, C.POINTER(ALMParamsC), C.POINTER(PANTRParamsC),
                                      C.POINTER(SteihaugParamsC), C.POINTER(NewtonTRParamsC), C.c_int,
                                      C.POINTER(vp)]
    lib.pantr_b200_destroy.argtypes = [vp]
    lib.pantr_b200_get_dims.argtypes = [vp, C.POINTER(i64)]
    lib.pantr_b200_get_box_C.argtypes = [vp, vp, vp]
    lib.pantr_b200_get_box_D.argtypes = [vp, vp, vp]
    lib.pantr_b200_get_initial.argtypes = [vp, vp, vp]
    lib.pantr_b200_solve_batch.argtypes = [vp, i64, vp, vp, vp, vp]
    lib.pantr_b200_solve_batch_device.argtypes = [vp, i64, vp, vp, vp, vp, vp]
    lib.pantr_b200_eval_batch.argtypes = [vp, i64] + [vp] * 9
    lib.pantr_b200_simulate_dynamics_batch.argtypes = [vp, i64, vp, vp, vp]
    lib.pantr_b200_simulate_dynamics_batch_device.argtypes = [vp, i64, vp, vp, i64, vp, vp]
    lib.pantr_b200_shift_warm_start_device.argtypes = [vp, i64, vp, vp, vp]
    lib.pantr_b200_set_trace.argtypes = [vp, i64, C.c_int]
    lib.pantr_b200_get_trace.argtypes = [vp, vp, C.c_int, C.POINTER(C.c_int)]
    lib.pantr_b200_kernel_launches.argtypes = [vp]
    lib.pantr_b200_kernel_launches.restype = i64
    lib.pantr_b200_grid_size.argtypes = [vp]
    lib.pantr_b200_set_exec_mode.argtypes = [vp, C.c_int]
    lib.pantr_b200_set_debug.argtypes = [vp, C.c_int, C.c_int]
    lib.pantr_b200_check_guards.argtypes = [vp, C.POINTER(i64)]
    lib.pantr_b200_set_order_device.argtypes = [vp, vp, i64]
    lib.pantr_b200_measure_fp64_peak.argtypes = [C.c_int, C.c_int, dp]
    f64 = C.c_double
    lib.pantr_b200_eval_f_batch.argtypes = [vp, i64, vp, vp, vp]
    lib.pantr_b200_eval_grad_f_batch.argtypes = [vp, i64, vp, vp, vp]
    lib.pantr_b200_eval_f_grad_f_batch.argtypes = [vp, i64, vp, vp, vp, vp]
    lib.pantr_b200_eval_g_batch.argtypes = [vp, i64, vp, vp, vp]
    lib.pantr_b200_eval_grad_g_prod_batch.argtypes = [vp, i64, vp, vp, vp, vp]
    lib.pantr_b200_eval_grad_L_batch.argtypes = [vp, i64, vp, vp, vp, vp]
    lib.pantr_b200_eval_hess_L_prod_batch.argtypes = [vp, i64, vp, vp, vp, f64, vp, vp]
    lib.pantr_b200_eval_hess_psi_prod_batch.argtypes = [vp, i64, vp, vp, vp, vp, f64, vp, vp]
    lib.pantr_b200_eval_psi_batch.argtypes = [vp, i64, vp, vp, vp, vp, vp, vp]
    lib.pantr_b200_eval_grad_psi_batch.argtypes = [vp, i64, vp, vp, vp, vp, vp]
    lib.pantr_b200_eval_psi_grad_psi_batch.argtypes = [vp, i64, vp, vp, vp, vp, vp, vp]
    lib.pantr_b200_eval_prox_grad_step_batch.argtypes = [vp, i64, f64, vp, vp, vp, vp]
    lib.pantr_b200_eval_proj_diff_g_batch.argtypes = [vp, i64, vp, vp]
    lib.pantr_b200_eval_proj_multipliers_batch.argtypes = [vp, i64, vp, f64]
    lib.pantr_b200_eval_inactive_indices_res_lna_batch.argtypes = [vp, i64, f64, vp, vp, vp, vp]
    lib.pantr_b200_eval_psi_grad_psi_noalloc_device.argtypes = [C.c_int, C.c_int, i64, vp, vp, vp, vp, vp, vp, vp]
    _lib = lib
    return lib


class PantrB200Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"pantr_b200 error {code}: {msg}")
        self.code = code


def check(rc):
    if rc != 0:
        code = rc
        msg = load().pantr_b200_last_error().decode()
        # mirror the reference's exception types (std::invalid_argument / std::out_of_range)
        if code == -1:
            raise ValueError(msg)
        if code == -4:
            raise IndexError(msg)
        raise PantrB200Error(code, msg)
