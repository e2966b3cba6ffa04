"""Option sets the reference benchmarks PANTR with, and solver construction from key=value options.

  option sets      new-benchmarks-paper/dodo.py:12-33,94-105
  driver defaults  drivers/mpc-driver.cpp:47-73,104-115
"""
from .params import (ALMParams, NewtonTRDirectionParams, PANTRParams, SteihaugCGParams, set_params)
from .solver import ALMSolver, NewtonTRDirection, PANTRSolver

# dodo.py:12-33 (common + PANTR-specific options)
_COMMON = [
    "alm.max_iter=300", "alm.max_time=30s", "alm.initial_penalty=1e4", "alm.penalty_update_factor=5",
    "alm.initial_tolerance=1e2", "alm.tolerance_update_factor=0.1",
    "solver.stop_crit=ProjGradUnitNorm", "alm.tolerance=1e-8", "alm.dual_tolerance=1e-8",
]
_PANTR = [
    "solver.radius_factor_rejected=0.35", "solver.radius_factor_acceptable=0.99",
    "solver.radius_factor_good=10", "solver.ratio_threshold_acceptable=0.3",
    "solver.ratio_threshold_good=0.5", "dir.hessian_vec_factor=1", "accel.tol_scale=1e-2",
]
# dodo.py:94-105
_PER_PROBLEM = {
    "hanging_chain": ["num_dist=5", "u_dist=-1,1,1"],
    "quadcopter": ["num_dist=0", "solver.max_iter=250"],
    "realistic_quadcopter": ["num_dist=0", "solver.max_iter=250", "accel.tol_scale=1",
                             "alm.initial_tolerance=1e-1"],
}


def benchmark_options(problem_name: str):
    """The key=value option list `dodo.py` passes to `benchmark-mpc-driver` for method=pantr."""
    return ["method=pantr"] + _COMMON + _PANTR + _PER_PROBLEM[problem_name]


def make_solver(opts):
    """drivers/mpc-driver.cpp:47-73,104-115: driver defaults, then alm./solver./dir./accel. keys."""
    used = [False] * len(opts)
    alm = ALMParams(max_iter=200, tolerance=1e-8, dual_tolerance=1e-8)          # mpc-driver.cpp:108-113
    solver = PANTRParams(max_iter=50000, stop_crit="FPRNorm")                   # mpc-driver.cpp:50-54
    accel, direction = SteihaugCGParams(), NewtonTRDirectionParams()
    set_params(alm, "alm", opts, used)
    set_params(solver, "solver", opts, used)
    set_params(direction, "dir", opts, used)
    set_params(accel, "accel", opts, used)
    return ALMSolver(alm, PANTRSolver(solver, NewtonTRDirection(accel, direction))), used
