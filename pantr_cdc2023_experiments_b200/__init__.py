"""pantr_b200 — B200-native batched PANTR+ALM solver behind alpaqa's solver/problem API surface.

The compute path is libpantr_b200.so (CUDA, sm_100a, FP64); see include/pantr_b200.h.
"""
from .params import (ALMParams, LipschitzEstimateParams, NewtonTRDirectionParams, PANOCStopCrit,
                     PANTRParams, SolverStatus, SteihaugCGParams, set_params)
from .solver import (ALMSolver, BatchedProblem, NewtonTRDirection, PANTRSolver, STATS_DTYPE,
                     status_name)
from .configs import benchmark_options, make_solver

__all__ = [
    "ALMParams", "LipschitzEstimateParams", "NewtonTRDirectionParams", "PANOCStopCrit", "PANTRParams",
    "SolverStatus", "SteihaugCGParams", "set_params", "ALMSolver", "BatchedProblem",
    "NewtonTRDirection", "PANTRSolver", "STATS_DTYPE", "status_name", "benchmark_options", "make_solver",
]
