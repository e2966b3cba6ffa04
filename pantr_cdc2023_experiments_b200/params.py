"""Parameter structs and the key=value option system of the reference, host side.

Mirrors (same field names, same defaults, same option keys):
  ALMParams                 alpaqa/src/alpaqa/include/alpaqa/outer/alm.hpp:19-90
  PANTRParams               alpaqa/src/alpaqa/include/alpaqa/inner/pantr.hpp:22-91
  LipschitzEstimateParams   alpaqa/src/alpaqa/include/alpaqa/inner/internal/lipschitz.hpp:8-20
  SteihaugCGParams          alpaqa/src/alpaqa/include/alpaqa/accelerators/steihaugcg.hpp:9-15
  NewtonTRDirectionParams   alpaqa/src/alpaqa/include/alpaqa/inner/directions/pantr/newton-tr.hpp:16-23
  set_params / key parsing  alpaqa/src/alpaqa/src/params/params.cpp:68-76,113-142,157-340
"""
from dataclasses import dataclass, field, fields, is_dataclass
import math
import re

from . import _lib

EPS = 2.220446049250313e-16

# alpaqa/inner/internal/panoc-stop-crit.hpp:7-102
PANOCStopCrit = {name: i for i, name in enumerate((
    "ApproxKKT", "ApproxKKT2", "ProjGradNorm", "ProjGradNorm2", "ProjGradUnitNorm",
    "ProjGradUnitNorm2", "FPRNorm", "FPRNorm2", "Ipopt", "LBFGSBpp"))}
# alpaqa/inner/internal/solverstatus.hpp:11-20
SolverStatus = {name: i for i, name in enumerate((
    "Busy", "Converged", "MaxTime", "MaxIter", "NotFinite", "NoProgress", "Interrupted",
    "Exception"))}
SolverStatusName = {v: k for k, v in SolverStatus.items()}


@dataclass
class LipschitzEstimateParams:
    L_0: float = 0.0
    ε: float = 1e-6
    δ: float = 1e-12
    Lγ_factor: float = 0.95


@dataclass
class PANTRParams:
    Lipschitz: LipschitzEstimateParams = field(default_factory=LipschitzEstimateParams)
    max_iter: int = 100
    max_time: float = 300.0  # seconds (pantr.hpp:27: 5 min); enforced on the device clock
    L_min: float = 1e-5
    L_max: float = 1e20
    stop_crit: str = "ApproxKKT"
    max_no_progress: int = 10
    print_interval: int = 0
    print_precision: int = 8
    quadratic_upperbound_tolerance_factor: float = 10 * EPS
    TR_tolerance_factor: float = 10 * EPS
    ratio_threshold_acceptable: float = 0.2
    ratio_threshold_good: float = 0.8
    radius_factor_rejected: float = 0.35
    radius_factor_acceptable: float = 0.999
    radius_factor_good: float = 2.5
    initial_radius: float = math.nan
    min_radius: float = 100 * EPS
    compute_ratio_using_new_stepsize: bool = False
    update_direction_on_prox_step: bool = True
    recompute_last_prox_step_after_direction_reset: bool = False
    disable_acceleration: bool = False
    ratio_approx_fbe_quadratic_model: bool = True

    def to_c(self):
        c = _lib.PANTRParamsC()
        c.Lipschitz = _lib.LipschitzParamsC(self.Lipschitz.L_0, self.Lipschitz.ε, self.Lipschitz.δ,
                                            self.Lipschitz.Lγ_factor)
        if self.stop_crit not in PANOCStopCrit:
            raise IndexError(f"Invalid PANOCStopCrit '{self.stop_crit}'")
        c.stop_crit = PANOCStopCrit[self.stop_crit]
        for f in ("max_iter", "max_no_progress", "L_min", "L_max",
                  "quadratic_upperbound_tolerance_factor", "TR_tolerance_factor",
                  "ratio_threshold_acceptable", "ratio_threshold_good", "radius_factor_rejected",
                  "radius_factor_acceptable", "radius_factor_good", "initial_radius", "min_radius",
                  "compute_ratio_using_new_stepsize", "update_direction_on_prox_step",
                  "recompute_last_prox_step_after_direction_reset", "disable_acceleration",
                  "ratio_approx_fbe_quadratic_model", "max_time"):
            setattr(c, f, getattr(self, f))
        return c


@dataclass
class ALMParams:
    tolerance: float = 1e-5
    dual_tolerance: float = 1e-5
    penalty_update_factor: float = 10.0
    penalty_update_factor_lower: float = 0.8
    min_penalty_update_factor: float = 1.1
    initial_penalty: float = 1.0
    initial_penalty_factor: float = 20.0
    initial_penalty_lower: float = 0.6
    initial_tolerance: float = 1.0
    initial_tolerance_increase: float = 1.1
    tolerance_update_factor: float = 1e-1
    ρ_increase: float = 2.0
    ρ_max: float = 0.5
    rel_penalty_increase_threshold: float = 0.1
    max_multiplier: float = 1e9
    max_penalty: float = 1e9
    min_penalty: float = 1e-9
    max_iter: int = 100
    max_time: float = 300.0  # seconds (alm.hpp:66: 5 min); enforced on the device clock
    max_num_initial_retries: int = 0
    max_num_retries: int = 0
    max_total_num_retries: int = 0
    print_interval: int = 0
    print_precision: int = 8
    single_penalty_factor: bool = False

    def to_c(self):
        c = _lib.ALMParamsC()
        for f, _ in _lib.ALMParamsC._fields_:
            setattr(c, f, getattr(self, {"rho_increase": "ρ_increase", "rho_max": "ρ_max"}.get(f, f)))
        return c


@dataclass
class SteihaugCGParams:
    tol_scale: float = 1.0
    tol_scale_root: float = 0.5
    tol_max: float = math.inf
    max_iter_factor: float = 1.0

    def to_c(self):
        return _lib.SteihaugParamsC(self.tol_scale, self.tol_scale_root, self.tol_max, self.max_iter_factor)


@dataclass
class NewtonTRDirectionParams:
    rescale_on_step_size_changes: bool = False
    hessian_vec_factor: float = 0.5
    finite_diff: bool = False
    finite_diff_stepsize: float = math.sqrt(EPS)

    def to_c(self):
        return _lib.NewtonTRParamsC(self.rescale_on_step_size_changes, self.hessian_vec_factor,
                                    self.finite_diff, self.finite_diff_stepsize)


_DURATION_UNITS = {"": 1.0, "s": 1.0, "ms": 1e-3, "us": 1e-6, "µs": 1e-6, "ns": 1e-9, "min": 60.0}


def _parse_value(current, value: str, key: str):
    """Scalar parsing rules of params.cpp:78-184."""
    if isinstance(current, bool):
        if value in ("0", "false"):
            return False
        if value in ("1", "true"):
            return True
        raise ValueError(f"Invalid value '{value}' for type 'bool' in '{key}'")
    if key.endswith("max_time"):  # std::chrono duration with unit suffix, params.cpp:113-142
        mt = re.fullmatch(r"([-+0-9.eE]+)\s*(\D*)", value)
        if not mt or mt.group(2) not in _DURATION_UNITS:
            raise ValueError(f"Invalid units '{value}' in '{key}'")
        return float(mt.group(1)) * _DURATION_UNITS[mt.group(2)]
    if isinstance(current, int):
        try:
            return int(value)
        except ValueError:
            raise ValueError(f"Invalid value '{value}' for type 'int' in '{key}'")
    if isinstance(current, float):
        try:
            return float(value)
        except ValueError:
            raise ValueError(f"Invalid value '{value}' for type 'double' in '{key}'")
    if isinstance(current, str):  # enum by name, params.cpp:157-184
        if key.endswith("stop_crit") and value not in PANOCStopCrit:
            raise ValueError(f"Invalid value '{value}' for type 'PANOCStopCrit' in '{key}'")
        return value
    raise ValueError(f"Unsupported option type for '{key}'")


def set_params(obj, prefix: str, opts, used=None):
    """`set_params(obj, "alm", opts)` of alpaqa/params/params.hpp:48-64: apply every
    `prefix.key[.subkey]=value` string of `opts` to the dataclass `obj`; unknown keys raise
    (params.tpp:60-80).  `used[i]` is set for consumed options."""
    names = {f.name for f in fields(obj)}
    for i, opt in enumerate(opts):
        keyval = opt.split("=", 1)
        key, value = keyval[0], (keyval[1] if len(keyval) > 1 else "")
        parts = key.split(".")
        if parts[0] != prefix:
            continue
        if used is not None:
            used[i] = True
        target, path = obj, parts[1:]
        while len(path) > 1:
            if not is_dataclass(target) or path[0] not in {f.name for f in fields(target)}:
                raise ValueError(f"Invalid key '{path[0]}' in '{key}'")
            target = getattr(target, path[0])
            path = path[1:]
        if not path or path[0] not in {f.name for f in fields(target)}:
            raise ValueError(f"Invalid key '{path[0] if path else ''}' for type "
                             f"'{type(target).__name__}' in '{key}', possible keys are: "
                             + ", ".join(sorted(names)))
        setattr(target, path[0], _parse_value(getattr(target, path[0]), value, key))
