"""Results writer in the reference's format.

  write_results            drivers/results.hpp:227-251   (a Python module defining `results = {...}`)
  write_evaluations        drivers/results.hpp:63-87     (19 EvalCounter keys, in this order)
  python_literal           drivers/output.hpp:14-55
  float_to_str / print_python  alpaqa/src/alpaqa/include/alpaqa/implementation/util/print.tpp:21-33,128-146

The files are consumed unchanged by new-benchmarks-paper/benchmark_util.py:4-9 (`load_results`).
"""
import math
import secrets
import time

import numpy as np

# drivers/results.hpp:66-84
EVAL_KEYS = ["proj_diff_g", "proj_multipliers", "prox_grad_step", "f", "grad_f", "f_grad_f", "f_g",
             "grad_f_grad_g_prod", "g", "grad_g_prod", "grad_gi", "grad_L", "hess_L_prod", "hess_L",
             "hess_ψ_prod", "hess_ψ", "ψ", "grad_ψ", "ψ_grad_ψ"]
# device counter -> EvalCounter key
STAT_TO_EVAL = {"ev_proj_multipliers": "proj_multipliers", "ev_prox_grad_step": "prox_grad_step",
                "ev_f": "f", "ev_g": "g", "ev_grad_L": "grad_L", "ev_hess_psi_prod": "hess_ψ_prod",
                "ev_psi": "ψ", "ev_grad_psi": "grad_ψ", "ev_psi_grad_psi": "ψ_grad_ψ"}


def float_to_str(v: float) -> str:
    """print.tpp:21-33: explicit '+' unless the sign bit is set, scientific, 17 significant digits."""
    v = float(v)
    if math.isnan(v):
        return "-nan" if math.copysign(1.0, v) < 0 else "+nan"
    if math.isinf(v):
        return "+inf" if v > 0 else "-inf"
    return f"{v:+.17e}"


def _quoted(s: str) -> str:  # std::quoted
    return '"' + str(s).replace("\\", "\\\\").replace('"', '\\"') + '"'


def python_literal(v) -> str:
    if isinstance(v, (bool, np.bool_)):
        return "True" if v else "False"
    if isinstance(v, (int, np.integer)):
        return str(int(v))
    if isinstance(v, (float, np.floating)):
        return float_to_str(v)
    if isinstance(v, str):
        return _quoted(v)
    if isinstance(v, (list, tuple)) and all(isinstance(e, str) for e in v):
        return "[" + "".join(_quoted(e) + ", " for e in v) + "]"          # output.hpp:49-54
    a = np.asarray(v, dtype=np.float64)
    if a.ndim == 1:                                                           # print.tpp:130-131
        return "np.array([" + ", ".join(float_to_str(e) for e in a) + "])"
    if a.ndim == 2:                                                           # print.tpp:132-144
        rows = ["".join(float_to_str(e) + ", " for e in r) for r in a]
        return "np.array([[" + "],\n [".join(rows) + "]])"
    raise TypeError(f"no python literal for {type(v)}")


def results_name(opts=()):
    """mpc-driver.cpp:312-318: results_<timestamp_ms>_<8 hex>, overridable by results_name=<name>."""
    ts = int(time.time() * 1000)
    name = f"results_{ts}_{secrets.token_hex(4)}"
    for o in opts:
        if o.startswith("results_name="):
            name = o.split("=", 1)[1]
    return name, ts


def write_results(fh, *, opts, timestamp, runtimes, states, inputs, solver, evaluations, problem, nvar,
                  ncon, path, horizon, second_order, formulation, extra=()):
    """states: nx x (num_sim+1), inputs: n x num_sim (column per MPC step), runtimes: num_sim
    (negative = the solve did not converge, mpc-driver.cpp:219-221)."""
    w = fh.write
    w("from numpy import nan, inf\nimport numpy as np\n__all__ = ['results']\nresults = {\n")

    def elem(k, v):
        w("    " + _quoted(k) + ": " + python_literal(v) + ",\n")
    elem("opts", list(opts))
    elem("time_utc_ms", int(timestamp))
    elem("runtimes", np.asarray(runtimes, dtype=np.float64))
    elem("states", np.asarray(states, dtype=np.float64))
    elem("inputs", np.asarray(inputs, dtype=np.float64))
    elem("solver", solver)
    w(_quoted("evaluations") + ": {\n")
    for k in EVAL_KEYS:
        elem(k, int(evaluations.get(k, 0)))
    w("},\n")
    for k, v in extra:
        elem(k, v)
    elem("problem", problem)
    elem("nvar", int(nvar))
    elem("ncon", int(ncon))
    w("    " + _quoted("path") + ": " + _quoted(path) + ",\n")
    elem("horizon", int(horizon))
    elem("second_order", second_order)
    elem("formulation", formulation)
    w("}\n")
