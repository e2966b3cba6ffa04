"""Algorithmic FP64 operation counts of the problem functions (roofline numerator).

Convention: add = sub = mul = div = sqrt = 1 flop, counted on the scalar algorithm
(30 position lanes / 27 velocity lanes / 10 links; per-link scalars counted once, not per lane),
i.e. the operations of oracle/hanging_chain.hpp.  O(n) solver vector work (prox steps, dots, CG
updates: < 2 % of a solve) is NOT counted, so the reported FLOP/s is a lower bound.

hanging chain (oracle/hanging_chain.hpp, oracle/hanging_chain_sweeps.inc):
  fc      27 (dist) + 30 (sq) + 10*10 (n2, sqrt, 1/w, c, iw2, e) + 30 (F) + 81 (kV)          = 268
  rk4     4 fc + 3*114 (stage points) + 57*7 (combination)                                  = 1813
  f       N*(102 stage cost + rk4) + 93 terminal + 31 reduction                             = 1915 N + 124
  vjp     60 (Fbar) + 30 + 20 (dot3) + 120 (db) + 27                                         = 257
  adjoint N*(4 vjp + 57 + 3*171 (w) + 228 + 9 (sums) + 69 (stage cost)) + 33                 = 1904 N + 33
  jvp     27 + 30 + 20 + 120 + 54                                                            = 251
  hvjp    60 + 60 + 40 + 30*18 + 27                                                          = 727
  hvp     forward: the nonlinear stage evaluations 1414 + tangent 4 jvp + 342 + 399 = 3159 N ;
          reverse: 1140 (w, dw) + 4 vjp + 4 hvjp + 465 (sums) + 129 (stage cost) = 5670 N ; + 63  = 8829 N + 63
These are the operations of ONE from-scratch evaluation of each function with nothing computed
twice inside the call and nothing reused across calls -- what the reference's CasADi-generated
f / psi_grad_psi / hess_psi_prod do per call.  The kernels execute more in some places (the reverse
sweeps rebuild the stage points from the cached 1/|dist|: +1414 N, and the tangent stage data: +1346 N)
and less in others (the products reuse the adjoint data of the last gradient: -4 vjp - 570 N); neither
changes the algorithmic count.  (Until r01_v3 the count included the in-call recomputation:
5233 N + 157 and 11589 N + 63.)
"""


_GEN = None


def _gen_counts():
    """Exact operation counts of the sympy-generated model functions (tools/gen_models.py writes them)."""
    global _GEN
    if _GEN is None:
        import json
        import os
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc", "generated", "flop_counts.json")) as fh:
            _GEN = json.load(fh)
    return _GEN


def _quadcopter(N, ns):
    """The two quadcopter models: one from-scratch evaluation of each function in the REFERENCE formulation
    (rollout, discrete adjoint, forward-over-adjoint: what CasADi's AD generates and oracle/quadcopter.hpp restates),
    built from the exact counts of the generated f_c / c functions (generated/flop_counts.json) plus the hand-written
    RK4 / cost / augmented-Lagrangian arithmetic around them; sin and cos count 1 flop each.
      rk4 step        4 (f_c + 6 sincos) + 3*18 (stage points) + 9*7 (combination)
      stage cost 40, terminal cost 24 (incl. the accumulation)
      AL terms        per constraint: div, add, fmin, fmax, sub, 2 mul, add = 8  -> 32 per node, + c(x), + 6 sincos at node N
      adjoint step    4 vjp + c_vjp + 182 (stage weights 9 + 3*27, sums 9 + 3*13, gradient 8, node 36)
      tangent step    4 jvp + 117
      2nd-order step  4 vjp + 4 hvjp + c_hvjp + c_jvp + c_vjp + 238
    The device runs the CONDENSED formulation of the same derivatives (csrc/quadcopter_condensed.cuh), which executes
    more operations (13 unit tangents per step for the step Jacobians) in exchange for lane parallelism; the roofline
    numerator stays the algorithmic count."""
    g = _gen_counts()[ns]
    fc, jvp, vjp, hvjp = g["fc"], g["fc_jvp"], g["fc_vjp"], g["fc_hvjp"]
    con, cj, cv, ch = g["con"], g["con_jvp"], g["con_vjp"], g["con_hvjp"]
    rk4 = 4 * (fc + 6) + 3 * 18 + 9 * 7
    f = N * (rk4 + 40) + 24
    al = (N + 1) * (con + 32) + 6 + 2
    adj = (cv + 27) + N * (4 * vjp + cv + 182)
    node2 = ch + cj + cv + 31
    hvp = (N * rk4 + al + N * (4 * vjp + 90)          # stage points, multipliers, first-order adjoint stage weights
           + N * (4 * jvp + 117) + node2 + N * (4 * vjp + 4 * hvjp + node2 + 207))
    return {"f": f, "psi": f + al, "psi_grad_psi": f + al + adj, "grad_psi": f + al + adj, "hess_psi_prod": hvp}


def quadcopter(N):
    return _quadcopter(N, "quadcopter_gen")


def realistic_quadcopter(N):
    return _quadcopter(N, "realistic_quadcopter_gen")


def hanging_chain(N):
    f = 1915 * N + 124
    adj = 1904 * N + 33
    return {"f": f, "psi": f, "psi_grad_psi": f + adj, "grad_psi": f + adj, "hess_psi_prod": 8829 * N + 63}


MODELS = {"hanging_chain": hanging_chain, "quadcopter": quadcopter, "realistic_quadcopter": realistic_quadcopter}


def solve_flops(problem_name, N, stats):
    """Total algorithmic flops of the solves described by the stats array (per-instance counters:
    ProblemWithCounters, alpaqa/problem/problem-with-counters.hpp:24-112).  grad_L is counted by
    the reference but its result is dead (pantr.tpp:306-314) and is not executed on the device."""
    F = MODELS[problem_name](N)
    tot = 0.0
    for key, stat in (("psi", "ev_psi"), ("psi_grad_psi", "ev_psi_grad_psi"), ("grad_psi", "ev_grad_psi"),
                      ("hess_psi_prod", "ev_hess_psi_prod"), ("f", "ev_f")):
        tot += float(stats[stat].astype("f8").sum()) * F[key]
    return tot
