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


def hanging_chain(N):
    f = 1915 * N + 124
    adj = 1904 * N + 33
    return {"f": f, "psi": f, "psi_grad_psi": f + adj, "grad_psi": f + adj, "hess_psi_prod": 8829 * N + 63}


MODELS = {"hanging_chain": hanging_chain}


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
