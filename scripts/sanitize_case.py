"""One small workload per invocation, meant to run under compute-sanitizer (scripts/gpu_sanitize.sh).

usage: sanitize_case.py <problem> <N> <B> <auto|warp|thread|sliced> [face]
Runs a batched ALM+PANTR solve (iteration counts capped so that the 10-100x slower instrumented run
stays short) and, with `face`, every batched member of the problem plugin table once.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import pantr_cdc2023_experiments_b200 as pb

name, N, B, mode = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
face = len(sys.argv) > 5 and sys.argv[5] == "face"
solver, _ = pb.make_solver(pb.benchmark_options(name) + ["alm.max_iter=4", "solver.max_iter=12"])
prob = pb.BatchedProblem(name, N)
solver.set_exec_mode(prob, mode)
rng = np.random.default_rng(7)
x0 = np.tile(prob.initial_state(), (B, 1))
if name == "hanging_chain":
    for _ in range(5):
        x0 = prob.simulate_dynamics(x0, rng.uniform(-1, 1, (B, 3)))
else:
    x0 += 0.02 * rng.standard_normal((B, prob.nx))
u = np.tile(prob.initial_guess(), (B, 1))
y = np.zeros((B, prob.m)) if prob.m else None
st = solver(prob, u, y, x0=x0)
print(name, N, B, mode, "status counts", np.bincount(st["status"], minlength=8).tolist(),
      "inner iterations", int(st["inner_iterations"].sum()))
if face:
    S = np.full((B, prob.m), 1e3)
    yy = rng.standard_normal((B, prob.m))
    v = rng.standard_normal((B, prob.n))
    prob.eval_f(u, x0=x0); prob.eval_grad_f(u, x0=x0); prob.eval_f_grad_f(u, x0=x0); prob.eval_g(u, x0=x0)
    prob.eval_grad_g_prod(u, yy, x0=x0); prob.eval_grad_L(u, yy, x0=x0); prob.eval_hess_L_prod(u, yy, 1.0, v, x0=x0)
    prob.eval_hess_psi_prod(u, yy, S, 1.0, v, x0=x0); prob.eval_psi(u, yy, S, x0=x0); prob.eval_grad_psi(u, yy, S, x0=x0)
    prob.eval_psi_grad_psi(u, yy, S, x0=x0)
    g = rng.standard_normal((B, prob.n))
    prob.eval_prox_grad_step(0.3, u, g); prob.eval_inactive_indices_res_lna(0.3, u, g)
    if prob.m:
        prob.eval_proj_diff_g(yy); prob.eval_proj_multipliers(yy, 1e9)
    print("problem face ok")
