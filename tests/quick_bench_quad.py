"""Quick cold-start throughput of the quadcopter problems (SURVEY.md §8d.3 distribution)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import pantr_cdc2023_experiments_b200 as pb
name = sys.argv[1] if len(sys.argv) > 1 else "quadcopter"
N = int(sys.argv[2]) if len(sys.argv) > 2 else 60
B = int(sys.argv[3]) if len(sys.argv) > 3 else 4096
solver, _ = pb.make_solver(pb.benchmark_options(name))
prob = pb.BatchedProblem(name, N)
rng = np.random.default_rng(0)
x0 = np.zeros((B, 9))
x0[:, :3] = np.array([-0.2, -0.25, 0.5]) + rng.uniform(-0.1, 0.1, (B, 3))
x0[:, 3:6] = rng.uniform(-0.5, 0.5, (B, 3)); x0[:, 6:9] = rng.uniform(-0.2, 0.2, (B, 3))
for rep in range(2):
    u = np.tile(prob.initial_guess(), (B, 1)); y = np.zeros((B, prob.m))
    t = time.time(); st = solver(prob, u, y, x0=x0); t = time.time() - t
    print(f"{name} cold N={N} B={B}: {t:.3f}s  {B/t:.1f} solves/s  converged {np.mean(st['status']==1):.3f} "
          f"outer mean {st['outer_iterations'].mean():.1f} inner mean {st['inner_iterations'].mean():.1f} max {st['inner_iterations'].max()} "
          f"hvp mean {st['ev_hess_psi_prod'].mean():.1f} grid {solver.grid_size(prob)} statuses {np.unique(st['status'], return_counts=True)}")
