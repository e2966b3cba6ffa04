"""Quick cold-start throughput of the quadcopter problems.
usage: quick_bench_quad.py <name> <N> <B> <auto|warp|thread> <scale>
x0 = nominal + scale * (SURVEY.md §8d.3 perturbation: p +-0.1, v +-0.5, theta +-0.2)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import pantr_cdc2023_experiments_b200 as pb
name = sys.argv[1] if len(sys.argv) > 1 else "quadcopter"
N = int(sys.argv[2]) if len(sys.argv) > 2 else 60
B = int(sys.argv[3]) if len(sys.argv) > 3 else 4096
mode = sys.argv[4] if len(sys.argv) > 4 else "auto"
scale = float(sys.argv[5]) if len(sys.argv) > 5 else 0.2
extra = sys.argv[6:]  # option overrides, e.g. alm.max_iter=10 solver.max_iter=100 (bounds the tail)
solver, _ = pb.make_solver(pb.benchmark_options(name) + extra)
prob = pb.BatchedProblem(name, N)
solver.set_exec_mode(prob, mode)
rng = np.random.default_rng(0)
x0 = np.tile(prob.initial_state(), (B, 1))
x0[:, :3] += scale * rng.uniform(-0.1, 0.1, (B, 3))
x0[:, 3:6] += scale * rng.uniform(-0.5, 0.5, (B, 3)); x0[:, 6:9] += scale * rng.uniform(-0.2, 0.2, (B, 3))
for rep in range(2):
    u = np.tile(prob.initial_guess(), (B, 1)); y = np.zeros((B, prob.m))
    t = time.time(); st = solver(prob, u, y, x0=x0); t = time.time() - t
    it = st['inner_iterations']
    print(f"{name} {mode} scale {scale} cold N={N} B={B}: {t:.3f}s  {B/t:.1f} solves/s  converged {np.mean(st['status']==1):.3f} "
          f"outer mean {st['outer_iterations'].mean():.1f} inner mean {it.mean():.1f} p50 {np.median(it):.0f} p99 {np.percentile(it, 99):.0f} max {it.max()} "
          f"hvp mean {st['ev_hess_psi_prod'].mean():.1f} -> {st['ev_hess_psi_prod'].sum()/t/1e6:.2f} M hvp/s ; "
          f"per-slot {st['ev_hess_psi_prod'].sum()/(st['cycles'].sum()/1.965e9):.0f} hvp/s while busy", flush=True)
