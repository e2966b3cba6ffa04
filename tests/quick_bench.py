import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import pantr_cdc2023_experiments_b200 as pb
N = int(sys.argv[1]) if len(sys.argv) > 1 else 60
B = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
solver, _ = pb.make_solver(pb.benchmark_options("hanging_chain"))
prob = pb.BatchedProblem("hanging_chain", N)
solver.set_exec_mode(prob, os.environ.get("PANTR_EXEC_MODE", "auto"))
rng = np.random.default_rng(0)
x0 = np.tile(prob.initial_state(), (B, 1))
for _ in range(5):
    x0 = prob.simulate_dynamics(x0, rng.uniform(-1, 1, (B, 3)))
for rep in range(2):
    u = np.zeros((B, prob.n))
    t = time.time(); st = solver(prob, u, None, x0=x0); t = time.time() - t
    print(f"cold N={N} B={B}: {t:.3f}s  {B/t:.1f} solves/s  converged {np.mean(st['status']==1):.3f} "
          f"iters mean {st['inner_iterations'].mean():.1f} max {st['inner_iterations'].max()} hvp mean {st['ev_hess_psi_prod'].mean():.1f} "
          f"grid {solver.grid_size(prob)} cycles mean {st['cycles'].mean():.3e} max {st['cycles'].max():.3e}")
# warm: plant step + shift
xn = prob.simulate_dynamics(x0, u[:, :3])
uw = np.concatenate([u[:, 3:], u[:, -3:]], axis=1).copy()
t = time.time(); st = solver(prob, uw, None, x0=xn); t = time.time() - t
print(f"warm N={N} B={B}: {t:.3f}s  {B/t:.1f} solves/s  converged {np.mean(st['status']==1):.3f} "
      f"iters mean {st['inner_iterations'].mean():.1f} hvp mean {st['ev_hess_psi_prod'].mean():.1f} "
      f"SM clock in kernel {st['cycles'].sum()/max(1,st['elapsed_us'].sum()):.0f} MHz, slot busy "
      f"{st['elapsed_us'].sum()/1e6/(t*min(B, solver.grid_size(prob))):.3f} of host-timed wall")
