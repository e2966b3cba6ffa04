"""CPU arm of the quadcopter configs on this box's host cores: the oracle port (reference build flags,
forward-over-adjoint formulation = what CasADi generates) on a bounded sample of the bench distribution.
Test infrastructure / baseline only."""
import ctypes as C, os, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import oracle_util as ou
import pantr_cdc2023_experiments_b200 as pb
odir = os.path.join(ROOT, "oracle")
subprocess.check_call(["make", "-C", odir, "-B", "liboracle_fast.so"], stdout=subprocess.DEVNULL)
lib = C.CDLL(os.path.join(odir, "liboracle_fast.so"))
cores = os.cpu_count()
P = lambda a: a.ctypes.data_as(C.c_void_p)
for name, pid in (("quadcopter", 1), ("realistic_quadcopter", 2)):
    for N in (60,) if pid == 1 else (20, 60):
        solver, _ = pb.make_solver(pb.benchmark_options(name))
        par = ou.pack_params(solver)
        for scale in (0.1, 0.2):
            S = 8 * cores
            rng = np.random.default_rng(0)
            x0 = np.zeros((S, 9)); x0[:, :3] = (-0.2, -0.25, 0.5)
            x0[:, :3] += scale * rng.uniform(-0.1, 0.1, (S, 3))
            x0[:, 3:6] += scale * rng.uniform(-0.5, 0.5, (S, 3)); x0[:, 6:9] += scale * rng.uniform(-0.2, 0.2, (S, 3))
            u = np.tile(np.tile([9.81, 0, 0, 0], N), (S, 1)).astype(np.float64); y = np.zeros((S, 4 * (N + 1))); st = np.zeros((S, 24))
            t = time.perf_counter()
            assert lib.oracle_solve_batch(pid, N, P(par), C.c_long(S), P(x0), P(u), P(y), P(st), cores) == 0
            t = time.perf_counter() - t
            print(f"{name} N={N} scale {scale}: {S} instances on {cores} threads in {t:.2f} s = {S/t:.1f} solves/s "
                  f"({S/t/cores:.2f}/core), converged {np.mean(st[:,0]==1):.3f}, inner mean {st[:,2].mean():.0f} "
                  f"max {st[:,2].max():.0f}, hvp mean {st[:,15].mean():.0f} -> {st[:,15].sum()/t/1e6:.3f} M hvp/s", flush=True)
