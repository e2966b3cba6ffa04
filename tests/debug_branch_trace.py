"""Debug aid: find the first PANTR iteration where the device and the oracle part ways for one option set.
usage: debug_branch_trace.py <problem> <N> <seed> key=value ..."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import numpy as np
import oracle_util as ou
import pantr_cdc2023_experiments_b200 as pb
import test_gpu_branches as tb

name, N, seed, extra = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), sys.argv[4:]
pid = tb.MODELS[name]
solver, _ = pb.make_solver(pb.benchmark_options(name) + extra)
prob, x0 = tb.instances(pb, name, N, 3, seed)
par = ou.pack_params(solver)
cols = "k psi fbe gamma eps radius rho q_model nJ cg_iters accepted".split()
for b in range(3):
    u = np.tile(prob.initial_guess(), (3, 1)); y = np.zeros((3, prob.m))
    solver.set_trace(prob, b, 20000)
    st = solver(prob, u, y, x0=x0)
    tr = solver.get_trace(prob, 20000)
    uo, yo, so, tro = ou.solve(pid, N, par, x0[b], prob.initial_guess(), trace_rows=20000)
    n = min(len(tr), len(tro))
    same = np.all((tr[:n] == tro[:n]) | (np.isnan(tr[:n]) & np.isnan(tro[:n])), axis=1)
    first = int(np.argmin(same)) if not same.all() else -1
    print(f"instance {b}: device rows {len(tr)} oracle rows {len(tro)} status {st['status'][b]}/{int(so['status'])} outer {st['outer_iterations'][b]}/{int(so['outer_iterations'])} "
          f"inner {st['inner_iterations'][b]}/{int(so['inner_iterations'])} ev_psi {st['ev_psi'][b]}/{int(so['ev_psi'])} first differing row {first}")
    if first >= 0:
        for r in range(max(0, first - 2), min(n, first + 3)):
            print("   row", r)
            for c, a, o in zip(cols, tr[r], tro[r]):
                print(f"      {c:9s} device {a!r:26} oracle {o!r:26} {'' if (a == o or (a != a and o != o)) else '<<<'}")
