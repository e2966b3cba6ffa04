"""Bring-up check of the condensed quadcopter path (exec mode 'warp') on a GPU: bit parity with the oracle mirror
(oracle problem ids 5 / 6), then throughput next to the thread-per-instance path."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import numpy as np
import oracle_util as ou
import pantr_cdc2023_experiments_b200 as pb

what = sys.argv[1] if len(sys.argv) > 1 else "all"
rng = np.random.default_rng(0)
ok = True
def check(name, a, b):
    global ok
    a, b = np.asarray(a), np.asarray(b)
    same = np.array_equal(a, b)
    if not same:
        ok = False
        err = np.abs(a - b).max() / max(1e-300, np.abs(b).max())
        print(f"  MISMATCH {name}: max rel err {err:.3e}")
    return same

if what in ("all", "parity"):
    for name, pid in (("quadcopter", 5), ("realistic_quadcopter", 6)):
        for N in (1, 2, 7, 33, 60):
            prob = pb.BatchedProblem(name, N)
            B = 3
            x0 = np.tile(prob.initial_state(), (B, 1)) + 0.05 * rng.standard_normal((B, 9))
            u = np.tile(prob.initial_guess(), (B, 1)) + 0.05 * rng.standard_normal((B, prob.n))
            y = rng.standard_normal((B, prob.m)) * 3; S = 10.0 ** rng.uniform(0, 3, (B, prob.m)); v = rng.standard_normal((B, prob.n))
            psi, grad, yhat, Hv = prob.eval(x0, u, y=y, Sigma=S, v=v, want_yhat=True)
            f = prob.eval_f(u, x0=x0); g = prob.eval_g(u, x0=x0); gL = prob.eval_grad_L(u, y, x0=x0)
            ggy = prob.eval_grad_g_prod(u, y, x0=x0); HL = prob.eval_hess_L_prod(u, y, -2.5, v, x0=x0)
            Hp = prob.eval_hess_psi_prod(u, y, S, 0.5, v, x0=x0)
            xn = prob.simulate_dynamics(x0, u[:, :4])
            good = True
            for b in range(B):
                po, go, yo, Ho = ou.eval_(pid, N, x0[b], u[b], y=y[b], Sigma=S[b], v=v[b], want_yhat=True)
                good &= check("psi", psi[b], po) & check("grad", grad[b], go) & check("yhat", yhat[b], yo) & check("Hv", Hv[b], Ho)
                good &= check("f", f[b], ou.problem_eval(pid, N, "f", x0[b], u[b])["s"])
                good &= check("g", g[b], ou.problem_eval(pid, N, "g", x0[b], u[b])["m"])
                good &= check("grad_L", gL[b], ou.problem_eval(pid, N, "grad_L", x0[b], u[b], y=y[b])["n"])
                good &= check("grad_g_prod", ggy[b], ou.problem_eval(pid, N, "grad_g_prod", x0[b], u[b], y=y[b])["n"])
                good &= check("hess_L_prod", HL[b], ou.problem_eval(pid, N, "hess_L_prod", x0[b], u[b], y=y[b], v=v[b], scalar=-2.5)["n"])
                good &= check("hess_psi_prod(0.5)", Hp[b], ou.problem_eval(pid, N, "hess_psi_prod", x0[b], u[b], y=y[b], Sigma=S[b], v=v[b], scalar=0.5)["n"])
                good &= check("plant", xn[b], ou.simulate(pid, x0[b], u[b, :4]))
            print(f"{name} N={N}: eval / face parity {'bit-exact' if good else 'FAILED'}", flush=True)
        for N, B in ((10, 6), (30, 3)):
            solver, _ = pb.make_solver(pb.benchmark_options(name) + ["alm.max_time=600min"])  # the oracle has no clock
            prob = pb.BatchedProblem(name, N)
            solver.set_exec_mode(prob, "warp")
            x0 = np.tile(prob.initial_state(), (B, 1)); x0[1:] += 0.03 * rng.standard_normal((B - 1, 9))
            u = np.tile(prob.initial_guess(), (B, 1)); y = np.zeros((B, prob.m))
            solver.set_trace(prob, 0, 8192)
            st = solver(prob, u, y, x0=x0)
            trace = solver.get_trace(prob, 8192)
            par = ou.pack_params(solver)
            good = True
            for b in range(B):
                uo, yo, so, tro = ou.solve(pid, N, par, x0[b], prob.initial_guess(), trace_rows=8192 if b == 0 else 0)
                good &= check("u", u[b], uo) & check("y", y[b], yo) & check("status", st["status"][b], int(so["status"]))
                good &= check("iters", st["inner_iterations"][b], int(so["inner_iterations"]))
                good &= check("hvps", st["ev_hess_psi_prod"][b], int(so["ev_hess_psi_prod"]))
                if b == 0:
                    good &= check("trace", trace, tro)
            print(f"{name} N={N} B={B}: solve parity {'bit-exact' if good else 'FAILED'}; status {st['status'].tolist()} iters {st['inner_iterations'].tolist()}", flush=True)
    print("PARITY", "OK" if ok else "FAILED", flush=True)

if what in ("all", "speed", "speed2", "speed3", "speed4"):
    scale = 0.1
    CAP = ["alm.max_iter=10", "solver.max_iter=100"]
    RUNS = ((("quadcopter", 60, 4736, "warp", CAP), ("realistic_quadcopter", 60, 4736, "warp", CAP), ("quadcopter", 60, 16384, "warp", []))
            if what == "speed2" else
            (("quadcopter", 60, 4736, "warp", CAP), ("realistic_quadcopter", 60, 4736, "warp", CAP)) if what == "speed3" else (("quadcopter", 60, 14208, "warp", CAP), ("realistic_quadcopter", 60, 14208, "warp", CAP)) if what == "speed4" else None)
    for name, N, B, mode, extra in RUNS or (("quadcopter", 60, 2368, "warp", CAP), ("quadcopter", 60, 4736, "warp", CAP),
                                    ("quadcopter", 60, 37888, "thread", CAP),
                                    ("quadcopter", 60, 4736, "warp", []),
                                    ("realistic_quadcopter", 60, 4736, "warp", []),
                                    ("quadcopter", 60, 16384, "warp", [])):
        solver, _ = pb.make_solver(pb.benchmark_options(name) + extra)
        prob = pb.BatchedProblem(name, N)
        solver.set_exec_mode(prob, mode)
        r = np.random.default_rng(0)
        x0 = np.tile(prob.initial_state(), (B, 1))
        x0[:, :3] += scale * r.uniform(-0.1, 0.1, (B, 3)); x0[:, 3:6] += scale * r.uniform(-0.5, 0.5, (B, 3)); x0[:, 6:9] += scale * r.uniform(-0.2, 0.2, (B, 3))
        for rep in range(2 if B <= 5000 else 1):
            u = np.tile(prob.initial_guess(), (B, 1)); y = np.zeros((B, prob.m))
            t = time.time(); st = solver(prob, u, y, x0=x0); t = time.time() - t
            it = st["inner_iterations"]
            print(f"{name} {mode} {' '.join(extra)} N={N} B={B}: {t:.3f}s {B/t:.1f} solves/s converged {np.mean(st['status']==1):.3f} "
                  f"inner mean {it.mean():.0f} max {it.max()} hvp/s {st['ev_hess_psi_prod'].sum()/t/1e6:.2f}M it/s {it.sum()/t/1e6:.3f}M "
                  f"busy-slot it/s {it.sum()/(st['elapsed_us'].sum()*1e-6):.0f} grid {solver.grid_size(prob)}", flush=True)
