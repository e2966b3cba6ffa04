"""Edge cases of the CUDA path against the oracle: empty batches, non-finite inputs, iteration caps,
the largest horizon of the realistic-quadcopter sweep (N = 120), cold closed loop, quadcopter closed loop."""
import importlib.util

import numpy as np
import pytest

import oracle_util as ou

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pb():
    import pantr_cdc2023_experiments_b200 as pb
    return pb


def test_empty_batch_is_a_no_op(pb):
    solver, _ = pb.make_solver(pb.benchmark_options("quadcopter"))
    prob = pb.BatchedProblem("quadcopter", 5)
    st = solver(prob, np.zeros((0, prob.n)), np.zeros((0, prob.m)), x0=np.zeros((0, 9)))
    assert st.shape == (0,)
    assert prob.eval_f(np.zeros((0, prob.n)), x0=np.zeros((0, 9))).shape == (0,)
    assert prob.simulate_dynamics(np.zeros((0, 9)), np.zeros((0, 4))).shape == (0, 9)


@pytest.mark.parametrize("mode", ["warp", "thread"])
def test_non_finite_initial_state_is_handled_like_the_oracle(pb, mode):
    name, N, B = "quadcopter", 6, 34
    pid = {"warp": 5, "thread": 1}[mode]  # oracle formulation of each device path
    solver, _ = pb.make_solver(pb.benchmark_options(name))
    prob = pb.BatchedProblem(name, N)
    solver.set_exec_mode(prob, mode)
    x0 = np.tile(prob.initial_state(), (B, 1))
    x0[3, 1] = np.nan; x0[33, 7] = np.inf                     # two poisoned instances, the rest are normal
    u = np.tile(prob.initial_guess(), (B, 1)); y = np.zeros((B, prob.m))
    st = solver(prob, u, y, x0=x0)
    solver.set_exec_mode(prob, "auto")
    par = ou.pack_params(solver)
    for b in (0, 3, 33):
        uo, yo, so, _ = ou.solve(pid, N, par, x0[b], prob.initial_guess())
        assert int(so["status"]) == st["status"][b], (b, so["status"], st["status"][b])
        assert int(so["inner_iterations"]) == st["inner_iterations"][b]
        np.testing.assert_array_equal(u[b], uo)
    assert st["status"][0] == pb.SolverStatus["Converged"] and st["status"][3] != pb.SolverStatus["Converged"]
    assert (st["status"][[1, 2, 4, 32]] == pb.SolverStatus["Converged"]).all()   # neighbours of a poisoned lane unaffected


def test_iteration_caps_report_max_iter_like_the_oracle(pb):
    N, B = 20, 4
    opts = pb.benchmark_options("hanging_chain") + ["solver.max_iter=3"]
    solver, _ = pb.make_solver(opts)
    prob = pb.BatchedProblem("hanging_chain", N)
    rng = np.random.default_rng(1)
    x0 = np.tile(prob.initial_state(), (B, 1))
    for _ in range(5):
        x0 = prob.simulate_dynamics(x0, rng.uniform(-1, 1, (B, 3)))
    u = np.zeros((B, prob.n)); st = solver(prob, u, None, x0=x0)
    par = ou.pack_params(solver)
    for b in range(B):
        uo, _, so, _ = ou.solve(0, N, par, x0[b], np.zeros(prob.n))
        assert int(so["status"]) == st["status"][b] == pb.SolverStatus["MaxIter"]
        assert st["inner_iterations"][b] == 3 == int(so["inner_iterations"])
        np.testing.assert_array_equal(u[b], uo)               # always_overwrite_results for m = 0 (alm.tpp:29-52)


def test_largest_horizon_of_the_realistic_quadcopter_sweep(pb):
    """BASELINE configs[3]: N = 10..120; n = 480, m = 484 at the top end."""
    name, pid, N, B = "realistic_quadcopter", 6, 120, 3
    prob = pb.BatchedProblem(name, N)
    assert (prob.n, prob.m) == (480, 484)
    rng = np.random.default_rng(12)
    x0 = np.tile(prob.initial_state(), (B, 1)) + 0.02 * rng.standard_normal((B, 9))
    u = np.tile(prob.initial_guess(), (B, 1)) + 0.05 * rng.standard_normal((B, prob.n))
    y = rng.standard_normal((B, prob.m)); S = 10.0 ** rng.uniform(0, 3, (B, prob.m)); v = rng.standard_normal((B, prob.n))
    psi, grad, yhat, Hv = prob.eval(x0, u, y=y, Sigma=S, v=v, want_yhat=True)
    for b in range(B):
        po, go, yo, Ho = ou.eval_(pid, N, x0[b], u[b], y=y[b], Sigma=S[b], v=v[b], want_yhat=True)
        assert po == psi[b]
        np.testing.assert_array_equal(go, grad[b]); np.testing.assert_array_equal(Ho, Hv[b])
    # a capped solve in both mappings, each against its own oracle formulation
    solver, _ = pb.make_solver(pb.benchmark_options(name) + ["alm.max_iter=2", "solver.max_iter=15"])
    par = ou.pack_params(solver)
    for mode, opid in (("warp", 6), ("thread", 2)):
        solver.set_exec_mode(prob, mode)
        uu = np.tile(prob.initial_guess(), (B, 1)); yy = np.zeros((B, prob.m))
        st = solver(prob, uu, yy, x0=x0)
        uo, yo, so, _ = ou.solve(opid, N, par, x0[1], prob.initial_guess())
        np.testing.assert_array_equal(uu[1], uo); np.testing.assert_array_equal(yy[1], yo)
        assert int(so["status"]) == st["status"][1] and int(so["inner_iterations"]) == st["inner_iterations"][1]
    solver.set_exec_mode(prob, "auto")


def test_full_solve_at_the_largest_horizon(pb):
    """One uncapped ALM+PANTR solve at N = 120 (the reference itself stops at N = 60, problems/CMakeLists.txt:33):
    device vs oracle, bit for bit, for the nominal initial state and a perturbed one."""
    name, pid, N = "realistic_quadcopter", 6, 120
    solver, _ = pb.make_solver(pb.benchmark_options(name) + ["alm.max_time=600min"])
    prob = pb.BatchedProblem(name, N)
    x0 = np.tile(prob.initial_state(), (2, 1)); x0[1] += 0.01 * np.random.default_rng(120).standard_normal(9)
    u = np.tile(prob.initial_guess(), (2, 1)); y = np.zeros((2, prob.m))
    st = solver(prob, u, y, x0=x0)
    par = ou.pack_params(solver)
    for b in range(2):
        uo, yo, so, _ = ou.solve(pid, N, par, x0[b], prob.initial_guess())
        assert int(so["status"]) == st["status"][b]
        assert int(so["inner_iterations"]) == st["inner_iterations"][b] and int(so["outer_iterations"]) == st["outer_iterations"][b]
        np.testing.assert_array_equal(u[b], uo); np.testing.assert_array_equal(y[b], yo)
    assert st["status"][0] == 1


def _run_driver(pb, tmp_path, monkeypatch, name, N, extra):
    from pantr_cdc2023_experiments_b200 import mpc_driver
    monkeypatch.chdir(tmp_path)
    opts = pb.benchmark_options(name) + extra + ["results_name=results_x"]
    assert mpc_driver.run([f"/d/{name}", str(N), "ss2p"] + opts) == 0
    spec = importlib.util.spec_from_file_location("", str(tmp_path / "results_x.py"))
    mod = importlib.util.module_from_spec(spec); spec.loader.exec_module(mod)
    return mod.results, opts


def test_quadcopter_closed_loop_cold_start_matches_oracle(pb, tmp_path, monkeypatch):
    """warm=0 branch of the driver (mpc-driver.cpp:192-197) with general constraints: u and y are reset every step."""
    name, pid, N, num_sim = "quadcopter", 5, 8, 3
    res, opts = _run_driver(pb, tmp_path, monkeypatch, name, N, [f"num_sim={num_sim}", "warm=0", "batch=2"])
    assert res["ncon"] == 4 * N + 4 and res["states"].shape == (9, num_sim + 1)
    solver, _ = pb.make_solver(opts)
    par = ou.pack_params(solver)
    x = np.array([-0.2, -0.25, 0.5, 0, 0, 0, 0, 0, 0.])            # num_dist=0 for the quadcopters (dodo.py:94-105)
    guess = np.tile([9.81, 0, 0, 0], N)
    u, _, _, _ = ou.solve(pid, N, par, x, guess)
    x = ou.simulate(pid, x, u[:4])
    for i in range(num_sim):
        u, _, so, _ = ou.solve(pid, N, par, x, guess)
        assert (so["status"] == 1) == (res["runtimes"][i] > 0)
        np.testing.assert_array_equal(res["states"][:, i], x)
        np.testing.assert_array_equal(res["inputs"][:, i], u)
        x = ou.simulate(pid, x, u[:4])
    np.testing.assert_array_equal(res["states"][:, num_sim], x)
