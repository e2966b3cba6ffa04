"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle on the same inputs."""
import numpy as np
import pytest

import oracle_util as ou

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pb():
    import pantr_cdc2023_experiments_b200 as pb
    return pb


def disturbed_states(prob, B, rng, steps=5):
    x0 = np.tile(prob.initial_state(), (B, 1))
    for _ in range(steps):
        x0 = prob.simulate_dynamics(x0, rng.uniform(-1, 1, (B, 3)))
    return x0


@pytest.mark.parametrize("N", [1, 2, 3, 7, 60])  # 1..3: the clamped look-ahead of the product's cp.async rings
def test_hanging_chain_eval_bit_exact(pb, N):
    prob = pb.BatchedProblem("hanging_chain", N)
    rng = np.random.default_rng(1 + N)
    B = 6
    x0 = disturbed_states(prob, B, rng)
    u = rng.uniform(-1, 1, (B, prob.n)); v = rng.standard_normal((B, prob.n))
    psi, grad, _, Hv = prob.eval(x0, u, v=v)
    for b in range(B):
        po, go, _, Ho = ou.eval_(0, N, x0[b], u[b], v=v[b])
        assert po == psi[b]
        np.testing.assert_array_equal(go, grad[b])
        np.testing.assert_array_equal(Ho, Hv[b])


def test_hanging_chain_plant_step_bit_exact(pb):
    prob = pb.BatchedProblem("hanging_chain", 5)
    rng = np.random.default_rng(3)
    x = disturbed_states(prob, 5, rng, steps=3)
    u = rng.uniform(-1, 1, (5, 3))
    xn = prob.simulate_dynamics(x, u)
    for b in range(5):
        np.testing.assert_array_equal(ou.simulate(0, x[b], u[b]), xn[b])


@pytest.mark.parametrize("N,B", [(10, 8), (30, 4), (60, 3)])
def test_hanging_chain_solve_parity(pb, N, B):
    solver, _ = pb.make_solver(pb.benchmark_options("hanging_chain"))
    prob = pb.BatchedProblem("hanging_chain", N)
    rng = np.random.default_rng(100 + N)
    x0 = disturbed_states(prob, B, rng)
    u = np.zeros((B, prob.n))
    solver.set_trace(prob, 0)
    stats = solver(prob, u, None, x0=x0)
    trace = solver.get_trace(prob)
    par = ou.pack_params(solver)
    for b in range(B):
        uo, _, so, tro = ou.solve(0, N, par, x0[b], np.zeros(prob.n), trace_rows=4096 if b == 0 else 0)
        assert int(so["status"]) == stats["status"][b] == 1
        # north_star tolerance: iterations within +-1, iterate and objective within 1e-8 relative
        assert abs(int(so["inner_iterations"]) - int(stats["inner_iterations"][b])) <= 1
        np.testing.assert_allclose(u[b], uo, rtol=1e-8, atol=1e-10)
        np.testing.assert_allclose(stats["final_psi"][b], so["final_psi"], rtol=1e-8)
        for k in ("ev_psi", "ev_psi_grad_psi", "ev_grad_L", "ev_hess_psi_prod", "ev_prox_grad_step",
                  "ev_grad_psi", "ev_inactive_indices_res_lna"):
            assert abs(int(so[k]) - int(stats[k][b])) <= max(2, 0.02 * so[k]), k
        if b == 0:  # iteration-level parity (the arithmetic is designed to be bit-identical)
            assert len(tro) == len(trace)
            np.testing.assert_array_equal(tro, trace)


def test_closed_loop_driver_matches_oracle_closed_loop(pb, tmp_path, monkeypatch):
    """Batched mpc_driver (device-resident closed loop) vs the reference's closed loop
    (drivers/mpc-driver.cpp:123-210) executed with the CPU oracle, instance 0 = the reference run."""
    import importlib.util
    from pantr_cdc2023_experiments_b200 import mpc_driver
    N, num_sim = 10, 4
    monkeypatch.chdir(tmp_path)
    opts = pb.benchmark_options("hanging_chain") + [f"num_sim={num_sim}", "warm=1", "batch=3", "results_name=results_cl"]
    assert mpc_driver.run(["/some/dir/hanging_chain", str(N), "ss2p"] + opts) == 0
    spec = importlib.util.spec_from_file_location("", str(tmp_path / "results_cl.py"))
    mod = importlib.util.module_from_spec(spec); spec.loader.exec_module(mod)
    res = mod.results
    assert res["states"].shape == (57, num_sim + 1) and res["inputs"].shape == (3 * N, num_sim)
    assert res["batch_size"] == 3 and res["nvar"] == 3 * N and res["ncon"] == 0
    # oracle closed loop
    solver, _ = pb.make_solver(opts)
    par = ou.pack_params(solver)
    x = np.zeros(57); x[0:30:3] = (np.arange(10) + 1) / 10
    for _ in range(5):
        x = ou.simulate(0, x, np.array([-1.0, 1.0, 1.0]))            # num_dist=5 u_dist=-1,1,1
    u, _, _, _ = ou.solve(0, N, par, x, np.zeros(3 * N))
    x = ou.simulate(0, x, u[:3])
    evals_hvp = 0
    for i in range(num_sim):
        u = np.concatenate([u[3:], u[-3:]])                             # mpc-driver.cpp:170
        u, _, so, _ = ou.solve(0, N, par, x, u)
        assert so["status"] == 1 and res["runtimes"][i] > 0
        np.testing.assert_array_equal(res["states"][:, i], x)
        np.testing.assert_array_equal(res["inputs"][:, i], u)
        evals_hvp += int(so["ev_hess_psi_prod"])
        x = ou.simulate(0, x, u[:3])
    np.testing.assert_array_equal(res["states"][:, num_sim], x)
    assert res["evaluations"]["hess_ψ_prod"] == evals_hvp
    batch = np.load(tmp_path / "results_cl.npz")
    assert batch["states"].shape == (3, 57, num_sim + 1) and (batch["runtimes"] > 0).all()


# ---- quadcopter / realistic quadcopter: m = 4(N+1) > 0, the ALM loop runs on the device ----
# oracle ids: 5 / 6 = condensed formulation (the device default, one warp per instance); 1 / 2 = forward-over-adjoint
# sweeps (the thread-per-instance path)
QUADS = [("quadcopter", 5), ("realistic_quadcopter", 6)]


def quad_states(B, rng):
    """x0 = (p0 + U[-0.1,0.1]^3, v ~ U[-0.5,0.5]^3, theta ~ U[-0.2,0.2]^3)."""
    x0 = np.zeros((B, 9))
    x0[:, :3] = np.array([-0.2, -0.25, 0.5]) + rng.uniform(-0.1, 0.1, (B, 3))
    x0[:, 3:6] = rng.uniform(-0.5, 0.5, (B, 3))
    x0[:, 6:9] = rng.uniform(-0.2, 0.2, (B, 3))
    return x0


@pytest.mark.parametrize("name,pid", QUADS)
@pytest.mark.parametrize("N", [1, 7, 60])
def test_quadcopter_eval_bit_exact(pb, name, pid, N):
    prob = pb.BatchedProblem(name, N)
    assert (prob.n, prob.m, prob.nx, prob.nu, prob.nc) == (4 * N, 4 * N + 4, 9, 4, 4)
    rng = np.random.default_rng(7 * pid + N)
    B = 5
    x0 = quad_states(B, rng)
    u = np.tile(prob.initial_guess(), (B, 1)) + rng.uniform(-0.1, 0.1, (B, prob.n)) * np.tile([30, 1, 1, 1], N)
    y = rng.standard_normal((B, prob.m)) * 5
    S = 10.0 ** rng.uniform(0, 3, (B, prob.m))
    v = rng.standard_normal((B, prob.n))
    psi, grad, yhat, Hv = prob.eval(x0, u, y=y, Sigma=S, v=v, want_yhat=True)
    for b in range(B):
        po, go, yo, Ho = ou.eval_(pid, N, x0[b], u[b], y=y[b], Sigma=S[b], v=v[b], want_yhat=True)
        assert po == psi[b]
        np.testing.assert_array_equal(go, grad[b])
        np.testing.assert_array_equal(yo, yhat[b])
        np.testing.assert_array_equal(Ho, Hv[b])
    # Hessian-vector product alone (no gradient requested): the model refreshes its adjoint cache
    _, _, _, Hv2 = prob.eval(x0, u, y=y, Sigma=S, v=v, want_grad=False)
    np.testing.assert_array_equal(Hv2, Hv)


@pytest.mark.parametrize("name,pid", QUADS)
def test_quadcopter_plant_step_and_boxes(pb, name, pid):
    prob = pb.BatchedProblem(name, 5)
    rng = np.random.default_rng(pid)
    x = quad_states(4, rng)
    u = np.tile([9.81, 0, 0, 0], (4, 1)) + rng.uniform(-0.1, 0.1, (4, 4))
    xn = prob.simulate_dynamics(x, u)
    for b in range(4):
        np.testing.assert_array_equal(ou.simulate(pid, x[b], u[b]), xn[b])
    lb, ub = prob.get_box_C()
    np.testing.assert_array_equal(lb, np.tile([0, -0.1, -0.1, -0.1], 5))
    np.testing.assert_array_equal(ub, np.tile([9.81 * 5, 0.1, 0.1, 0.1], 5))
    zl, zu = prob.get_box_D()
    np.testing.assert_array_equal(zl, np.tile([-np.pi / 2, -np.pi / 2, np.cos(1.1 / 2), -np.inf], 6))
    np.testing.assert_array_equal(zu, np.tile([np.pi / 2, np.pi / 2, np.inf, 0], 6))
    np.testing.assert_array_equal(prob.initial_state(), [-0.2, -0.25, 0.5, 0, 0, 0, 0, 0, 0])


@pytest.mark.parametrize("name,pid", QUADS)
@pytest.mark.parametrize("N,B", [(10, 6), (30, 3), (60, 3)])  # 60 = the horizon of BASELINE configs[2]
def test_quadcopter_solve_parity(pb, name, pid, N, B):
    # alm.max_time=30s of the option set is enforced on the device clock; the oracle has no clock, and the
    # full-scale perturbations take up to 5 * 10^4 iterations
    solver, _ = pb.make_solver(pb.benchmark_options(name) + ["alm.max_time=600min"])
    prob = pb.BatchedProblem(name, N)
    rng = np.random.default_rng(200 + 10 * pid + N)
    x0 = quad_states(B, rng)
    x0[0] = prob.initial_state()  # instance 0 = the reference's own cold start
    u = np.tile(prob.initial_guess(), (B, 1))
    y = np.zeros((B, prob.m))
    solver.set_trace(prob, 0, 8192)
    stats = solver(prob, u, y, x0=x0)
    trace = solver.get_trace(prob, 8192)
    par = ou.pack_params(solver)
    for b in range(B):
        uo, yo, so, tro = ou.solve(pid, N, par, x0[b], prob.initial_guess(), trace_rows=8192 if b == 0 else 0)
        assert int(so["status"]) == stats["status"][b]
        assert abs(int(so["outer_iterations"]) - int(stats["outer_iterations"][b])) <= 1
        assert abs(int(so["inner_iterations"]) - int(stats["inner_iterations"][b])) <= 1
        np.testing.assert_allclose(u[b], uo, rtol=1e-8, atol=1e-10)
        np.testing.assert_allclose(y[b], yo, rtol=1e-8, atol=1e-8)
        np.testing.assert_allclose(stats["final_psi"][b], so["final_psi"], rtol=1e-8)
        np.testing.assert_allclose(stats["delta"][b], so["delta"], rtol=1e-6, atol=1e-12)
        np.testing.assert_allclose(stats["norm_penalty"][b], so["norm_penalty"], rtol=1e-12)
        for k in ("ev_psi", "ev_psi_grad_psi", "ev_grad_L", "ev_hess_psi_prod", "ev_prox_grad_step",
                  "ev_grad_psi", "ev_inactive_indices_res_lna", "ev_proj_multipliers"):
            assert abs(int(so[k]) - int(stats[k][b])) <= max(2, 0.02 * so[k]), k
        if b == 0:  # same arithmetic on both sides: iteration-level parity
            assert len(tro) == len(trace)
            np.testing.assert_array_equal(tro, trace)
            np.testing.assert_array_equal(u[b], uo)
            np.testing.assert_array_equal(y[b], yo)
    assert (stats["status"] == 1).all()


@pytest.mark.parametrize("name,pid", [("quadcopter", 1), ("realistic_quadcopter", 2)])
@pytest.mark.parametrize("N,B", [(10, 40), (30, 33)])
def test_quadcopter_thread_per_instance_path_matches_its_oracle(pb, name, pid, N, B):
    """ThreadExec (32 instances per warp, interleaved vectors, forward-over-adjoint sweeps) against the oracle's
    forward-over-adjoint formulation: same bits; against the condensed default path: same solution to rounding."""
    solver, _ = pb.make_solver(pb.benchmark_options(name) + ["alm.max_time=600min"])
    prob = pb.BatchedProblem(name, N)
    rng = np.random.default_rng(300 + 10 * pid + N)
    x0 = quad_states(B, rng) * 0.2 + 0.8 * prob.initial_state()
    x0[0] = prob.initial_state()
    res = {}
    for mode in ("warp", "thread"):
        solver.set_exec_mode(prob, mode)
        u = np.tile(prob.initial_guess(), (B, 1)); y = np.zeros((B, prob.m))
        if mode == "thread":
            solver.set_trace(prob, 0, 8192)
        st = solver(prob, u, y, x0=x0)
        res[mode] = (u, y, st, solver.get_trace(prob, 8192) if mode == "thread" else None)
    solver.set_exec_mode(prob, "auto")
    uw, yw, sw, _ = res["warp"]; ut, yt, stt, trace = res["thread"]
    both = (sw["status"] == 1) & (stt["status"] == 1)  # (an instance may run into alm.max_iter in one formulation only)
    assert both.sum() >= B - 2
    np.testing.assert_allclose(ut[both], uw[both], rtol=0, atol=1e-5 * np.abs(uw).max())  # two formulations, one minimiser
    par = ou.pack_params(solver)
    for b in (0, B - 1):
        uo, yo, so, tro = ou.solve(pid, N, par, x0[b], prob.initial_guess(), trace_rows=8192 if b == 0 else 0)
        assert int(so["status"]) == stt["status"][b]
        np.testing.assert_array_equal(ut[b], uo)
        np.testing.assert_array_equal(yt[b], yo)
        assert int(so["inner_iterations"]) == int(stt["inner_iterations"][b])
        if b == 0:
            np.testing.assert_array_equal(tro, trace)


def test_hanging_chain_warm_start_parity_at_the_bench_horizon(pb):
    """BASELINE configs[1] is the WARM solve at N = 60: cold solve, plant step, shift (mpc-driver.cpp:158-175), then the
    timed warm solve -- compared with the oracle doing the same, bit for bit."""
    N, B = 60, 4
    solver, _ = pb.make_solver(pb.benchmark_options("hanging_chain"))
    prob = pb.BatchedProblem("hanging_chain", N)
    rng = np.random.default_rng(60)
    x0 = disturbed_states(prob, B, rng)
    u = np.zeros((B, prob.n))
    st0 = solver(prob, u, None, x0=x0)
    x1 = prob.simulate_dynamics(x0, u[:, :3])
    uw = np.concatenate([u[:, 3:], u[:, -3:]], axis=1).copy()
    uw_in = uw.copy()
    st = solver(prob, uw, None, x0=x1)
    par = ou.pack_params(solver)
    for b in range(B):
        uo, _, so0, _ = ou.solve(0, N, par, x0[b], np.zeros(prob.n))
        np.testing.assert_array_equal(u[b], uo)
        x1o = ou.simulate(0, x0[b], uo[:3])
        np.testing.assert_array_equal(x1[b], x1o)
        np.testing.assert_array_equal(uw_in[b], np.concatenate([uo[3:], uo[-3:]]))
        uwo, _, so, _ = ou.solve(0, N, par, x1o, np.concatenate([uo[3:], uo[-3:]]))
        assert int(so["status"]) == st["status"][b] == 1
        assert int(so["inner_iterations"]) == st["inner_iterations"][b]
        assert int(so["ev_hess_psi_prod"]) == st["ev_hess_psi_prod"][b]
        np.testing.assert_array_equal(uw[b], uwo)
    assert (st["inner_iterations"] < st0["inner_iterations"]).any()  # the warm start helps


def test_instance_order_hint_does_not_change_results(pb):
    import torch
    N, B = 12, 50
    solver, _ = pb.make_solver(pb.benchmark_options("hanging_chain"))
    prob = pb.BatchedProblem("hanging_chain", N)
    rng = np.random.default_rng(5)
    x0 = disturbed_states(prob, B, rng)
    u0 = np.zeros((B, prob.n)); st0 = solver(prob, u0, None, x0=x0)
    order = torch.from_numpy(rng.permutation(B).astype(np.int32)).cuda()
    solver.set_order(prob, order)
    u1 = np.zeros((B, prob.n)); st1 = solver(prob, u1, None, x0=x0)
    solver.set_order(prob, None)
    np.testing.assert_array_equal(u1, u0)
    for k in st0.dtype.names:
        if k not in ("cycles", "elapsed_us"):
            np.testing.assert_array_equal(st1[k], st0[k], err_msg=k)
    # the hint carries its length: another batch size is refused, an out-of-range entry is skipped (status Busy)
    solver.set_order(prob, order)
    with pytest.raises(ValueError):
        solver(prob, np.zeros((B - 1, prob.n)), None, x0=x0[:B - 1])
    bad = order.clone(); bad[3] = B + 7
    skipped = int(order[3])
    solver.set_order(prob, bad)
    u2 = np.zeros((B, prob.n)); st2 = solver(prob, u2, None, x0=x0)
    solver.set_order(prob, None)
    assert st2["status"][skipped] == 0 and not u2[skipped].any()
    keep = np.arange(B) != skipped
    np.testing.assert_array_equal(u2[keep], u0[keep])
