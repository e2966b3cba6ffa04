"""Own memory / race checks of the device code.  compute-sanitizer is closed on the GPU pool (every invocation answers
"compute-sanitizer is closed on this pool and stays closed", profiles/r02_a_compute_sanitizer_closed.log), so the
analogue of the reference's Debug-mode checks (NaN-poisoned stale iterates pantr.tpp:460-467, allocation-free inner loop
pantr.tpp:242, ASan/UBSan builds alpaqa/cmake/Sanitizers.cmake:1-15) is done with the library's own facility
(pantr_b200_set_debug / pantr_b200_check_guards):
  initcheck analogue  the whole per-block workspace is set to NaN before every launch: a read of stale or
                      uninitialised workspace poisons the result, so results must stay bit-identical;
  memcheck analogue   a 32-double guard band behind every block's workspace carries a canary: it must survive;
  racecheck analogue  results must not depend on how many blocks are resident (1 block vs the full grid: blocks share
                      only the instance queue) nor on the run (three repeats bit-identical: a missing __syncwarp
                      between lanes exchanging data through memory shows up as run-to-run differences).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
CASES = [("hanging_chain", 7, 40), ("hanging_chain", 33, 6), ("quadcopter", 7, 40), ("quadcopter", 33, 6),
         ("realistic_quadcopter", 7, 40), ("realistic_quadcopter", 60, 4)]


@pytest.fixture(scope="module")
def pb():
    import pantr_cdc2023_experiments_b200 as pb
    return pb


def _instances(pb, name, N, B):
    prob = pb.BatchedProblem(name, N)
    rng = np.random.default_rng(N + B)
    x0 = np.tile(prob.initial_state(), (B, 1))
    if name == "hanging_chain":
        for _ in range(5):
            x0 = prob.simulate_dynamics(x0, rng.uniform(-1, 1, (B, 3)))
    else:
        x0[1:] += 0.02 * rng.standard_normal((B - 1, prob.nx))
    return prob, x0


def _solve(solver, prob, x0):
    B = x0.shape[0]
    u = np.tile(prob.initial_guess(), (B, 1)); y = np.zeros((B, prob.m))
    st = solver(prob, u, y if prob.m else None, x0=x0)
    return u, y, st


@pytest.mark.parametrize("name,N,B", CASES)
def test_poisoned_workspace_guard_bands_grid_size_and_repeats(pb, name, N, B):
    solver, _ = pb.make_solver(pb.benchmark_options(name) + ["alm.max_iter=6", "solver.max_iter=40"])
    prob, x0 = _instances(pb, name, N, B)
    ref = _solve(solver, prob, x0)
    for flags, limit in ((3, 0), (3, 1), (0, 3), (3, 0), (0, 0)):
        solver.set_debug(prob, flags, limit)
        got = _solve(solver, prob, x0)
        if flags & 2:
            assert solver.check_guards(prob) == 0
        np.testing.assert_array_equal(got[0], ref[0]); np.testing.assert_array_equal(got[1], ref[1])
        for k in ref[2].dtype.names:
            if k not in ("cycles", "elapsed_us"):
                np.testing.assert_array_equal(got[2][k], ref[2][k], err_msg=f"{k} flags={flags} grid_limit={limit}")
    solver.set_debug(prob, 0, 0)


@pytest.mark.parametrize("name,N,B", [("hanging_chain", 9, 5), ("quadcopter", 9, 5), ("realistic_quadcopter", 9, 5)])
def test_problem_functions_with_poisoned_workspace(pb, name, N, B):
    prob, x0 = _instances(pb, name, N, B)
    rng = np.random.default_rng(1)
    x = np.tile(prob.initial_guess(), (B, 1)) + 0.05 * rng.standard_normal((B, prob.n))
    y = rng.standard_normal((B, prob.m)); S = 10.0 ** rng.uniform(0, 3, (B, prob.m)); v = rng.standard_normal((B, prob.n))

    def everything():
        out = [prob.eval_f(x, x0=x0), prob.eval_grad_f(x, x0=x0), prob.eval_g(x, x0=x0), prob.eval_grad_g_prod(x, y, x0=x0),
               prob.eval_grad_L(x, y, x0=x0), prob.eval_hess_L_prod(x, y, -1.5, v, x0=x0),
               prob.eval_hess_psi_prod(x, y, S, 1.0, v, x0=x0), prob.eval_hess_psi_prod(x, y, S, 0.25, v, x0=x0)]
        out += list(prob.eval_psi(x, y, S, x0=x0)) + list(prob.eval_psi_grad_psi(x, y, S, x0=x0))
        out += [a for a in prob.eval(x0, x, y=y if prob.m else None, Sigma=S if prob.m else None, v=v, want_yhat=bool(prob.m)) if a is not None]
        return out

    ref = everything()
    prob.set_debug(3)
    got = everything()
    assert prob.check_guards() == 0
    prob.set_debug(0)
    for a, b in zip(got, ref):
        np.testing.assert_array_equal(a, b)
