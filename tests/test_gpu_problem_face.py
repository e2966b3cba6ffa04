"""GPU parity of the batched problem plugin face (pantr_b200_eval_*_batch, the members of
alpaqa_problem_functions_t) against the CPU oracle: bit-exact, every function, all three models."""
import numpy as np
import pytest

import oracle_util as ou

pytestmark = pytest.mark.gpu
# oracle ids of the formulations the device runs: 5 / 6 = the condensed quadcopters (oracle/quadcopter_condensed.hpp)
PROBLEMS = [("hanging_chain", 0), ("quadcopter", 5), ("realistic_quadcopter", 6)]


@pytest.fixture(scope="module")
def pb():
    import pantr_cdc2023_experiments_b200 as pb
    return pb


def _inputs(prob, pid, N, B, rng):
    if pid == 0:
        x0 = np.tile(prob.initial_state(), (B, 1)) + 0.01 * rng.standard_normal((B, prob.nx))
        x = rng.uniform(-0.8, 0.8, (B, prob.n))
    else:
        x0 = np.tile(prob.initial_state(), (B, 1)) + 0.05 * rng.standard_normal((B, prob.nx))
        x = np.tile(prob.initial_guess(), (B, 1)) + 0.05 * rng.standard_normal((B, prob.n))
    y = rng.standard_normal((B, prob.m)) * 3
    S = 10.0 ** rng.uniform(0, 3, (B, prob.m))
    v = rng.standard_normal((B, prob.n))
    return x0, x, y, S, v


@pytest.mark.parametrize("name,pid", PROBLEMS)
@pytest.mark.parametrize("N", [1, 9, 60])
def test_smooth_functions_bit_exact(pb, name, pid, N):
    prob = pb.BatchedProblem(name, N)
    rng = np.random.default_rng(1000 + 10 * pid + N)
    B = 4
    x0, x, y, S, v = _inputs(prob, pid, N, B, rng)
    f = prob.eval_f(x, x0=x0)
    gf = prob.eval_grad_f(x, x0=x0)
    f2, gf2 = prob.eval_f_grad_f(x, x0=x0)
    g = prob.eval_g(x, x0=x0)
    ggy = prob.eval_grad_g_prod(x, y, x0=x0)
    gL = prob.eval_grad_L(x, y, x0=x0)
    HL = prob.eval_hess_L_prod(x, y, 1.0, v, x0=x0)
    HL2 = prob.eval_hess_L_prod(x, y, -2.5, v, x0=x0)
    Hp = prob.eval_hess_psi_prod(x, y, S, 1.0, v, x0=x0)
    psi, yhat = prob.eval_psi(x, y, S, x0=x0)
    gpsi = prob.eval_grad_psi(x, y, S, x0=x0)
    psi2, gpsi2 = prob.eval_psi_grad_psi(x, y, S, x0=x0)
    np.testing.assert_array_equal(f2, f); np.testing.assert_array_equal(gf2, gf)
    np.testing.assert_array_equal(gpsi2, gpsi)
    for b in range(B):
        yb, Sb = (y[b], S[b]) if prob.m else (None, None)
        assert ou.problem_eval(pid, N, "f", x0[b], x[b])["s"] == f[b]
        np.testing.assert_array_equal(ou.problem_eval(pid, N, "grad_f", x0[b], x[b])["n"], gf[b])
        np.testing.assert_array_equal(ou.problem_eval(pid, N, "g", x0[b], x[b])["m"], g[b])
        np.testing.assert_array_equal(ou.problem_eval(pid, N, "grad_g_prod", x0[b], x[b], y=yb)["n"], ggy[b])
        np.testing.assert_array_equal(ou.problem_eval(pid, N, "grad_L", x0[b], x[b], y=yb)["n"], gL[b])
        np.testing.assert_array_equal(ou.problem_eval(pid, N, "hess_L_prod", x0[b], x[b], y=yb, v=v[b])["n"], HL[b])
        np.testing.assert_array_equal(
            ou.problem_eval(pid, N, "hess_L_prod", x0[b], x[b], y=yb, v=v[b], scalar=-2.5)["n"], HL2[b])
        np.testing.assert_array_equal(
            ou.problem_eval(pid, N, "hess_psi_prod", x0[b], x[b], y=yb, Sigma=Sb, v=v[b])["n"], Hp[b])
        po, go, yo, _ = ou.eval_(pid, N, x0[b], x[b], y=yb, Sigma=Sb, want_yhat=True)
        assert po == psi[b] and po == psi2[b]
        np.testing.assert_array_equal(go, gpsi[b])
        if prob.m:
            np.testing.assert_array_equal(yo, yhat[b])


@pytest.mark.parametrize("name,pid", PROBLEMS)
def test_box_functions_bit_exact(pb, name, pid):
    N, B = 11, 5
    prob = pb.BatchedProblem(name, N)
    rng = np.random.default_rng(77 + pid)
    lb, ub = prob.get_box_C()
    x = rng.uniform(lb, ub, (B, prob.n)); x[0] = lb; x[1] = ub  # on the bounds as well
    g = rng.standard_normal((B, prob.n)) * 4; g[2] = 0
    gamma = 0.21
    h, xh, p = prob.eval_prox_grad_step(gamma, x, g)
    J = prob.eval_inactive_indices_res_lna(gamma, x, g)
    assert not h.any()
    for b in range(B):
        r = ou.problem_eval(pid, N, "prox_grad_step", x=x[b], v=g[b], scalar=gamma)
        np.testing.assert_array_equal(r["n"], xh[b]); np.testing.assert_array_equal(r["n2"], p[b])
        np.testing.assert_array_equal(ou.problem_eval(pid, N, "inactive_indices", x=x[b], v=g[b], scalar=gamma)["J"], J[b])
    z = rng.standard_normal((B, prob.m)) * 2
    yy = rng.standard_normal((B, prob.m)) * 10
    e = prob.eval_proj_diff_g(z)
    yp = prob.eval_proj_multipliers(yy, 4.0)
    assert e.shape == (B, prob.m) and yp.shape == (B, prob.m)
    for b in range(B if prob.m else 0):
        np.testing.assert_array_equal(ou.problem_eval(pid, N, "proj_diff_g", y=z[b])["m"], e[b])
        np.testing.assert_array_equal(ou.problem_eval(pid, N, "proj_multipliers", y=yy[b], scalar=4.0)["m"], yp[b])


def test_null_buffers_are_invalid_argument(pb):
    import ctypes as C
    from pantr_cdc2023_experiments_b200 import _lib
    prob = pb.BatchedProblem("quadcopter", 3)
    lib = _lib.load()
    assert lib.pantr_b200_eval_f_batch(prob._probe.h, 1, None, None, None) == -1
    assert lib.pantr_b200_eval_grad_L_batch(prob._probe.h, 1, None, None, None, None) == -1
    with pytest.raises(ValueError):
        prob.eval_f(np.zeros((2, prob.n + 1)), x0=np.zeros((2, 9)))


def test_benchmark_problem_files_hold_the_reference_model_data(pb, tmp_path):
    """<name>_ss2p_<N>.csv / .guess.csv as problems/generate_problem.py writes them, from the device library's model."""
    from pantr_cdc2023_experiments_b200 import problem_data as pd
    prob = pb.BatchedProblem("quadcopter", 4)
    base = pd.write_benchmark_problem_files(prob, str(tmp_path))
    assert base.endswith("quadcopter_ss2p_4")
    d = pd.read_problem_data(base + ".csv", n=16, m=20, p=9)
    np.testing.assert_array_equal(d["C"][0], np.tile([0, -0.1, -0.1, -0.1], 4))       # quadcopter.py:56-59, ss.py:28-29
    np.testing.assert_array_equal(d["C"][1], np.tile([9.81 * 5, 0.1, 0.1, 0.1], 4))
    np.testing.assert_array_equal(d["D"][0], np.tile([-np.pi / 2, -np.pi / 2, np.cos(1.1 / 2), -np.inf], 5))
    np.testing.assert_array_equal(d["D"][1], np.tile([np.pi / 2, np.pi / 2, np.inf, 0], 5))
    np.testing.assert_array_equal(d["param"], [-0.2, -0.25, 0.5, 0, 0, 0, 0, 0, 0])   # quadcopter.py:78-83
    g, split = pd.read_guess(base + ".guess.csv")
    np.testing.assert_array_equal(g, np.tile([9.81, 0, 0, 0], 4)); assert split == 0  # quadcopter.py:112
    hc = pb.BatchedProblem("hanging_chain", 3)
    d = pd.read_problem_data(pd.write_benchmark_problem_files(hc, str(tmp_path)) + ".csv", n=9, m=0, p=57)
    np.testing.assert_array_equal(d["C"][0], -np.ones(9)); assert d["D"][0].size == 0
    np.testing.assert_array_equal(d["param"][0:30:3], (np.arange(10) + 1) / 10)       # hanging_chain.py:92-96
