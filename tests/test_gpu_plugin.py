"""GPU test of the reference's plugin ABI: the plugin is loaded the way alpaqa's DLProblem and the benchmark driver
load it (tests/plugin_harness.cpp), every non-null member of alpaqa_problem_functions_t and every extra function is
called, and the results are compared bit for bit with the CPU oracle.
  alpaqa/src/interop/dl-api/include/alpaqa/dl/dl-problem.h:13-169,339-447
  problems/problems/hanging_chain.cpp:139-170, drivers/casadi-dll-wrapper.cpp:69-129"""
import numpy as np
import pytest

import oracle_util as ou
import plugin_util as pu

pytestmark = pytest.mark.gpu
CASES = [("hanging_chain", 0, 12), ("quadcopter", 5, 9), ("realistic_quadcopter", 6, 9)]
# what the reference's single-shooting plugins provide (problems/problems/formulation.hpp:104-123) minus the sparse
# matrices (eval_jac_g, eval_hess_L and their nonzero counts), plus three members of the table they leave out
PROVIDED = {"eval_f", "eval_grad_f", "eval_g", "eval_grad_g_prod", "eval_psi", "eval_psi_grad_psi", "eval_hess_psi_prod",
            "eval_grad_L", "initialize_box_C", "initialize_box_D", "eval_hess_L_prod", "eval_f_grad_f", "eval_grad_psi"}


@pytest.fixture(scope="module", autouse=True)
def built():
    if not pu.available():
        pytest.fail("plugins / harness not built: run __graft_entry__.build() where /root/reference is mounted")


@pytest.mark.parametrize("name,pid,N", CASES)
def test_every_member_of_the_function_table(name, pid, N):
    prob = pu.LoadedProblem(name, N)
    try:
        n, m, nx, nu = ou.dims(pid, N)
        assert (prob.n, prob.m) == (n, m)
        assert prob.provided() == PROVIDED  # null exactly where "not provided" (dl-problem.cpp:99-119)
        rng = np.random.default_rng(pid)
        # extra functions (casadi-dll-wrapper.cpp:100-126)
        for key, val in (("get_nx", nx), ("get_nu", nu), ("get_nc", 0 if pid == 0 else 4), ("get_nc_N", 0 if pid == 0 else 4)):
            assert prob.extra(key)[1] == val
        guess, _ = prob.extra("initial_guess", out_len=n)
        x_init, ln = prob.extra("get_initial_state_ptr", out_len=nx)
        assert ln == nx
        if pid == 0:
            np.testing.assert_array_equal(guess, np.zeros(n))
            np.testing.assert_array_equal(x_init[0:30:3], (np.arange(10) + 1) / 10)
        else:
            np.testing.assert_array_equal(guess, np.tile([9.81, 0, 0, 0], N))
            np.testing.assert_array_equal(x_init, [-0.2, -0.25, 0.5, 0, 0, 0, 0, 0, 0])
        # the driver re-parameterises the NLP by writing through the span (drivers/mpc-driver.cpp:126,148-162)
        x0 = x_init + 0.02 * rng.standard_normal(nx)
        back, _ = prob.extra("get_initial_state_ptr", in1=x0, out_len=nx)
        np.testing.assert_array_equal(back, x0)
        np.testing.assert_array_equal(prob.extra("get_initial_state", out_len=nx)[0], x0)
        uu = rng.uniform(-0.05, 0.05, nu) + (0 if pid == 0 else np.array([9.81, 0, 0, 0]))
        xn, _ = prob.extra("simulate_dynamics", in1=x0, in2=uu, out_len=nx)
        np.testing.assert_array_equal(xn, ou.simulate(pid, x0, uu))
        # boxes
        r = prob.call("initialize_box_C"); lbC, ubC = r["n"], r["work_n"]
        assert (lbC < ubC).all() and lbC[0] == (-1 if pid == 0 else 0)
        r = prob.call("initialize_box_D"); zl, zu = r["m"], r["work_m"]
        # function table vs the oracle
        x = guess + 0.05 * rng.standard_normal(n)
        y = rng.standard_normal(m) * 3 if m else None
        S = 10.0 ** rng.uniform(0, 3, m) if m else None
        v = rng.standard_normal(n)
        pe = lambda fn, **kw: ou.problem_eval(pid, N, fn, x0, x, **kw)
        assert prob.call("eval_f", x)["s"] == pe("f")["s"]
        np.testing.assert_array_equal(prob.call("eval_grad_f", x)["n"], pe("grad_f")["n"])
        r = prob.call("eval_f_grad_f", x)
        assert r["s"] == pe("f")["s"]
        np.testing.assert_array_equal(r["n"], pe("grad_f")["n"])
        np.testing.assert_array_equal(prob.call("eval_grad_L", x, y)["n"], pe("grad_L", y=y)["n"])
        for scale in (1.0, -2.5):
            np.testing.assert_array_equal(prob.call("eval_hess_L_prod", x, y, v=v, scale=scale)["n"],
                                          pe("hess_L_prod", y=y, v=v, scalar=scale)["n"])
            np.testing.assert_array_equal(prob.call("eval_hess_psi_prod", x, y, S, v=v, scale=scale, zl=zl, zu=zu)["n"],
                                          pe("hess_psi_prod", y=y, Sigma=S, v=v, scalar=scale)["n"])
        po, go, yo, _ = ou.eval_(pid, N, x0, x, y=y, Sigma=S, want_yhat=bool(m))
        r = prob.call("eval_psi", x, y, S, zl=zl, zu=zu)
        assert r["s"] == po
        r2 = prob.call("eval_psi_grad_psi", x, y, S, zl=zl, zu=zu)
        assert r2["s"] == po
        np.testing.assert_array_equal(r2["n"], go)
        np.testing.assert_array_equal(prob.call("eval_grad_psi", x, y, S, zl=zl, zu=zu)["n"], go)
        if m:
            np.testing.assert_array_equal(r["m"], yo)
            np.testing.assert_array_equal(prob.call("eval_g", x)["m"], pe("g")["m"])
            np.testing.assert_array_equal(prob.call("eval_grad_g_prod", x, y)["n"], pe("grad_g_prod", y=y)["n"])
            np.testing.assert_array_equal(zl, np.tile([-np.pi / 2, -np.pi / 2, np.cos(1.1 / 2), -np.inf], N + 1))
        else:
            np.testing.assert_array_equal(prob.call("eval_grad_g_prod", x, y)["n"], np.zeros(n))
    finally:
        prob.close()


def test_default_horizon_and_two_instances():
    """size <= 0 keeps the plugin's default horizon (hanging_chain.cpp:29,147-148); instances are independent."""
    a, b = pu.LoadedProblem("hanging_chain", 0), pu.LoadedProblem("hanging_chain", 7)
    try:
        assert a.n == 3 * 30 and b.n == 3 * 7
        xa, _ = a.extra("get_initial_state_ptr", out_len=57)
        b.extra("get_initial_state_ptr", in1=xa + 1.0, out_len=57)
        np.testing.assert_array_equal(a.extra("get_initial_state", out_len=57)[0], xa)
    finally:
        a.close(); b.close()
