"""CPU checks of the oracle's problem plugin face (the functions of alpaqa_problem_functions_t that PANTR
itself does not call: grad_f, grad_g_prod, grad_L, hess_L_prod, the box projections) against finite
differences and the reference's formulas (box-constr-problem.hpp:59-185, box.hpp:34-53)."""
import numpy as np
import pytest

import oracle_util as ou


def _point(pid, N, seed):
    n, m, nx, nu = ou.dims(pid, N)
    rng = np.random.default_rng(seed)
    if pid == 0:
        x0 = np.zeros(nx); x0[0:30:3] = (np.arange(10) + 1) / 10
        x0 += 0.01 * rng.standard_normal(nx)
        x = rng.uniform(-0.5, 0.5, n)
    else:
        x0 = np.array([-0.2, -0.25, 0.5, 0, 0, 0, 0, 0, 0.]) + 0.05 * rng.standard_normal(nx)
        x = np.tile([9.81, 0, 0, 0], N) + 0.05 * rng.standard_normal(n)
    return n, m, x0, x, rng.standard_normal(m), rng.standard_normal(n)


@pytest.mark.parametrize("pid", [0, 1, 2])
def test_first_order_face_matches_finite_differences(pid):
    N = 6
    n, m, x0, x, y, v = _point(pid, N, 10 + pid)
    f = ou.problem_eval(pid, N, "f", x0, x)["s"]
    fg = ou.problem_eval(pid, N, "f_grad_f", x0, x)
    gf = ou.problem_eval(pid, N, "grad_f", x0, x)["n"]
    assert fg["s"] == f and np.array_equal(fg["n"], gf)
    h = 1e-6
    d = v / np.linalg.norm(v)
    fd = (ou.problem_eval(pid, N, "f", x0, x + h * d)["s"] - ou.problem_eval(pid, N, "f", x0, x - h * d)["s"]) / (2 * h)
    assert abs(fd - gf @ d) <= 1e-6 * max(1.0, abs(fd))
    ggy = ou.problem_eval(pid, N, "grad_g_prod", x0, x, y=y)["n"]
    gL = ou.problem_eval(pid, N, "grad_L", x0, x, y=y)["n"]
    np.testing.assert_allclose(gL, gf + ggy, rtol=1e-11, atol=1e-11 * np.abs(gf).max())
    if m:
        gp = ou.problem_eval(pid, N, "g", x0, x + h * d)["m"]
        gm = ou.problem_eval(pid, N, "g", x0, x - h * d)["m"]
        fdg = (gp - gm) @ y / (2 * h)
        assert abs(fdg - ggy @ d) <= 1e-6 * max(1.0, abs(fdg))
    else:
        assert not ggy.any()


@pytest.mark.parametrize("pid", [0, 1, 2])
def test_hess_L_prod_matches_finite_differences_and_scale(pid):
    N = 6
    n, m, x0, x, y, v = _point(pid, N, 20 + pid)
    Hv = ou.problem_eval(pid, N, "hess_L_prod", x0, x, y=y, v=v)["n"]
    h = 1e-6
    fd = (ou.problem_eval(pid, N, "grad_L", x0, x + h * v, y=y)["n"] -
          ou.problem_eval(pid, N, "grad_L", x0, x - h * v, y=y)["n"]) / (2 * h)
    np.testing.assert_allclose(Hv, fd, rtol=2e-5, atol=2e-5 * np.abs(fd).max())
    Hv3 = ou.problem_eval(pid, N, "hess_L_prod", x0, x, y=y, v=v, scalar=3.0)["n"]
    np.testing.assert_array_equal(Hv3, Hv * 3.0)
    if m:  # hess psi = hess L(y_hat) + J' W J : with nothing active outside D and Sigma tiny they agree
        S = np.full(m, 1e-300)
        Hp = ou.problem_eval(pid, N, "hess_psi_prod", x0, x, y=np.zeros(m), Sigma=S, v=v)["n"]
        H0 = ou.problem_eval(pid, N, "hess_L_prod", x0, x, y=np.zeros(m), v=v)["n"]
        np.testing.assert_allclose(Hp, H0, rtol=1e-9, atol=1e-9 * np.abs(H0).max())


@pytest.mark.parametrize("pid", [0, 1])
def test_box_functions_follow_reference_formulas(pid):
    N = 5
    n, m, nx, nu = ou.dims(pid, N)
    rng = np.random.default_rng(5)
    lbC, ubC = (np.full(n, -1.0), np.full(n, 1.0)) if pid == 0 else \
        (np.tile([0, -0.1, -0.1, -0.1], N), np.tile([9.81 * 5, 0.1, 0.1, 0.1], N))
    x = rng.uniform(lbC, ubC); g = rng.standard_normal(n) * 3; gamma = 0.37
    r = ou.problem_eval(pid, N, "prox_grad_step", x=x, v=g, scalar=gamma)
    p = np.minimum(np.maximum(-gamma * g, lbC - x), ubC - x)
    np.testing.assert_array_equal(r["n2"], p); np.testing.assert_array_equal(r["n"], x + p)
    J = ou.problem_eval(pid, N, "inactive_indices", x=x, v=g, scalar=gamma)["J"]
    xfw = x - gamma * g
    np.testing.assert_array_equal(J, np.nonzero((lbC < xfw) & (xfw < ubC))[0])
    if m:
        hp = np.pi / 2
        lbD = np.tile([-hp, -hp, np.cos(1.1 / 2), -np.inf], N + 1); ubD = np.tile([hp, hp, np.inf, 0.0], N + 1)
        z = rng.standard_normal(m) * 2
        np.testing.assert_array_equal(ou.problem_eval(pid, N, "proj_diff_g", y=z)["m"], z - np.clip(z, lbD, ubD))
        yy = rng.standard_normal(m) * 10; M = 4.0
        want = np.clip(yy, np.where(lbD == -np.inf, 0, -M), np.where(ubD == np.inf, 0, M))
        np.testing.assert_array_equal(ou.problem_eval(pid, N, "proj_multipliers", y=yy, scalar=M)["m"], want)
