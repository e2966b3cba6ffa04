"""Checks the oracle's two quadcopter models, in both formulations -- problem ids 1 / 2: forward-over-adjoint
sweeps (oracle/quadcopter.hpp), ids 5 / 6: the condensed formulation the CUDA library uses
(oracle/quadcopter_condensed.hpp) -- against an independent numpy transcription of the reference's
Python models and against finite differences, the augmented-Lagrangian compositions against their
definition, and the two formulations against each other.

  python/alpaqa_mpc_benchmarks/problems/quadcopter.py:5-121
  python/alpaqa_mpc_benchmarks/problems/realistic_quadcopter.py:5-127
  python/alpaqa_mpc_benchmarks/formulations/ss.py:20-55
  alpaqa/python/alpaqa/casadi_generator/__init__.py:64-77 (psi), :181-189 (hess_psi_prod)

The reference has no golden vectors for these functions (parity unpinned for the model
arithmetic), hence the independent transcription.
"""
import ctypes as C

import numpy as np
import pytest

import oracle_util as ou

TS = 0.1
PF = np.array([0.25, 0.25, 0.5])


def f_c_simple(x, u):  # quadcopter.py:30-47
    v, th, at, om = x[3:6], x[6:9], u[0], u[1:4]
    cr, sr, cp, sp, cy, sy = np.cos(th[0]), np.sin(th[0]), np.cos(th[1]), np.sin(th[1]), np.cos(th[2]), np.sin(th[2])
    R = np.array([[cy * cp, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr],
                  [sy * cp, sy * sp * sr + cy * cr, sy * sp * cr - cy * sr],
                  [-sp, cp * sr, cp * cr]])
    a = R @ np.array([0, 0, at]) + np.array([0, 0, -9.81])
    return np.concatenate([v, a, om])


def f_c_realistic(x, u):  # realistic_quadcopter.py:30-52
    v, (phi, th, psi), at, om = x[3:6], x[6:9], u[0], u[1:4]
    cpsi, cphi, cth, spsi, sphi, sth = np.cos(psi), np.cos(phi), np.cos(th), np.sin(psi), np.sin(phi), np.sin(th)
    ARB = np.array([[cpsi * cth - sphi * spsi * sth, -cphi * spsi, cpsi * sth + cth * sphi * spsi],
                    [cth * spsi + cpsi * sphi * sth, cphi * cpsi, spsi * sth - cpsi * cth * sphi],
                    [-cphi * sth, sphi, cphi * cth]])
    Om = np.array([[cth, 0, -cphi * sth], [0, 1, sphi], [sth, 0, cphi * cth]])
    a = ARB @ np.array([0, 0, at]) + np.array([0, 0, -9.81])
    return np.concatenate([v, a, np.linalg.solve(Om, om)])


F_C = {1: f_c_simple, 2: f_c_realistic, 5: f_c_simple, 6: f_c_realistic}
PIDS = [1, 2, 5, 6]


def f_d(pid, x, u):  # quadcopter.py:49-55
    fc = F_C[pid]
    k1 = fc(x, u); k2 = fc(x + TS * k1 / 2, u); k3 = fc(x + TS * k2 / 2, u); k4 = fc(x + TS * k3, u)
    return x + (TS / 6) * (k1 + 2 * k2 + 2 * k3 + k4)


def con(x):  # quadcopter.py:62-69
    return np.array([x[6], x[7], np.cos(x[6]) * np.cos(x[7]), 0.1 ** 2 - x[0] ** 2 - x[1] ** 2])


def stage_cost(x, u):  # quadcopter.py:87-103
    return 1e-4 * u @ u + 10 * u[1:] @ u[1:] + x[3:6] @ x[3:6] + 10 * (x[:3] - PF) @ (x[:3] - PF) + x[6:9] @ x[6:9]


def term_cost(x):  # quadcopter.py:104-110
    return 25 * x[3:6] @ x[3:6] + 250 * (x[:3] - PF) @ (x[:3] - PF) + 10 * x[6:9] @ x[6:9]


def f_g(pid, x0, u, N):  # ss.py:24-40
    x, f, g = x0.copy(), 0.0, []
    for k in range(N):
        uk = u[4 * k:4 * k + 4]
        f += stage_cost(x, uk)
        g.append(con(x))
        x = f_d(pid, x, uk)
    g.append(con(x))
    return f + term_cost(x), np.concatenate(g)


def boxes(N):
    zl = np.tile([-np.pi / 2, -np.pi / 2, np.cos(1.1 / 2), -np.inf], N + 1)
    zu = np.tile([np.pi / 2, np.pi / 2, np.inf, 0.0], N + 1)
    return zl, zu


def psi_ref(pid, x0, u, y, S, N):  # casadi_generator/__init__.py:64-72
    f, g = f_g(pid, x0, u, N)
    zl, zu = boxes(N)
    zeta = g + y / S
    d = zeta - np.maximum(zl, np.minimum(zeta, zu))
    return f + 0.5 * (S * d) @ d, S * d


def random_point(rng, N, spread=1.0):
    x0 = np.concatenate([[-0.2, -0.25, 0.5] + rng.uniform(-0.1, 0.1, 3), rng.uniform(-0.5, 0.5, 3),
                         rng.uniform(-0.2, 0.2, 3)])
    u = np.tile([9.81, 0, 0, 0], N) + spread * np.concatenate([np.r_[rng.uniform(-3, 3), rng.uniform(-0.1, 0.1, 3)]
                                                               for _ in range(N)])
    m = 4 * (N + 1)
    # multipliers / penalties that put some constraints on either side of their bounds
    y = rng.standard_normal(m) * 5
    S = 10.0 ** rng.uniform(0, 3, m)
    return x0, u, y, S


@pytest.mark.parametrize("fn", ["oracle_sincos", "oracle_sincos_fma"])
def test_sincos_restatement(fn):
    """csrc/pantr_hd.h pantr_sincos (forward-over-adjoint formulation) and pantr_sincos_fma (condensed formulation)
    vs libm over the angle ranges the models see."""
    lib = ou.load()
    sincos = getattr(lib, fn)
    rng = np.random.default_rng(0)
    x = np.concatenate([rng.uniform(-4, 4, 20000), rng.uniform(-1000, 1000, 5000), [0.0, np.pi / 2, -np.pi, 1e-300]])
    s, c = np.zeros_like(x), np.zeros_like(x)
    sincos(C.c_long(x.size), ou.P(x), ou.P(s), ou.P(c))
    assert np.abs(s - np.sin(x)).max() <= 4e-16
    assert np.abs(c - np.cos(x)).max() <= 4e-16
    bad = np.array([np.inf, np.nan, 1e9])
    sb, cb = np.zeros(3), np.zeros(3)
    sincos(C.c_long(3), ou.P(bad), ou.P(sb), ou.P(cb))
    assert np.isnan(sb).all() and np.isnan(cb).all()


@pytest.mark.parametrize("pid", PIDS)
def test_dims_boxes(pid):
    assert ou.dims(pid, 60) == (240, 244, 9, 4)
    assert ou.dims(pid, 10) == (40, 44, 9, 4)


@pytest.mark.parametrize("pid", PIDS)
def test_plant_step_matches_independent_model(pid):
    rng = np.random.default_rng(pid)
    for _ in range(5):
        x0, u, _, _ = random_point(rng, 1)
        np.testing.assert_allclose(ou.simulate(pid, x0, u[:4]), f_d(pid, x0, u[:4]), rtol=1e-13, atol=1e-14)


@pytest.mark.parametrize("pid", PIDS)
@pytest.mark.parametrize("N", [1, 5, 20])
def test_f_g_psi_match_independent_model(pid, N):
    rng = np.random.default_rng(100 * pid + N)
    x0, u, y, S = random_point(rng, N)
    f = C.c_double()
    g = np.zeros(4 * (N + 1))
    assert ou.load().oracle_eval_f_g(pid, N, ou.P(x0), ou.P(u), C.byref(f), ou.P(g)) == 0
    f_ref, g_ref = f_g(pid, x0, u, N)
    assert abs(f.value - f_ref) <= 1e-12 * abs(f_ref)
    np.testing.assert_allclose(g, g_ref, rtol=1e-12, atol=1e-13)
    psi, _, yhat, _ = ou.eval_(pid, N, x0, u, y=y, Sigma=S, want_yhat=True)
    psi_r, yhat_r = psi_ref(pid, x0, u, y, S, N)
    assert abs(psi - psi_r) <= 1e-12 * abs(psi_r)
    np.testing.assert_allclose(yhat, yhat_r, rtol=1e-11, atol=1e-11)
    assert (yhat_r != 0).any() and (yhat_r == 0).any()  # both clamp branches are exercised


@pytest.mark.parametrize("pid", PIDS)
@pytest.mark.parametrize("N", [1, 4, 12])
def test_gradient_and_hvp_finite_differences(pid, N):
    rng = np.random.default_rng(10 * pid + N)
    x0, u, y, S = random_point(rng, N)
    v = rng.standard_normal(4 * N) * np.tile([1.0, 0.05, 0.05, 0.05], N)
    _, grad, _, Hv = ou.eval_(pid, N, x0, u, y=y, Sigma=S, v=v)
    h = 1e-6
    fd = np.array([(psi_ref(pid, x0, u + h * e, y, S, N)[0] - psi_ref(pid, x0, u - h * e, y, S, N)[0]) / (2 * h)
                   for e in np.eye(4 * N)])
    np.testing.assert_allclose(grad, fd, rtol=5e-6, atol=1e-5 * np.abs(fd).max())
    gp = ou.eval_(pid, N, x0, u + h * v, y=y, Sigma=S)[1]
    gm = ou.eval_(pid, N, x0, u - h * v, y=y, Sigma=S)[1]
    np.testing.assert_allclose(Hv, (gp - gm) / (2 * h), rtol=5e-6, atol=1e-5 * np.abs(Hv).max())


@pytest.mark.parametrize("pid", PIDS)
def test_hvp_is_symmetric_and_linear(pid):
    N = 10
    rng = np.random.default_rng(3 + pid)
    x0, u, y, S = random_point(rng, N)
    a, b = rng.standard_normal(4 * N), rng.standard_normal(4 * N)
    Ha, Hb, Hab = (ou.eval_(pid, N, x0, u, y=y, Sigma=S, v=w)[3] for w in (a, b, 2 * a - 3 * b))
    assert abs(a @ Hb - b @ Ha) <= 1e-10 * max(1.0, abs(a @ Hb))
    np.testing.assert_allclose(Hab, 2 * Ha - 3 * Hb, rtol=1e-9, atol=1e-9 * np.abs(Hab).max())


@pytest.mark.parametrize("pid", [1, 2])
@pytest.mark.parametrize("N", [1, 5, 60])
def test_condensed_formulation_agrees_with_forward_over_adjoint(pid, N):
    """Same exact derivatives, different association of the sums: agreement to rounding."""
    rng = np.random.default_rng(40 * pid + N)
    x0, u, y, S = random_point(rng, N)
    v = rng.standard_normal(4 * N)
    pa, ga, ya, Ha = ou.eval_(pid, N, x0, u, y=y, Sigma=S, v=v, want_yhat=True)
    pc, gc, yc, Hc = ou.eval_(pid + 4, N, x0, u, y=y, Sigma=S, v=v, want_yhat=True)
    assert abs(pa - pc) <= 1e-14 * abs(pa)
    # (the two formulations use the two sin/cos restatements: <= 1 ulp apart, not bit-equal)
    np.testing.assert_allclose(yc, ya, rtol=0, atol=2e-14 * np.abs(ya).max())
    np.testing.assert_allclose(gc, ga, rtol=0, atol=2e-14 * np.abs(ga).max())
    np.testing.assert_allclose(Hc, Ha, rtol=0, atol=2e-14 * np.abs(Ha).max())
    np.testing.assert_allclose(ou.simulate(pid + 4, x0, u[:4]), ou.simulate(pid, x0, u[:4]), rtol=0, atol=1e-15)


@pytest.mark.parametrize("pid", [5, 6])
def test_scaled_hessian_products(pid):
    """hess_L_prod / hess_psi_prod take a scale s that multiplies the COST only: s L = s f + y' g,
    s psi = s f + 1/2 y_hat' d (alpaqa/python/alpaqa/casadi_generator/__init__.py:65,69,146-189)."""
    N, s = 6, -2.5
    rng = np.random.default_rng(pid)
    x0, u, y, S = random_point(rng, N)
    v = rng.standard_normal(4 * N) * np.tile([1.0, 0.05, 0.05, 0.05], N)
    h = 1e-6

    def sL(uu):
        f, g = f_g(pid, x0, uu, N)
        return s * f + y @ g

    def spsi(uu):
        f, g = f_g(pid, x0, uu, N)
        zl, zu = boxes(N)
        zeta = g + y / S
        d = zeta - np.maximum(zl, np.minimum(zeta, zu))
        return s * f + 0.5 * (S * d) @ d

    def grad_fd(fun, uu):
        return np.array([(fun(uu + h * e) - fun(uu - h * e)) / (2 * h) for e in np.eye(4 * N)])

    for fn, fun, kw in (("hess_L_prod", sL, {}), ("hess_psi_prod", spsi, {"Sigma": S})):
        Hv = ou.problem_eval(pid, N, fn, x0, u, y=y, v=v, scalar=s, **kw)["n"]
        fd = (grad_fd(fun, u + 1e-4 * v) - grad_fd(fun, u - 1e-4 * v)) / 2e-4
        np.testing.assert_allclose(Hv, fd, rtol=2e-4, atol=2e-4 * np.abs(fd).max())
        H1 = ou.problem_eval(pid, N, fn, x0, u, y=y, v=v, scalar=1.0, **kw)["n"]
        assert np.abs(H1 - Hv).max() > 1e-3 * np.abs(H1).max()  # the scale matters
