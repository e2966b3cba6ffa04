"""Checks the oracle's hanging-chain model (oracle/hanging_chain.hpp) against an independent numpy
transcription of the reference's Python model and against finite differences.

  python/alpaqa_mpc_benchmarks/problems/hanging_chain.py:7-98   dynamics, RK4, costs, box, x0
  python/alpaqa_mpc_benchmarks/formulations/ss.py:20-55          single shooting

The reference has no golden vectors for the model functions (parity unpinned for the
model arithmetic; CasADi is not available to generate any), hence the independent transcription.
"""
import numpy as np
import pytest

import oracle_util as ou

NB, DIM = 9, 3
PAR = dict(alpha=25.0, beta=1.0, gamma=0.01, m=0.03, D=0.1, L=0.033, v_max=1.0, g=9.81, Ts=0.1)


def f_c(x, u):  # hanging_chain.py:47-56, state = [vec(x) (3x10, column major), vec(v) (3x9)]
    X = x[:DIM * (NB + 1)].reshape(NB + 1, DIM)
    V = x[DIM * (NB + 1):].reshape(NB, DIM)
    dist = np.vstack([X[0:1], X[1:] - X[:-1]])
    nrm = np.sqrt((dist ** 2).sum(axis=1, keepdims=True))
    F = dist * PAR["D"] * (1 - PAR["L"] / nrm)
    a = (F[1:] - F[:-1]) / PAR["m"] + np.array([0, 0, -PAR["g"]])
    return np.concatenate([V.ravel(), u, a.ravel()])


def f_d(x, u):  # hanging_chain.py:58-66
    h = PAR["Ts"]
    k1 = f_c(x, u); k2 = f_c(x + h / 2 * k1, u); k3 = f_c(x + h / 2 * k2, u); k4 = f_c(x + h * k3, u)
    return x + h / 6 * (k1 + 2 * k2 + 2 * k3 + k4)


def stage_cost(x, u=None):  # hanging_chain.py:68-80
    xe = x[DIM * NB:DIM * (NB + 1)] - np.array([1.0, 0, 0])
    v = x[DIM * (NB + 1):]
    c = PAR["alpha"] * xe @ xe + PAR["beta"] * v @ v
    return c if u is None else c + PAR["gamma"] * u @ u


def cost(x0, u, N):  # ss.py:24-37
    x, f = x0.copy(), 0.0
    for k in range(N):
        uk = u[3 * k:3 * k + 3]
        f += stage_cost(x, uk)
        x = f_d(x, uk)
    return f + stage_cost(x)


def rest_state():
    x = np.zeros(57)
    x[0:30:3] = (np.arange(10) + 1) / 10
    return x


def disturbed(rng, steps=5):
    x = rest_state()
    for _ in range(steps):
        x = f_d(x, rng.uniform(-1, 1, 3))
    return x


def test_dims_and_physics():
    assert ou.dims(0, 60) == (180, 0, 57, 3)
    k = f_c(rest_state(), np.zeros(3))
    acc = k[30:].reshape(9, 3)
    np.testing.assert_allclose(acc[:, 2], -9.81)          # gravity on all balls
    np.testing.assert_allclose(acc[:, :2], 0, atol=1e-12)  # spring forces cancel at the rest state


def test_plant_step_matches_independent_model():
    rng = np.random.default_rng(0)
    for _ in range(5):
        x, u = disturbed(rng), rng.uniform(-1, 1, 3)
        np.testing.assert_allclose(ou.simulate(0, x, u), f_d(x, u), rtol=1e-13, atol=1e-14)


@pytest.mark.parametrize("N", [1, 5, 20])
def test_cost_matches_independent_model(N):
    rng = np.random.default_rng(N)
    x0, u = disturbed(rng), rng.uniform(-1, 1, 3 * N)
    psi, _, _, _ = ou.eval_(0, N, x0, u)
    assert abs(psi - cost(x0, u, N)) <= 1e-12 * abs(psi)


@pytest.mark.parametrize("N", [1, 4, 12])
def test_gradient_and_hvp_finite_differences(N):
    rng = np.random.default_rng(10 + N)
    x0, u, v = disturbed(rng), rng.uniform(-1, 1, 3 * N), rng.standard_normal(3 * N)
    _, grad, _, Hv = ou.eval_(0, N, x0, u, v=v)
    h = 1e-6
    fd = np.array([(cost(x0, u + h * e, N) - cost(x0, u - h * e, N)) / (2 * h) for e in np.eye(3 * N)])
    np.testing.assert_allclose(grad, fd, rtol=2e-6, atol=1e-7)
    gp = ou.eval_(0, N, x0, u + h * v)[1]
    gm = ou.eval_(0, N, x0, u - h * v)[1]
    np.testing.assert_allclose(Hv, (gp - gm) / (2 * h), rtol=2e-6, atol=1e-6)


def test_hvp_is_symmetric_and_linear():
    N = 10
    rng = np.random.default_rng(3)
    x0, u = disturbed(rng), rng.uniform(-1, 1, 3 * N)
    a, b = rng.standard_normal(3 * N), rng.standard_normal(3 * N)
    Ha, Hb, Hab = (ou.eval_(0, N, x0, u, v=w)[3] for w in (a, b, 2 * a - 3 * b))
    assert abs(a @ Hb - b @ Ha) <= 1e-10 * max(1.0, abs(a @ Hb))
    np.testing.assert_allclose(Hab, 2 * Ha - 3 * Hb, rtol=1e-10, atol=1e-9)
