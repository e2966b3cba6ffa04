"""Pins the CPU oracle (oracle/) with the reference's own solver-independent known-answer tests.

  alpaqa/test/outer/test-alm.cpp:7-81    singleshooting1D   u* = -0.454545 (1e-4)
  alpaqa/test/outer/test-alm.cpp:83-179  multipleshooting1D u* = -0.454545 (1e-4)
  alpaqa/test/outer/test-alm.cpp:181-304 multipleshooting8D u, x, y references (6e-4), A/B from
                                         matrices/test-alm.mat.ipp (fixture: tests/golden/test_alm_8d.json)
  alpaqa/test/outer/test-casadi.cpp:9-85 Rosenbrock through the committed CasADi-generated C
                                         (compiled into oracle/_ref/), x, y references (rel 1e-6)
The reference runs them with PANOC; the minimisers are solver independent, here they are reached by the
oracle's PANTR+ALM restatement.
"""
import ctypes as C
import json
import os
import subprocess

import numpy as np
import pytest

import oracle_util as ou

HERE = os.path.dirname(os.path.abspath(__file__))
ROSEN_SO = os.path.join(ou.ORACLE_DIR, "_ref", "rosenbrock_functions_test.so")
INF = np.inf


def pantr_params(alm, max_iter=200, stop_crit=4, tol_scale=1.0):
    """params[48] layout of oracle_capi.cpp:make_solver; PANTR defaults of inner/pantr.hpp:22-91."""
    p = np.zeros(48)
    p[0:13] = [alm["tolerance"], alm["dual_tolerance"], alm["penalty_update_factor"], alm["initial_penalty"],
               alm["initial_tolerance"], alm.get("tolerance_update_factor", 0.1),
               alm.get("rel_penalty_increase_threshold", 0.1), 1e9, 1e9, 1e-9, alm["max_iter"], 20, 0]
    p[16:37] = [max_iter, stop_crit, 1e-5, 1e20, 10 * ou.EPS, 10 * ou.EPS, 0.2, 0.8, 0.35, 0.999, 2.5,
                np.nan, 100 * ou.EPS, 0, 1e-6, 1e-12, 0.95, 0, 0, 0, 1]
    p[40] = 1.0
    p[44:48] = [tol_scale, 0.5, INF, 1]
    return p


def solve_qp(H, c, f0, G, g0, Clb, Cub, Dlb, Dub, par, x, y):
    n, m = len(c), len(g0)
    arr = lambda a: np.ascontiguousarray(np.asarray(a, dtype=np.float64))
    H, c, G, g0, Clb, Cub, Dlb, Dub = map(arr, (H, c, G, g0, Clb, Cub, Dlb, Dub))
    x, y, st = arr(x).copy(), arr(y).copy(), np.zeros(ou.NSTATS)
    lib = ou.load()
    lib.oracle_solve_qp.argtypes = [C.c_long, C.c_long, C.c_void_p, C.c_void_p, C.c_double] + [C.c_void_p] * 10
    rc = lib.oracle_solve_qp(n, m, ou.P(H), ou.P(c), f0, ou.P(G), ou.P(g0), ou.P(Clb), ou.P(Cub), ou.P(Dlb),
                             ou.P(Dub), ou.P(par), ou.P(x), ou.P(y), ou.P(st))
    assert rc == 0
    return x, y, dict(zip(ou.STAT_KEYS, st))


def test_alm_singleshooting1D():
    # f(u) = 10 (0.5 + u)^2 + u^2, u in [-1, 1]: test-alm.cpp:10-45
    alm = dict(tolerance=1e-4, dual_tolerance=1e-4, penalty_update_factor=100, initial_penalty=20,
               initial_tolerance=1, rel_penalty_increase_threshold=0.25, max_iter=10)
    x, _, st = solve_qp([[22.0]], [10.0], 2.5, np.zeros((0, 1)), [], [-1], [1], [], [], pantr_params(alm, 100), [1.0], [])
    assert st["status"] == 1
    assert abs(x[0] - (-0.454545)) < 1e-4


def test_alm_multipleshooting1D():
    # variables (u, x); f = 10 x^2 + u^2; g = 0.5*1 + u - x = 0: test-alm.cpp:87-132
    alm = dict(tolerance=1e-4, dual_tolerance=1e-4, penalty_update_factor=5, initial_penalty=1,
               initial_tolerance=1e-4, rel_penalty_increase_threshold=0.25, max_iter=10)
    H = np.diag([2.0, 20.0])
    x, y, st = solve_qp(H, [0, 0], 0.0, [[1.0, -1.0]], [0.5], [-1, -INF], [1, INF], [0], [0],
                        pantr_params(alm, 100), [0.5, 0.5], [1.0])
    assert st["status"] == 1
    assert abs(x[0] - (-0.454545)) < 1e-4


@pytest.mark.parametrize("stop_crit", range(10))
def test_alm_multipleshooting8D(stop_crit):
    """All ten PANOCStopCrit values (panoc-helpers.tpp:62-137) reach the same minimiser; 0 = ApproxKKT is the
    reference's default (the one test-alm.cpp itself runs with), 0 / 1 / 8 need grad psi(x_hat)."""
    fix = json.load(open(os.path.join(HERE, "golden", "test_alm_8d.json")))
    A = np.array(fix["A_rowmajor_8x8"]).reshape(8, 8)
    B = np.array(fix["B_rowmajor_8x8"]).reshape(8, 8)
    nx = nu = 8
    H = np.diag([2.0] * nu + [20.0] * nx)            # f = x'Qx + u'Ru, Q = 10 I, R = I
    G = np.hstack([B, -np.eye(nx)])                  # g = A x0 + B u - x
    g0 = A @ np.ones(nx)
    alm = dict(tolerance=1e-4, dual_tolerance=1e-4, penalty_update_factor=5, initial_penalty=1,
               initial_tolerance=1e-4, rel_penalty_increase_threshold=0.25, max_iter=20)
    x, y, st = solve_qp(H, np.zeros(16), 0.0, G, g0, [-1] * nu + [-INF] * nx, [1] * nu + [INF] * nx,
                        np.zeros(nx), np.zeros(nx), pantr_params(alm, 200, stop_crit=stop_crit), np.full(16, 5.0), np.ones(nx))
    assert st["status"] == 1
    if stop_crit in (2, 3):  # ProjGradNorm(2) measure the step p itself, which scales with the step size gamma:
        return               # at the same tolerance they stop far earlier than the gamma-free criteria
    tol = fix["tolerance"]
    np.testing.assert_allclose(x[:nu], fix["u_ref"], atol=tol)
    np.testing.assert_allclose(x[nu:], fix["xs_ref"], atol=tol)
    np.testing.assert_allclose(y, fix["y_ref"], atol=tol)


def _rosenbrock_so():
    if not os.path.exists(ROSEN_SO) and os.path.isdir("/root/reference"):
        subprocess.check_call(["make", "-C", ou.ORACLE_DIR, "ref"])
    if not os.path.exists(ROSEN_SO):
        pytest.skip("oracle/_ref/rosenbrock_functions_test.so not built (reference not mounted)")
    return ROSEN_SO.encode()


ROSEN = dict(n=2, m=1, param=[2.0], Clb=[0.0, 0.0], Cub=[5.0, 5.0], Dlb=[0.0], Dub=[1.0])  # test-casadi.cpp:16-22


@pytest.mark.parametrize("composed", [0, 1])
def test_casadi_rosenbrock_minimiser(composed):
    so = _rosenbrock_so()
    alm = dict(tolerance=1e-10, dual_tolerance=1e-10, penalty_update_factor=10, initial_penalty=1,
               initial_tolerance=1, max_iter=20)                        # test-casadi.cpp:30-35 (+ defaults)
    par = pantr_params(alm, 500)
    arr = lambda a: np.array(a, dtype=np.float64)
    x, y, st = arr([2.0, 1.0]), arr([1.0]), np.zeros(ou.NSTATS)        # test-casadi.cpp:52-55
    lib = ou.load()
    lib.oracle_solve_casadi.argtypes = [C.c_char_p, C.c_long, C.c_long, C.c_void_p, C.c_long] + [C.c_void_p] * 4 + \
                                       [C.c_int] + [C.c_void_p] * 4
    rc = lib.oracle_solve_casadi(so, 2, 1, ou.P(arr(ROSEN["param"])), 1, ou.P(arr(ROSEN["Clb"])),
                                 ou.P(arr(ROSEN["Cub"])), ou.P(arr(ROSEN["Dlb"])), ou.P(arr(ROSEN["Dub"])),
                                 composed, ou.P(par), ou.P(x), ou.P(y), ou.P(st))
    assert rc == 0
    assert st[0] == 1  # Converged
    np.testing.assert_allclose(x, [+7.54681194348482909e-01, +4.63926877264482063e-01], rtol=1e-6)  # :78-84
    np.testing.assert_allclose(y, [+1.13829175485413714e-01], rtol=1e-6)


def test_casadi_al_formulas_match_generated_code():
    """psi, grad psi, yhat composed by the oracle (type-erased-problem.tpp defaults) == CasADi's symbolic
    psi / psi_grad_psi on random points, covering inside / below / above D."""
    so = _rosenbrock_so()
    lib = ou.load()
    lib.oracle_eval_casadi.argtypes = [C.c_char_p, C.c_long, C.c_long, C.c_void_p, C.c_long, C.c_void_p,
                                       C.c_void_p, C.c_int] + [C.c_void_p] * 9
    rng = np.random.default_rng(7)
    arr = lambda a: np.array(a, dtype=np.float64)
    seen = set()
    for _ in range(200):
        x, y, S, v = rng.uniform(-2, 3, 2), rng.uniform(-20, 5, 1), rng.uniform(0.1, 5, 1), rng.standard_normal(2)
        out = []
        for composed in (0, 1):
            psi, psi2 = C.c_double(), C.c_double()
            grad, yhat, Hv = np.zeros(2), np.zeros(1), np.zeros(2)
            rc = lib.oracle_eval_casadi(so, 2, 1, ou.P(arr(ROSEN["param"])), 1, ou.P(arr(ROSEN["Dlb"])),
                                        ou.P(arr(ROSEN["Dub"])), composed, ou.P(x), ou.P(y), ou.P(S), ou.P(v),
                                        C.byref(psi), ou.P(grad), ou.P(yhat), ou.P(Hv), C.byref(psi2))
            assert rc == 0
            out.append((psi.value, psi2.value, grad, yhat))
        (p0, q0, g0, y0), (p1, q1, g1, y1) = out
        seen.add(0 if y0[0] == 0 else (1 if y0[0] > 0 else -1))
        assert abs(p0 - p1) <= 1e-12 * max(1, abs(p0)) and abs(q0 - q1) <= 1e-12 * max(1, abs(q0))
        np.testing.assert_allclose(g1, g0, rtol=1e-11, atol=1e-11)
        np.testing.assert_allclose(y1, y0, rtol=1e-12, atol=1e-12)
    assert seen == {-1, 0, 1}
