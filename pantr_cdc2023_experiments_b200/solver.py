"""Host-side mirror of the reference's solver / problem interface for the batched GPU path.

  ALMSolver(ALMParams, PANTRSolver(PANTRParams, NewtonTRDirection(SteihaugCGParams,
            NewtonTRDirectionParams)))         drivers/mpc-driver.cpp:47-73,104-115
  stats = solver(problem, x, y)                alpaqa/outer/alm.hpp:157

Here `problem` is a `BatchedProblem` (B independent instances of one of the three benchmark MPC
problems, each with its own initial state = NLP parameter), `x` is [B, n] and `y` is [B, m].
All compute happens in libpantr_b200.so (CUDA); this file only marshals buffers.
"""
import ctypes as C

import numpy as np

from . import _lib
from .params import (ALMParams, NewtonTRDirectionParams, PANTRParams, SolverStatusName,
                     SteihaugCGParams)

PROBLEM_IDS = {"hanging_chain": 0, "quadcopter": 1, "realistic_quadcopter": 2}


class NewtonTRDirection:
    """inner/directions/pantr/newton-tr.hpp:26-190"""

    def __init__(self, accelerator: SteihaugCGParams = None, direction: NewtonTRDirectionParams = None):
        self.accelerator = accelerator or SteihaugCGParams()
        self.direction = direction or NewtonTRDirectionParams()

    def get_name(self):
        return "NewtonTRDirection<EigenConfigd>"


class PANTRSolver:
    """inner/pantr.hpp:141-200"""

    def __init__(self, params: PANTRParams = None, direction: NewtonTRDirection = None):
        self.params = params or PANTRParams()
        self.direction = direction or NewtonTRDirection()

    def get_name(self):
        return f"PANTRSolver<{self.direction.get_name()}>"


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f64(a, shape, name):
    a = np.ascontiguousarray(a, dtype=np.float64)
    if a.shape != tuple(shape):
        raise ValueError(f"{name}: expected shape {tuple(shape)}, got {a.shape}")
    return a


class BatchedProblem:
    """B instances of `<name>_ss2p_<horizon>` (drivers/casadi-dll-wrapper.cpp:25-67); holds the
    per-instance initial states (the NLP parameter, `initial_state()` in the driver)."""

    def __init__(self, name: str, horizon: int, x0=None, device: int = 0):
        if name not in PROBLEM_IDS:
            raise ValueError(f"unknown problem '{name}'")
        self.name, self.horizon, self.device = name, int(horizon), int(device)
        self.problem_id = PROBLEM_IDS[name]
        self._probe = _Handle(self.problem_id, self.horizon, None, None, None, None, self.device)
        self.n, self.m, self.nx, self.nu, self.nc = self._probe.dims
        self.x0 = None if x0 is None else _f64(x0, (np.shape(x0)[0], self.nx), "x0")

    def set_debug(self, flags, grid_limit=0):
        """Memory checks for the problem-function launches of this object (pantr_b200_set_debug)."""
        _lib.check(_lib.load().pantr_b200_set_debug(self._probe.h, flags, grid_limit))

    def check_guards(self):
        bad = C.c_int64(-1)
        _lib.check(_lib.load().pantr_b200_check_guards(self._probe.h, C.byref(bad)))
        return bad.value

    # TypeErasedProblem::get_n / get_m / get_box_C / get_box_D
    def get_n(self): return self.n
    def get_m(self): return self.m

    def get_box_C(self):
        lb, ub = np.empty(self.n), np.empty(self.n)
        _lib.check(_lib.load().pantr_b200_get_box_C(self._probe.h, _ptr(lb), _ptr(ub)))
        return lb, ub

    def get_box_D(self):
        lb, ub = np.empty(self.m), np.empty(self.m)
        _lib.check(_lib.load().pantr_b200_get_box_D(self._probe.h, _ptr(lb), _ptr(ub)))
        return lb, ub

    def initial_state(self):
        x0 = np.empty(self.nx)
        _lib.check(_lib.load().pantr_b200_get_initial(self._probe.h, _ptr(x0), None))
        return x0

    def initial_guess(self):
        u0 = np.empty(self.n)
        _lib.check(_lib.load().pantr_b200_get_initial(self._probe.h, None, _ptr(u0)))
        return u0

    # batched TypeErasedProblem::eval_psi_grad_psi / eval_psi / eval_hess_psi_prod
    def eval(self, x0, u, y=None, Sigma=None, v=None, want_grad=True, want_yhat=False):
        B = np.shape(u)[0]
        x0 = _f64(x0, (B, self.nx), "x0"); u = _f64(u, (B, self.n), "u")
        y = _f64(y, (B, self.m), "y") if self.m else None
        Sigma = _f64(Sigma, (B, self.m), "Sigma") if self.m else None
        v = None if v is None else _f64(v, (B, self.n), "v")
        psi = np.empty(B)
        grad = np.empty((B, self.n)) if want_grad else None
        yhat = np.empty((B, self.m)) if (want_yhat and self.m) else None
        Hv = np.empty((B, self.n)) if v is not None else None
        _lib.check(_lib.load().pantr_b200_eval_batch(self._probe.h, B, _ptr(x0), _ptr(u), _ptr(y),
                                                     _ptr(Sigma), _ptr(v), _ptr(psi), _ptr(grad),
                                                     _ptr(yhat), _ptr(Hv)))
        return psi, grad, yhat, Hv

    # ---- TypeErasedProblem API surface, batched (problem/type-erased-problem.hpp:228-571): every method
    # takes the per-instance initial states x0 [B, nx] first (the NLP parameter), then the reference's
    # arguments with a leading batch axis, and returns the reference's outputs. ----
    def _xy(self, x0, x, y=None, Sigma=None, v=None):
        B = np.shape(x)[0]
        x0 = self.x0 if x0 is None else x0
        out = [B, _f64(x0, (B, self.nx), "x0"), _f64(x, (B, self.n), "x")]
        for a, shape, nm in ((y, (B, self.m), "y"), (Sigma, (B, self.m), "Sigma"), (v, (B, self.n), "v")):
            out.append(None if (a is None or shape[1] == 0) else _f64(a, shape, nm))
        return out

    def eval_f(self, x, x0=None):
        B, x0, x, *_ = self._xy(x0, x); f = np.empty(B)
        _lib.check(_lib.load().pantr_b200_eval_f_batch(self._probe.h, B, _ptr(x0), _ptr(x), _ptr(f)))
        return f

    def eval_grad_f(self, x, x0=None):
        B, x0, x, *_ = self._xy(x0, x); g = np.empty((B, self.n))
        _lib.check(_lib.load().pantr_b200_eval_grad_f_batch(self._probe.h, B, _ptr(x0), _ptr(x), _ptr(g)))
        return g

    def eval_f_grad_f(self, x, x0=None):
        B, x0, x, *_ = self._xy(x0, x); f = np.empty(B); g = np.empty((B, self.n))
        _lib.check(_lib.load().pantr_b200_eval_f_grad_f_batch(self._probe.h, B, _ptr(x0), _ptr(x), _ptr(f), _ptr(g)))
        return f, g

    def eval_g(self, x, x0=None):
        B, x0, x, *_ = self._xy(x0, x); g = np.empty((B, self.m))
        _lib.check(_lib.load().pantr_b200_eval_g_batch(self._probe.h, B, _ptr(x0), _ptr(x), _ptr(g)))
        return g

    def eval_grad_g_prod(self, x, y, x0=None):
        B, x0, x, y, *_ = self._xy(x0, x, y); out = np.empty((B, self.n))
        _lib.check(_lib.load().pantr_b200_eval_grad_g_prod_batch(self._probe.h, B, _ptr(x0), _ptr(x), _ptr(y), _ptr(out)))
        return out

    def eval_grad_L(self, x, y, x0=None):
        B, x0, x, y, *_ = self._xy(x0, x, y); out = np.empty((B, self.n))
        _lib.check(_lib.load().pantr_b200_eval_grad_L_batch(self._probe.h, B, _ptr(x0), _ptr(x), _ptr(y), _ptr(out)))
        return out

    def eval_hess_L_prod(self, x, y, scale, v, x0=None):
        B, x0, x, y, _, v = self._xy(x0, x, y, None, v); Hv = np.empty((B, self.n))
        _lib.check(_lib.load().pantr_b200_eval_hess_L_prod_batch(self._probe.h, B, _ptr(x0), _ptr(x), _ptr(y),
                                                                 float(scale), _ptr(v), _ptr(Hv)))
        return Hv

    def eval_hess_psi_prod(self, x, y, Sigma, scale, v, x0=None):
        B, x0, x, y, Sigma, v = self._xy(x0, x, y, Sigma, v); Hv = np.empty((B, self.n))
        _lib.check(_lib.load().pantr_b200_eval_hess_psi_prod_batch(self._probe.h, B, _ptr(x0), _ptr(x), _ptr(y),
                                                                   _ptr(Sigma), float(scale), _ptr(v), _ptr(Hv)))
        return Hv

    def eval_psi(self, x, y, Sigma, x0=None):
        B, x0, x, y, Sigma, _ = self._xy(x0, x, y, Sigma); psi = np.empty(B)
        yhat = np.empty((B, self.m))
        _lib.check(_lib.load().pantr_b200_eval_psi_batch(self._probe.h, B, _ptr(x0), _ptr(x), _ptr(y), _ptr(Sigma),
                                                         _ptr(yhat) if self.m else None, _ptr(psi)))
        return psi, yhat

    def eval_grad_psi(self, x, y, Sigma, x0=None):
        B, x0, x, y, Sigma, _ = self._xy(x0, x, y, Sigma); g = np.empty((B, self.n))
        _lib.check(_lib.load().pantr_b200_eval_grad_psi_batch(self._probe.h, B, _ptr(x0), _ptr(x), _ptr(y),
                                                              _ptr(Sigma), _ptr(g)))
        return g

    def eval_psi_grad_psi(self, x, y, Sigma, x0=None):
        B, x0, x, y, Sigma, _ = self._xy(x0, x, y, Sigma); psi = np.empty(B); g = np.empty((B, self.n))
        _lib.check(_lib.load().pantr_b200_eval_psi_grad_psi_batch(self._probe.h, B, _ptr(x0), _ptr(x), _ptr(y),
                                                                  _ptr(Sigma), _ptr(psi), _ptr(g)))
        return psi, g

    def eval_prox_grad_step(self, gamma, x, grad_psi):
        B = np.shape(x)[0]
        x = _f64(x, (B, self.n), "x"); g = _f64(grad_psi, (B, self.n), "grad_psi")
        xh, p = np.empty((B, self.n)), np.empty((B, self.n))
        _lib.check(_lib.load().pantr_b200_eval_prox_grad_step_batch(self._probe.h, B, float(gamma), _ptr(x), _ptr(g),
                                                                    _ptr(xh), _ptr(p)))
        return np.zeros(B), xh, p  # h(x_hat) = 0: no l1 term

    def eval_proj_diff_g(self, z):
        B = np.shape(z)[0]
        z = _f64(z, (B, self.m), "z"); e = np.empty((B, self.m))
        _lib.check(_lib.load().pantr_b200_eval_proj_diff_g_batch(self._probe.h, B, _ptr(z), _ptr(e)))
        return e

    def eval_proj_multipliers(self, y, M):
        B = np.shape(y)[0]
        y = np.array(_f64(y, (B, self.m), "y"))
        _lib.check(_lib.load().pantr_b200_eval_proj_multipliers_batch(self._probe.h, B, _ptr(y), float(M)))
        return y

    def eval_inactive_indices_res_lna(self, gamma, x, grad_psi):
        B = np.shape(x)[0]
        x = _f64(x, (B, self.n), "x"); g = _f64(grad_psi, (B, self.n), "grad_psi")
        J = np.zeros((B, self.n), dtype=np.int32); nJ = np.zeros(B, dtype=np.int32)
        _lib.check(_lib.load().pantr_b200_eval_inactive_indices_res_lna_batch(self._probe.h, B, float(gamma), _ptr(x),
                                                                              _ptr(g), _ptr(J), _ptr(nJ)))
        return [J[b, :nJ[b]].copy() for b in range(B)]

    # "simulate_dynamics" extra function (problems/problems/formulation.hpp:258-261)
    def simulate_dynamics(self, x, u):
        B = np.shape(x)[0]
        x = _f64(x, (B, self.nx), "x"); u = _f64(u, (B, self.nu), "u")
        xn = np.empty_like(x)
        _lib.check(_lib.load().pantr_b200_simulate_dynamics_batch(self._probe.h, B, _ptr(x), _ptr(u), _ptr(xn)))
        return xn


class _Handle:
    def __init__(self, problem_id, horizon, alm, pantr, cg, dirp, device):
        lib = _lib.load()
        self.h = C.c_void_p()
        ref = lambda p: None if p is None else C.byref(p)
        _lib.check(lib.pantr_b200_create(problem_id, horizon, ref(alm), ref(pantr), ref(cg), ref(dirp),
                                         device, C.byref(self.h)))
        dims = (C.c_int64 * 5)()
        _lib.check(lib.pantr_b200_get_dims(self.h, dims))
        self.dims = tuple(int(d) for d in dims)

    def __del__(self):
        try:
            if self.h:
                _lib.load().pantr_b200_destroy(self.h)
                self.h = None
        except Exception:
            pass


STATS_DTYPE = np.dtype([(k, {C.c_int32: "i4", C.c_uint32: "u4", C.c_double: "f8", C.c_uint64: "u8"}[t])
                        for k, t in _lib.StatsC._fields_], align=True)
assert STATS_DTYPE.itemsize == C.sizeof(_lib.StatsC)


class ALMSolver:
    """outer/alm.hpp:94-189 for batches.  `solver(problem, x, y)` solves every instance of the
    batch; x [B, n] and y [B, m] are updated in place and a structured array of per-instance
    stats (ALMSolver::Stats with the accumulated inner stats and evaluation counters) is returned."""

    def __init__(self, params: ALMParams = None, inner_solver: PANTRSolver = None):
        self.params = params or ALMParams()
        self.inner_solver = inner_solver or PANTRSolver()
        self._handles = {}

    def get_name(self):
        return f"ALMSolver<{self.inner_solver.get_name()}>"

    def _handle(self, problem: BatchedProblem):
        key = (problem.problem_id, problem.horizon, problem.device)
        if key not in self._handles:
            inner = self.inner_solver
            self._handles[key] = _Handle(problem.problem_id, problem.horizon, self.params.to_c(),
                                         inner.params.to_c(), inner.direction.accelerator.to_c(),
                                         inner.direction.direction.to_c(), problem.device)
        return self._handles[key]

    def invalidate(self):
        """Call after changing params (handles capture params at creation)."""
        self._handles.clear()

    def __call__(self, problem: BatchedProblem, x, y=None, x0=None):
        x0 = problem.x0 if x0 is None else x0
        if x0 is None:
            raise ValueError("problem has no initial states")
        B = np.shape(x)[0]
        if not (isinstance(x, np.ndarray) and x.dtype == np.float64 and x.flags.c_contiguous):
            raise ValueError("x must be a C-contiguous float64 array (updated in place)")
        x0 = _f64(x0, (B, problem.nx), "x0")
        if x.shape != (B, problem.n):
            raise ValueError(f"x: expected shape {(B, problem.n)}, got {x.shape}")
        if problem.m:
            if y is None or y.shape != (B, problem.m) or y.dtype != np.float64 or not y.flags.c_contiguous:
                raise ValueError(f"y must be a C-contiguous float64 array of shape {(B, problem.m)}")
        stats = np.zeros(B, dtype=STATS_DTYPE)
        h = self._handle(problem)
        _lib.check(_lib.load().pantr_b200_solve_batch(h.h, B, _ptr(x0), _ptr(x), _ptr(y) if problem.m else None,
                                                      _ptr(stats)))
        return stats

    # ---- device-resident variant (torch tensors on the handle's CUDA device) ----
    def solve_device(self, problem: BatchedProblem, x0_t, x_t, y_t, stats_t, stream=None):
        """x0_t [B,nx], x_t [B,n], y_t [B,m] or None: float64 CUDA tensors; stats_t: uint8 CUDA
        tensor of B*sizeof(stats) bytes.  Asynchronous on `stream` (torch.cuda.Stream or None)."""
        h = self._handle(problem)
        B = x_t.shape[0]
        sp = C.c_void_p(stream.cuda_stream) if stream is not None else None
        _lib.check(_lib.load().pantr_b200_solve_batch_device(
            h.h, B, C.c_void_p(x0_t.data_ptr()), C.c_void_p(x_t.data_ptr()),
            C.c_void_p(y_t.data_ptr()) if (y_t is not None and problem.m) else None,
            C.c_void_p(stats_t.data_ptr()), sp))

    def set_order(self, problem, order_t):
        """Scheduling hint: int32 CUDA tensor with a permutation of range(B) (kept alive here), or None."""
        self._order = order_t
        if order_t is not None and (str(order_t.dtype) != "torch.int32" or not order_t.is_contiguous() or not order_t.is_cuda):
            raise ValueError("order must be a contiguous int32 CUDA tensor")
        ptr = None if order_t is None else C.c_void_p(order_t.data_ptr())
        _lib.check(_lib.load().pantr_b200_set_order_device(self._handle(problem).h, ptr, 0 if order_t is None else order_t.numel()))

    def set_exec_mode(self, problem, mode):
        """'auto' = 'warp' (one warp per instance; quadcopters: condensed formulation) | 'thread' (one thread per
        instance, forward-over-adjoint sweeps: quadcopters only, the round-1 path kept for A/B runs)."""
        code = {"auto": 0, "warp": 1, "thread": 2}[mode]
        _lib.check(_lib.load().pantr_b200_set_exec_mode(self._handle(problem).h, code))

    def set_debug(self, problem, flags, grid_limit=0):
        """flags: 1 = NaN-poison the workspace before every launch, 2 = guard bands (see check_guards)."""
        _lib.check(_lib.load().pantr_b200_set_debug(self._handle(problem).h, flags, grid_limit))

    def check_guards(self, problem):
        """Number of guard-band doubles overwritten since the last launch (0 = no workspace overrun)."""
        bad = C.c_int64(-1)
        _lib.check(_lib.load().pantr_b200_check_guards(self._handle(problem).h, C.byref(bad)))
        return bad.value

    def set_trace(self, problem, instance, max_rows=4096):
        _lib.check(_lib.load().pantr_b200_set_trace(self._handle(problem).h, instance, max_rows))

    def get_trace(self, problem, max_rows=4096):
        rows = np.zeros((max_rows, 11)); n = C.c_int()
        _lib.check(_lib.load().pantr_b200_get_trace(self._handle(problem).h, _ptr(rows), max_rows, C.byref(n)))
        return rows[:min(n.value, max_rows)]

    def kernel_launches(self, problem):
        return int(_lib.load().pantr_b200_kernel_launches(self._handle(problem).h))

    def grid_size(self, problem):
        return int(_lib.load().pantr_b200_grid_size(self._handle(problem).h))


def status_name(code):
    return SolverStatusName.get(int(code), "Unknown")
