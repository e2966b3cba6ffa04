"""ctypes access to tests/plugin_harness.so (the stand-in for alpaqa's DLProblem loader) -- test infrastructure."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HARNESS = os.path.join(ROOT, "tests", "plugin_harness.so")
PLUGIN_DIR = os.path.join(ROOT, "pantr_cdc2023_experiments_b200", "plugins")
# order of the bits harness_provided() returns = order of the members in alpaqa_problem_functions_t (dl-problem.h:18-153)
MEMBERS = ["eval_f", "eval_grad_f", "eval_g", "eval_grad_g_prod", "eval_jac_g", "get_jac_g_num_nonzeros", "eval_grad_gi",
           "eval_hess_L_prod", "eval_hess_L", "get_hess_L_num_nonzeros", "eval_hess_psi_prod", "eval_hess_psi",
           "get_hess_psi_num_nonzeros", "eval_f_grad_f", "eval_f_g", "eval_grad_f_grad_g_prod", "eval_grad_L", "eval_psi",
           "eval_grad_psi", "eval_psi_grad_psi", "eval_prox_grad_step", "initialize_box_C", "initialize_box_D",
           "initialize_l1_reg"]


def available():
    return os.path.exists(HARNESS) and all(os.path.exists(plugin_path(n)) for n in ("hanging_chain", "quadcopter", "realistic_quadcopter"))


def plugin_path(name):
    return os.path.join(PLUGIN_DIR, name + ".so")


_h = None


def harness():
    global _h
    if _h is None:
        _h = C.CDLL(HARNESS)
        _h.harness_open.restype = C.c_void_p
        _h.harness_open.argtypes = [C.c_char_p, C.c_int, C.c_int]
        _h.harness_close.argtypes = [C.c_void_p]
        _h.harness_last_error.restype = C.c_char_p
        for f in ("harness_n", "harness_m"):
            getattr(_h, f).restype = C.c_long
            getattr(_h, f).argtypes = [C.c_void_p]
        _h.harness_provided.restype = C.c_uint
        _h.harness_provided.argtypes = [C.c_void_p]
        _h.harness_call.argtypes = [C.c_void_p, C.c_char_p] + [C.c_void_p] * 4 + [C.c_double] + [C.c_void_p] * 7
        _h.harness_extra.argtypes = [C.c_void_p, C.c_char_p] + [C.c_void_p] * 4
    return _h


def P(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class LoadedProblem:
    """One plugin instance, loaded the way DLProblem loads it."""

    def __init__(self, name, size, null_user_data=False):
        self.p = harness().harness_open(plugin_path(name).encode(), size, int(null_user_data))
        if not self.p:
            raise RuntimeError(harness().harness_last_error().decode())
        self.n, self.m = harness().harness_n(self.p), harness().harness_m(self.p)

    def close(self):
        if self.p:
            harness().harness_close(self.p)
            self.p = None

    def provided(self):
        mask = harness().harness_provided(self.p)
        return {nm for i, nm in enumerate(MEMBERS) if mask >> i & 1}

    def call(self, name, x=None, y=None, S=None, v=None, scale=1.0, zl=None, zu=None):
        s = C.c_double(np.nan)
        on, om, wn, wm = np.full(self.n, np.nan), np.full(max(self.m, 1), np.nan), np.zeros(self.n), np.zeros(max(self.m, 1))
        rc = harness().harness_call(self.p, name.encode(), P(x), P(y), P(S), P(v), scale, P(zl), P(zu), C.byref(s), P(on), P(om),
                                    P(wn), P(wm))
        if rc:
            raise RuntimeError(harness().harness_last_error().decode())
        return {"s": s.value, "n": on, "m": om[:self.m], "work_n": wn, "work_m": wm[:self.m]}

    def extra(self, name, in1=None, in2=None, out_len=0):
        out = np.full(max(out_len, 1), np.nan)
        ln = C.c_long(-1)
        rc = harness().harness_extra(self.p, name.encode(), P(in1), P(in2), P(out), C.byref(ln))
        if rc:
            raise RuntimeError(harness().harness_last_error().decode())
        return out[:out_len], ln.value
