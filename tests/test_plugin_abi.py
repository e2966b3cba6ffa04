"""The reference's problem plugin ABI (alpaqa_problem_functions_t + benchmark_problem_register,
alpaqa/src/interop/dl-api/include/alpaqa/dl/dl-problem.h:13-169) as exported by
pantr_cdc2023_experiments_b200/plugins/<problem>.so.  CPU part: the objects exist, export exactly the symbol the
loader looks up, and fail the way the reference's plugins fail; the calls themselves need a GPU (test_gpu_plugin.py)."""
import os
import subprocess

import pytest

import plugin_util as pu

needs_build = pytest.mark.skipif(not pu.available(), reason="plugins / harness not built (they need the reference headers: "
                                 "make -C pantr_cdc2023_experiments_b200/csrc/plugin where /root/reference is mounted)")


@needs_build
@pytest.mark.parametrize("name", ["hanging_chain", "quadcopter", "realistic_quadcopter"])
def test_plugin_exports_only_the_registration_symbol(name):
    out = subprocess.check_output(["nm", "-D", "--defined-only", pu.plugin_path(name)], text=True)
    syms = [ln.split()[-1] for ln in out.splitlines() if " T " in ln]
    assert syms == ["benchmark_problem_register"]  # dl-problem.cpp:19-29: dlsym("<prefix>_register")


@needs_build
def test_missing_user_data_throws_invalid_argument():
    """problems/problems/hanging_chain.cpp:141-142"""
    with pytest.raises(RuntimeError, match="invalid_argument: Missing user data"):
        pu.LoadedProblem("hanging_chain", 10, null_user_data=True)


@needs_build
def test_no_device_is_an_exception_not_a_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(RuntimeError, match="no CUDA device"):
        pu.LoadedProblem("quadcopter", 10)
