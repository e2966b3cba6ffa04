"""The results file is a Python module `results = {...}` in the reference's format
(drivers/results.hpp:227-251), loadable by new-benchmarks-paper/benchmark_util.py:4-9."""
import importlib.util
import io
import re

import numpy as np

from pantr_cdc2023_experiments_b200.results import (EVAL_KEYS, float_to_str, python_literal, write_results)


def test_float_to_str_matches_print_tpp():
    assert float_to_str(1.0) == "+1.00000000000000000e+00"           # 17 digits, explicit sign
    assert float_to_str(-0.1) == "-1.00000000000000006e-01"
    assert float_to_str(0.0) == "+0.00000000000000000e+00" and float_to_str(-0.0).startswith("-0.0")
    assert float_to_str(float("inf")) == "+inf" and float_to_str(float("-inf")) == "-inf"
    assert float_to_str(float("nan")) == "+nan"
    assert float(float_to_str(np.pi)) == np.pi                        # max_digits10 round trip


def test_python_literals():
    assert python_literal(True) == "True" and python_literal(7) == "7" and python_literal("a\"b") == '"a\\"b"'
    assert python_literal(["x=1", "y"]) == '["x=1", "y", ]'            # output.hpp:49-54
    assert python_literal(np.array([1.0, -2.0])) == "np.array([+1.00000000000000000e+00, -2.00000000000000000e+00])"
    m = python_literal(np.array([[1.0, 2.0], [3.0, 4.0]]))
    assert m.startswith("np.array([[+1.0") and "],\n [" in m and m.endswith(", ]])")  # print.tpp:132-144


def test_results_module_round_trip(tmp_path):
    rng = np.random.default_rng(0)
    runtimes = np.array([0.01, -0.02, np.nan]); states = rng.standard_normal((5, 4)); inputs = rng.standard_normal((6, 3))
    buf = io.StringIO()
    write_results(buf, opts=["method=pantr", "warm=1"], timestamp=1234567890123, runtimes=runtimes, states=states,
                  inputs=inputs, solver="ALMSolver<PANTRSolver<NewtonTRDirection<EigenConfigd>>>",
                  evaluations={"ψ": 3, "hess_ψ_prod": 9}, problem="hanging_chain", nvar=6, ncon=0,
                  path="/p/hanging_chain_ss2p_2.so", horizon=2, second_order=False, formulation="ss2p",
                  extra=[("batch_size", 3)])
    text = buf.getvalue()
    assert text.startswith("from numpy import nan, inf\nimport numpy as np\n__all__ = ['results']\nresults = {\n")
    assert re.search(r'"runtimes": np\.array\(\[\+1\.0{16}2e-02, -2\.0{16}4e-02, \+nan\]\),', text)
    f = tmp_path / "results_x.py"
    f.write_text(text)
    spec = importlib.util.spec_from_file_location("", str(f))     # benchmark_util.load_results
    mod = importlib.util.module_from_spec(spec); spec.loader.exec_module(mod)
    r = mod.results
    assert list(r.keys()) == ["opts", "time_utc_ms", "runtimes", "states", "inputs", "solver", "evaluations",
                              "batch_size", "problem", "nvar", "ncon", "path", "horizon", "second_order", "formulation"]
    np.testing.assert_array_equal(r["states"], states); np.testing.assert_array_equal(r["inputs"], inputs)
    np.testing.assert_array_equal(r["runtimes"][:2], runtimes[:2]); assert np.isnan(r["runtimes"][2])
    assert list(r["evaluations"].keys()) == EVAL_KEYS and r["evaluations"]["hess_ψ_prod"] == 9
    assert r["opts"] == ["method=pantr", "warm=1"] and r["second_order"] is False and r["nvar"] == 6
