"""The reference's problem data files: the reader against the reference's own committed data file
(tests/golden/rosenbrock_functions_test.csv, see make_problem_data_fixture.py), the writer byte for byte."""
import os

import numpy as np
import pytest

from pantr_cdc2023_experiments_b200 import problem_data as pd

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "rosenbrock_functions_test.csv")


def test_reader_on_the_reference_data_file():
    d = pd.read_problem_data(GOLDEN, n=2, m=1, p=1)
    np.testing.assert_array_equal(d["C"][0], [0, 0]); np.testing.assert_array_equal(d["C"][1], [5, 5])
    np.testing.assert_array_equal(d["D"][0], [0]); np.testing.assert_array_equal(d["D"][1], [1])
    np.testing.assert_array_equal(d["param"], [2])
    assert d["l1_reg"].size == 0 and d["penalty_alm_split"] == 0


def test_writer_reproduces_the_reference_file_byte_for_byte(tmp_path):
    out = tmp_path / "rosenbrock.csv"
    pd.write_problem_data(out, C=([0, 0], [5, 5]), D=([0], [1]), param=[2], l1_reg=[], penalty_alm_split=0)
    assert open(out, "rb").read() == open(GOLDEN, "rb").read()


def test_round_trip_with_infinite_bounds_and_guess(tmp_path):
    rng = np.random.default_rng(0)
    C = (rng.standard_normal(7), rng.standard_normal(7) + 5)
    D = (np.array([-np.inf, 0.25, -1e-300]), np.array([0.0, np.inf, 1e300]))
    pd.write_problem_data(tmp_path / "p.csv", C=C, D=D, param=rng.standard_normal(4), l1_reg=[0.5], penalty_alm_split=2)
    d = pd.read_problem_data(tmp_path / "p.csv", n=7, m=3, p=4)
    np.testing.assert_array_equal(d["C"][0], C[0]); np.testing.assert_array_equal(d["C"][1], C[1])
    np.testing.assert_array_equal(d["D"][0], D[0]); np.testing.assert_array_equal(d["D"][1], D[1])
    np.testing.assert_array_equal(d["l1_reg"], [0.5]); assert d["penalty_alm_split"] == 2
    g = rng.standard_normal(12)
    pd.write_guess(tmp_path / "p.guess.csv", g, 0)
    g2, split = pd.read_guess(tmp_path / "p.guess.csv")
    np.testing.assert_array_equal(g2, g); assert split == 0


def test_errors_like_the_reference(tmp_path):
    with pytest.raises(RuntimeError, match="Unable to open data file"):
        pd.read_problem_data(tmp_path / "missing.csv")
    with pytest.raises(RuntimeError, match="Failed to open guess file"):
        pd.read_guess(tmp_path / "missing.guess.csv")
    (tmp_path / "bad.csv").write_text("1,2\n3,x\n\n\n\n\n0")
    with pytest.raises(RuntimeError, match="C.upperbound"):
        pd.read_problem_data(tmp_path / "bad.csv")
    (tmp_path / "short.csv").write_text("1,2\n3,4\n\n\n\n\n")
    with pytest.raises(RuntimeError, match="penalty_alm_split"):
        pd.read_problem_data(tmp_path / "short.csv")
    with pytest.raises(RuntimeError, match="expected 3 values"):
        pd.read_problem_data(GOLDEN, n=3)
