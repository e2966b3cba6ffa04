"""The reference's on-disk problem data files.

  <name>.csv        rows: C.lowerbound, C.upperbound, D.lowerbound, D.upperbound, param, l1_reg (may be an
                    empty line), then the integer penalty_alm_split.
                    writer: alpaqa/python/alpaqa/casadi_generator/__init__.py:520-540 (np.savetxt, '%.18e', ',')
                    reader: CasADiProblem::load_numerical_data,
                            alpaqa/src/interop/casadi/include/alpaqa/implementation/casadi/CasADiProblem.tpp:145-192
  <name>.guess.csv  row: initial guess, then penalty_alm_split (problems/generate_problem.py:71-73,
                    drivers/casadi-dll-wrapper.cpp:57-61)

`write_benchmark_problem_files` emits exactly what `problems/generate_problem.py <name> ss2p <N>` writes
next to `<name>_ss2p_<N>.so`, from the device library's own model data (boxes, initial state, guess), so
the reference's loader can be pointed at them; `read_problem_data` is the loader's mirror.
"""
import os

import numpy as np


def _row(fh, x):
    np.savetxt(fh, np.reshape(np.asarray(x, dtype=np.float64), (1, -1)), delimiter=",", newline="\n")


def write_problem_data(path, C=None, D=None, param=None, l1_reg=None, penalty_alm_split=0):
    """casadi_generator.write_casadi_problem_data.  C, D: (lowerbound, upperbound) pairs."""
    C = ([], []) if C is None else C
    D = ([], []) if D is None else D
    with open(path, "w") as fh:
        for v in (C[0], C[1], D[0], D[1], [] if param is None else param, [] if l1_reg is None else l1_reg):
            _row(fh, v)
        fh.write(str(int(penalty_alm_split)))


def _floats(line, name, path, lineno):
    line = line.strip()
    if not line:
        return np.zeros(0)
    try:
        return np.array([float(t) for t in line.split(",")])
    except ValueError as e:
        raise RuntimeError(f'Unable to read {name} from data file "{path}:{lineno}": {e}') from None


def read_problem_data(path, n=None, m=None, p=None):
    """CasADiProblem::load_numerical_data: fixed-size rows (an empty line leaves the default), dynamic
    l1_reg, one integer.  n, m, p (optional) are checked like the reference's fixed-size reads."""
    if not os.path.exists(path):
        raise RuntimeError(f'Unable to open data file "{path}"')
    with open(path) as fh:
        lines = fh.read().split("\n")
    names = ["C.lowerbound", "C.upperbound", "D.lowerbound", "D.upperbound", "param", "l1_reg"]
    sizes = [n, n, m, m, p, None]
    rows = []
    for i, (nm, sz) in enumerate(zip(names, sizes)):
        v = _floats(lines[i] if i < len(lines) else "", nm, path, i + 1)
        if sz is not None and v.size not in (0, sz):
            raise RuntimeError(f'Unable to read {nm} from data file "{path}:{i + 1}": expected {sz} values, got {v.size}')
        rows.append(v)
    rest = "\n".join(lines[6:]).strip()
    try:
        split = int(rest.split()[0]) if rest else None
    except ValueError:
        split = None
    if split is None:
        raise RuntimeError(f'Unable to read penalty_alm_split from data file "{path}:7"')
    return {"C": (rows[0], rows[1]), "D": (rows[2], rows[3]), "param": rows[4], "l1_reg": rows[5],
            "penalty_alm_split": split}


def write_guess(path, initial_guess, penalty_alm_split=0):
    with open(path, "w") as fh:
        _row(fh, initial_guess)
        _row(fh, [penalty_alm_split])


def read_guess(path):
    """drivers/casadi-dll-wrapper.cpp:57-61"""
    if not os.path.exists(path):
        raise RuntimeError("Failed to open guess file")
    with open(path) as fh:
        lines = fh.read().split("\n")
    guess = _floats(lines[0], "initial_guess", path, 1)
    return guess, int(float(lines[1].strip())) if len(lines) > 1 and lines[1].strip() else 0


def write_benchmark_problem_files(problem, directory, formulation="ss2p"):
    """<dir>/<name>_<formulation>_<N>.csv and .guess.csv for a BatchedProblem (needs the device library)."""
    base = os.path.join(directory, f"{problem.name}_{formulation}_{problem.horizon}")
    write_problem_data(base + ".csv", C=problem.get_box_C(), D=problem.get_box_D(), param=problem.initial_state(),
                       l1_reg=[], penalty_alm_split=0)
    write_guess(base + ".guess.csv", problem.initial_guess(), 0)
    return base
