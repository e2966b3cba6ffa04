"""Host-side logic and the C-ABI surface (no GPU): option parsing identical to the reference's
key=value system, the benchmark option sets, library exports, error behaviour without a device."""
import ctypes as C
import math
import os
import re

import numpy as np
import pytest

import pantr_cdc2023_experiments_b200 as pb
from pantr_cdc2023_experiments_b200 import _lib, flops, sharding

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    header = open(os.path.join(ROOT, "include", "pantr_b200.h")).read()
    declared = set(re.findall(r"\b(pantr_b200_[A-Za-z0-9_]+)\s*\(", header))
    assert declared == set(_lib.SYMBOLS), declared ^ set(_lib.SYMBOLS)
    lib = C.CDLL(_lib.LIB_PATH)
    for sym in declared:
        assert hasattr(lib, sym), sym


def test_default_params_match_reference_defaults():
    lib = _lib.load()
    a, p, c, d = _lib.ALMParamsC(), _lib.PANTRParamsC(), _lib.SteihaugParamsC(), _lib.NewtonTRParamsC()
    lib.pantr_b200_default_params(C.byref(a), C.byref(p), C.byref(c), C.byref(d))
    pa, pp, pc, pd = pb.ALMParams(), pb.PANTRParams(), pb.SteihaugCGParams(), pb.NewtonTRDirectionParams()
    assert (a.tolerance, a.penalty_update_factor, a.max_iter, a.max_penalty) == (1e-5, 10, 100, 1e9)  # alm.hpp:19-90
    assert (p.max_iter, p.stop_crit, p.L_max) == (100, 0, 1e20)                                       # pantr.hpp:22-91
    assert p.Lipschitz.Lgamma_factor == pp.Lipschitz.Lγ_factor == 0.95                                # lipschitz.hpp:20
    assert math.isnan(p.initial_radius) and p.min_radius == 100 * 2.220446049250313e-16
    assert (c.tol_scale, c.tol_scale_root, c.max_iter_factor) == (pc.tol_scale, pc.tol_scale_root, 1)
    assert math.isinf(c.tol_max) and d.hessian_vec_factor == pd.hessian_vec_factor == 0.5
    for f, _ in _lib.ALMParamsC._fields_:
        assert getattr(a, f) == getattr(pa.to_c(), f), f


def test_no_device_is_an_error_not_a_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(_lib.PantrB200Error) as e:
        pb.BatchedProblem("hanging_chain", 10)
    assert e.value.code == -3 and "no CPU fallback" in str(e.value)


def test_benchmark_option_set_and_solver_construction():
    opts = pb.benchmark_options("hanging_chain")  # dodo.py:12-33,94-105
    solver, used = pb.make_solver(opts)
    a, p = solver.params, solver.inner_solver.params
    assert (a.max_iter, a.initial_penalty, a.penalty_update_factor, a.tolerance) == (300, 1e4, 5, 1e-8)
    assert a.max_time == 30.0 and a.initial_tolerance == 100 and a.dual_tolerance == 1e-8
    assert p.stop_crit == "ProjGradUnitNorm" and p.max_iter == 50000          # mpc-driver.cpp:50-54 default kept
    assert (p.radius_factor_rejected, p.radius_factor_acceptable, p.radius_factor_good) == (0.35, 0.99, 10)
    assert (p.ratio_threshold_acceptable, p.ratio_threshold_good) == (0.3, 0.5)
    assert solver.inner_solver.direction.direction.hessian_vec_factor == 1
    assert solver.inner_solver.direction.accelerator.tol_scale == 1e-2
    assert [o for o, u in zip(opts, used) if not u] == ["method=pantr", "num_dist=5", "u_dist=-1,1,1"]
    assert solver.get_name() == "ALMSolver<PANTRSolver<NewtonTRDirection<EigenConfigd>>>"  # alm.hpp:163-165
    q, _ = pb.make_solver(pb.benchmark_options("realistic_quadcopter"))
    assert q.inner_solver.params.max_iter == 250 and q.inner_solver.direction.accelerator.tol_scale == 1
    assert q.params.initial_tolerance == 0.1


def test_set_params_rules():
    a = pb.ALMParams()
    pb.set_params(a, "alm", ["alm.max_time=250ms", "alm.single_penalty_factor=true", "solver.max_iter=3"])
    assert a.max_time == 0.25 and a.single_penalty_factor is True           # params.cpp:113-142, :99-111
    p = pb.PANTRParams()
    pb.set_params(p, "solver", ["solver.Lipschitz.ε=1e-3", "solver.max_iter=7", "solver.stop_crit=FPRNorm2"])
    assert p.Lipschitz.ε == 1e-3 and p.max_iter == 7 and p.stop_crit == "FPRNorm2"
    for bad in (["alm.nonexistent=1"], ["alm.max_iter=abc"], ["alm.max_time=3 parsecs"],
                ["alm.single_penalty_factor=maybe"]):
        with pytest.raises(ValueError):
            pb.set_params(pb.ALMParams(), "alm", bad)
    with pytest.raises(ValueError):
        pb.set_params(pb.PANTRParams(), "solver", ["solver.stop_crit=Nope"])


def test_enum_values_match_reference():
    assert pb.SolverStatus["Converged"] == 1 and pb.SolverStatus["MaxIter"] == 3 and pb.SolverStatus["NotFinite"] == 4
    assert pb.PANOCStopCrit["ProjGradUnitNorm"] == 4 and pb.PANOCStopCrit["FPRNorm"] == 6


def test_flop_model():
    F = flops.hanging_chain(60)
    assert F["f"] == 1915 * 60 + 124 and F["psi_grad_psi"] == 3819 * 60 + 157 and F["hess_psi_prod"] == 8829 * 60 + 63
    st = np.zeros(3, dtype=pb.STATS_DTYPE)
    st["ev_psi"], st["ev_psi_grad_psi"], st["ev_grad_psi"], st["ev_hess_psi_prod"] = 2, 3, 1, 10
    assert flops.solve_flops("hanging_chain", 60, st) == 3 * (2 * F["f"] + 4 * F["psi_grad_psi"] + 10 * F["hess_psi_prod"])


def test_shard_ranges_cover_batch():
    for B, W in ((16384, 8), (10, 4), (3, 8), (0, 2), (1000003, 8)):
        r = [sharding.shard_range(B, k, W) for k in range(W)]
        assert r[0][0] == 0 and r[-1][1] == B
        assert all(a[1] == b[0] for a, b in zip(r[:-1], r[1:])) and all(lo <= hi for lo, hi in r)


def test_traffic_json_is_keyed_by_library_and_source_hash():
    """profiles/traffic.json (tools/make_traffic.py): every capture names the library AND the sources it was taken from;
    bench.py reports roofline.traffic only for a matching build."""
    import json
    import os
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, os.path.join(root, "tools"))
    from source_hash import source_sha256
    caps = json.load(open(os.path.join(root, "profiles", "traffic.json")))["captures"]
    assert caps
    for c in caps:
        assert len(c["lib_sha256"]) == 64 and len(c["src_sha256"]) == 64
        assert c["dram_bytes_per_launch"] == c["dram_bytes_read"] + c["dram_bytes_write"]
        assert {"problem", "batch", "horizon", "start", "source"} <= set(c)
    assert len(source_sha256()) == 64 and source_sha256() == source_sha256()


def test_bench_config_of_a_globally_sharded_batch():
    import argparse
    import os
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    a = argparse.Namespace(config="hanging_chain", batch=32768, horizon=60, horizons="", scale=0.1, shard="global")
    d = bench.config_dict(a, 2)
    assert d["global_batch"] == 32768 and d["batch_per_gpu"] == 16384 and "ONE global batch" in d["workload"]
    a.shard = "weak"
    d = bench.config_dict(a, 2)
    assert d["global_batch"] == 65536 and d["batch_per_gpu"] == 32768 and "per GPU" in d["workload"]
