"""How far do two correct evaluations of the same problem drift apart?  

The north_star asks for "same status, iterations within +-1, objective and iterate within 1e-8 relative" against the
reference, whose CasADi-generated arithmetic (operation order, contraction) cannot be reproduced here.  This script
measures what such a tolerance can mean, on the CPU oracle only:
  (a) hanging chain N = 60, the bench's warm instances: strict build (-ffp-contract=off, the device's arithmetic)
      vs the reference's flag policy (-O3 -march=native -ffp-contract=fast): same algorithm, same operation order,
      only the contraction of a*b+c differs;
  (b) quadcopters N = 60, the bench's cold instances: forward-over-adjoint formulation (what CasADi's AD generates)
      vs the condensed formulation the device runs: same exact derivatives, different association of the sums.
Writes profiles/r02_fp_sensitivity.json.  Test infrastructure (oracle only).
usage: python scripts/fp_sensitivity.py [instances_hc] [instances_quad]
"""
import ctypes as C, json, os, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import oracle_util as ou
import bench

S_HC = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
S_Q = int(sys.argv[2]) if len(sys.argv) > 2 else 256
cores = os.cpu_count()
odir = os.path.join(ROOT, "oracle")
subprocess.check_call(["make", "-C", odir, "liboracle.so", "liboracle_fast.so"], stdout=subprocess.DEVNULL)
strict, fast = ou.load(), C.CDLL(os.path.join(odir, "liboracle_fast.so"))


def compare(tag, a, b):
    (ua, sa), (ub, sb) = a, b
    both = (sa[:, 0] == 1) & (sb[:, 0] == 1)
    du = np.abs(ua - ub).max(axis=1) / np.maximum(1.0, np.abs(ua).max(axis=1))
    dpsi = np.abs(sa[:, 6] - sb[:, 6]) / np.maximum(1e-300, np.abs(sa[:, 6]))
    dit = np.abs(sa[:, 2] - sb[:, 2])
    q = lambda v: {p: float(np.percentile(v, p)) for p in (50, 90, 99, 100)}
    return {"pair": tag, "instances": int(len(sa)), "status_differs": int((sa[:, 0] != sb[:, 0]).sum()),
            "converged_both": int(both.sum()),
            "inner_iterations_equal": int((dit == 0).sum()), "inner_iterations_within_1": int((dit <= 1).sum()),
            "inner_iterations_abs_diff_percentiles": q(dit),
            "inner_iterations_rel_diff_percentiles": q(dit / np.maximum(1, sa[:, 2])),
            "u_rel_err_percentiles_converged": q(du[both]) if both.any() else None,
            "psi_rel_err_percentiles_converged": q(dpsi[both]) if both.any() else None,
            "u_within_1e-8": int((du[both] <= 1e-8).sum()), "psi_within_1e-8": int((dpsi[both] <= 1e-8).sum()),
            "mean_inner_iterations": [float(sa[:, 2].mean()), float(sb[:, 2].mean())]}


class A:  # the bench's argument set
    pass


out = {"cores": cores, "cases": []}
t0 = time.time()
# (a) hanging chain, warm bench instances
a = A(); a.config = "hanging_chain"; a.batch = 16384; a.horizon = 60; a.scale = 0.1; a.horizons = ""
par, x1, uw, m = bench.cpu_segment_inputs(strict, a, 0, "hanging_chain", 60, S_HC, cores)
res = {}
for nm, lib in (("strict", strict), ("fast", fast)):
    u, st, t = bench.cpu_solve_batch(lib, 0, 60, par, x1, uw, 0, cores)
    res[nm] = (u, st)
out["cases"].append(dict(compare("hanging chain N=60 warm: -ffp-contract=off vs -O3 -march=native -ffp-contract=fast", res["strict"], res["fast"]),
                         problem="hanging_chain"))
print(json.dumps(out["cases"][-1]), flush=True)
# (b) quadcopters, cold bench instances: formulation pair (strict build) and contraction pair (condensed formulation)
for cfg, name, ref_id, dev_id in (("quadcopter", "quadcopter", 1, 5), ("realistic_sweep", "realistic_quadcopter", 2, 6)):
    a = A(); a.config = cfg; a.batch = 65536; a.horizon = 60; a.scale = 0.1; a.horizons = "60"
    par, x0, u0, m = bench.cpu_segment_inputs(strict, a, 0, name, 60, S_Q, cores)
    r = {}
    for nm, lib, pid in (("foa_strict", strict, ref_id), ("condensed_strict", strict, dev_id), ("condensed_fast", fast, dev_id)):
        u, st, t = bench.cpu_solve_batch(lib, pid, 60, par, x0, u0, m, cores)
        r[nm] = (u, st)
    out["cases"].append(dict(compare(f"{name} N=60 cold: forward-over-adjoint vs condensed formulation (both strict)", r["foa_strict"], r["condensed_strict"]), problem=name))
    print(json.dumps(out["cases"][-1]), flush=True)
    out["cases"].append(dict(compare(f"{name} N=60 cold, condensed formulation: -ffp-contract=off vs fast", r["condensed_strict"], r["condensed_fast"]), problem=name))
    print(json.dumps(out["cases"][-1]), flush=True)
out["seconds"] = time.time() - t0
with open(os.path.join(ROOT, "profiles", "r02_fp_sensitivity.json"), "w") as fh:
    json.dump(out, fh, indent=1)
