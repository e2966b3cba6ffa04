"""Batched analogue of the reference's MPC benchmark driver.

  python -m pantr_cdc2023_experiments_b200.mpc_driver <path>/<name> <horizon> <formulation> [key=value ...]

Same command line, option keys and results file as drivers/mpc-driver.cpp:
  option parsing / solver construction   mpc-driver.cpp:47-73,104-115,326-370
  closed loop (do_mpc_experiment)        mpc-driver.cpp:123-210
  results                                mpc-driver.cpp:298-323, drivers/results.hpp:227-251
plus `batch=<B>` (number of independent closed loops, default 1), `seed=<s>`, `device=<d>` and
`record=<K>` (closed loops whose full input trajectories are written out, default min(B, 4)).

Closed loop b = 0 reproduces the reference run exactly (disturbance input `u_dist` applied
`num_dist` times); loops b >= 1 use disturbance inputs drawn uniformly from [-|u_dist|, |u_dist|]
(SURVEY.md §8d.2).  All B loops advance in lock step: one batched solve, one batched plant step and
one batched warm-start shift per MPC step; iterates stay resident on the device between steps.
"""
import ctypes as C
import os
import sys
import time

import numpy as np

from . import _lib
from .configs import make_solver
from .results import STAT_TO_EVAL, results_name, write_results
from .solver import PROBLEM_IDS, STATS_DTYPE, BatchedProblem


def _get_opt(opts, used, key, default, conv):
    val = default
    for i, o in enumerate(opts):
        if o.startswith(key + "="):
            val = conv(o.split("=", 1)[1])
            used[i] = True
    return val


def _bool(s):
    if s in ("0", "false"):
        return False
    if s in ("1", "true"):
        return True
    raise ValueError(f"Invalid value '{s}' for type 'bool'")


def run(argv):
    if len(argv) < 3:
        print("Usage: mpc_driver <path>/<name> <horizon> <formulation> [key=value ...]")
        return -1 if argv else 0
    path, name = os.path.split(argv[0])
    horizon, formulation, opts = int(argv[1]), argv[2], list(argv[3:])
    if name.startswith("dl:"):
        name = name[3:]
    if name not in PROBLEM_IDS:
        raise ValueError(f"unknown problem '{name}'")
    if formulation != "ss2p":  # new-benchmarks-paper/dodo.py:76: the only formulation PANTR is run with
        raise ValueError("only the single-shooting formulation 'ss2p' is on the device path")
    used = [False] * len(opts)
    method = _get_opt(opts, used, "method", "pantr", str)
    if method != "pantr":
        raise ValueError("only method=pantr is on the device path")          # mpc-driver.cpp:360-373
    solver, used_s = make_solver(opts)
    used = [a or b for a, b in zip(used, used_s)]
    B = _get_opt(opts, used, "batch", 1, int)
    seed = _get_opt(opts, used, "seed", 0xC0FFEE, int)
    device = _get_opt(opts, used, "device", 0, int)
    problem = BatchedProblem(name, horizon, device=device)
    n, m, nx, nu, nc = problem.n, problem.m, problem.nx, problem.nu, problem.nc
    ub = problem.get_box_C()[1][:nu]
    u_dist = np.where(np.isfinite(ub), ub, 0.0)                                # mpc-driver.cpp:140-143
    u_dist = _get_opt(opts, used, "u_dist", u_dist, lambda s: np.array([float(t) for t in s.split(",")]))
    if u_dist.size != nu:
        raise IndexError("Invalid u_dist size")                                # mpc-driver.cpp:145-146
    num_dist = _get_opt(opts, used, "num_dist", 3, int)
    num_sim = _get_opt(opts, used, "num_sim", horizon, int)
    warm = _get_opt(opts, used, "warm", True, _bool)
    record = min(B, _get_opt(opts, used, "record", 4, int))
    batch_npz = _get_opt(opts, used, "batch_npz", B <= 4096, _bool)  # per-instance arrays of all B closed loops (large)
    _get_opt(opts, used, "results_name", None, str)
    unused = [o for o, u in zip(opts, used) if not u]
    if unused:                                                                 # mpc-driver.cpp:403-411
        print("Unused options:", " ".join(unused), file=sys.stderr)
    print("u_dist =", " ".join(f"{v:g}" for v in u_dist), flush=True)

    import torch
    lib = _lib.load()
    dev = torch.device("cuda", device)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    sp = C.c_void_p(stream.cuda_stream)
    vp = lambda t: C.c_void_p(t.data_ptr())
    h = solver._handle(problem).h
    f64 = dict(dtype=torch.float64, device=dev)

    rng = np.random.default_rng(seed)
    dist = rng.uniform(-1.0, 1.0, (B, nu)) * np.abs(u_dist)
    dist[0] = u_dist
    x0 = torch.from_numpy(np.tile(problem.initial_state(), (B, 1))).to(dev)
    xn = torch.empty_like(x0)
    ud = torch.from_numpy(dist).to(dev)
    for _ in range(num_dist):                                                  # mpc-driver.cpp:152-154
        _lib.check(lib.pantr_b200_simulate_dynamics_batch_device(h, B, vp(x0), vp(ud), nu, vp(xn), sp))
        x0, xn = xn, x0
    u_initial = torch.from_numpy(np.tile(problem.initial_guess(), (B, 1))).to(dev)
    u = u_initial.clone()
    y = torch.zeros(B, max(m, 1), **f64)
    stats_t = torch.zeros(B * STATS_DTYPE.itemsize, dtype=torch.uint8, device=dev)
    yarg = y if m else None

    def solve():
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        solver.solve_device(problem, x0, u, yarg, stats_t, stream)
        e1.record(stream)
        e1.synchronize()
        return stats_t.cpu().numpy().view(STATS_DTYPE).copy(), e0.elapsed_time(e1) * 1e-3

    def plant():
        nonlocal x0, xn
        _lib.check(lib.pantr_b200_simulate_dynamics_batch_device(h, B, vp(x0), vp(u), n, vp(xn), sp))
        x0, xn = xn, x0

    solve()                                                                    # initial solve, :158
    plant()                                                                    # :160-161
    # The closed loop itself is enqueued on the stream in one go: shift -> solve -> record -> plant for
    # all num_sim steps with no host round trip; per-step results land in device arrays that are read
    # back once at the end (the reference's loop, mpc-driver.cpp:163-206, per instance).
    isz = STATS_DTYPE.itemsize
    stats_all = torch.zeros(num_sim, B * isz, dtype=torch.uint8, device=dev)
    states_all = torch.empty(num_sim + 1, B, nx, **f64)
    first_all = torch.empty(num_sim, B, nu, **f64)
    rec_all = torch.empty(num_sim, record, n, **f64)
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(num_sim)]
    t_begin = time.perf_counter()
    for i in range(num_sim):
        if warm:                                                               # :168-175
            _lib.check(lib.pantr_b200_shift_warm_start_device(h, B, vp(u), vp(y) if m else None, sp))
        else:                                                                  # :192-197
            u.copy_(u_initial)
            y.zero_()
        evs[i][0].record(stream)
        solver.solve_device(problem, x0, u, yarg, stats_all[i], stream)        # :199
        evs[i][1].record(stream)
        states_all[i].copy_(x0)                                                # :202
        rec_all[i].copy_(u[:record])                                           # :203
        first_all[i].copy_(u[:, :nu])
        plant()                                                                # :205
    states_all[num_sim].copy_(x0)
    torch.cuda.synchronize()
    total = time.perf_counter() - t_begin
    st_all = stats_all.cpu().numpy().view(STATS_DTYPE).reshape(num_sim, B)
    wall = np.array([a.elapsed_time(b) * 1e-3 for a, b in evs])
    t = st_all["elapsed_us"].T * 1e-6                                          # device wall time of each solve (%globaltimer)
    runtimes = np.where(st_all["status"].T == 1, t, -t)                        # :216-222
    iters = st_all["inner_iterations"].T.astype(np.int64)
    evals = np.zeros(B, dtype=[(k, "i8") for k in STAT_TO_EVAL])
    for k in STAT_TO_EVAL:
        evals[k] = st_all[k].sum(axis=0)
    states = states_all.cpu().numpy().transpose(1, 2, 0).copy()               # [B, nx, num_sim + 1]
    inputs = rec_all.cpu().numpy().transpose(1, 2, 0).copy()                   # [record, n, num_sim]
    first_inputs = first_all.cpu().numpy().transpose(1, 2, 0).copy()           # [B, nu, num_sim]

    conv = (runtimes > 0).mean()
    print(f"\nsolver:  {solver.get_name()}\nproblem: {name} {horizon} (from {path!r})\n"
          f"num var: {n}\nnum con: {m}\nbatch:   {B} closed loops x {num_sim} MPC steps\n"
          f"time: {total:.3f} s wall, {B * num_sim / max(wall.sum(), 1e-12):.1f} solves/s on device, "
          f"converged {100 * conv:.2f} %, mean PANTR iterations {iters.mean():.2f}, "
          f"mean Hessian-vector products {evals['ev_hess_psi_prod'].sum() / (B * num_sim):.1f} per solve\n", flush=True)
    rname, ts = results_name(opts)
    extra = [("batch_size", B), ("batch_wall_time_per_step", wall), ("batch_converged_fraction", float(conv)),
             ("batch_mean_inner_iterations", float(iters.mean()))]
    if B > 1 and batch_npz:
        npz = rname + ".npz"
        np.savez_compressed(npz, runtimes=runtimes, states=states, inputs=inputs, first_inputs=first_inputs,
                            inner_iterations=iters, disturbance_inputs=dist,
                            **{"evaluations_" + v: evals[k] for k, v in STAT_TO_EVAL.items()})
        extra.append(("batch_npz", npz))
    with open(rname + ".py", "w") as fh:
        write_results(fh, opts=opts, timestamp=ts, runtimes=runtimes[0], states=states[0], inputs=inputs[0],
                      solver=solver.get_name(), evaluations={v: int(evals[k][0]) for k, v in STAT_TO_EVAL.items()},
                      problem=name, nvar=n, ncon=m,
                      path=os.path.join(path, f"{name}_{formulation}_{horizon}.so"), horizon=horizon,
                      second_order=formulation.endswith("2"), formulation=formulation, extra=extra)
    print("results:", rname[len("results_"):] if rname.startswith("results_") else rname)
    return 0


if __name__ == "__main__":
    sys.exit(run(sys.argv[1:]))
