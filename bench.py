#!/usr/bin/env python
"""bench.py — converged MPC solves/s of the batched PANTR+ALM hot path (hanging chain N=60).

  python bench.py --gpus N --steps K --warmup W            (our arm, one rank per GPU under torchrun)
  python bench.py --impl reference --gpus N --steps K ...  (CPU arm: the oracle port on all host cores)

Workload (BASELINE.json configs[1]): B = 16384 hanging-chain instances per GPU,
horizon 60.  Instance b: rest state, 5 plant steps with a random disturbance input in [-1,1]^3, one
untimed cold solve from u = 0, one plant step with the first input, warm-start shift.  A timed "step"
= one warm solve of the whole batch (ALMSolver::operator() for every instance).  Multi-GPU: weak
scaling, every rank owns its own B instances (seeded by rank), no data-path collective.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

PROBLEM, PID, HORIZON = "hanging_chain", 0, 60
METRIC = "converged MPC solves/sec (hanging chain N=60, batched, warm-started)"
SEED = 0xC0FFEE + 1


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=16384, help="instances per GPU")
    ap.add_argument("--horizon", type=int, default=HORIZON)
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--order", default="index", choices=["history", "index"],
                    help="instance scheduling inside the launch: by the previous (cold) solve's cost, longest first, or by index")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------
# clocks: nvidia-smi sampled during the timed region
# ---------------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "200",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); smax.append(float(f[1])); power.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(smax)), "power_w_max": float(max(power)),
                "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------
# CPU oracle (checker / baseline only)
# ---------------------------------------------------------------------------------------------
def load_cpu_oracle():
    """The oracle built with the reference's flag policy (-O3 -march=native, default contraction).
    Built on this machine if g++ is here; otherwise the strict in-tree build."""
    import oracle_util as ou
    odir = os.path.join(ROOT, "oracle")
    fast = os.path.join(odir, "liboracle_fast.so")
    try:
        subprocess.check_call(["make", "-C", odir, "-B", "liboracle_fast.so"], stdout=subprocess.DEVNULL,
                              stderr=subprocess.DEVNULL)
        lib, flags = C.CDLL(fast), "-O3 -march=native -ffp-contract=fast"
    except Exception:
        lib, flags = ou.load(), "-O3 -march=x86-64-v3 -ffp-contract=off (prebuilt)"
    return lib, flags


def cpu_solve_batch(lib, N, par, x0, u, nthreads):
    P = lambda a: a.ctypes.data_as(C.c_void_p)
    B = x0.shape[0]
    u = np.array(u, dtype=np.float64)
    y = np.zeros((B, 1))
    st = np.zeros((B, 24))
    x0 = np.ascontiguousarray(x0)
    t = time.perf_counter()
    rc = lib.oracle_solve_batch(PID, N, P(par), C.c_long(B), P(x0), P(u), P(y), P(st), nthreads)
    t = time.perf_counter() - t
    assert rc == 0
    return u, st, t


def cpu_simulate(lib, x, u):
    P = lambda a: a.ctypes.data_as(C.c_void_p)
    xn = np.zeros_like(x)
    for b in range(x.shape[0]):
        xb, ub, ob = np.ascontiguousarray(x[b]), np.ascontiguousarray(u[b]), np.zeros(x.shape[1])
        assert lib.oracle_simulate_dynamics(PID, P(xb), P(ub), P(ob)) == 0
        xn[b] = ob
    return xn


def rest_state():
    x = np.zeros(57)
    x[0:30:3] = (np.arange(10) + 1) / 10  # hanging_chain.py:92-96
    return x


def disturbances(B, rank):
    rng = np.random.default_rng([SEED, rank])
    return rng.uniform(-1.0, 1.0, (5, B, 3))


def solver_and_params():
    import oracle_util as ou
    import pantr_cdc2023_experiments_b200 as pb
    solver, _ = pb.make_solver(pb.benchmark_options(PROBLEM))
    return pb, solver, ou.pack_params(solver)


def cpu_warm_inputs(lib, N, par, dist, S, cores):
    """Warm-start inputs of the first S instances, prepared on the CPU only."""
    x0 = np.tile(rest_state(), (S, 1))
    for s in range(dist.shape[0]):
        x0 = cpu_simulate(lib, x0, dist[s, :S])
    u, st, t_cold = cpu_solve_batch(lib, N, par, x0, np.zeros((S, 3 * N)), cores)
    x1 = cpu_simulate(lib, x0, u[:, :3])
    uw = np.concatenate([u[:, 3:], u[:, -3:]], axis=1).copy()
    return x1, uw, t_cold, st


# ---------------------------------------------------------------------------------------------
# reference arm: the CPU implementation of the path on all host cores
# ---------------------------------------------------------------------------------------------
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    _, _, par = solver_and_params()
    lib, flags = load_cpu_oracle()
    cores = os.cpu_count() or 1
    N = args.horizon
    S = min(args.batch, max(cores * 8, 32))
    dist = disturbances(args.batch, 0)
    x1, uw, t_cold, _ = cpu_warm_inputs(lib, N, par, dist, S, cores)
    for _ in range(args.warmup):
        cpu_solve_batch(lib, N, par, x1, uw, cores)
    tt, conv = 0.0, 0
    for _ in range(args.steps):
        _, st, t = cpu_solve_batch(lib, N, par, x1, uw, cores)
        tt += t
        conv += int((st[:, 0] == 1).sum())
    value = conv / tt
    sample = (f"first {S} of the {args.batch} warm instances per step, {cores} threads, one instance per thread "
              f"at a time; oracle port built {flags}")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "solves/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"{PROBLEM} N={N} batch {args.batch} warm-started (bounded sample of {S})",
                   "problem": PROBLEM, "horizon": N, "batch_per_gpu": args.batch, "sample": S},
        "cpu_baseline": {"value": value, "unit": "solves/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "solves/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "reference binary unavailable (needs Eigen + CasADi); CPU restatement (oracle/) timed instead",
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------
# our arm: batched CUDA solver, one rank per GPU
# ---------------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist_

    pb, solver, par = solver_and_params()
    from pantr_cdc2023_experiments_b200 import _lib, flops
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (pantr_b200 has no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # stdout carries exactly one JSON line: NCCL prints its version banner with printf when the first
        # communicator comes up, so file descriptor 1 points at stderr until that has happened
        sys.stdout.flush()
        saved_fd = os.dup(1)
        os.dup2(2, 1)
        try:
            dist_.init_process_group("nccl", device_id=dev)
            warm = torch.zeros(1, device=dev)
            dist_.all_reduce(warm)
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_fd, 1)
            os.close(saved_fd)

    def barrier():
        if world > 1:
            dist_.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist_.all_reduce(t, op=dist_.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist_.all_reduce(t, op=dist_.ReduceOp.SUM)
        return float(t.item())

    N, B = args.horizon, args.batch
    lib = _lib.load()
    prob = pb.BatchedProblem(PROBLEM, N, device=local)
    n, nx, nu = prob.n, prob.nx, prob.nu
    h = solver._handle(prob).h
    # A non-default stream: the C ABI treats a NULL stream as "the handle's own stream", and the
    # legacy default stream's handle is 0.  All torch work below is issued on this stream as well.
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0
    sp = C.c_void_p(stream.cuda_stream)
    vp = lambda t: C.c_void_p(t.data_ptr())
    f64 = dict(dtype=torch.float64, device=dev)

    # ---- synthetic instances, prepared on the device (untimed) ----
    dist = disturbances(B, rank)
    x0 = torch.from_numpy(np.tile(rest_state(), (B, 1))).to(dev)
    xn = torch.empty_like(x0)
    for s in range(dist.shape[0]):
        ud = torch.from_numpy(np.ascontiguousarray(dist[s])).to(dev)
        _lib.check(lib.pantr_b200_simulate_dynamics_batch_device(h, B, vp(x0), vp(ud), 3, vp(xn), sp))
        x0, xn = xn, x0
    u = torch.zeros(B, n, **f64)
    stats_t = torch.zeros(B * pb.STATS_DTYPE.itemsize, dtype=torch.uint8, device=dev)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    solver.solve_device(prob, x0, u, None, stats_t, stream)  # cold solve (mpc-driver.cpp:158)
    ev1.record(stream)
    torch.cuda.synchronize()
    cold_ms = ev0.elapsed_time(ev1)
    cold_stats = stats_t.cpu().numpy().view(pb.STATS_DTYPE).copy()
    x1 = torch.empty_like(x0)
    _lib.check(lib.pantr_b200_simulate_dynamics_batch_device(h, B, vp(x0), vp(u), n, vp(x1), sp))
    _lib.check(lib.pantr_b200_shift_warm_start_device(h, B, vp(u), None, sp))
    torch.cuda.synchronize()
    u_warm = u.clone()
    u_work = torch.empty_like(u_warm)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def step(events=None):
        flush.zero_()
        u_work.copy_(u_warm)
        if events is not None:
            events[0].record(stream)
        solver.solve_device(prob, x1, u_work, None, stats_t, stream)
        if events is not None:
            events[1].record(stream)

    # Scheduling hint (pantr_b200_set_order_device): in a closed loop the previous solve's cost of each
    # instance is known; starting the expensive ones first shortens the tail of the launch.  The
    # predictor here is the cycle count of the untimed cold solve, never of the timed solve itself.
    order_t = torch.argsort(torch.from_numpy(cold_stats["cycles"].astype(np.int64)), descending=True).to(torch.int32).to(dev)
    other_ms = None
    def timed_once():
        a_, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        step((a_, b_)); torch.cuda.synchronize()
        return a_.elapsed_time(b_)
    solver.set_order(prob, None if args.order == "history" else order_t)   # the OTHER order, for information
    step(); other_ms = min(timed_once(), timed_once())
    solver.set_order(prob, order_t if args.order == "history" else None)
    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = solver.kernel_launches(prob)
    kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t0.record(stream)
    for k in range(args.steps):
        step(kev[k])
    t1.record(stream)
    barrier()
    total_ms = max_over_ranks(t0.elapsed_time(t1))
    clocks = sampler.stop() if rank == 0 else None
    launches = solver.kernel_launches(prob) - launches0
    kernel_ms = float(np.mean([a.elapsed_time(b) for a, b in kev]))
    stats = stats_t.cpu().numpy().view(pb.STATS_DTYPE).copy()
    grid_now = solver.grid_size(prob)
    busy_us = float(stats["elapsed_us"].astype("f8").sum())                 # globaltimer, per instance
    slot_util = busy_us / (kernel_ms * 1e3 * min(grid_now, B))
    sm_mhz_in_kernel = float(stats["cycles"].astype("f8").sum() / max(busy_us, 1.0))   # clock64 / globaltimer
    conv_local = int((stats["status"] == 1).sum())
    conv_total = sum_over_ranks(conv_local)
    value = conv_total * args.steps / (total_ms * 1e-3)
    u_dev_result = u_work.cpu().numpy()

    # ---- e2e: the public host-buffer API (pantr_b200_solve_batch via ALMSolver.__call__) ----
    x1_h = torch.empty(B, nx, dtype=torch.float64).pin_memory()
    uw_h = torch.empty(B, n, dtype=torch.float64).pin_memory()
    u_h = torch.empty(B, n, dtype=torch.float64).pin_memory()
    x1_h.copy_(x1); uw_h.copy_(u_warm)
    torch.cuda.synchronize()
    e2e_steps = max(1, min(args.steps, 3))
    u_np, x1_np = u_h.numpy(), x1_h.numpy()

    def e2e_step():
        flush.zero_()
        torch.cuda.synchronize()
        u_np[...] = uw_h.numpy()
        t = time.perf_counter()
        st = solver(prob, u_np, None, x0=x1_np)  # H2D x0,u ; kernel ; D2H u,stats ; synchronous
        return time.perf_counter() - t, st

    e2e_step()  # warm-up (allocates the staging buffers)
    barrier()
    e2e_t, e2e_conv = 0.0, 0
    for _ in range(e2e_steps):
        t, st = e2e_step()
        e2e_t += t
        e2e_conv += int((st["status"] == 1).sum())
    e2e_t = max_over_ranks(e2e_t)
    e2e_value = sum_over_ranks(e2e_conv) / e2e_t
    same_as_device = bool(np.array_equal(u_np, u_dev_result))
    h2d = B * (nx + n) * 8
    d2h = B * n * 8 + B * pb.STATS_DTYPE.itemsize

    # ---- roofline of the solve kernel: algorithmic FP64 flops / kernel time vs measured DFMA peak ----
    F_local = flops.solve_flops(PROBLEM, N, stats)
    peak = C.c_double()
    _lib.check(lib.pantr_b200_measure_fp64_peak(local, 3, C.byref(peak)))
    achieved = F_local / (kernel_ms * 1e-3) / 1e12
    state_bytes = B * ((nx + 2 * n) * 8 + pb.STATS_DTYPE.itemsize)
    peaks_file = os.path.join(ROOT, "MEASURED_PEAKS.json")
    hbm_peak = json.load(open(peaks_file))["hbm_gbs"] if os.path.exists(peaks_file) else 6650.0
    # DRAM traffic of this launch shape from the committed `ncu --set full` capture (never measured live)
    traffic, traffic_src = None, None
    tfile = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tfile):
        tj = json.load(open(tfile))
        if tj.get("batch") == B and tj.get("horizon") == N:
            traffic, traffic_src = tj["dram_bytes_per_launch"], tj.get("source")
    roofline = {
        "bound": "fp64", "achieved": achieved, "peak": peak.value, "unit": "TFLOP/s",
        "frac": achieved / peak.value if peak.value > 0 else None, "traffic": traffic,
        "traffic_unit": "bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum)", "traffic_source": traffic_src,
        "kernel": "solve_kernel<HangingChain>", "kernel_ms": kernel_ms,
        "flops_per_launch": F_local, "flops_per_solve": F_local / B,
        "peak_source": "FP64 DFMA peak measured live on this GPU (pantr_b200_measure_fp64_peak); "
                       "MEASURED_PEAKS.json has no FP64 entry",
        "hbm": {"achieved_gbs": state_bytes / (kernel_ms * 1e-3) / 1e9, "peak_gbs": hbm_peak,
                "peak_source": "measured" if os.path.exists(peaks_file) else "fallback",
                "algorithmic_bytes_per_launch": state_bytes,
                # the kernel's own cache streaming (per-block sweep caches do not fit in L2): committed
                # ncu traffic of this launch shape over the live kernel time, the second roofline
                "cache_stream_gbs": traffic / (kernel_ms * 1e-3) / 1e9 if traffic else None,
                "cache_stream_frac_of_peak": traffic / (kernel_ms * 1e-3) / 1e9 / hbm_peak if traffic else None},
    }

    # ---- CPU baseline (rank 0, N = 1 only): the oracle port on this box's host cores ----
    cpu_baseline = None
    parity = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        olib, oflags = load_cpu_oracle()
        cores = os.cpu_count() or 1
        x1c, uwc = x1_h.numpy(), uw_h.numpy()
        S0 = min(B, cores)
        _, _, tcal = cpu_solve_batch(olib, N, par, x1c[:S0], uwc[:S0], cores)
        S = int(min(B, max(S0, cores * max(1, int(args.cpu_seconds / max(tcal, 1e-3))))))
        uo, sto, t = cpu_solve_batch(olib, N, par, x1c[:S], uwc[:S], cores)
        cpu_baseline = {
            "value": float((sto[:, 0] == 1).sum()) / t, "unit": "solves/s", "cores": cores, "kind": "port",
            "sample": f"first {S} of the {B} warm instances of this run, {cores} threads (one instance per "
                      f"thread at a time), {t:.1f} s; oracle port built {oflags}"}
        # checker: device results of the first instances vs the strict oracle build (same arithmetic
        # order, no contraction): north_star tolerance 1e-8 relative, same status, iterations +-1
        import oracle_util as ou
        Sp = min(B, 4 * cores)
        us, sts, _ = cpu_solve_batch(ou.load(), N, par, x1c[:Sp], uwc[:Sp], cores)
        rel = np.abs(us - u_dev_result[:Sp]).max() / max(1.0, np.abs(us).max())
        parity = {"instances": Sp, "max_rel_err_u": float(rel),
                  "bitwise_equal_u": bool(np.array_equal(us, u_dev_result[:Sp])),
                  "status_equal": bool((sts[:, 0] == stats["status"][:Sp]).all()),
                  "max_iter_diff": int(np.abs(sts[:, 2] - stats["inner_iterations"][:Sp]).max()),
                  "fast_vs_strict_oracle_max_rel_err_u": float(np.abs(uo[:Sp] - us).max() / max(1.0, np.abs(us).max()))}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "solves/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"{PROBLEM} N={N} batch {B} per GPU, random disturbed initial states, "
                                   f"warm-started (BASELINE.json configs[1])",
                       "problem": PROBLEM, "horizon": N, "batch_per_gpu": B, "global_batch": B * world,
                       "parallelism": f"instances sharded over {world} GPU(s), no collective on the data path",
                       "l2": "256 MiB flush write between timed steps (inside the timed region)",
                       "order": ("instances taken longest-first by the cycle count of the previous (cold) solve"
                                 if args.order == "history" else "instances taken by index"),
                       "options": "new-benchmarks-paper/dodo.py PANTR option set, tolerance 1e-8"},
            "converged_fraction": conv_total / (B * world),
            "mean_inner_iterations": float(stats["inner_iterations"].mean()),
            "mean_hess_psi_prod": float(stats["ev_hess_psi_prod"].mean()),
            "cold_solve": {"ms": cold_ms, "solves_per_s": B / (cold_ms * 1e-3),
                           "converged_fraction": float((cold_stats["status"] == 1).mean()),
                           "mean_inner_iterations": float(cold_stats["inner_iterations"].mean()),
                           "mean_hess_psi_prod": float(cold_stats["ev_hess_psi_prod"].mean())},
            "e2e": {"value": e2e_value, "unit": "solves/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "steps": e2e_steps, "bitwise_equal_to_device_path": same_as_device},
            "gpu_launches": int(launches), "grid": solver.grid_size(prob),
            "scheduling": {"order": args.order, "slot_utilisation": slot_util, "sm_mhz_in_kernel": sm_mhz_in_kernel,
                           "kernel_ms_other_order": other_ms,
                           "cycles_mean": float(stats["cycles"].mean()), "cycles_max": float(stats["cycles"].max())},
            "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu_baseline, "parity_check": parity,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist_.destroy_process_group()


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
