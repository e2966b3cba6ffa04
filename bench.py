#!/usr/bin/env python
"""bench.py — converged MPC solves/s of the batched PANTR+ALM hot path.

  python bench.py --gpus N --steps K --warmup W [--config C]     (our arm, one rank per GPU under torchrun)
  python bench.py --impl reference --gpus N --steps K ...         (CPU arm: the oracle port on all host cores)

--config selects the BASELINE.json workload (default = configs[1], the one the headline metric is quoted on):
  hanging_chain    configs[1]  hanging chain N=60, 16384 instances per GPU, random disturbed initial states, WARM start:
                   rest state, 5 plant steps with a disturbance input in [-1,1]^3, one untimed cold solve from u = 0,
                   one plant step with the first input, warm-start shift (mpc-driver.cpp:148-175); a timed "step" =
                   one warm solve of the whole batch.
  quadcopter       configs[2]  simplified quadcopter N=60, 65536 instances per GPU, cold start from the reference's
                   guess; x0 = nominal + scale * (p +-0.1, v +-0.5, theta +-0.2) (--scale,
                   default 0.1: at scale 1 the tilt perturbation alone makes 18 % of the instances run into
                   alm.max_iter x solver.max_iter = 75000 iterations).
  realistic_sweep  configs[3]  realistic quadcopter, horizons 10, 20, ..., 120, --batch instances per horizon and GPU;
                   a step = one cold solve of every horizon's batch, value = all solves / total time.
  mixed            configs[4]  one third each of the three problems at N=60, cold start, sorted by problem type per
                   GPU (one launch per type); --batch = instances per GPU (1M / 8 = 131072 on the 8-GPU box).
Multi-GPU: instances are sharded by index, no collective on the data path.  --shard weak (default): every rank owns
--batch instances of its own (seeded by rank); --shard global: all ranks derive the SAME global batch of --batch
instances from one seed, rank r solves shard_range(B, r, world) and the per-instance stats are gathered at the end.
"""
import argparse
import ctypes as C
import hashlib
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

SEED = 0xC0FFEE
# name -> (problem id of the C ABI, oracle id of the device formulation, oracle id of the reference formulation)
PROBLEMS = {"hanging_chain": (0, 0, 0), "quadcopter": (1, 5, 1), "realistic_quadcopter": (2, 6, 2)}
CONFIGS = {
    "hanging_chain": dict(index=1, metric="converged MPC solves/sec (hanging chain N=60, batched, warm-started)", batch=16384),
    "quadcopter": dict(index=2, metric="converged MPC solves/sec (simplified quadcopter N=60, batched, cold start)", batch=65536),
    "realistic_sweep": dict(index=3, metric="converged MPC solves/sec (realistic quadcopter, horizon sweep N=10..120, batched, cold start)",
                            batch=65536),
    "mixed": dict(index=4, metric="converged MPC solves/sec (mixed scenario: hanging chain + quadcopter + realistic quadcopter, N=60, cold start)",
                  batch=131072),
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="hanging_chain", choices=list(CONFIGS))
    ap.add_argument("--batch", type=int, default=None, help="instances per GPU (per horizon for realistic_sweep); default: the config's")
    ap.add_argument("--horizon", type=int, default=60)
    ap.add_argument("--horizons", default="10,20,30,40,50,60,70,80,90,100,110,120", help="realistic_sweep only")
    ap.add_argument("--scale", type=float, default=0.1, help="scale of the quadcopter x0 perturbation (SURVEY.md 8d.3 = 1)")
    ap.add_argument("--shard", default="weak", choices=["weak", "global"])
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-closed-loop", action="store_true", help="skip the closed-loop leg (hanging_chain config, N = 1)")
    ap.add_argument("--closed-loop-steps", type=int, default=60)
    ap.add_argument("--e2e-steps", type=int, default=None)
    a = ap.parse_args()
    if a.batch is None:
        a.batch = CONFIGS[a.config]["batch"]
    return a


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
        return {"sm_mhz": float(np.median(sm)), "sm_min_mhz": float(min(sm)), "sm_max_mhz": float(max(smax)),
                "power_w_max": float(max(power)), "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------
# workload: a list of segments (problem, horizon, instances); a step solves every segment once
# ---------------------------------------------------------------------------------------------
def segments_of(args):
    if args.config == "hanging_chain":
        return [("hanging_chain", args.horizon, args.batch)]
    if args.config == "quadcopter":
        return [("quadcopter", args.horizon, args.batch)]
    if args.config == "realistic_sweep":
        return [("realistic_quadcopter", int(h), args.batch) for h in args.horizons.split(",")]
    B = args.batch  # mixed: thirds, sorted by problem type (349526 / 349525 / 349525 of 1M)
    third = B // 3
    return [("hanging_chain", args.horizon, B - 2 * third), ("quadcopter", args.horizon, third),
            ("realistic_quadcopter", args.horizon, third)]


def rest_state():
    x = np.zeros(57)
    x[0:30:3] = (np.arange(10) + 1) / 10  # hanging_chain.py:92-96
    return x


def instance_inputs(args, seg_index, name, B, stream_id):
    """Deterministic synthetic instances of one segment: numpy PCG64 streams keyed by (seed + config index, segment,
    stream_id); stream_id = rank for weak sharding, 0 for the global batch."""
    rng = np.random.default_rng([SEED + CONFIGS[args.config]["index"], seg_index, stream_id])
    if name == "hanging_chain":
        return {"dist": rng.uniform(-1.0, 1.0, (5, B, 3))}
    s = args.scale
    x0 = np.zeros((B, 9)); x0[:, :3] = (-0.2, -0.25, 0.5)  # quadcopter.py:78-83
    x0[:, :3] += s * rng.uniform(-0.1, 0.1, (B, 3))
    x0[:, 3:6] += s * rng.uniform(-0.5, 0.5, (B, 3))
    x0[:, 6:9] += s * rng.uniform(-0.2, 0.2, (B, 3))
    # c(x_0) must lie in D, i.e. p_x^2 + p_y^2 >= 0.01 (always true for scale <= 1: |p_x| >= 0.1)
    assert (x0[:, 0] ** 2 + x0[:, 1] ** 2 >= 0.01).all()
    return {"x0": x0}


def workload_text(args):
    segs = segments_of(args)
    per = ", ".join(f"{n} N={N} x {B}" for n, N, B in segs) if len(segs) <= 3 else \
        f"{segs[0][0]} N={segs[0][1]}..{segs[-1][1]} ({len(segs)} horizons) x {segs[0][2]} each"
    start = "warm-started (cold solve, plant step, shift; the warm solve is timed)" if args.config == "hanging_chain" else "cold start"
    quad = "" if args.config == "hanging_chain" else f"; quadcopter x0 = nominal + {args.scale} * (p +-0.1, v +-0.5, theta +-0.2)"
    scope = "per GPU" if args.shard == "weak" else "in ONE global batch sharded over the GPUs by instance index"
    return f"{per} {scope}, {start}{quad} (BASELINE.json configs[{CONFIGS[args.config]['index']}])"


def config_dict(args, world, sample=None):
    segs = segments_of(args)
    d = {"workload": workload_text(args), "config": args.config,
         "segments": [{"problem": n, "horizon": N, "batch_per_gpu": B if args.shard == "weak" else round(B / world, 1)} for n, N, B in segs],
         "batch_per_gpu": sum(B for _, _, B in segs) if args.shard == "weak" else round(sum(B for _, _, B in segs) / world, 1),
         "global_batch": sum(B for _, _, B in segs) * (world if args.shard == "weak" else 1),
         "x0_scale": None if args.config == "hanging_chain" else args.scale,
         "parallelism": f"instances sharded by index over {world} GPU(s) ({args.shard}), no collective on the data path",
         "l2": "256 MiB flush write between timed steps (inside the timed region)",
         "options": "new-benchmarks-paper/dodo.py PANTR option set per problem, tolerance 1e-8"}
    if sample is not None:
        d["sample"] = sample
    return d


# ---------------------------------------------------------------------------------------------
# CPU oracle (checker / baseline only)
# ---------------------------------------------------------------------------------------------
def load_cpu_oracle():
    """The oracle built with the reference's flag policy (-O3 -march=native, default contraction,
    scripts/build-benchmarks.sh:19-22).  Built on this machine if g++ is here; otherwise the strict in-tree build."""
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


def cpu_solve_batch(lib, pid, N, par, x0, u, m, nthreads):
    P = lambda a: a.ctypes.data_as(C.c_void_p)
    B = x0.shape[0]
    u = np.array(u, dtype=np.float64)
    y = np.zeros((B, max(m, 1)))
    st = np.zeros((B, 24))
    x0 = np.ascontiguousarray(x0)
    t = time.perf_counter()
    rc = lib.oracle_solve_batch(pid, N, P(par), C.c_long(B), P(x0), P(u), P(y), P(st), nthreads)
    t = time.perf_counter() - t
    assert rc == 0
    return u, st, t


def cpu_simulate(lib, pid, x, u):
    P = lambda a: a.ctypes.data_as(C.c_void_p)
    xn = np.zeros_like(x)
    for b in range(x.shape[0]):
        xb, ub, ob = np.ascontiguousarray(x[b]), np.ascontiguousarray(u[b]), np.zeros(x.shape[1])
        assert lib.oracle_simulate_dynamics(pid, P(xb), P(ub), P(ob)) == 0
        xn[b] = ob
    return xn


def solver_for(name):
    import oracle_util as ou
    import pantr_cdc2023_experiments_b200 as pb
    solver, _ = pb.make_solver(pb.benchmark_options(name))
    return solver, ou.pack_params(solver)


def cpu_segment_inputs(lib, args, seg_index, name, N, S, cores):
    """Timed-solve inputs of the first S instances of a segment (rank 0's stream), prepared on the CPU only."""
    _, par = solver_for(name)
    inp = instance_inputs(args, seg_index, name, segments_of(args)[seg_index][2], 0)
    if name == "hanging_chain":
        x0 = np.tile(rest_state(), (S, 1))
        for s in range(5):
            x0 = cpu_simulate(lib, 0, x0, inp["dist"][s, :S])
        if args.config != "hanging_chain":  # cold start
            return par, x0, np.zeros((S, 3 * N)), 0
        u, _, _ = cpu_solve_batch(lib, 0, N, par, x0, np.zeros((S, 3 * N)), 0, cores)
        x1 = cpu_simulate(lib, 0, x0, u[:, :3])
        return par, x1, np.concatenate([u[:, 3:], u[:, -3:]], axis=1).copy(), 0
    return par, inp["x0"][:S], np.tile(np.tile([9.81, 0.0, 0.0, 0.0], N), (S, 1)), 4 * (N + 1)


# ---------------------------------------------------------------------------------------------
# reference arm: the CPU implementation of the path on all host cores
# ---------------------------------------------------------------------------------------------
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    lib, flags = load_cpu_oracle()
    cores = os.cpu_count() or 1
    segs = segments_of(args)
    # bounded sample, the same rule as the cpu_baseline leg of the b200 arm: the first S instances of every segment,
    # S large enough (>= 256 per core for the cheap hanging-chain solves) that the step is not limited by stragglers
    per_seg = {"hanging_chain": max(4096, 256 * cores), "quadcopter": 8 * cores, "realistic_quadcopter": 8 * cores}
    prepared = []
    for i, (name, N, B) in enumerate(segs):
        S = min(B, max(cores, per_seg[name] // (1 if len(segs) <= 3 else 4)))
        par, x0, u, m = cpu_segment_inputs(lib, args, i, name, N, S, cores)
        prepared.append((name, N, S, par, x0, u, m))

    def step():
        tt, conv, tot = 0.0, 0, 0
        for name, N, S, par, x0, u, m in prepared:
            _, st, t = cpu_solve_batch(lib, PROBLEMS[name][2], N, par, x0, u, m, cores)
            tt += t; conv += int((st[:, 0] == 1).sum()); tot += S
        return tt, conv, tot

    for _ in range(args.warmup):
        step()
    tt, conv, tot = 0.0, 0, 0
    for _ in range(args.steps):
        t, c, n = step()
        tt += t; conv += c; tot += n
    value = conv / tt
    sample = (f"first {[p[2] for p in prepared]} instances of the segment(s) per step, {cores} threads, one instance per thread "
              f"at a time off a shared queue; oracle port (reference formulation: forward-over-adjoint sweeps) built {flags}")
    line = {
        "impl": "reference", "metric": CONFIGS[args.config]["metric"], "value": value, "unit": "solves/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tt / args.steps,
        "higher_is_better": True, "scaling": "weak" if args.shard == "weak" else "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": config_dict(args, args.gpus, sample=[p[2] for p in prepared]),
        "converged_fraction": conv / max(tot, 1),
        "cpu_baseline": {"value": value, "unit": "solves/s", "cores": cores, "solves_per_s_per_core": value / cores,
                         "kind": "port", "sample": sample},
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

    import oracle_util as ou
    import pantr_cdc2023_experiments_b200 as pb
    from pantr_cdc2023_experiments_b200 import _lib, flops, sharding
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

    def reduce_(x, op):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist_.all_reduce(t, op=getattr(dist_.ReduceOp, op))
        return float(t.item())

    def gather_scalars(vals):
        """a few floats per rank -> list over ranks (per-rank attribution of the scaling loss)"""
        t = torch.tensor(vals, dtype=torch.float64, device=dev)
        if world == 1:
            return [list(map(float, vals))]
        out = [torch.zeros_like(t) for _ in range(world)]
        dist_.all_gather(out, t)
        return [list(map(float, o.tolist())) for o in out]

    lib = _lib.load()
    # A non-default stream: the C ABI treats a NULL stream as "the handle's own stream", and the
    # legacy default stream's handle is 0.  All torch work below is issued on this stream as well.
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0
    sp = C.c_void_p(stream.cuda_stream)
    vp = lambda t: C.c_void_p(t.data_ptr())
    f64 = dict(dtype=torch.float64, device=dev)
    isz = pb.STATS_DTYPE.itemsize

    # ---- synthetic instances, prepared on the device (untimed) ----
    class Seg:
        pass
    segs = []
    cold_info = None
    for i, (name, N, Bg) in enumerate(segments_of(args)):
        s = Seg()
        s.index, s.name, s.N = i, name, N
        if args.shard == "global":  # one global batch, this rank's contiguous shard
            lo, hi = sharding.shard_range(Bg, rank, world)
            inp = instance_inputs(args, i, name, Bg, 0)
            inp = {k: np.ascontiguousarray(v[:, lo:hi] if k == "dist" else v[lo:hi]) for k, v in inp.items()}
            s.B = hi - lo
        else:
            inp = instance_inputs(args, i, name, Bg, rank)
            s.B = Bg
        s.solver, s.par = solver_for(name)
        s.prob = pb.BatchedProblem(name, N, device=local)
        s.h = s.solver._handle(s.prob).h
        n, m = s.prob.n, s.prob.m
        B = s.B
        s.stats_t = torch.zeros(max(B, 1) * isz, dtype=torch.uint8, device=dev)
        s.y_init = torch.zeros(B, max(m, 1), **f64)
        if name == "hanging_chain":
            x0 = torch.from_numpy(np.tile(rest_state(), (B, 1))).to(dev)
            xn = torch.empty_like(x0)
            for d in range(5):
                ud = torch.from_numpy(np.ascontiguousarray(inp["dist"][d])).to(dev)
                _lib.check(lib.pantr_b200_simulate_dynamics_batch_device(s.h, B, vp(x0), vp(ud), 3, vp(xn), sp))
                x0, xn = xn, x0
            u = torch.zeros(B, n, **f64)
            if args.config == "hanging_chain":  # warm start: untimed cold solve, plant step, shift
                ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                ev0.record(stream)
                s.solver.solve_device(s.prob, x0, u, None, s.stats_t, stream)  # mpc-driver.cpp:158
                ev1.record(stream)
                torch.cuda.synchronize()
                cs = s.stats_t.cpu().numpy().view(pb.STATS_DTYPE)[:B].copy()
                cold_ms = ev0.elapsed_time(ev1)
                cold_info = {"ms": cold_ms, "solves_per_s": B / (cold_ms * 1e-3),
                             "converged_fraction": float((cs["status"] == 1).mean()),
                             "mean_inner_iterations": float(cs["inner_iterations"].mean()),
                             "mean_hess_psi_prod": float(cs["ev_hess_psi_prod"].mean())}
                x1 = torch.empty_like(x0)
                _lib.check(lib.pantr_b200_simulate_dynamics_batch_device(s.h, B, vp(x0), vp(u), n, vp(x1), sp))
                _lib.check(lib.pantr_b200_shift_warm_start_device(s.h, B, vp(u), None, sp))
                torch.cuda.synchronize()
                x0 = x1
            s.x0, s.u_init = x0, u.clone()
        else:
            s.x0 = torch.from_numpy(inp["x0"]).to(dev)
            s.u_init = torch.from_numpy(np.tile(s.prob.initial_guess(), (B, 1))).to(dev)
        s.u_work, s.y_work = torch.empty_like(s.u_init), torch.empty_like(s.y_init)
        segs.append(s)
    torch.cuda.synchronize()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def step(events=None):
        flush.zero_()
        for k, s in enumerate(segs):
            s.u_work.copy_(s.u_init)
            if s.prob.m:
                s.y_work.copy_(s.y_init)
            if events is not None:
                events[k][0].record(stream)
            if s.B:
                s.solver.solve_device(s.prob, s.x0, s.u_work, s.y_work if s.prob.m else None, s.stats_t, stream)
            if events is not None:
                events[k][1].record(stream)

    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = sum(s.solver.kernel_launches(s.prob) for s in segs)
    mk = lambda: [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in segs]
    kev = [mk() for _ in range(args.steps)]
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t0.record(stream)
    for k in range(args.steps):
        step(kev[k])
    t1.record(stream)
    barrier()
    my_ms = t0.elapsed_time(t1)
    total_ms = reduce_(my_ms, "MAX")
    clocks = sampler.stop()
    launches = sum(s.solver.kernel_launches(s.prob) for s in segs) - launches0
    for k, s in enumerate(segs):
        s.kernel_ms = float(np.mean([kev[j][k][0].elapsed_time(kev[j][k][1]) for j in range(args.steps)]))
        s.stats = s.stats_t.cpu().numpy().view(pb.STATS_DTYPE)[:s.B].copy()
        s.u_dev = s.u_work.cpu().numpy()
        s.conv = int((s.stats["status"] == 1).sum())
        s.flops = flops.solve_flops(s.name, s.N, s.stats)
        grid_now = s.solver.grid_size(s.prob)
        busy_us = float(s.stats["elapsed_us"].astype("f8").sum())                 # globaltimer, per instance
        s.slot_util = busy_us / max(s.kernel_ms * 1e3 * min(grid_now, max(s.B, 1)), 1e-9)
        s.sm_mhz = float(s.stats["cycles"].astype("f8").sum() / max(busy_us, 1.0))   # clock64 / globaltimer
    conv_local = sum(s.conv for s in segs)
    inst_local = sum(s.B for s in segs)
    conv_total = reduce_(conv_local, "SUM")
    inst_total = reduce_(inst_local, "SUM")
    value = conv_total * args.steps / (total_ms * 1e-3)
    dom = max(segs, key=lambda s: s.kernel_ms)  # the dominant kernel launch of a step
    kernel_ms_sum = sum(s.kernel_ms for s in segs)
    # per-rank attribution of the scaling loss (the slowest rank sets the step time)
    per_rank = gather_scalars([my_ms / args.steps, kernel_ms_sum, dom.sm_mhz, dom.slot_util,
                               clocks.get("sm_mhz") or 0.0, clocks.get("sm_min_mhz") or 0.0, float(conv_local)])
    gathered_ok = None
    if args.shard == "global" and world > 1:  # the one inter-rank transfer of a sharded solve: the final stats gather
        allst = sharding.gather_stats(np.concatenate([s.stats for s in segs]), world)
        gathered_ok = bool(len(allst) == int(inst_total) and int((allst["status"] == 1).sum()) == int(conv_total))

    # ---- e2e: the public host-buffer API (pantr_b200_solve_batch via ALMSolver.__call__) ----
    host = []
    for s in segs:
        x0_h = torch.empty(s.B, s.prob.nx, dtype=torch.float64).pin_memory(); x0_h.copy_(s.x0)
        ui_h = torch.empty(s.B, s.prob.n, dtype=torch.float64).pin_memory(); ui_h.copy_(s.u_init)
        u_h = torch.empty(s.B, s.prob.n, dtype=torch.float64).pin_memory()
        y_h = torch.zeros(s.B, max(s.prob.m, 1), dtype=torch.float64).pin_memory()
        host.append((x0_h.numpy(), ui_h.numpy(), u_h.numpy(), y_h.numpy()))
    torch.cuda.synchronize()
    e2e_steps = args.e2e_steps if args.e2e_steps is not None else max(1, min(args.steps, 3))

    def e2e_step():
        flush.zero_()
        torch.cuda.synchronize()
        conv, tt = 0, 0.0
        for s, (x0n, uin, un, yn) in zip(segs, host):
            if not s.B:
                continue
            un[...] = uin
            yn[...] = 0
            t = time.perf_counter()
            st = s.solver(s.prob, un, yn if s.prob.m else None, x0=x0n)  # H2D x0, u, y ; kernel ; D2H u, y, stats ; synchronous
            tt += time.perf_counter() - t
            conv += int((st["status"] == 1).sum())
        return tt, conv

    e2e_warmup = total_ms / args.steps < 5000.0  # a step of several seconds does not notice the one-off staging allocation
    if e2e_warmup:
        e2e_step()  # warm-up (allocates the staging buffers)
    barrier()
    e2e_t, e2e_conv = 0.0, 0
    for _ in range(e2e_steps):
        t, c = e2e_step()
        e2e_t += t
        e2e_conv += c
    e2e_t = reduce_(e2e_t, "MAX")
    e2e_value = reduce_(e2e_conv, "SUM") / e2e_t
    same_as_device = bool(all(np.array_equal(h[2], s.u_dev) for s, h in zip(segs, host)))
    h2d = sum(s.B * (s.prob.nx + s.prob.n + s.prob.m) * 8 for s in segs)
    d2h = sum(s.B * (s.prob.n + s.prob.m) * 8 + s.B * isz for s in segs)

    # ---- roofline of the dominant solve kernel: algorithmic FP64 flops / kernel time vs measured DFMA peak ----
    peak = C.c_double()
    _lib.check(lib.pantr_b200_measure_fp64_peak(local, 3, C.byref(peak)))
    peaks_file = os.path.join(ROOT, "MEASURED_PEAKS.json")
    hbm_peak = json.load(open(peaks_file))["hbm_gbs"] if os.path.exists(peaks_file) else 6650.0
    achieved = dom.flops / (dom.kernel_ms * 1e-3) / 1e12
    state_bytes = dom.B * ((dom.prob.nx + 2 * dom.prob.n + 2 * dom.prob.m) * 8 + isz)
    # DRAM traffic of this launch shape from a committed `ncu --set full` capture: only valid for the library it was
    # taken from (sha256 of libpantr_b200.so recorded in profiles/traffic.json) and for this launch shape; else null
    traffic, traffic_src = None, None
    tfile = os.path.join(ROOT, "profiles", "traffic.json")
    lib_sha = hashlib.sha256(open(_lib.LIB_PATH, "rb").read()).hexdigest()
    # a rebuild of the same sources may not reproduce the .so byte for byte: the capture also records a hash of the sources
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    try:
        from source_hash import source_sha256
        src_sha = source_sha256()
    except Exception:
        src_sha = None
    if os.path.exists(tfile):
        for tj in json.load(open(tfile)).get("captures", []):
            if ((tj.get("lib_sha256") == lib_sha or tj.get("src_sha256") == src_sha) and tj.get("problem") == dom.name and tj.get("batch") == dom.B
                    and tj.get("horizon") == dom.N and tj.get("start") == ("warm" if args.config == "hanging_chain" else "cold")):
                traffic, traffic_src = tj["dram_bytes_per_launch"], tj.get("source")
    kname = {"hanging_chain": "solve_kernel<HangingChain>", "quadcopter": "solve_kernel<QuadCondensedModel<SimpleFns>>",
             "realistic_quadcopter": "solve_kernel<QuadCondensedModel<RealisticFns>>"}[dom.name]
    roofline = {
        "bound": "fp64", "achieved": achieved, "peak": peak.value, "unit": "TFLOP/s",
        "frac": achieved / peak.value if peak.value > 0 else None, "traffic": traffic,
        "traffic_unit": "bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum)", "traffic_source": traffic_src,
        "lib_sha256": lib_sha[:16],
        "kernel": kname, "kernel_ms": dom.kernel_ms, "kernel_share_of_step": dom.kernel_ms / max(my_ms / args.steps, 1e-9),
        "flops_per_launch": dom.flops, "flops_per_solve": dom.flops / max(dom.B, 1),
        "flops_convention": "algorithmic: per-instance evaluation counters x one from-scratch evaluation of each problem function "
                            "in the reference formulation (pantr_cdc2023_experiments_b200/flops.py)",
        "peak_source": "FP64 DFMA peak measured live on this GPU (pantr_b200_measure_fp64_peak); "
                       "MEASURED_PEAKS.json has no FP64 entry",
        "hbm": {"achieved_gbs": state_bytes / (dom.kernel_ms * 1e-3) / 1e9, "peak_gbs": hbm_peak,
                "peak_source": "measured" if os.path.exists(peaks_file) else "fallback",
                "algorithmic_bytes_per_launch": state_bytes,
                "cache_stream_gbs": traffic / (dom.kernel_ms * 1e-3) / 1e9 if traffic else None,
                "cache_stream_frac_of_peak": traffic / (dom.kernel_ms * 1e-3) / 1e9 / hbm_peak if traffic else None},
        "segments": [{"problem": s.name, "horizon": s.N, "batch": s.B, "kernel_ms": s.kernel_ms,
                      "tflops": s.flops / (s.kernel_ms * 1e-3) / 1e12 if s.kernel_ms > 0 else None,
                      "frac": s.flops / (s.kernel_ms * 1e-3) / 1e12 / peak.value if s.kernel_ms > 0 and peak.value > 0 else None,
                      "solves_per_s": s.conv / (s.kernel_ms * 1e-3) if s.kernel_ms > 0 else None,
                      "converged_fraction": s.conv / max(s.B, 1),
                      "mean_inner_iterations": float(s.stats["inner_iterations"].mean()) if s.B else None,
                      "max_inner_iterations": int(s.stats["inner_iterations"].max()) if s.B else None,
                      "mean_hess_psi_prod": float(s.stats["ev_hess_psi_prod"].mean()) if s.B else None,
                      "slot_utilisation": s.slot_util, "sm_mhz_in_kernel": s.sm_mhz,
                      "cycles_max_over_mean": float(s.stats["cycles"].max() / max(s.stats["cycles"].mean(), 1.0)) if s.B else None}
                     for s in segs],
    }

    # ---- CPU baseline + parity check (rank 0, N = 1 only): the oracle port on this box's host cores ----
    cpu_baseline = None
    parity = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        olib, oflags = load_cpu_oracle()
        cores = os.cpu_count() or 1
        tt, cc, samples = 0.0, 0, []
        budget = args.cpu_seconds / len(segs)
        for s, (x0n, uin, _, _) in zip(segs, host):
            pid_ref = PROBLEMS[s.name][2]
            S0 = min(s.B, cores)
            _, _, tcal = cpu_solve_batch(olib, pid_ref, s.N, s.par, x0n[:S0], uin[:S0], s.prob.m, cores)
            S = int(min(s.B, max(S0, cores * max(1, int(budget / max(tcal, 1e-3))))))
            _, sto, t = cpu_solve_batch(olib, pid_ref, s.N, s.par, x0n[:S], uin[:S], s.prob.m, cores)
            tt += t; cc += int((sto[:, 0] == 1).sum()); samples.append(S)
        cpu_baseline = {
            "value": cc / tt, "unit": "solves/s", "cores": cores, "solves_per_s_per_core": cc / tt / cores, "kind": "port",
            "sample": f"first {samples} instances of the segment(s) of this run, {cores} threads (one instance per "
                      f"thread at a time off a shared queue), {tt:.1f} s; oracle port (reference formulation: forward-over-adjoint "
                      f"sweeps) built {oflags}"}
        # checker: device results of the first instances vs the strict oracle build of the device formulation (same
        # arithmetic order, no contraction): north_star tolerance 1e-8 relative, same status, iterations +-1
        parity = {"segments": []}
        for s, (x0n, uin, _, _) in zip(segs, host):
            Sp = min(s.B, cores if s.name != "hanging_chain" else 4 * cores)
            us, sts, _ = cpu_solve_batch(ou.load(), PROBLEMS[s.name][1], s.N, s.par, x0n[:Sp], uin[:Sp], s.prob.m, cores)
            # instances the device ended with MaxTime (alm.max_time on the device clock; the oracle has no clock) are
            # timing dependent and cannot be compared: counted, not checked
            keep = np.asarray(s.stats["status"][:Sp]) != 2
            ud, st_d, it_d = s.u_dev[:Sp][keep], np.asarray(s.stats["status"][:Sp])[keep], np.asarray(s.stats["inner_iterations"][:Sp])[keep]
            uo, st_o, it_o = us[keep], sts[keep, 0], sts[keep, 2]
            rel = np.abs(uo - ud).max() / max(1.0, np.abs(uo).max()) if keep.any() else 0.0
            parity["segments"].append({
                "problem": s.name, "horizon": s.N, "instances": int(keep.sum()), "excluded_max_time": int((~keep).sum()),
                "max_rel_err_u": float(rel), "bitwise_equal_u": bool(np.array_equal(uo, ud)),
                "status_equal": bool((st_o == st_d).all()),
                "max_iter_diff": int(np.abs(it_o - it_d).max()) if keep.any() else 0})
        parity["all_bitwise_equal"] = all(p["bitwise_equal_u"] and p["status_equal"] and p["max_iter_diff"] == 0
                                          for p in parity["segments"])

    # ---- closed loop, the reference's own metric (mean over num_sim MPC steps; perfplot-cold-warm.py:52-61) ----
    closed_loop = None
    if args.config == "hanging_chain" and world == 1 and rank == 0 and not args.no_closed_loop:
        s = segs[0]
        n, B = s.prob.n, s.B
        x, xn = s.x0.clone(), torch.empty_like(s.x0)
        u = s.u_init.clone()
        stats_all = torch.zeros(args.closed_loop_steps, max(B, 1) * isz, dtype=torch.uint8, device=dev)
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0.record(stream)
        for i in range(args.closed_loop_steps):  # warm = 1: shift, solve, plant (mpc-driver.cpp:163-206), no host sync
            if i > 0:
                _lib.check(lib.pantr_b200_shift_warm_start_device(s.h, B, vp(u), None, sp))
            s.solver.solve_device(s.prob, x, u, None, stats_all[i], stream)
            _lib.check(lib.pantr_b200_simulate_dynamics_batch_device(s.h, B, vp(x), vp(u), n, vp(xn), sp))
            x, xn = xn, x
        c1.record(stream)
        torch.cuda.synchronize()
        sa = stats_all.cpu().numpy().view(pb.STATS_DTYPE).reshape(args.closed_loop_steps, -1)[:, :B]
        ms = c0.elapsed_time(c1)
        closed_loop = {"warm": True, "num_sim": args.closed_loop_steps, "batch": B,
                       "solves_per_s": float((sa["status"] == 1).sum() / (ms * 1e-3)),
                       "converged_fraction": float((sa["status"] == 1).mean()),
                       "mean_inner_iterations": float(sa["inner_iterations"].mean()),
                       "mean_hess_psi_prod": float(sa["ev_hess_psi_prod"].mean()),
                       "note": "shift + solve + plant step for every MPC step enqueued without host synchronisation "
                               "(drivers/mpc-driver.cpp:163-206 per instance); the first step is the timed bench step"}

    if rank == 0:
        names = ["ms_per_step", "kernel_ms", "sm_mhz_in_kernel", "slot_utilisation", "smi_sm_mhz_median", "smi_sm_mhz_min", "converged"]
        ranks = {nm: [r[i] for r in per_rank] for i, nm in enumerate(names)}
        ranks["slowest_rank"] = int(np.argmax(ranks["ms_per_step"]))
        line = {
            "metric": CONFIGS[args.config]["metric"], "value": value, "unit": "solves/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "weak" if args.shard == "weak" else "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_dict(args, world),
            "converged_fraction": conv_total / max(inst_total, 1),
            "mean_inner_iterations": float(np.mean(np.concatenate([s.stats["inner_iterations"] for s in segs]))),
            "mean_hess_psi_prod": float(np.mean(np.concatenate([s.stats["ev_hess_psi_prod"] for s in segs]))),
            "cold_solve": cold_info,
            "e2e": {"value": e2e_value, "unit": "solves/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "steps": e2e_steps, "warmup_steps": int(e2e_warmup),
                    "bitwise_equal_to_device_path": same_as_device},
            "gpu_launches": int(launches), "grid": dom.solver.grid_size(dom.prob),
            "per_rank": ranks, "stats_gather_ok": gathered_ok,
            "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu_baseline, "parity_check": parity,
            "closed_loop": closed_loop,
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
