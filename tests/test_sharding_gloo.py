"""N > 1 host path on CPU: two gloo ranks shard a batch by instance index, each "solves" its shard
(here: the CPU oracle stands in for the device so the test runs without a GPU) and the per-instance
stats are gathered; the result must equal the single-process result, instance for instance."""
import os
import sys

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _inputs(B, N):
    rng = np.random.default_rng(5)
    x0 = np.zeros((B, 57))
    x0[:, 0:30:3] = (np.arange(10) + 1) / 10
    x0[:, 27:30] += rng.uniform(-0.05, 0.05, (B, 3))
    return x0, np.zeros((B, 3 * N))


def _worker(rank, world, port, B, N, out_dir):
    sys.path.insert(0, ROOT); sys.path.insert(0, HERE)
    import oracle_util as ou
    import pantr_cdc2023_experiments_b200 as pb
    from pantr_cdc2023_experiments_b200 import sharding
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    solver, _ = pb.make_solver(pb.benchmark_options("hanging_chain"))
    par = ou.pack_params(solver)
    x0, u = _inputs(B, N)
    lo, hi = sharding.shard_range(B, rank, world)
    ul, _, st = ou.solve_batch(0, N, par, x0[lo:hi], u[lo:hi])
    local = np.zeros(hi - lo, dtype=[("index", "i8"), ("status", "i4"), ("iters", "u4"), ("psi", "f8")])
    local["index"], local["status"], local["iters"], local["psi"] = np.arange(lo, hi), st[:, 0], st[:, 2], st[:, 6]
    allst = sharding.gather_stats(local, world)
    if rank == 0:
        np.save(os.path.join(out_dir, "gathered.npy"), allst)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharded_solve_equals_single_process(tmp_path):
    import oracle_util as ou
    import pantr_cdc2023_experiments_b200 as pb
    B, N, world = 5, 6, 2   # odd batch: ragged shards (3 + 2)
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(world, port, B, N, str(tmp_path)), nprocs=world, join=True)
    got = np.load(os.path.join(str(tmp_path), "gathered.npy"))
    solver, _ = pb.make_solver(pb.benchmark_options("hanging_chain"))
    x0, u = _inputs(B, N)
    _, _, st = ou.solve_batch(0, N, ou.pack_params(solver), x0, u)
    assert list(got["index"]) == list(range(B))
    np.testing.assert_array_equal(got["status"], st[:, 0].astype("i4"))
    np.testing.assert_array_equal(got["iters"], st[:, 2].astype("u4"))
    np.testing.assert_array_equal(got["psi"], st[:, 6])
