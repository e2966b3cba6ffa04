"""One batch, really sharded over two GPUs (north_star: "sharded ... by instance index, ... no inter-GPU traffic beyond a
final stats gather"): both ranks derive the same global batch from one seed, rank r solves shard_range(B, r, world) on
its own device, the per-instance stats travel through sharding.gather_stats over NCCL, and the result must equal the
single-GPU solve of the whole batch, bit for bit.  Skipped on boxes with fewer than two devices."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _batch(name, N, B):
    sys.path.insert(0, ROOT)
    import pantr_cdc2023_experiments_b200 as pb
    rng = np.random.default_rng(9)
    prob = pb.BatchedProblem(name, N)
    x0 = np.tile(prob.initial_state(), (B, 1))
    if name == "hanging_chain":
        x0[:, 27:30] += rng.uniform(-0.05, 0.05, (B, 3))
    else:
        x0 += 0.01 * rng.standard_normal((B, 9))
    return x0, np.tile(prob.initial_guess(), (B, 1))


def _worker(rank, world, port, name, N, B, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import pantr_cdc2023_experiments_b200 as pb
    from pantr_cdc2023_experiments_b200 import sharding
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world,
                            device_id=torch.device("cuda", rank))
    x0, u0 = _batch(name, N, B)
    lo, hi = sharding.shard_range(B, rank, world)
    solver, _ = pb.make_solver(pb.benchmark_options(name))
    prob = pb.BatchedProblem(name, N, device=rank)
    u = u0[lo:hi].copy(); y = np.zeros((hi - lo, prob.m))
    st = solver(prob, u, y if prob.m else None, x0=x0[lo:hi])
    allst = sharding.gather_stats(st, world)                 # NCCL: staged through the device
    packed = np.concatenate([u, y], axis=1)
    allu = sharding.gather_stats(packed.reshape(-1), world)  # (test only: the iterates, to compare them too)
    if rank == 0:
        np.save(os.path.join(out_dir, "stats.npy"), allst)
        np.save(os.path.join(out_dir, "uy.npy"), allu)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("name,N,B", [("hanging_chain", 12, 37), ("quadcopter", 10, 21)])
def test_two_gpu_sharded_solve_equals_single_gpu(tmp_path, name, N, B):
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    sys.path.insert(0, ROOT)
    import pantr_cdc2023_experiments_b200 as pb
    world = 2
    port = 29600 + os.getpid() % 2000
    mp.spawn(_worker, args=(world, port, name, N, B, str(tmp_path)), nprocs=world, join=True)
    got = np.load(os.path.join(str(tmp_path), "stats.npy"))
    uy = np.load(os.path.join(str(tmp_path), "uy.npy"))
    x0, u0 = _batch(name, N, B)
    solver, _ = pb.make_solver(pb.benchmark_options(name))
    prob = pb.BatchedProblem(name, N, device=0)
    u = u0.copy(); y = np.zeros((B, prob.m))
    st = solver(prob, u, y if prob.m else None, x0=x0)
    assert len(got) == B
    for k in st.dtype.names:
        if k not in ("cycles", "elapsed_us"):
            np.testing.assert_array_equal(got[k], st[k], err_msg=k)
    np.testing.assert_array_equal(uy.reshape(B, -1), np.concatenate([u, y], axis=1))
