"""Instance sharding across GPUs: instances are independent, so the batch is split
by instance index into contiguous chunks, one per rank; there is no collective on the data path.
The reference's only parallelism is one process per experiment (`doit -n`,
new-benchmarks-paper/run-single-mpc-exp.sh:19); this is its batched analogue.
"""
import numpy as np


def shard_range(B: int, rank: int, world: int):
    """[lo, hi) of the instances rank owns: ceil(B/world) per rank, the last ranks may be short/empty."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    per = -(-B // world)
    lo = min(B, rank * per)
    return lo, min(B, lo + per)


def gather_stats(local_stats: np.ndarray, world: int, group=None):
    """Final per-instance stats gather (the only inter-rank traffic of a sharded solve): returns the
    concatenation over ranks on every rank.  gloo gathers host tensors; NCCL only moves device memory, so under
    an NCCL group the bytes are staged through the current CUDA device."""
    if world == 1:
        return local_stats
    import torch
    import torch.distributed as dist
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")
    raw = np.ascontiguousarray(local_stats).view(np.uint8).reshape(-1)
    sizes = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(sizes, torch.tensor([raw.size], dtype=torch.int64, device=dev), group=group)
    sizes = [int(s.item()) for s in sizes]
    mx = max(sizes)
    buf = torch.zeros(mx, dtype=torch.uint8, device=dev)
    buf[:raw.size] = torch.from_numpy(raw.copy()).to(dev)
    out = [torch.zeros(mx, dtype=torch.uint8, device=dev) for _ in range(world)]
    dist.all_gather(out, buf, group=group)
    parts = [o.cpu().numpy()[:sz].view(local_stats.dtype) for o, sz in zip(out, sizes)]
    return np.concatenate(parts)
