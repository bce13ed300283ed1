"""Multi-GPU host logic: robots are independent, so the batch is partitioned into contiguous index ranges,
one osc_b200.Batch per GPU / rank, with NO collective on the tick.  The only cross-rank steps are bookkeeping:
a barrier around timed regions, a MAX over ranks of the device time, and an optional gather of torques on the host.
torch.distributed is used for those (nccl on GPUs, gloo in the CPU tests); the tick itself never calls it."""
import numpy as np


def shard_range(n_robots: int, world: int, rank: int):
    """contiguous range [lo, hi) of rank `rank`: sizes differ by at most one, every robot exactly once"""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad world/rank")
    base, rem = divmod(int(n_robots), world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard(arrays: dict, n_robots: int, world: int, rank: int) -> dict:
    """slice every per-robot array [n_robots, ...] of a workload dict to this rank's range"""
    lo, hi = shard_range(n_robots, world, rank)
    return {k: (v[lo:hi] if isinstance(v, np.ndarray) and v.shape[:1] == (n_robots,) else v) for k, v in arrays.items()}


def max_over_ranks(x: float, dist=None, device=None) -> float:
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return float(x)
    import torch
    t = torch.tensor([x], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def gather_rows(local: np.ndarray, n_robots: int, dist=None):
    """host-side join of per-rank results into one [n_robots, k] array on rank 0 (None elsewhere)"""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return local
    import torch
    world, rank = dist.get_world_size(), dist.get_rank()
    k = local.shape[1]
    sizes = [shard_range(n_robots, world, r) for r in range(world)]
    pad = max(hi - lo for lo, hi in sizes)
    buf = torch.zeros((pad, k), dtype=torch.float64)
    buf[:local.shape[0]] = torch.from_numpy(np.ascontiguousarray(local, dtype=np.float64))
    out = [torch.zeros_like(buf) for _ in range(world)] if rank == 0 else None
    dist.gather(buf, out, dst=0)
    if rank != 0:
        return None
    return np.concatenate([out[r][:hi - lo].numpy() for r, (lo, hi) in enumerate(sizes)], axis=0)
