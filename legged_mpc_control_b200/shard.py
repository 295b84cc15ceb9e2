"""Batch-index sharding across the GPUs of one box.

Every QP is independent, so rank g owns the contiguous range [g*B/G, (g+1)*B/G) and the data
path has no collective.  The only exchange is the final result gather to rank 0, done with
torch.distributed (NCCL on GPUs, gloo in the CPU tests)."""
import numpy as np


def shard_range(total: int, rank: int, world: int):
    """Contiguous balanced partition; the first (total % world) ranks get one extra problem."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    base, rem = divmod(int(total), world)
    start = rank * base + min(rank, rem)
    return start, start + base + (1 if rank < rem else 0)


def shard_batch(batch: dict, rank: int, world: int) -> dict:
    n = len(batch["x0"])
    lo, hi = shard_range(n, rank, world)
    return {k: v[lo:hi] for k, v in batch.items() if isinstance(v, np.ndarray) and len(v) == n}


def gather_results(local, total: int, rank: int, world: int, dst: int = 0):
    """Gather per-rank result rows (a torch tensor [n_local, ...]) into one [total, ...] tensor on
    rank `dst` (None elsewhere).  Uneven shards are padded to the largest shard for the gather."""
    import torch
    import torch.distributed as dist
    if world == 1:
        return local
    sizes = [shard_range(total, r, world) for r in range(world)]
    nmax = max(hi - lo for lo, hi in sizes)
    pad = torch.zeros((nmax,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    bufs = [torch.empty_like(pad) for _ in range(world)] if rank == dst else None
    dist.gather(pad, bufs, dst=dst)
    if rank != dst:
        return None
    return torch.cat([bufs[r][: hi - lo] for r, (lo, hi) in enumerate(sizes)], dim=0)
