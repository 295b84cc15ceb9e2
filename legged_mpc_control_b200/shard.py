"""Batch-index sharding across the GPUs of one box.

Every QP is independent, so rank g owns the contiguous range [g*B/G, (g+1)*B/G) and the data
path has no collective.  The only exchange is the final result gather to rank 0, which goes through
HOST memory (a gloo group over the ranks' CPU copies): NCCL is not used for it — north_star: "NCCL is
not used, because the work does not reduce across GPUs".  (A single process that drives several GPUs
needs no gather at all: lmpc_solve_sharded writes every shard into its slice of one host array.)"""
import numpy as np

_HOST_GROUP = None


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
    """Gather per-rank result rows (a torch tensor [n_local, ...], host or device) into one
    [total, ...] HOST tensor on rank `dst` (None elsewhere) through a gloo group: device results are
    copied to host memory first, so no NCCL collective is issued.  Uneven shards are padded to the
    largest shard for the gather."""
    import torch
    import torch.distributed as dist
    if world == 1:
        return local.cpu()
    global _HOST_GROUP
    if _HOST_GROUP is None:
        _HOST_GROUP = dist.new_group(backend="gloo") if dist.get_backend() != "gloo" else dist.group.WORLD
    local = local.cpu()
    sizes = [shard_range(total, r, world) for r in range(world)]
    nmax = max(hi - lo for lo, hi in sizes)
    pad = torch.zeros((nmax,) + tuple(local.shape[1:]), dtype=local.dtype)
    pad[: local.shape[0]] = local
    bufs = [torch.empty_like(pad) for _ in range(world)] if rank == dst else None
    dist.gather(pad, bufs, dst=dst, group=_HOST_GROUP)
    if rank != dst:
        return None
    return torch.cat([bufs[r][: hi - lo] for r, (lo, hi) in enumerate(sizes)], dim=0)
