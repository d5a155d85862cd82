"""Frame-pair sharding across the GPUs of one box.

The path shards by frame pair: pair k = (frame k, frame k+1) is independent of every other
pair once inputs are read-only (quirk Q1 dropped), a frame is never split, and there is no
data-path collective.  The only exchange is the end-of-batch gather of per-pair gesture
statistics (32 B per pair, ParallelVideoPyr.cpp:845-888) -- or, on request, flow fields -- to
rank 0, done with torch.distributed (NCCL over NVLink on the GPU box, gloo in CPU tests).
"""
from __future__ import annotations


def partition_pairs(n_pairs: int, world: int, rank: int):
    """Contiguous block of pair indices [start, stop) owned by ``rank``.  Contiguous (not
    round-robin) so that a rank uploads each interior frame once and reuses its pyramid for
    both pairs that share it.  Rank r needs frames start .. stop (inclusive)."""
    if world < 1 or not (0 <= rank < world) or n_pairs < 0:
        raise ValueError("bad partition arguments")
    base, rem = divmod(n_pairs, world)
    start = rank * base + min(rank, rem)
    stop = start + base + (1 if rank < rem else 0)
    return start, stop


def gather_to_rank0(local, counts, dst: int = 0, group=None):
    """Gathers per-pair rows (stats ``(n_local, 4)`` or flows ``(n_local, H, W)``) to ``dst``.
    ``counts[r]`` = rows of rank r (ragged allowed).  Returns the concatenated tensor on
    ``dst`` (rows in pair order) and None elsewhere."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    if world == 1:
        return local
    n_max = max(counts)
    pad = local
    if local.shape[0] < n_max:
        pad = torch.zeros((n_max,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        pad[:local.shape[0]] = local
    bufs = [torch.empty_like(pad) for _ in range(world)] if rank == dst else None
    dist.gather(pad.contiguous(), bufs, dst=dst, group=group)
    if rank != dst:
        return None
    return torch.cat([bufs[r][:counts[r]] for r in range(world)], dim=0)
