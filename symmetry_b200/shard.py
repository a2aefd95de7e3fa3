"""Sharding of connections over GPUs / ranks.

Connections are independent (each `peer` is handled in isolation,
/root/reference/src/provider.ts:173-193), so the path shards by connection with
NO data-path collective: inside one context slot % n_gpus picks the GPU; across
processes (bench.py under torchrun) rank r owns a contiguous block of streams.
The only inter-rank traffic is the benchmark's barrier and the max over ranks
of the elapsed time."""
from __future__ import annotations


def owner_gpu(slot: int, n_gpus: int) -> int:
    return slot % n_gpus


def shard_streams(total: int, world: int, rank: int):
    """Contiguous block [lo, hi) of `total` streams for `rank`; blocks differ by at most one."""
    base, extra = divmod(total, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def max_over_ranks(value: float, dist=None, device=None) -> float:
    """Elapsed time of the job = the slowest rank's."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return value
    import torch

    t = torch.tensor([value], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def aggregate_throughput(units_per_rank: float, seconds: float, dist=None, device=None) -> float:
    """Whole-job units/s: sum of every rank's units over the slowest rank's time."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return units_per_rank / seconds
    import torch

    t = torch.tensor([units_per_rank], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item()) / max_over_ranks(seconds, dist, device)
