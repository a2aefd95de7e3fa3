"""N>1 host logic on CPU: two gloo ranks shard the streams with no data-path collective; the only
inter-rank operations are the barrier and the max/sum reductions bench.py uses for its JSON line."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from symmetry_b200 import shard, synth


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, total):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = shard.shard_streams(total, world, rank)
    keys, hdrs = synth.stream_keys(hi - lo, first=lo)
    # every rank derives its own streams; gather only the *ranges* and a digest to prove disjoint coverage
    mine = torch.tensor([lo, hi, int(keys.astype("int64").sum())], dtype=torch.int64)
    got = [torch.zeros(3, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(got, mine)
    ranges = sorted((int(g[0]), int(g[1])) for g in got)
    assert ranges[0][0] == 0 and ranges[-1][1] == total
    for a, b in zip(ranges, ranges[1:]):
        assert a[1] == b[0], "shards must tile the stream set without overlap"
    all_keys, _ = synth.stream_keys(total)
    assert sum(int(g[2]) for g in got) == int(all_keys.astype("int64").sum())
    # timing reductions: slowest rank's time, summed units
    secs = 1.0 + rank
    assert shard.max_over_ranks(secs, dist) == float(world)
    assert abs(shard.aggregate_throughput(hi - lo, secs, dist) - total / world) < 1e-9
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharding_with_gloo():
    world, total = 2, 1025
    mp.spawn(_worker, args=(world, _free_port(), total), nprocs=world, join=True)


def test_slot_to_gpu_and_block_sharding():
    for n in (1, 2, 4, 8):
        owners = [shard.owner_gpu(s, n) for s in range(64)]
        assert set(owners) == set(range(n))
        covered = []
        for r in range(n):
            lo, hi = shard.shard_streams(65536 + 3, n, r)
            covered.append((lo, hi))
        assert covered[0][0] == 0 and covered[-1][1] == 65539
        assert all(a[1] == b[0] for a, b in zip(covered, covered[1:]))
        sizes = [hi - lo for lo, hi in covered]
        assert max(sizes) - min(sizes) <= 1
