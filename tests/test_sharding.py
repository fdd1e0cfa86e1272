"""Host-side multi-GPU plumbing on CPU: world_size-2 gloo process group (the N > 1 path of bench.py without GPUs).
Scans shard by rank with no data-path collective; the collectives that exist are the timing MAX and the count gather."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gndnet_b200.pipeline import gather_rates, gather_shard_sizes, max_over_ranks, rate_shards, shard_range


def test_shard_range_partitions_exactly():
    for world in (1, 2, 3, 4, 8):
        for n in (0, 1, 7, 8, 64, 65, 511, 512):
            parts = [shard_range(n, r, world) for r in range(world)]
            assert parts[0][0] == 0 and parts[-1][1] == n
            assert all(parts[i][1] == parts[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in parts]
            assert max(sizes) - min(sizes) <= 1 and sorted(sizes, reverse=True) == sizes
    with pytest.raises(ValueError):
        shard_range(8, 2, 2)


def test_rate_shards_follow_the_link_rates():
    # the measured 8-GPU box: 20.5 GB/s to GPUs 0-3, 36 GB/s to GPUs 4-7
    s = rate_shards([20.5] * 4 + [36.0] * 4, 512, 32, 128)
    assert sum(s) == 512 and max(s[:4]) - min(s[:4]) <= 1 and max(s[4:]) - min(s[4:]) <= 1 and s[0] in (46, 47) and s[4] in (81, 82)
    assert abs(s[0] / s[4] - 20.5 / 36.0) < 0.02
    assert rate_shards([55.4, 55.4], 128, 32, 128) == [64, 64]           # equal links: equal shards
    assert rate_shards([1.0, 100.0], 128, 32, 128) == [32, 96]           # the bounds hold
    assert rate_shards([3.0], 64, 8, 128) == [64]
    # a skewed but feasible request: the fast rank clips at the upper bound, the slow one takes the rest
    assert rate_shards([1.0, 100.0], 128, 16, 100) == [28, 100]
    assert rate_shards([1.0, 1.0, 100.0], 200, 10, 100) == [50, 50, 100]
    assert rate_shards([1.0, 50.0, 50.0], 100, 20, 60) == [20, 40, 40]
    for rates, total, lo, hi in ([[1.0, 2.0, 4.0, 8.0, 16.0, 32.0, 64.0, 128.0], 512, 32, 128], [[5.0, 1.0, 1.0], 96, 32, 32],
                                 [[1e-3, 1e3], 65, 1, 64]):
        s = rate_shards(rates, total, lo, hi)
        assert sum(s) == total and min(s) >= lo and max(s) <= hi
        assert all(s[i] <= s[j] for i in range(len(s)) for j in range(len(s)) if rates[i] < rates[j])   # monotone in the rate
    for bad in ([], [1.0, 0.0], [1.0, float("nan")]):
        with pytest.raises(ValueError):
            rate_shards(bad, 128, 32, 128)
    with pytest.raises(ValueError):
        rate_shards([1.0, 1.0], 300, 32, 128)


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lo, hi = shard_range(129, rank, world)
        sizes = gather_shard_sizes(hi - lo)
        ms = max_over_ranks(10.0 + 5.0 * rank)
        # link-rate-proportional e2e shards: every rank gathers the rates and computes the same split
        shards = rate_shards(gather_rates(20.0 + 16.0 * rank), 128, 32, 128)
        assert sum(shards) == 128 and shards == [46, 82], shards
        # every scan is owned by exactly one rank: gather the ownership masks
        own = torch.zeros(129, dtype=torch.int32)
        own[lo:hi] = 1
        dist.all_reduce(own)
        dist.barrier()
        q.put((rank, sizes, ms, bool((own == 1).all())))
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, sizes, ms, ok in res:
        assert sizes == [65, 64] and ms == 15.0 and ok
