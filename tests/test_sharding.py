"""Host-side multi-GPU plumbing on CPU: world_size-2 gloo process group (the N > 1 path of bench.py without GPUs).
Scans shard by rank with no data-path collective; the collectives that exist are the timing MAX and the count gather."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gndnet_b200.pipeline import gather_shard_sizes, max_over_ranks, shard_range


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


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lo, hi = shard_range(129, rank, world)
        sizes = gather_shard_sizes(hi - lo)
        ms = max_over_ranks(10.0 + 5.0 * rank)
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
