"""N > 1 host logic on CPU: world-size-2 gloo processes exercise the shard arithmetic, the
gradient-bucket all-reduce (+ 1/world scaling, identical parameters on every rank afterwards)
and the max-over-ranks timing reduction that bench.py uses."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from hanabi_agents_b200 import dist as hd


def test_shard_partition_properties():
    for n_total in (0, 1, 7, 65536, 1048576 + 3):
        for world in (1, 2, 3, 8):
            spans = [hd.shard(n_total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and sum(n for _, n in spans) == n_total
            for (b0, n0), (b1, _) in zip(spans, spans[1:]):
                assert b0 + n0 == b1
            sizes = [n for _, n in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        hd.shard(10, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        assert hd.env_world() == (rank, rank, world)
        # each rank computes gradients on its own replay shard ...
        g = torch.Generator().manual_seed(100 + rank)
        bucket = torch.randn(1000, generator=g)
        mine = bucket.clone()
        hd.allreduce_sum_(bucket)
        # ... and every rank ends with the same averaged gradient, hence the same parameters
        theta = torch.zeros(1000) - 0.1 * bucket / world
        gathered = [torch.zeros(1000) for _ in range(world)]
        dist.all_gather(gathered, theta)
        assert all(torch.equal(gathered[0], t) for t in gathered)
        others = [torch.randn(1000, generator=torch.Generator().manual_seed(100 + r)) for r in range(world)]
        torch.testing.assert_close(bucket, sum(others))
        assert not torch.equal(mine, bucket)
        # slowest rank defines the step time; throughput sums over ranks
        assert hd.max_over_ranks(10.0 + rank) == 10.0 + world - 1
        assert hd.sum_over_ranks(1.0 + rank) == sum(1.0 + r for r in range(world))
        base, n = hd.shard(65536 * world + 1, rank, world)
        spans = [None] * world
        dist.all_gather_object(spans, (base, n))
        assert spans[0][0] == 0 and all(a[0] + a[1] == b[0] for a, b in zip(spans, spans[1:]))
        out.put((rank, "ok"))
    finally:
        dist.destroy_process_group()


def test_world2_gloo():
    ctx = mp.get_context("spawn")
    out = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    got = sorted(out.get() for _ in range(2))
    assert got == [(0, "ok"), (1, "ok")]


def test_single_process_helpers_are_noops():
    t = torch.ones(4)
    assert hd.allreduce_sum_(t) is t and torch.equal(t, torch.ones(4))
    assert hd.max_over_ranks(3.5) == 3.5 and hd.sum_over_ranks(2.0) == 2.0
