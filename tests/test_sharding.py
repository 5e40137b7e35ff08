"""Multi-GPU host logic on CPU: frame sharding is a partition with no data-path collective; the only
cross-rank traffic is the harness (barrier, max-over-ranks time, count reduction). world_size-2 gloo."""
import os
import socket

import numpy as np
import pytest

from orb_slam2_2021_b200.shard import chunks, shard_range


@pytest.mark.parametrize("n,world", [(4096, 1), (4096, 2), (4096, 8), (7, 4), (3, 8), (0, 2)])
def test_shard_partition(n, world):
    ranges = [shard_range(n, r, world) for r in range(world)]
    assert ranges[0][0] == 0 and ranges[-1][1] == n
    assert all(a[1] == b[0] for a, b in zip(ranges, ranges[1:]))
    sizes = [e - b for b, e in ranges]
    assert max(sizes) - min(sizes) <= 1
    for b, e in ranges:
        cs = chunks(b, e, 256)
        assert sum(y - x for x, y in cs) == e - b and all(y - x <= 256 for x, y in cs)


def _worker(rank, world, port, n_frames, q):
    import torch
    import torch.distributed as dist

    from oracle import oracle as O
    from orb_slam2_2021_b200 import synth

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    b, e = shard_range(n_frames, rank, world)
    ex = O.OracleExtractor(300, 1.2, 4, 20, 7)  # CPU stand-in for the per-rank engine: same sharding code path
    count = 0
    for f in range(b, e):
        k, _ = ex(synth.synth(320, 240, 500 + 2 * f))
        count += len(k)
    t = torch.tensor([count, e - b], dtype=torch.int64)
    dist.all_reduce(t)  # harness-only reduction of counts
    elapsed = torch.tensor([float(rank + 1)])
    dist.all_reduce(elapsed, op=dist.ReduceOp.MAX)  # max-over-ranks timing rule
    dist.barrier()
    if rank == 0:
        q.put((int(t[0]), int(t[1]), float(elapsed[0])))
    dist.destroy_process_group()


def test_two_rank_gloo_sharding():
    import torch.multiprocessing as mp

    from oracle import oracle as O
    from orb_slam2_2021_b200 import synth

    n_frames, world = 5, 2
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_frames, q)) for r in range(world)]
    [p.start() for p in procs]
    total, frames, tmax = q.get(timeout=120)
    [p.join(60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    ex = O.OracleExtractor(300, 1.2, 4, 20, 7)
    ref = sum(len(ex(synth.synth(320, 240, 500 + 2 * f))[0]) for f in range(n_frames))
    assert (total, frames, tmax) == (ref, n_frames, float(world))
