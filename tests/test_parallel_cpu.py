"""world_size-2 gloo tests (CPU) of the multi-GPU plumbing: arena sharding covers the arena set exactly once, and the
statistics reductions reproduce the single-process values of the union of the ranks' data."""
import importlib
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    par = importlib.import_module("reinforcement-learning_b200.parallel")
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(1234)
        full = torch.randn(10007, generator=g) * 3.0 + 0.7           # every rank draws the same union
        start, count = par.shard_arenas(full.numel(), world, rank)
        mine = full[start:start + count]
        mean, std = par.global_advantage_stats(mine)
        x = mine.double().numpy()
        n, gm, m2 = par.welford_merge_across_ranks(float(x.size), float(x.mean()), float(((x - x.mean()) ** 2).sum()))
        tmax = par.max_over_ranks(1.0 + rank)
        tsum = par.sum_over_ranks(float(count))
        q.put((rank, start, count, mean, std, n, gm, m2, tmax, tsum))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_two_rank_reductions_match_single_process():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=100) for _ in range(world))
    for p in procs:
        p.join(30)
        assert p.exitcode == 0
    g = torch.Generator().manual_seed(1234)
    full = (torch.randn(10007, generator=g) * 3.0 + 0.7).double()
    assert res[0][1] == 0 and res[0][1] + res[0][2] == res[1][1] and res[1][1] + res[1][2] == 10007
    for r in res:
        assert abs(r[3] - float(full.mean())) < 1e-9
        assert abs(r[4] - float(full.std(unbiased=True))) < 1e-9
        assert r[5] == 10007 and abs(r[6] - float(full.mean())) < 1e-9
        assert abs(r[7] - float(((full - full.mean()) ** 2).sum())) < 1e-6
        assert r[8] == 2.0 and r[9] == 10007.0


def test_shards_partition_the_arena_set():
    par = importlib.import_module("reinforcement-learning_b200.parallel")
    for total in (0, 1, 7, 4096, 65536, 65537):
        for world in (1, 2, 3, 4, 8):
            seen = np.zeros(total, np.int32)
            sizes = []
            for r in range(world):
                s, c = par.shard_arenas(total, world, r)
                seen[s:s + c] += 1
                sizes.append(c)
            assert (seen == 1).all() and max(sizes) - min(sizes) <= 1
    assert par.shard_arenas(65536, 8, 3) == (3 * 8192, 8192)
    assert len({par.rank_seed(123, r) for r in range(8)}) == 8
    with pytest.raises(ValueError):
        par.shard_arenas(8, 2, 2)
    assert par.welford_std(1, 0.0) == 1.0 and abs(par.welford_std(5, 16.0) - 2.0) < 1e-12
