"""Host logic of the N>1 path (stream sharding + the timing protocol) on CPU with gloo,
world_size 2 - the same code bench.py runs under torchrun with NCCL on the GPU box."""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "voxtral.c_b200"))

import streams  # noqa: E402


def test_shards_cover_all_streams_once():
    for n in (0, 1, 7, 64, 65):
        for world in (1, 2, 3, 4, 8):
            got = []
            for r in range(world):
                a, b = streams.shard_streams(n, r, world)
                assert 0 <= a <= b <= n
                got += list(range(a, b))
            assert got == list(range(n))
            sizes = [streams.shard_streams(n, r, world) for r in range(world)]
            lens = [b - a for a, b in sizes]
            assert max(lens) - min(lens) <= 1
    with pytest.raises(ValueError):
        streams.shard_streams(4, 2, 2)


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    a, b = streams.shard_streams(5, rank, world)           # 5 streams of 60 s: ranks get 3 and 2
    dist.barrier()
    total, slowest = streams.aggregate(dist, "cpu", 60.0 * (b - a), 1.0 + rank)
    dist.barrier()
    q.put((rank, a, b, total, slowest))
    dist.destroy_process_group()


def test_gloo_world2_aggregation():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert [(r[1], r[2]) for r in res] == [(0, 3), (3, 5)]
    for r in res:
        assert r[3] == 300.0 and r[4] == 2.0          # SUM of audio seconds, MAX of elapsed
