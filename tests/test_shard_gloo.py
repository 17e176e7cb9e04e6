"""N>1 host logic on CPU: two gloo ranks each take a contiguous shard (+32 KiB halo),
compress it with the CPU twin of the kernels (test infrastructure standing in for the
GPU), and rank 0 concatenates and folds CRCs.  No collective carries payload bytes
other than the final gather of results, as on the real multi-GPU path."""
import os
import socket
import sys
import zlib

import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

import synthdata
from helpers import twin_deflate

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, q):
    sys.path.insert(0, os.path.join(ROOT, "python-zlib-ng_b200"))
    from zlib_ng_b200 import _native, sharding
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    data = synthdata.make("mixed", n, 12).tobytes()          # every rank can regenerate the stream
    start, end = sharding.plan_shards(n, world)[rank]
    zdict = sharding.shard_dictionary(data, start)
    blob = twin_deflate(data[start:end], 6, zdict) if end > start else b""
    part = (blob, zlib.crc32(data[start:end]), end - start)
    gathered = [None] * world
    dist.all_gather_object(gathered, part)
    if rank == 0:
        stream, crc, total = sharding.fold(gathered, _native.lib().b2d_crc32_combine)
        ok = (zlib.decompressobj(-15).decompress(stream) == data and crc == zlib.crc32(data) and total == n)
        # sharding must not change the bytes: same stream as one unsharded job
        ok = ok and stream == twin_deflate(data, 6)
        q.put(ok)
    dist.destroy_process_group()


@pytest.mark.parametrize("n", [1_000_000, 131072 * 3, 5])
def test_two_rank_shard_and_fold(n):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True


def test_plan_shards_properties():
    sys.path.insert(0, os.path.join(ROOT, "python-zlib-ng_b200"))
    from zlib_ng_b200 import sharding
    for n in (0, 1, 131071, 131072, 131073, 10 * 131072 + 7, (1 << 30) + 3):
        for world in (1, 2, 3, 8):
            plan = sharding.plan_shards(n, world)
            assert len(plan) == world and plan[0][0] == 0 and plan[-1][1] == n
            for (a, b), (c, d) in zip(plan, plan[1:]):
                assert b == c and a <= b
            for a, b in plan:
                assert a % sharding.BLOCK == 0 or a == n
            sizes = [(-(-(b - a) // sharding.BLOCK)) for a, b in plan]
            assert max(sizes) - min(sizes) <= 1
