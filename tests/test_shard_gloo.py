"""CPU tests of the multi-GPU host logic with the gloo backend, world size 2 (and the ragged 3-rank case):
partitioning and the all-gather of the 128-byte result records."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from laser_slam_b200 import shard
from laser_slam_b200.engine import RESULT_DTYPE


def test_partition_covers_everything():
    for total in (0, 1, 7, 1000, 1_000_000):
        for world in (1, 2, 3, 8):
            blocks = [shard.partition(total, world, r) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == total
            for a, b in zip(blocks, blocks[1:]):
                assert a[1] == b[0]
            sizes = [hi - lo for lo, hi in blocks]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard.partition(10, 2, 2)


def _records(lo, hi):
    r = np.zeros(hi - lo, RESULT_DTYPE)
    r["x"][:, 0] = np.arange(lo, hi)
    r["valid"] = 1
    r["iterations"] = np.arange(lo, hi) % 11
    r["corr_hash"] = (np.arange(lo, hi) * 2654435761) % (1 << 31)
    return r


def _worker(rank, world, port, total, ok):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lo, hi = shard.partition(total, world, rank)
        local = torch.from_numpy(_records(lo, hi).view(np.uint8).reshape(-1, 128).copy())
        out = shard.all_gather_records(local, total)
        got = out.numpy().view(RESULT_DTYPE).reshape(-1)
        want = _records(0, total)
        good = np.array_equal(got.view(np.uint8), want.view(np.uint8))
        if total % world == 0:
            # the overlapped (double-buffered) gather of bench.py: three steps through two slots, each step's records
            # differ, every result must be that step's
            g = shard.OverlappedGather(total, local, depth=2)
            for step in range(3):
                k = g.slot()
                assert k == step % 2
                loc = local.clone()
                loc[:, 100] = step + 1          # a byte of the record's reserved area marks the step
                g.submit(loc)
                res = g.result(k).numpy()
                w = want.view(np.uint8).reshape(-1, 128).copy()
                w[:, 100] = step + 1
                good = good and np.array_equal(res, w)
        ok[rank] = int(good)
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world,total", [(2, 1000), (2, 1001), (3, 10)])
def test_all_gather_records_gloo(world, total):
    ctx = mp.get_context("spawn")
    ok = ctx.Array("i", [0] * world)
    mp.spawn(_worker, args=(world, _free_port(), total, ok), nprocs=world, join=True)
    assert list(ok) == [1] * world


def test_best_of_starts():
    r = np.zeros(8, RESULT_DTYPE)
    r["valid"] = [1, 1, 0, 1, 0, 0, 0, 0]
    r["nvalid"] = 100
    r["error"] = [5, 3, 1, 3, 9, 9, 9, 9]
    best, idx = shard.best_of_starts(r, 4)
    assert idx.tolist() == [1, 0] and best["error"][0] == 3 and best["valid"][1] == 0
