"""N>1 host logic on the CPU with gloo, world_size 2: contiguous batch sharding and the
rank-ordered pose all-gather.  No CUDA kernels run here."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from scflow_b200.refiner import PoseGatherer, gather_poses, pack_pose, shard_range, unpack_pose


def test_shard_range_covers_everything():
    for total in (1, 7, 64, 512, 513):
        for world in (1, 2, 3, 4, 8):
            spans = [shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def test_pack_unpack_roundtrip():
    R = torch.randn(5, 3, 3)
    t = torch.randn(5, 3)
    R2, t2 = unpack_pose(pack_pose(R, t))
    assert torch.equal(R, R2) and torch.equal(t, t2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, total, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = shard_range(total, rank, world)
    full = torch.arange(total * 12, dtype=torch.float32).reshape(total, 12)
    got = gather_poses(full[lo:hi].clone(), total)
    ok = torch.equal(got, full)
    pg = PoseGatherer(total, "cpu")                    # the pipelined form: three steps through two slots
    for step in range(3):
        slot = pg.submit(full[lo:hi] + step)
        ok = ok and torch.equal(pg.result(slot), full + step)
    q.put((rank, ok))
    dist.destroy_process_group()


@pytest.mark.parametrize("total", [8, 7])
def test_gather_poses_world2(total):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, total, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(ok for _, ok in res), res
