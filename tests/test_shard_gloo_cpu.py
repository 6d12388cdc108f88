"""N > 1 host logic on CPU: two gloo ranks shard a batch of anchor-target work items, each computes its slice
with the CPU oracle (standing in for the per-rank GPU work), and rank 0 reassembles the batch in image order.
Also checks the max-over-ranks timing reduction bench.py relies on."""
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


def _worker(rank, world, port, tmpdir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import anchors as oa, overlap as oo
        from pyrapose_b200.utils import shard
        total = 5
        anchors = oa.anchors_for_shape((64, 96, 3))
        begin, end = shard.shard_range(total, rank, world)
        pos, arg = [], []
        for i in range(begin, end):
            rng = np.random.default_rng(1000 + i)
            x1 = rng.uniform(0, 60, 6)
            y1 = rng.uniform(0, 40, 6)
            q = np.stack([x1, y1, x1 + rng.uniform(8, 40, 6), y1 + rng.uniform(8, 30, 6)], axis=1)
            p, _, a, _ = oo.gt_assign(anchors, q)
            pos.append(p)
            arg.append(a)
        out = shard.gather_arrays([np.stack(pos), np.stack(arg)])
        slowest = shard.max_over_ranks(10.0 + rank)
        assert slowest == 10.0 + world - 1
        if rank == 0:
            np.savez(os.path.join(tmpdir, "out.npz"), pos=out[0], arg=out[1])
        else:
            assert out is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_two_rank_sharded_assignment_matches_single_process(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    from oracle import anchors as oa, overlap as oo
    got = np.load(str(tmp_path / "out.npz"))
    anchors = oa.anchors_for_shape((64, 96, 3))
    for i in range(5):
        rng = np.random.default_rng(1000 + i)
        x1 = rng.uniform(0, 60, 6)
        y1 = rng.uniform(0, 40, 6)
        q = np.stack([x1, y1, x1 + rng.uniform(8, 40, 6), y1 + rng.uniform(8, 30, 6)], axis=1)
        p, _, a, _ = oo.gt_assign(anchors, q)
        assert np.array_equal(got["pos"][i], p) and np.array_equal(got["arg"][i], a)
