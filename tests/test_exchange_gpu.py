"""Peer-to-peer gradient exchange (ppb_p2p_*, utils/exchange.py): the data-parallel step's alternative to the
overlapped NCCL all-reduce (compile(exchange='p2p')). The sum kernel runs on one GPU; the exchange itself needs two (skipped otherwise — the
driver-run bench covers it at N > 1 through train.dp_check)."""
import ctypes
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_sum_kernel_adds_in_rank_order(cuda_device):
    import torch
    from pyrapose_b200 import _native as nat
    lib = nat.load()
    rng = np.random.default_rng(0)
    for n, world in ((1 << 20, 2), (1000003, 8), (5, 3)):
        srcs = [torch.from_numpy(rng.standard_normal(n).astype(np.float32)).to(cuda_device) for _ in range(world)]
        want = srcs[0].clone()
        for t in srcs[1:]:
            want = want + t          # left to right, fp32: the kernel's order
        out = srcs[1]                 # in place on one of the sources, as the training step does
        ptrs = (ctypes.c_void_p * world)(*[t.data_ptr() for t in srcs])
        nat.check(lib.ppb_p2p_sum_f32(ctypes.c_void_p(out.data_ptr()), ptrs, world, ctypes.c_longlong(n),
                                      nat.current_stream_ptr()))
        torch.cuda.synchronize()
        assert torch.equal(out, want)
    with pytest.raises(ValueError):
        nat.check(lib.ppb_p2p_sum_f32(ctypes.c_void_p(srcs[0].data_ptr()), ptrs, 0, ctypes.c_longlong(4), None))


def _replica(rank, world, port, out_dir):
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank), LOCAL_WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    from pyrapose_b200 import models, optimizers
    from pyrapose_b200.utils.anchors import anchor_targets_bbox, anchors_for_shape
    from synth import pose_annotations
    C, H, W = 3, 96, 128
    rng = np.random.default_rng(10 + rank)
    img = rng.integers(0, 256, (1, H, W, 3)).astype(np.float32) - np.array([103.939, 116.779, 123.68], np.float32)
    y = anchor_targets_bbox(anchors_for_shape((H, W, 3)), [np.zeros((H, W, 3), np.float32)], [pose_annotations(70 + rank, C)], C)
    res = {}
    for mode in ("nccl", "p2p"):
        m = models.backbone("resnet50").retinanet(num_classes=C, seed=1)
        m.compile(optimizer=optimizers.Adam(lr=1e-5, clipnorm=0.001), exchange=mode, bucket_bytes=8 << 20)
        for _ in range(3):   # three steps: both staging parities and the reuse of the first are exercised
            m.train_on_batch(img, y)
        eng = m._train_engine
        assert bool(eng.__dict__.get("_p2p")) == (mode == "p2p")
        res[mode] = (eng.train_read("grads", 0, eng.grads_tensor().numel()), eng.train_read("weights", 0, eng.grads_tensor().numel()),
                     m.last_step["grad_norm"])
    np.savez(os.path.join(out_dir, "r%d.npz" % rank), g_nccl=res["nccl"][0], g_p2p=res["p2p"][0], w_nccl=res["nccl"][1],
             w_p2p=res["p2p"][1], n=np.array([res["nccl"][2], res["p2p"][2]]))
    dist.barrier()
    dist.destroy_process_group()


def test_two_replicas_p2p_equals_nccl(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    mp.spawn(_replica, args=(2, 29533, str(tmp_path)), nprocs=2, join=True)
    r = [np.load(str(tmp_path / ("r%d.npz" % i))) for i in range(2)]
    for k in ("g_p2p", "w_p2p"):
        assert np.array_equal(r[0][k], r[1][k]), "replicas diverged on %s" % k
    # two replicas: a + b in either order, so the copy-engine exchange and NCCL's all-reduce agree to the last bit
    # in the gradient of the first step; after three steps allow fp32 noise from NCCL's own reduction order
    g0, g1 = r[0]["g_nccl"].astype(np.float64), r[0]["g_p2p"].astype(np.float64)
    assert np.linalg.norm(g0 - g1) <= 1e-5 * np.linalg.norm(g0)
    assert abs(r[0]["n"][0] - r[0]["n"][1]) <= 1e-5 * r[0]["n"][0]
    d = r[0]["w_nccl"].astype(np.float64) - r[0]["w_p2p"]
    assert np.abs(d).max() <= 1e-6
