#!/usr/bin/env python
"""Training-step benchmark (BASELINE.json config 4): YCB-Video 21 classes, PFPN forward + losses + backward +
gradient all-reduce (NCCL) + Adam, batch 16 per GPU at 480x640, data parallel (one process per GPU).

  python tools/bench_train.py [--steps K] [--warmup W] [--batch B]
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/bench_train.py ...

Prints ONE JSON line on rank 0: images/s over all ranks (device events, max over ranks), the split of the step
(forward+backward graph / all-reduce / optimizer), the tcgen05 kernels' TFLOP/s (forward conv, dgrad, wgrad) from a
per-launch profile, and the losses of the last step. Targets are built once with anchor_targets_bbox from seeded
synthetic pose annotations (tests/synth.py) and stay resident; images are resident too (this is a kernel-level
number, the e2e number of the round is bench.py's)."""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

H, W, C = 480, 640, 21


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--profile", action="store_true", help="print the per-launch table")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    from pyrapose_b200 import models, optimizers
    from pyrapose_b200.utils.anchors import anchor_targets_bbox, anchors_for_shape
    from synth import pose_annotations

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    B = args.batch
    rng = np.random.default_rng(4 + rank)
    img = rng.integers(0, 256, (B, H, W, 3)).astype(np.float32) - np.array([103.939, 116.779, 123.68], np.float32)
    anchors = anchors_for_shape((H, W, 3))
    anns = [pose_annotations(1000 * rank + i, C) for i in range(B)]
    y = anchor_targets_bbox(anchors, [np.zeros((H, W, 3), np.float32)] * B, anns, C)
    model = models.backbone("resnet50").retinanet(num_classes=C, seed=1)  # same init on every rank
    model.compile(optimizer=optimizers.Adam(lr=1e-5, clipnorm=0.001))
    dev = torch.device("cuda", local)
    x_d = torch.from_numpy(img).to(dev)
    y_d = [torch.from_numpy(t).to(dev) for t in y]

    for _ in range(args.warmup):
        model.train_on_batch(x_d, y_d, sync_losses=False)
    eng = model._train_engine
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4 * args.steps + 1)]
    opt = model.optimizer
    buf = eng.buffers
    ev[0].record()
    for s in range(args.steps):
        buf["images"].copy_(x_d, non_blocking=True)
        buf["y_reg"].copy_(y_d[0], non_blocking=True)
        buf["y_cls"].copy_(y_d[1], non_blocking=True)
        buf["y_mask"].copy_(y_d[2], non_blocking=True)
        ev[4 * s + 1].record()
        eng.train_forward_backward()
        ev[4 * s + 2].record()
        if world > 1:
            dist.all_reduce(eng.grads_tensor())
        ev[4 * s + 3].record()
        eng.train_apply(opt.lr, opt.beta_1, opt.beta_2, opt.epsilon, opt.clipnorm, 1.0 / world)
        ev[4 * s + 4].record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    total_ms = ev[0].elapsed_time(ev[-1])
    parts = np.zeros(4)
    for s in range(args.steps):
        for k in range(4):
            parts[k] += ev[4 * s + k].elapsed_time(ev[4 * s + k + 1])
    parts /= args.steps
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    # ---- end to end through the public API: raw uint8 images from pinned host memory, targets built on the device from
    # the annotation lists (anchor_targets_bbox_device), train_on_batch, loss values read back every step
    raw = torch.from_numpy(rng.integers(0, 256, (B, H, W, 3)).astype(np.uint8)).pin_memory()
    from pyrapose_b200.utils.anchors import anchor_targets_bbox_device, anchors_for_shape_device
    anchors_dev = anchors_for_shape_device((H, W, 3))
    shapes = [np.zeros((H, W, 3), np.uint8)] * B
    for _ in range(2):
        model.train_on_batch(raw, anchor_targets_bbox_device(anchors_dev, shapes, anns, C))
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    import time
    t0 = time.perf_counter()
    for _ in range(args.steps):
        out = model.train_on_batch(raw, anchor_targets_bbox_device(anchors_dev, shapes, anns, C))
    torch.cuda.synchronize()
    loop_s = time.perf_counter() - t0
    # ---- the same through model.fit_generator (the reference's loop, bin/train.py:381-390): a Sequence whose
    # __getitem__ builds the batch (raw uint8 host images + device targets); the next batch is staged by a background
    # thread while the GPU runs the current step, every batch's losses are read back (one launch late)
    class Batches(object):
        def __len__(self):
            return args.steps

        def __getitem__(self, i):
            return raw, list(anchor_targets_bbox_device(anchors_dev, shapes, anns, C))

    class Count(object):
        n = 0

        def on_batch_end(self, batch, logs):
            Count.n += 1
            assert np.isfinite(logs["loss"])

    model.fit_generator(Batches(), steps_per_epoch=3, epochs=1, verbose=0)   # warm-up: thread, pinned slots, streams
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    model.fit_generator(Batches(), steps_per_epoch=args.steps, epochs=1, verbose=0, callbacks=[Count()], shuffle=False)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    assert Count.n == args.steps
    te = torch.tensor([e2e_s, loop_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_s, loop_s = float(te[0].item()), float(te[1].item())
    eng = model._train_engine
    losses = eng.train_losses()
    prof = eng.train_profile()
    groups = {}
    for name, fl, ms in prof:
        kind = name.split(":")[0] if ":" in name else ("fwd_conv" if fl > 0 else "fwd_other")
        g = groups.setdefault(kind, [0.0, 0.0, 0])
        g[0] += ms
        g[1] += fl
        g[2] += 1
    if rank == 0:
        if args.profile:
            for name, fl, ms in prof:
                print("%-44s %8.4f ms %8.1f TFLOP/s" % (name, ms, fl / ms * 1e-9 if ms > 0 else 0.0))
        n_params = int(eng.grads_tensor().numel())
        line = {
            "metric": "images_per_sec_pfpn_480x640_train_step", "value": B * world * args.steps / (total_ms * 1e-3),
            "unit": "images/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "dtype": "bf16",
            "data": "synthetic",
            "config": {"workload": "PFPN ResNet-50 training step (forward + focal/orthogonal_l1 losses + backward + "
                                   "NCCL gradient all-reduce + Adam clipnorm), YCB-Video 21 classes, batch %d per GPU, "
                                   "480x640" % B, "global_batch": B * world, "parallelism": "dp%d" % world,
                       "trainable_params": n_params},
            "split_ms": {"copy_inputs": parts[0], "forward_backward": parts[1], "allreduce": parts[2],
                         "optimizer": parts[3]},
            "kernels": {k: {"ms": v[0], "launches": v[2], "tflops": (v[1] / v[0] * 1e-9) if v[0] > 0 and v[1] > 0 else None}
                        for k, v in groups.items()},
            "e2e": {"value": B * world * args.steps / e2e_s, "unit": "images/s", "ms_per_step": 1e3 * e2e_s / args.steps,
                    "h2d_bytes_per_step": int(raw.numel()) + B * 8 * (4 + 16) * 8,
                    "d2h_bytes_per_step": 5 * 8 + 3 * 4,
                    "api": "model.fit_generator(Sequence -> (uint8 host images, anchor_targets_bbox_device(annotations))), "
                           "per-batch losses read back through the callbacks",
                    "train_on_batch_loop": {"value": B * world * args.steps / loop_s, "ms_per_step": 1e3 * loop_s / args.steps,
                                            "api": "anchor_targets_bbox_device(annotations) + model.train_on_batch(uint8 "
                                                   "host images) -> losses, synchronous"}},
            "losses": {k: losses[k] for k in ("3Dbox", "cls", "mask", "l2", "grad_norm")},
            "device_bytes": eng.device_bytes(),
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
