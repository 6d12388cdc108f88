"""Where the per-step time of model.fit_generator goes (config 4, batch 16, 480x640): the same loop fed with
(A) constant device tensors, (B) pinned uint8 host images + constant device targets, (C) the full batch builder
(uint8 host images + anchor_targets_bbox_device from the annotation lists). Prints ms per step of each."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
H, W, C, B = 480, 640, 21, 16


def main():
    import torch
    from pyrapose_b200 import models, optimizers
    from pyrapose_b200.utils.anchors import anchor_targets_bbox_device, anchors_for_shape_device
    from synth import pose_annotations
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 40
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(4)
    raw = torch.from_numpy(rng.integers(0, 256, (B, H, W, 3)).astype(np.uint8)).pin_memory()
    anchors_dev = anchors_for_shape_device((H, W, 3))
    anns = [pose_annotations(i, C) for i in range(B)]
    shapes = [np.zeros((H, W, 3), np.uint8)] * B
    y_const = list(anchor_targets_bbox_device(anchors_dev, shapes, anns, C))
    x_const = raw.to(dev)
    model = models.backbone("resnet50").retinanet(num_classes=C, seed=1)
    model.compile(optimizer=optimizers.Adam(lr=1e-5, clipnorm=0.001))

    class Seq(object):
        def __init__(self, fn):
            self.fn = fn

        def __len__(self):
            return steps

        def __getitem__(self, i):
            return self.fn()

    def alone(tag):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(10):
            y = anchor_targets_bbox_device(anchors_dev, shapes, anns, C)
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        print("%s: anchor_targets_bbox_device %.3f ms host per call (+%.3f ms drain)" %
              (tag, 1e2 * (t1 - t0), 1e3 * (time.perf_counter() - t1)))

    alone("before any fit")
    variants = [
        ("A constant device tensors", lambda: (x_const, y_const)),
        ("B pinned uint8 images + constant device targets", lambda: (raw, y_const)),
        ("C uint8 images + anchor_targets_bbox_device", lambda: (raw, list(anchor_targets_bbox_device(anchors_dev, shapes, anns, C)))),
    ]
    for name, fn in variants:
        model.fit_generator(Seq(fn), steps_per_epoch=3, epochs=1, verbose=0)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        model.fit_generator(Seq(fn), steps_per_epoch=steps, epochs=1, verbose=0, shuffle=False)
        torch.cuda.synchronize()
        print("%-55s %.3f ms/step" % (name, 1e3 * (time.perf_counter() - t0) / steps))
        alone("after " + name[:1])
    import cProfile
    import pstats
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(10):
        y = anchor_targets_bbox_device(anchors_dev, shapes, anns, C)
    torch.cuda.synchronize()
    pr.disable()
    pstats.Stats(pr).sort_stats("tottime").print_stats(8)
    print(torch.cuda.memory_summary(abbreviated=True))
    # the batch builder alone, host time and device time
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        y = anchor_targets_bbox_device(anchors_dev, shapes, anns, C)
    e1.record()
    torch.cuda.synchronize()
    print("anchor_targets_bbox_device alone: %.3f ms host wall, %.3f ms device span per batch" %
          (1e3 * (time.perf_counter() - t0) / 20, e0.elapsed_time(e1) / 20))
    # plain device loop for reference
    eng = model._train_engine
    opt = model.optimizer
    torch.cuda.synchronize()
    e0.record()
    for _ in range(steps):
        eng.train_forward_backward()
        eng.train_apply(opt.lr, opt.beta_1, opt.beta_2, opt.epsilon, opt.clipnorm, 1.0)
    e1.record()
    torch.cuda.synchronize()
    print("device loop (graph + optimizer, nothing else): %.3f ms/step" % (e0.elapsed_time(e1) / steps))


if __name__ == "__main__":
    main()
