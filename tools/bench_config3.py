"""BASELINE.json config 3 side by side: anchor-target IoU assignment for B = 64 images at
480x640, K = 100 GT boxes per image (seeds of SURVEY.md section 8d), N = 56 700 anchors (3 levels) and the 5-level
variant N = 57 600.

  GPU   ppb_compute_overlap_f64 (the materialised (N, K) matrices, 2.9 GB) and the fused ppb_gt_assign_f64
  CPU   the reference's OWN Cython compute_overlap (oracle/_ref, compiled from the reference's .pyx where it lies) +
        the numpy argmax / threshold lines of PyraPose/utils/anchors.py:310-316 — on one core (the reference has no
        threading) and sharded by image over all host cores with multiprocessing; when oracle/_ref is absent the C
        port (oracle/compute_overlap_port.c, bit-identical, ~100x faster than the Cython build's per-pair Python
        float call) is timed instead and labelled so.
Prints one JSON line. Usage: python tools/bench_config3.py [--images 64] [--cpu-images 16]"""
import argparse
import json
import multiprocessing as mp
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np


def gt_boxes(i, K=100):
    rng = np.random.default_rng(1000 + i)
    x1 = rng.uniform(0, 624, K); y1 = rng.uniform(0, 464, K)
    w = rng.uniform(16, 256, K); h = rng.uniform(16, 256, K)
    return np.stack([x1, y1, np.minimum(x1 + w, 640), np.minimum(y1 + h, 480)], axis=1)


_state = {}


def _anchors(levels):
    from oracle import anchors as oa
    prm = oa.AnchorParameters.default
    if len(levels) == 5:   # the reference's unused 5-level builder (models/retinanet.py:134-157): RetinaNet's sizes / strides
        prm = oa.AnchorParameters(sizes=[32, 64, 128, 256, 512], strides=[8, 16, 32, 64, 128], ratios=prm.ratios,
                                  scales=prm.scales)
    return oa.anchors_for_shape((480, 640, 3), pyramid_levels=levels, anchor_params=prm)


def _init(levels):
    from oracle import anchors as oa, build as obuild, overlap as oo
    fn = obuild.load_ref_compute_overlap()
    _state["kind"] = "reference" if fn is not None else "port"
    _state["fn"] = fn if fn is not None else oo.compute_overlap
    _state["anchors"] = _anchors(levels)


def _one(i):
    ov = _state["fn"](_state["anchors"].astype(np.float64), gt_boxes(i).astype(np.float64))   # anchors.py:310
    arg = np.argmax(ov, axis=1)                                                               # :311
    mx = ov[np.arange(ov.shape[0]), arg]                                                      # :312
    pos = mx >= 0.5                                                                           # :315
    ign = (mx > 0.4) & ~pos                                                                   # :316
    return int(pos.sum()), int(ign.sum())


def cpu_side(levels, n_images, procs):
    t0 = time.perf_counter()
    if procs == 1:
        _init(levels)
        out = [_one(i) for i in range(n_images)]
    else:
        with mp.get_context("fork").Pool(procs, initializer=_init, initargs=(levels,)) as pool:
            pool.map(_one, range(procs))   # warm the workers (imports, anchors)
            t0 = time.perf_counter()
            out = pool.map(_one, range(n_images))
    return time.perf_counter() - t0, out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--images", type=int, default=64)
    ap.add_argument("--cpu-images", type=int, default=16, help="images of the single-core sample")
    args = ap.parse_args()
    B, K = args.images, 100
    level_sets = ([3, 4, 5], [3, 4, 5, 6, 7])
    # CPU side first: the worker pool is forked before this process touches CUDA
    cpu = []
    for levels in level_sets:
        t1, out1 = cpu_side(levels, args.cpu_images, 1)
        tn, outn = cpu_side(levels, B, os.cpu_count())
        cpu.append((t1, out1, tn, outn))
    import torch
    from pyrapose_b200 import _kernels
    dev = torch.device("cuda:0")
    q = torch.from_numpy(np.stack([gt_boxes(i) for i in range(B)])).to(dev)
    res = {"config": "anchor-target assignment (compute_overlap IoU), 480x640, %d GT boxes per image, batch %d" % (K, B),
           "cores": os.cpu_count(), "runs": []}

    def ev(fn, reps=5):
        fn(); torch.cuda.synchronize()
        ts = []
        for _ in range(reps):
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        return float(np.median(ts))

    for levels, (t1, out1, tn, outn) in zip(level_sets, cpu):
        anchors = _anchors(levels)
        a_dev = torch.from_numpy(anchors).to(dev)
        N = anchors.shape[0]
        ms_mat = ev(lambda: _kernels.compute_overlap(a_dev, q))
        ms_fused = ev(lambda: _kernels.gt_assign(a_dev, q))
        pos, ign, arg = _kernels.gt_assign(a_dev, q)
        gpu_counts = [(int(pos[i].sum()), int(ign[i].sum())) for i in range(B)]
        assert outn == gpu_counts and out1 == gpu_counts[:args.cpu_images], "GPU and CPU assignments differ"
        res["runs"].append({
            "levels": levels, "anchors": N,
            "gpu_materialised_ms": ms_mat, "gpu_materialised_GBps": B * N * K * 8 / (ms_mat * 1e-3) / 1e9,
            "gpu_fused_ms": ms_fused, "gpu_fused_images_per_s": B / (ms_fused * 1e-3),
            "cpu_kind": _state.get("kind", "port"),
            "cpu_1core_s_per_image": t1 / args.cpu_images, "cpu_1core_images_per_s": args.cpu_images / t1,
            "cpu_1core_sample": "%d images" % args.cpu_images,
            "cpu_all_cores_s": tn, "cpu_all_cores_images_per_s": B / tn,
            "positive_ignore_counts_equal": True})
    print(json.dumps(res))


if __name__ == "__main__":
    main()
