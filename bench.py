#!/usr/bin/env python
"""Benchmark of the PyraPose hot path on B200: PFPN inference (ResNet-50 -> PFPN -> heads) + 3D-box-corner decode
+ FilterDetections, batch 32 per GPU at 480x640 (BASELINE.json configs[1]); metric images/s.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  torchrun --nproc-per-node N bench.py --gpus N ...      (one rank per GPU, weak scaling, no data-path collective)

Prints ONE JSON line (rank 0). `value` = device-resident throughput (inputs already in HBM), `e2e` = the same step
through model.detect_on_batch with pinned HOST images in and HOST detections out. `--impl reference` times the CPU
restatement of the reference path (oracle/, torch-CPU fp32 on all host cores; the reference's TF/Keras stack is
not installable here) on a bounded sample of the same workload.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

H, W, C, A = 480, 640, 15, 9
BATCH = 32
METRIC = "images_per_sec_pfpn_480x640_infer_decode_filter"


def conv_flops_per_image(h, w, num_classes, num_anchors=9):
    """Algorithmic conv FLOPs (2*MAC) of one forward pass, from the layer table (234.73 G @ C=15)."""
    from pyrapose_b200.models import spec
    lv = [((h + (8 << i) - 1) // (8 << i), (w + (8 << i) - 1) // (8 << i)) for i in range(3)]
    res = {}
    ch, cw = (h + 1) // 2, (w + 1) // 2
    res["conv1"] = (ch, cw)
    ch, cw = (ch + 1) // 2, (cw + 1) // 2
    for stage, blocks in enumerate(spec.RESNET50_BLOCKS):
        for b in range(blocks):
            tag = "%d%s" % (stage + 2, chr(ord("a") + b))
            if b == 0 and stage > 0:
                ch, cw = (ch + 1) // 2, (cw + 1) // 2
            for br in ("branch2a", "branch2b", "branch2c", "branch1"):
                res["res%s_%s" % (tag, br)] = (ch, cw)
    pf = {"P3_lateral": lv[0], "P4_lateral": lv[1], "P5_lateral": lv[2], "P4_mid": lv[1], "P3_mid": lv[0],
          "P3_down": lv[1], "P3": lv[0], "P4_down": lv[2], "P4": lv[1], "P5": lv[2]}
    total = 0.0
    all_locs = sum(a * b for a, b in lv)
    for l in spec.all_layers(num_classes, num_anchors):
        if l.name in res:
            px = res[l.name][0] * res[l.name][1]
        elif l.name in pf:
            px = pf[l.name][0] * pf[l.name][1]
        elif l.name.startswith("mask"):
            px = lv[0][0] * lv[0][1]
        else:
            px = all_locs
        total += 2.0 * px * l.ksize * l.ksize * l.cin * l.cout
    return total


def bench_weights(args):
    """Random-init weights of the reference architecture. The reference's own initialisers give every score
    sigmoid(-4.595) = 0.01 and nothing passes the 0.05 threshold, so the heads are drawn
    he-normal (logit std ~1.5) and the classification bias is set so that logits ~ N(cls_bias, 1.5^2):
    -6.0 -> ~2 %% of the 850 500 scores per image pass 0.05 (primary setting), -4.6 -> ~13 %% (dense stress)."""
    from pyrapose_b200.models import spec
    w = spec.init_weights(C, A, seed=1, head_init="spread")
    w["pyramid_classification/bias"] = np.full_like(w["pyramid_classification/bias"], args.cls_bias)
    return w


def make_images(seed, batch):
    rng = np.random.default_rng(seed)
    img = rng.integers(0, 256, (batch, H, W, 3)).astype(np.float32)
    img -= np.array([103.939, 116.779, 123.68], dtype=np.float32)
    return img


class ClockSampler(object):
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows = []
        self.proc = None
        self.idx = gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.QUERY,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
                for n, v in zip(names, r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def _build_info():
    """Which build of libpyrapose_b200.so ran: source digest, compiler, flags (pyrapose_b200/_lib/build_info.json)."""
    try:
        from pyrapose_b200 import build as native_build
        info = native_build.build_info()
        if info is None:
            return None
        return {"sources_sha256": info["sources_sha256"][:16], "lib_sha256": info["lib_sha256"][:16],
                "lib_unchanged_since_build": info["lib_sha256"] == info["lib_sha256_now"], "nvcc": info["nvcc"],
                "arch_flags": info["arch_flags"], "built_utc": info["built_utc"], "translation_units": len(info["sources"])}
    except Exception as e:
        return {"error": repr(e)}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d, "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


def cpu_step(weights, n_images, seed, batch=1):
    """One bounded CPU sample: oracle forward + decode + filter over n_images, `batch` images per oracle call. Both
    batch 1 and batch 4 are measured and reported (batch 4 is ~2x faster per image on an 8-core host, slower on the
    16-core GPU box); larger batches gain nothing."""
    from oracle import decode as od, filter as of, pfpn
    for i in range(0, n_images, batch):
        img = np.concatenate([make_images(seed + i + j, 1) for j in range(min(batch, n_images - i))], axis=0)
        b3, cls, _ = pfpn.forward_inference(img, weights, C)
        of.filter_detections_batch(od.hull_boxes(b3), b3, cls)


def cpu_reference(weights, n_images, threads, seed=100, batch=1):
    """Oracle forward + decode + filter on the host cores. Returns (images/s, seconds)."""
    import torch
    torch.set_num_threads(threads)
    cpu_step(weights, batch, seed, batch)  # warm-up (oneDNN primitive creation)
    t0 = time.perf_counter()
    cpu_step(weights, n_images, seed + 1, batch)
    dt = time.perf_counter() - t0
    return n_images / dt, dt


def run_reference(args):
    """The reference arm: the CPU restatement of the path on all host cores; each step is a bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    from pyrapose_b200.models import spec
    weights = bench_weights(args)
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    # both batchings once, then the timed steps in the faster one
    b1, _ = cpu_reference(weights, 4, threads, seed=900, batch=1)
    b4, _ = cpu_reference(weights, 4, threads, seed=950, batch=4)
    sample = 4 if b4 > b1 else 2
    cb = 4 if b4 > b1 else 1
    for i in range(max(args.warmup, 1)):
        cpu_step(weights, sample, 1000 + 10 * i, cb)
    t0 = time.perf_counter()
    for i in range(args.steps):
        cpu_step(weights, sample, 2000 + 10 * i, cb)
    dt = time.perf_counter() - t0
    ips = args.steps * sample / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": ips, "unit": "images/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1000.0 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        # the same workload as our arm's line (each step here is a bounded sample of it: see cpu_baseline.sample)
        "config": {"workload": "PFPN ResNet-50 inference + 3D-box-corner decode + filter_detections (thr 0.05, NMS 0.5, "
                               "max 300), batch %d per GPU, %dx%d, %d classes" % (args.batch, H, W, C),
                   "global_batch": args.gpus * args.batch,
                   "parallelism": "replicas x%d (no collective)" % args.gpus,
                   "weights": "random-init (seed 1), he-normal heads, cls bias %.1f" % args.cls_bias,
                   "note": "reference TF/Keras path unavailable on this box; this is the cited CPU restatement (oracle/) "
                           "on rank 0's host cores"},
        "cpu_baseline": {"value": ips, "unit": "images/s", "cores": threads, "kind": "port",
                         "sample": "%d images per step, %d per oracle call (the faster of batch 1: %.2f and batch 4: %.2f "
                                   "images/s on this host; a bounded sample of the batch-32 step), forward+decode+filter, "
                                   "torch-CPU fp32" % (sample, cb, b1, b4),
                         "batch1_images_per_s": b1, "batch4_images_per_s": b4},
        "e2e": {"value": ips, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    _emit(line)


def _max_over_ranks(value, dev, world):
    import torch
    import torch.distributed as dist
    t = torch.tensor([value], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def train_block(args, world, rank, dev, barrier):
    """BASELINE.json config 4, the one place on the path with a collective: YCB-Video 21 classes, batch 16 per GPU at
    480x640, forward + focal / orthogonal_l1 losses + backward + bucketed NCCL gradient all-reduce + Adam(clipnorm)
    through model.train_on_batch. Images and targets are resident (targets built once by anchor_targets_bbox from
    seeded synthetic pose annotations). Returns the `train` object of the JSON line."""
    import torch
    import torch.distributed as dist
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from synth import pose_annotations
    from pyrapose_b200 import models, optimizers
    from pyrapose_b200.utils.anchors import anchor_targets_bbox, anchors_for_shape
    TH, TW, TC, TB = 480, 640, 21, args.train_batch
    anchors = anchors_for_shape((TH, TW, 3))
    mean = np.array([103.939, 116.779, 123.68], dtype=np.float32)

    def make(batch, seed0):
        rng = np.random.default_rng(seed0)
        img = rng.integers(0, 256, (batch, TH, TW, 3)).astype(np.float32) - mean
        anns = [pose_annotations(seed0 * 1000 + i, TC) for i in range(batch)]
        y = anchor_targets_bbox(anchors, [np.zeros((TH, TW, 3), np.float32)] * batch, anns, TC)
        return torch.from_numpy(img).to(dev), [torch.from_numpy(t).to(dev) for t in y]

    x_d, y_d = make(TB, 4 + rank)
    model = models.backbone("resnet50").retinanet(num_classes=TC, seed=1)   # same initial weights on every rank
    model.compile(optimizer=optimizers.Adam(lr=1e-5, clipnorm=0.001))
    steps, warm = args.train_steps, 3
    for _ in range(warm):
        model.train_on_batch(x_d, y_d, sync_losses=False)
    eng = model._train_engine
    opt = model.optimizer
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        model.train_on_batch(x_d, y_d, sync_losses=False)
    e1.record()
    torch.cuda.synchronize()
    step_ms = _max_over_ranks(e0.elapsed_time(e1) / steps, dev, world)
    losses = eng.train_losses()
    barrier()
    # the same step without the exchange (segments back to back, no NCCL): the difference is the EXPOSED all-reduce
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3 * steps + 1)]
    n_seg = len(eng.train_segments())
    ev[0].record()
    for s_ in range(steps):
        eng.buffers["images"].copy_(x_d, non_blocking=True)
        for k_, t_ in zip(("y_reg", "y_cls", "y_mask"), y_d):
            eng.buffers[k_].copy_(t_, non_blocking=True)
        ev[3 * s_ + 1].record()
        eng.train_count()
        for i in range(n_seg):
            eng.train_run_segment(i)
        ev[3 * s_ + 2].record()
        eng.train_apply(opt.lr, opt.beta_1, opt.beta_2, opt.epsilon, opt.clipnorm, 1.0)
        ev[3 * s_ + 3].record()
    torch.cuda.synchronize()
    local_ms = _max_over_ranks(ev[0].elapsed_time(ev[-1]) / steps, dev, world)
    parts = [sum(ev[3 * s_ + k_].elapsed_time(ev[3 * s_ + k_ + 1]) for s_ in range(steps)) / steps for k_ in range(3)]
    n_params = int(eng.grads_tensor().numel())
    # tcgen05 kernel classes of one step, per-launch CUDA events (forward conv / dgrad / wgrad)
    groups = {}
    if rank == 0:
        for name, fl, ms in eng.train_profile():
            kind = name.split(":")[0] if ":" in name else ("fwd_conv" if fl > 0 else "fwd_other")
            g = groups.setdefault(kind, [0.0, 0.0, 0])
            g[0] += ms
            g[1] += fl
            g[2] += 1
    out = {
        "metric": "images_per_sec_pfpn_480x640_train_step", "value": TB * world / (step_ms * 1e-3), "unit": "images/s",
        "ms_per_step": step_ms, "steps": steps, "warmup": warm,
        "config": {"workload": "PFPN ResNet-50 training step (forward + focal/orthogonal_l1 + backward + gradient "
                               "exchange over NVLink + Adam clipnorm), YCB-Video %d classes, batch %d per GPU, %dx%d"
                               % (TC, TB, TH, TW), "global_batch": TB * world, "parallelism": "dp%d" % world,
                   "dp_normalizer": model.dp_normalizer},
        "split_ms": {"copy_inputs": parts[0], "forward_backward": parts[1], "optimizer": parts[2],
                     "allreduce_exposed": max(step_ms - local_ms, 0.0) if world > 1 else 0.0,
                     "step_without_exchange": local_ms},
        "allreduce_bytes": 4 * n_params if world > 1 else 0,
        "allreduce": (("%d buckets of ~%d MB; " % (sum(1 for _, c, _ in eng.train_segments() if c), model.bucket_bytes >> 20)) +
                      ("peer-to-peer exchange: every finished bucket is pushed into the peers' staging memory by copy "
                       "engines over NVLink while the backward pass runs (no SM taken), one kernel sums the replicas' "
                       "slices afterwards (utils/exchange.py, ppb_p2p_*)" if eng.__dict__.get("_p2p") else
                       "each all-reduced (NCCL, async) as soon as its last weight gradient has landed, overlapping the "
                       "rest of the backward pass")) if world > 1 else None,
        "exchange": ("p2p" if eng.__dict__.get("_p2p") else "nccl") if world > 1 else None,
        "kernels": {k: {"ms": v[0], "launches": v[2], "tflops": (v[1] / v[0] * 1e-9) if v[0] > 0 and v[1] > 0 else None}
                    for k, v in groups.items()},
        "losses": {k: losses[k] for k in ("3Dbox", "cls", "mask", "l2", "grad_norm")},
        "trainable_params": n_params,
    }
    # end to end through model.fit_generator (the reference's loop, bin/train.py:381-390): every step a Sequence item =
    # raw uint8 images in pinned host memory + the annotation lists, targets built on the device
    # (anchor_targets_bbox_device), the next batch staged by a background thread while the GPU runs the current step,
    # every batch's losses copied back and handed to a callback. Host wall clock, max over ranks.
    import time
    from pyrapose_b200.utils.anchors import anchor_targets_bbox_device, anchors_for_shape_device
    raw = torch.from_numpy(np.random.default_rng(40 + rank).integers(0, 256, (TB, TH, TW, 3)).astype(np.uint8)).pin_memory()
    anns = [pose_annotations((4 + rank) * 1000 + i, TC) for i in range(TB)]
    shapes = [np.zeros((TH, TW, 3), np.uint8)] * TB
    anchors_dev = anchors_for_shape_device((TH, TW, 3))

    e2e_steps = max(steps, 20)   # one staged batch of pipeline fill is part of the timed call: amortised over >= 20

    class Batches(object):
        def __len__(self):
            return e2e_steps

        def __getitem__(self, i):
            return raw, list(anchor_targets_bbox_device(anchors_dev, shapes, anns, TC))

    seen = []

    class Logs(object):
        def on_batch_end(self, batch, logs):
            seen.append(logs["loss"])

    model.fit_generator(Batches(), steps_per_epoch=3, epochs=1, verbose=0)   # warm-up (rebinds for uint8 input)
    torch.cuda.synchronize()
    barrier()
    t0 = time.perf_counter()
    model.fit_generator(Batches(), steps_per_epoch=e2e_steps, epochs=1, verbose=0, callbacks=[Logs()], shuffle=False)
    torch.cuda.synchronize()
    e2e_ms = _max_over_ranks(1e3 * (time.perf_counter() - t0) / e2e_steps, dev, world)
    assert len(seen) == e2e_steps and np.isfinite(seen).all()
    out["e2e"] = {"value": TB * world / (e2e_ms * 1e-3), "unit": "images/s", "ms_per_step": e2e_ms, "steps": e2e_steps,
                  "h2d_bytes_per_step": int(raw.numel()) + sum(a['bboxes'].shape[0] for a in anns) * 8 * (4 + 16)
                  + TB * TH * TW,
                  "d2h_bytes_per_step": 5 * 8,
                  "api": "model.fit_generator(Sequence -> (uint8 host images, anchor_targets_bbox_device(annotation "
                         "lists))), per-batch losses through on_batch_end"}
    del model, eng
    torch.cuda.empty_cache()
    if world > 1:
        out["dp_check"] = dp_check(world, rank, dev, make)
    return out


def dp_check(world, rank, dev, make):
    """An N-replica step on shards must equal the 1-replica step on the concatenated batch (what
    keras.utils.multi_gpu_model computes, PyraPose/bin/train.py:82-86): every rank steps on its own 2 images over
    NCCL, rank 0 repeats the step alone on all 2N images; compared: global gradient norm and every updated tensor."""
    import torch
    import torch.distributed as dist
    from pyrapose_b200 import models, optimizers
    TC, b = 21, 2
    x_d, y_d = make(b, 900 + rank)
    m = models.backbone("resnet50").retinanet(num_classes=TC, seed=1)
    m.compile(optimizer=optimizers.Adam(lr=1e-5, clipnorm=0.001))
    m.train_on_batch(x_d, y_d)
    g_dp = m.last_step["grad_norm"]
    w_dp = {k: v.copy() for k, v in m.get_weights_dict().items()}
    gathered = []
    for t in [x_d] + y_d:
        parts = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(parts, t.contiguous())
        gathered.append(torch.cat(parts, dim=0))
    res = None
    if rank == 0:
        one = models.backbone("resnet50").retinanet(num_classes=TC, seed=1)
        w0 = {k: v.copy() for k, v in one.weights.items()}
        one.compile(optimizer=optimizers.Adam(lr=1e-5, clipnorm=0.001), data_parallel=False)
        one.train_on_batch(gathered[0], gathered[1:])
        g_one = one.last_step["grad_norm"]
        w_one = one.get_weights_dict()
        # (1) the exchanged gradient: after train_apply the flat buffer holds the prepared gradient of the step
        e_dp, e_one = m._train_engine, one._train_engine
        worst_g = 0.0
        n_g = e_one.grads_tensor().numel()
        ga = e_one.train_read("grads", 0, n_g).astype(np.float64)
        whole_g = float(np.linalg.norm(e_dp.train_read("grads", 0, n_g).astype(np.float64) - ga) / np.linalg.norm(ga))
        for lyr in one.layers:
            for bias in (False, True):
                a = e_one.train_layer_tensor("grads", lyr.name, lyr.ksize, lyr.cin, lyr.cout, bias=bias)
                if a is None or np.linalg.norm(a) < 1e-4 * g_one:
                    continue
                g = e_dp.train_layer_tensor("grads", lyr.name, lyr.ksize, lyr.cin, lyr.cout, bias=bias)
                worst_g = max(worst_g, float(np.linalg.norm(g.astype(np.float64) - a) / np.linalg.norm(a)))
        # (2) the updated weights. At lr = 1e-5 / clipnorm = 1e-3 a step is a few hundred fp32 ulps of the weight
        # down to less than one, so two correct results differ by rounding: allow 2 * sqrt(sum ulp^2 / 12)
        worst, worst_raw = 0.0, 0.0
        for k, v0 in w0.items():
            d = w_one[k].astype(np.float64) - v0
            n = np.linalg.norm(d)
            if n < 1e-12:
                continue
            ulp = np.spacing(np.abs(v0).astype(np.float32)).astype(np.float64)
            floor = 2.0 * float(np.sqrt((ulp ** 2).sum() / 12.0))
            diff = float(np.linalg.norm(w_dp[k].astype(np.float64) - w_one[k]))
            worst = max(worst, max(diff - floor, 0.0) / n)
            worst_raw = max(worst_raw, diff / n)
        rel_norm = abs(g_dp - g_one) / g_one
        res = {"replicas": world, "images_per_replica": b, "grad_norm_dp": g_dp, "grad_norm_single": g_one,
               "grad_norm_rel": rel_norm, "max_rel": max(whole_g, worst), "whole_gradient_rel": whole_g,
               "max_rel_gradient_tensor": worst_g, "max_rel_step_beyond_fp32_rounding": worst,
               "max_rel_step_raw": worst_raw, "tolerance": 1e-3, "tolerance_single_tensor": 1e-2,
               "tolerance_single_tensor_step": 3e-3,
               "ok": bool(whole_g <= 1e-3 and worst <= 3e-3 and rel_norm <= 1e-3 and worst_g <= 1e-2),
               "what": "N replicas on shards over NCCL (global normalisers, summed gradients) vs rank 0 alone on the "
                       "concatenated batch, one Adam(lr=1e-5, clipnorm=1e-3) step: relative difference of the global "
                       "gradient norm and of the whole exchanged gradient (<= 1e-3); every tensor's weight "
                       "step <= 3e-3, single gradient tensors <= 1e-2 (the same bf16 gradient maps summed over another split-K "
                       "partition; the tensor core's fp32 accumulation is not associative, "
                       "tests/test_train_config4_gpu.py)"}
        del one
    del m
    torch.cuda.empty_cache()
    dist.barrier()
    return res


def config5_block(args, world, rank, dev, barrier):
    """BASELINE.json config 5: 960x1280, T-LESS 30 classes, GLOBAL batch 64 sharded over the N GPUs (strong scaling:
    64 / N images per GPU), inference + decode + FilterDetections, images resident."""
    import torch
    from pyrapose_b200 import _kernels, models
    from pyrapose_b200.models import spec
    h5, w5, c5, gb = 960, 1280, 30, 64
    per = gb // world
    w = spec.init_weights(c5, A, seed=1, head_init="spread")
    w["pyramid_classification/bias"] = np.full_like(w["pyramid_classification/bias"], args.cls_bias)
    train = models.backbone("resnet50").retinanet(num_classes=c5)
    train.set_weights_dict(w)
    model = models.convert_model(train, alias_arena=True)   # production arena (shared backbone buffers)
    rng = np.random.default_rng(500 + rank)
    img = torch.from_numpy(rng.integers(0, 256, (per, h5, w5, 3), dtype=np.uint8)).pin_memory()
    buf = model.forward_device(img, want_hull=True)
    eng = model._engines[(h5, w5, 'alias')]

    def step():
        eng.run()
        return _kernels.filter_detections(buf["hull"], buf["boxes3D"], buf["cls"])

    for _ in range(3):
        step()
    barrier()
    steps = 5
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = _max_over_ranks(e0.elapsed_time(e1) / steps, dev, world)
    flops = conv_flops_per_image(h5, w5, c5, A)
    out = {"metric": "images_per_sec_pfpn_960x1280_infer_decode_filter", "value": gb / (ms * 1e-3), "unit": "images/s",
           "ms_per_step": ms, "steps": steps, "scaling": "strong", "global_batch": gb, "batch_per_gpu": per,
           "classes": c5, "conv_tflops_per_gpu": flops * per / (ms * 1e-3) / 1e12,
           "device_bytes_per_gpu": eng.device_bytes()}
    del model, train, eng, buf
    torch.cuda.empty_cache()
    return out


def run_ours(args):
    import torch
    import torch.distributed as dist
    from pyrapose_b200 import _kernels, _native, models
    from pyrapose_b200.models import spec

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    B = args.batch
    weights = bench_weights(args)
    train = models.backbone("resnet50").retinanet(num_classes=C)
    train.set_weights_dict(weights)
    model = models.convert_model(train, alias_arena=True)   # production arena: backbone intermediates share buffers

    # ---- device-resident arm: images already in HBM
    host_img = make_images(10 + rank, B)
    pinned = torch.from_numpy(host_img).pin_memory()
    buf = model.forward_device(pinned, want_hull=True)        # builds the engine, binds, first run
    eng = model._engines[(H, W, 'alias')]
    torch.cuda.synchronize()

    def step_device():
        eng.run()
        return _kernels.filter_detections(buf["hull"], buf["boxes3D"], buf["cls"])

    for _ in range(max(args.warmup, 3)):
        step_device()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = _native.launch_count()
    e0 = torch.cuda.Event(enable_timing=True)
    e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step_device()
    e1.record()
    torch.cuda.synchronize()
    dev_ms = e0.elapsed_time(e1)
    launches = _native.launch_count() - launches0
    dev_ms = _max_over_ranks(dev_ms, dev, world)
    barrier()

    # ---- sustained: the same step for >= args.sustain_s seconds (the K-step region above is a burst of ~0.2 s; under
    # back-to-back replay the board sits on its power cap), with its own clock sample
    sustained = None
    if args.sustain_s > 0:
        n_sus = max(int(args.sustain_s * 1e3 / (dev_ms / args.steps)) + 1, args.steps)
        barrier()
        e0.record()
        for _ in range(n_sus):
            step_device()
        e1.record()
        torch.cuda.synchronize()
        sus_ms = _max_over_ranks(e0.elapsed_time(e1), dev, world)
        barrier()
        sustained = {"value": world * B * n_sus / (sus_ms * 1e-3), "unit": "images/s", "steps": n_sus,
                     "seconds": sus_ms * 1e-3, "ms_per_step": sus_ms / n_sus}
    # one nvidia-smi sampler (100 ms period) over the K timed steps AND the sustained loop of the same step
    clocks = sampler.stop() if rank == 0 else None
    if clocks is not None:
        clocks["window"] = "timed steps + sustained loop (same step, %.1f s)" % ((sustained or {}).get("seconds", 0.0))

    # ---- e2e arms: HOST images in, HOST results out, every step, through the public API. The real caller's input is
    # raw uint8 BGR (cv2.imread -> preprocess_image, PyraPose/utils/linemod_eval.py:283-303): preprocess_image('caffe')
    # runs fused in the stem loader, so the headline e2e ships uint8 (29.5 MB per batch-32 step); the float32
    # mean-subtracted form the reference hands to predict_on_batch (118 MB) is timed beside it.
    #   stream : model.detect_stream — pipelined batch loop (H2D of step i+1 overlaps the forward pass of step i)
    #   sync   : model.detect_on_batch — one blocking call per step
    #   predict_on_batch : the reference's own call; returns the three FULL arrays (234 MB D2H per step)
    rng8 = np.random.default_rng(70 + rank)
    raw = [torch.from_numpy(rng8.integers(0, 256, (B, H, W, 3), dtype=np.uint8)).pin_memory() for _ in range(2)]
    f32 = [pinned, torch.from_numpy(make_images(50 + rank, B)).pin_memory()]

    def time_stream(batches):
        def gen(n):
            for i in range(n):
                yield batches[i & 1]
        det = None
        for det in model.detect_stream(gen(3)):
            pass
        barrier()
        t0 = time.perf_counter()
        for det in model.detect_stream(gen(args.steps)):
            pass
        torch.cuda.synchronize()
        return _max_over_ranks(time.perf_counter() - t0, dev, world), det

    def time_calls(fn, x, steps):
        out = None
        for _ in range(2):
            out = fn(x)
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            out = fn(x)
        torch.cuda.synchronize()
        return _max_over_ranks(time.perf_counter() - t0, dev, world), out

    e2e_s, det = time_stream(raw)
    e2e_f32_s, _ = time_stream(f32)
    sync_s, _ = time_calls(model.detect_on_batch, raw[0], args.steps)
    sync_f32_s, _ = time_calls(model.detect_on_batch, f32[0], args.steps)
    pob_steps = max(args.steps // 2, 3)
    pob_s, pob_out = time_calls(model.predict_on_batch, raw[0], pob_steps)
    pob_f32_s, _ = time_calls(model.predict_on_batch, f32[0], pob_steps)
    h2d = int(raw[0].numel())
    d2h = int(sum(o.nbytes for o in det))
    pob_d2h = int(sum(o.nbytes for o in pob_out))
    del pob_out
    ips = lambda sec, n=args.steps: world * B * n / sec
    e2e = {"value": ips(e2e_s), "unit": "images/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
           "api": "model.detect_stream(uint8 host batches) -> host detections [boxes, boxes3D, scores, labels]",
           "input": "raw uint8 BGR (preprocess_image('caffe') fused into the stem loader)",
           "e2e_sync": ips(sync_s), "api_sync": "model.detect_on_batch(uint8 host batch)",
           "fp32_input": {"value": ips(e2e_f32_s), "e2e_sync": ips(sync_f32_s), "h2d_bytes_per_step": int(pinned.numel() * 4)},
           "predict_on_batch": {"value": ips(pob_s, pob_steps), "fp32_input": ips(pob_f32_s, pob_steps),
                                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": pob_d2h,
                                "api": "model.predict_on_batch(host batch) -> [boxes3D, cls, mask] full float32 arrays "
                                       "(the reference's call, linemod_eval.py:303), pinned buffers + async D2H"}}

    train_out = train_block(args, world, rank, dev, barrier) if args.train_steps > 0 else None
    cfg5_out = config5_block(args, world, rank, dev, barrier) if (world > 1 and 64 % world == 0 and args.config5) else None

    if rank == 0:
        # ---- roofline of the dominant kernel (tcgen05 implicit-GEMM conv), timed per launch with CUDA events
        buf = model.forward_device(pinned, want_hull=True)   # back to the plan of the timed region
        eng = model._engines[(H, W, 'alias')]
        prof = [eng.profile() for _ in range(3)][-1]
        conv_ms = sum(ms for k, ms in prof if k == 0)
        other_ms = sum(ms for k, ms in prof if k != 0)
        n_conv = sum(1 for k, _ in prof if k == 0)
        flops_img = conv_flops_per_image(H, W, C, A)
        peaks, peak_src = load_peaks()
        achieved = flops_img * B / (conv_ms * 1e-3) / 1e12
        # the K-step region is a ~0.2 s burst and the per-launch profile times every kernel alone: the BURST cuBLAS
        # figure is the denominator; the sustained one is reported beside it
        peak = float(peaks["bf16_tflops"])
        peak_sus = float(peaks.get("bf16_tflops_sustained", peaks["bf16_tflops"]))
        # the single heaviest launch (a 3x3 512->512 regression-tower convolution over the three pyramid levels)
        ops = eng.op_table()
        top_i = max(range(len(prof)), key=lambda i: prof[i][1] if prof[i][0] == 0 else -1.0)
        top = {"name": ops[top_i][0], "ms": prof[top_i][1], "tflops": ops[top_i][1] / (prof[top_i][1] * 1e-3) / 1e12}
        top["frac"] = top["tflops"] / peak
        top["frac_of_sustained"] = top["tflops"] / peak_sus
        groups = {}
        for (name, fl), (k, ms) in zip(ops, prof):
            if k != 0:
                grp = "elementwise"
            elif name == "conv1":
                grp = "stem"
            elif name.startswith("res"):
                grp = "stage" + name[3]
            elif name.startswith("pyramid_regression"):
                grp = "reg_head"
            elif name.startswith("pyramid_class"):
                grp = "cls_head"
            elif name.startswith("mask"):
                grp = "mask_head"
            else:
                grp = "pfpn"
            g = groups.setdefault(grp, [0.0, 0.0])
            g[0] += ms
            g[1] += fl
        by_group = {k: {"ms": v[0], "tflops": v[1] / (v[0] * 1e-3) / 1e12 if v[1] > 0 else None}
                    for k, v in groups.items()}
        # ---- secondary: HBM-bound score pass of the filter (reads B*N*C fp32 once)
        n_boxes = eng.num_boxes
        pass_rate = float((buf["cls"] > 0.05).float().mean())
        flush = torch.empty(1 << 30, dtype=torch.uint8, device=dev)  # L2 flush; also hides host launch latency
        # four distinct score tensors (435 MB > the 126 MB L2), launched back to back inside ONE event pair: the
        # per-launch time is then the kernel's own (a lone launch between two events carries ~6 us of launch floor
        # on this box, tools/probe_hbm.py); the lone-launch figure is reported beside it
        score_set = [buf["cls"]] + [buf["cls"].clone() for _ in range(3)]
        for sc in score_set:
            _kernels.score_mask(sc, 0.05, want_counts=False)
        ts, ts1 = [], []
        for _ in range(10):
            flush.fill_(1)
            s0 = torch.cuda.Event(enable_timing=True)
            s1 = torch.cuda.Event(enable_timing=True)
            s0.record()
            for sc in score_set:
                _kernels.score_mask(sc, 0.05, want_counts=False)
            s1.record()
            torch.cuda.synchronize()
            ts.append(s0.elapsed_time(s1) / len(score_set))
            flush.fill_(1)
            s0 = torch.cuda.Event(enable_timing=True)
            s1 = torch.cuda.Event(enable_timing=True)
            s0.record()
            _kernels.score_mask(buf["cls"], 0.05, want_counts=False)
            s1.record()
            torch.cuda.synchronize()
            ts1.append(s0.elapsed_time(s1))
        del flush, score_set
        mask_ms = statistics.median(ts)
        mask_ms_single = statistics.median(ts1)
        mask_gbs = B * n_boxes * C * 4 / (mask_ms * 1e-3) / 1e9
        # ---- CPU baseline on this box's host cores (bounded sample)
        threads = os.cpu_count() or 1
        if world == 1:
            cpu_b1, cpu_s1 = cpu_reference(weights, args.cpu_images, threads, batch=1)
            cpu_b4, cpu_s4 = cpu_reference(weights, args.cpu_images, threads, batch=4)
            cpu_ips, cpu_s = max(cpu_b1, cpu_b4), cpu_s1 + cpu_s4
        else:  # the CPU baseline is an N=1 leg: at N>1 the other ranks' barrier spinning would share its cores
            cpu_ips, cpu_s, cpu_b1, cpu_b4 = None, 0.0, None, None
        value = world * B * args.steps / (dev_ms * 1e-3)
        line = {
            "metric": METRIC, "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": "PFPN ResNet-50 inference + 3D-box-corner decode + filter_detections (thr 0.05, NMS 0.5, "
                                   "max 300), batch %d per GPU, %dx%d, %d classes" % (B, H, W, C),
                       "global_batch": world * B, "parallelism": "replicas x%d (no collective)" % world,
                       "weights": "random-init (seed 1), he-normal heads, cls bias %.1f: %.1f%% of scores pass 0.05"
                                  % (args.cls_bias, 100 * pass_rate),
                       "candidates_per_image": pass_rate * n_boxes * C,
                       "l2": "working set per step (activations ~%.1f GB) far exceeds the 126 MB L2; no flush needed"
                             % (eng.device_bytes() / 1e9)},
            "e2e": e2e,
            "build": _build_info(),
            "sustained": sustained,
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                         "frac": achieved / peak, "kernel": "conv_igemm_kernel",
                         # dram__bytes_read.sum + dram__bytes_write.sum of the heaviest launch from the committed
                         # ncu --set full capture (profiles/r01_full_heads_pair.txt, launch 0; B=32): 227.9 + 184.0 MB against
                         # 2 x 222k rows x 512 ch x 2 B = 455 MB of algorithmic input + output (the tail of the output
                         # is still in L2 when the kernel ends): no wasted re-reads
                         "traffic": 411518976 if (B, H, W, C) == (32, 480, 640, 15) else None,
                         "traffic_launch": "pyramid_regression3D_1 (CTA-pair + slab mode), dram read 228.2 MB + write 183.3 MB, profiles/r02_full_heads.txt launch 3", "top_launch": top,
                         "launches_per_step": n_conv, "conv_ms_per_step": conv_ms, "other_ms_per_step": other_ms,
                         "flops_per_image": flops_img, "peak_source": peak_src + " bf16_tflops (burst)",
                         "frac_of_burst": achieved / peak, "frac_of_sustained": achieved / peak_sus,
                         "by_group": by_group},
            "roofline_hbm": {"kernel": "score_mask_kernel", "bound": "hbm", "achieved": mask_gbs,
                             "peak": float(peaks["hbm_gbs"]), "unit": "GB/s", "frac": mask_gbs / float(peaks["hbm_gbs"]),
                             "bytes_per_launch": B * n_boxes * C * 4, "ms": mask_ms,
                             "timing": "4 launches back to back over 4 distinct score tensors (435 MB > L2), L2 flushed "
                                       "before, one CUDA-event pair, / 4",
                             "ms_single_launch": mask_ms_single,
                             "frac_single_launch": B * n_boxes * C * 4 / (mask_ms_single * 1e-3) / 1e9
                                                   / float(peaks["hbm_gbs"])},
            "cpu_baseline": {"value": cpu_ips, "unit": "images/s", "cores": threads, "kind": "port",
                             "sample": "%d images forward+decode+filter at batch 1 and again at batch 4 (value = the faster; "
                                       "a bounded sample of the batch-32 step), torch-CPU fp32 oracle, %.1f s"
                                       % (args.cpu_images, cpu_s),
                             "batch1_images_per_s": cpu_b1, "batch4_images_per_s": cpu_b4} if world == 1 else None,
            "train": train_out,
            "config5": cfg5_out,
        }
        _emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def _emit(line):
    """The ONE JSON line goes to the process's original stdout; everything else written to fd 1 by this process or by
    the libraries it loads (NCCL prints its version banner to stdout at communicator creation) lands on stderr."""
    os.write(_JSON_FD, (json.dumps(line) + "\n").encode())


_JSON_FD = 1


def main():
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=BATCH)
    ap.add_argument("--cpu-images", type=int, default=8)
    ap.add_argument("--sustain-s", type=float, default=5.0, help="seconds of back-to-back steps for the `sustained` figure (0: skip)")
    ap.add_argument("--train-steps", type=int, default=10, help="timed steps of the `train` block (config 4); 0: skip")
    ap.add_argument("--train-batch", type=int, default=16)
    ap.add_argument("--no-config5", dest="config5", action="store_false",
                    help="skip the `config5` block (960x1280, 30 classes, global batch 64 over N >= 2 GPUs)")
    ap.add_argument("--height", type=int, default=H, help="other BASELINE.json configs, e.g. config 5: 960 x 1280, 30 classes")
    ap.add_argument("--width", type=int, default=W)
    ap.add_argument("--classes", type=int, default=C)
    ap.add_argument("--cls-bias", type=float, default=-6.0,
                    help="classification output bias: -6.0 = ~2%% of scores > 0.05 (primary), -4.6 = ~13%% (dense stress)")
    args = ap.parse_args()
    globals().update(H=args.height, W=args.width, C=args.classes,
                     METRIC="images_per_sec_pfpn_%dx%d_infer_decode_filter" % (args.height, args.width))
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
