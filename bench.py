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


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d, "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


def cpu_step(weights, n_images, seed):
    """One bounded CPU sample: n_images x (oracle forward + decode + filter), batch 1 each."""
    from oracle import decode as od, filter as of, pfpn
    for i in range(n_images):
        img = make_images(seed + i, 1)
        b3, cls, _ = pfpn.forward_inference(img, weights, C)
        of.filter_detections_batch(od.hull_boxes(b3), b3, cls)


def cpu_reference(weights, n_images, threads, seed=100):
    """Oracle forward + decode + filter on the host cores. Returns (images/s, seconds)."""
    import torch
    torch.set_num_threads(threads)
    cpu_step(weights, 1, seed)  # warm-up (oneDNN primitive creation)
    t0 = time.perf_counter()
    cpu_step(weights, n_images, seed + 1)
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
    sample = 2
    for i in range(max(args.warmup, 1)):
        cpu_step(weights, sample, 1000 + 10 * i)
    t0 = time.perf_counter()
    for i in range(args.steps):
        cpu_step(weights, sample, 2000 + 10 * i)
    dt = time.perf_counter() - t0
    ips = args.steps * sample / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": ips, "unit": "images/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1000.0 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "PFPN ResNet-50 inference + 3D-box-corner decode + filter_detections (thr 0.05, NMS 0.5, "
                               "max 300), 480x640, 15 classes",
                   "note": "reference TF/Keras path unavailable on this box; this is the cited CPU restatement (oracle/)"},
        "cpu_baseline": {"value": ips, "unit": "images/s", "cores": threads, "kind": "port",
                         "sample": "%d images per step (batch 1 each), forward+decode+filter, torch-CPU fp32" % sample},
        "e2e": {"value": ips, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    _emit(line)


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
    model = models.convert_model(train)

    # ---- device-resident arm: images already in HBM
    host_img = make_images(10 + rank, B)
    pinned = torch.from_numpy(host_img).pin_memory()
    buf = model.forward_device(pinned, want_hull=True)        # builds the engine, binds, first run
    eng = model._engines[(H, W)]
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
    t = torch.tensor([dev_ms], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms = float(t.item())
    barrier()

    # ---- e2e arm: pinned host images in, host detections out, every step, through the public API.
    # model.detect_stream is the pipelined batch loop (copy of step i+1 overlaps the forward pass of step i); every
    # step still pays its own H2D of the images and D2H of the detections. The synchronous one-call-per-step
    # figure (model.detect_on_batch) is reported beside it as e2e_sync.
    pinned2 = torch.from_numpy(make_images(50 + rank, B)).pin_memory()

    def host_batches(n):
        for i in range(n):
            yield pinned if (i & 1) == 0 else pinned2

    for det in model.detect_stream(host_batches(3)):
        pass
    barrier()
    t0 = time.perf_counter()
    for det in model.detect_stream(host_batches(args.steps)):
        pass
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_s = float(t.item())
    for _ in range(2):
        model.detect_on_batch(pinned)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        model.detect_on_batch(pinned)
    torch.cuda.synchronize()
    sync_s = time.perf_counter() - t0
    t = torch.tensor([sync_s], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    sync_s = float(t.item())
    clocks = sampler.stop() if rank == 0 else None
    h2d = int(pinned.numel() * 4)
    d2h = int(sum(o.nbytes for o in det))

    if rank == 0:
        # ---- roofline of the dominant kernel (tcgen05 implicit-GEMM conv), timed per launch with CUDA events
        prof = [eng.profile() for _ in range(3)][-1]
        conv_ms = sum(ms for k, ms in prof if k == 0)
        other_ms = sum(ms for k, ms in prof if k != 0)
        n_conv = sum(1 for k, _ in prof if k == 0)
        flops_img = conv_flops_per_image(H, W, C, A)
        peaks, peak_src = load_peaks()
        achieved = flops_img * B / (conv_ms * 1e-3) / 1e12
        peak = float(peaks.get("bf16_tflops_sustained", peaks["bf16_tflops"]))
        # the single heaviest launch (a 3x3 512->512 regression-tower convolution over the three pyramid levels)
        ops = eng.op_table()
        top_i = max(range(len(prof)), key=lambda i: prof[i][1] if prof[i][0] == 0 else -1.0)
        top = {"name": ops[top_i][0], "ms": prof[top_i][1], "tflops": ops[top_i][1] / (prof[top_i][1] * 1e-3) / 1e12}
        top["frac"] = top["tflops"] / peak
        top["frac_of_burst"] = top["tflops"] / float(peaks["bf16_tflops"])
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
            cpu_ips, cpu_s = cpu_reference(weights, args.cpu_images, threads)
        else:  # the CPU baseline is an N=1 leg: at N>1 the other ranks' barrier spinning would share its cores
            cpu_ips, cpu_s = None, 0.0
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
                       "l2": "working set per step (activations ~%.1f GB) far exceeds the 126 MB L2; no flush needed"
                             % (eng.device_bytes() / 1e9)},
            "e2e": {"value": world * B * args.steps / e2e_s, "unit": "images/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "api": "model.detect_stream(host batches) -> host detections",
                    "e2e_sync": world * B * args.steps / sync_s, "api_sync": "model.detect_on_batch(host batch)"},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                         "frac": achieved / peak, "kernel": "conv_igemm_kernel",
                         # dram__bytes_read.sum + dram__bytes_write.sum of the heaviest launch from the committed
                         # ncu --set full capture (profiles/r01_full_heads_pair.txt, launch 0; B=32): 227.9 + 184.0 MB against
                         # 2 x 222k rows x 512 ch x 2 B = 455 MB of algorithmic input + output (the tail of the output
                         # is still in L2 when the kernel ends): no wasted re-reads
                         "traffic": 411939840 if (B, H, W, C) == (32, 480, 640, 15) else None,
                         "traffic_launch": "pyramid_regression3D_1 (CTA-pair + slab mode), profiles/r01_full_heads_pair.txt launch 0", "top_launch": top,
                         "launches_per_step": n_conv, "conv_ms_per_step": conv_ms, "other_ms_per_step": other_ms,
                         "flops_per_image": flops_img, "peak_source": peak_src + " bf16_tflops_sustained",
                         "frac_of_burst": achieved / float(peaks["bf16_tflops"])},
            "roofline_hbm": {"kernel": "score_mask_kernel", "bound": "hbm", "achieved": mask_gbs,
                             "peak": float(peaks["hbm_gbs"]), "unit": "GB/s", "frac": mask_gbs / float(peaks["hbm_gbs"]),
                             "bytes_per_launch": B * n_boxes * C * 4, "ms": mask_ms,
                             "timing": "4 launches back to back over 4 distinct score tensors (435 MB > L2), L2 flushed "
                                       "before, one CUDA-event pair, / 4",
                             "ms_single_launch": mask_ms_single,
                             "frac_single_launch": B * n_boxes * C * 4 / (mask_ms_single * 1e-3) / 1e9
                                                   / float(peaks["hbm_gbs"])},
            "cpu_baseline": {"value": cpu_ips, "unit": "images/s", "cores": threads, "kind": "port",
                             "sample": "%d images (batch 1 each) forward+decode+filter, torch-CPU fp32 oracle, %.1f s"
                                       % (args.cpu_images, cpu_s)} if world == 1 else None,
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
