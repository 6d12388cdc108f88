"""Cycle-stamped trace of ONE CTA of selected convolution launches (VERDICT r01 task 2b: "commit a globaltimer trace
of one CTA"): where a tile's time goes between the TMA producer, the MMA issuer and the epilogue groups.

  python tools/trace_conv.py [B] [layer ...]      default layers: the shallow-K stage 4-5 layers and one of each kind

For each layer: launch shape, then per steady-state tile (median over the CTA's tiles, first and last excluded) in SM
cycles: tile period (MMA issuer tile start -> next tile start), the issuer's wait for the accumulator, its wait for
operand stages (sum over k-iterations), issue time; the producer's wait for free stages; per epilogue group the wait
for a full accumulator, for ring slots, the tcgen05.ld time, compute, and store issue."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pyrapose_b200 import models
from pyrapose_b200.models import spec

args = sys.argv[1:]
B = int(args[0]) if args and args[0].isdigit() else 32
want = [a for a in args if not a.isdigit()] or ["res2b_branch2a", "res2b_branch2b", "res2b_branch2c", "res3b_branch2a",
                                                "res3b_branch2c", "res4b_branch2a", "res4b_branch2b", "res4b_branch2c",
                                                "res5b_branch2a", "res5b_branch2b", "res5b_branch2c", "P3_lateral",
                                                "pyramid_classification_1", "pyramid_regression3D_1"]
H, W, C = 480, 640, 15
m = models.convert_model(models.backbone("resnet50").retinanet(num_classes=C))
m.set_weights_dict(spec.init_weights(C, seed=1, head_init="spread"))
img = torch.randn(B, H, W, 3).mul_(60).pin_memory()
m.forward_device(img, want_hull=True)
eng = m._engines[(H, W)]
for _ in range(2):
    eng.run()
torch.cuda.synchronize()
prof = np.median(np.array([[t for _, t in eng.profile()] for _ in range(3)]), axis=0)
ops = eng.op_table()


def d32(a, b):
    return (a - b) & 0xFFFFFFFF


def per_tile(ev, start_ev, end_ev):
    """sum over a tile of clock(end_ev) - clock(preceding start_ev) for matching (tile, sub)"""
    out = {}
    last = {}
    for e, t, s, c in ev:
        if e == start_ev:
            last[(t, s)] = c
        elif e == end_ev and (t, s) in last:
            out[t] = out.get(t, 0) + d32(c, last[(t, s)])
    return out


def med(d, tiles):
    v = [d[t] for t in tiles if t in d]
    return float(np.median(v)) if v else float("nan")


print("B=%d %dx%d, SM cycles per steady-state tile of CTA 0 (median)" % (B, H, W))
for idx, (name, fl) in enumerate(ops):
    if name not in want:
        continue
    eng.run()  # inputs of the layer are those of a normal forward pass
    info, roles = eng.trace_op(idx, cta=0)
    mma = roles.get("mma")
    if mma is None or len(mma) < 4:
        print("%s: no trace" % name)
        continue
    starts = {int(t): int(c) for e, t, s, c in mma if e == 3}
    n_t = len(starts)
    tiles = sorted(starts)[1:-1] if n_t > 3 else sorted(starts)
    period = {t: d32(starts[t + 1], starts[t]) for t in tiles if t + 1 in starts}
    # MMA issuer: wait for accumulator = start(3) - previous tile commit(6); operand waits = sum(4 - previous event)
    commit = {int(t): int(c) for e, t, s, c in mma if e == 6}
    acc_wait = {t: d32(starts[t], commit[t - 1]) for t in tiles if t - 1 in commit}
    opw, issue = {}, {}
    prev = None
    for e, t, s, c in mma:
        e, t, c = int(e), int(t), int(c)
        if e == 4 and prev is not None:
            opw[t] = opw.get(t, 0) + d32(c, prev)
        if e == 5 and prev is not None:
            issue[t] = issue.get(t, 0) + d32(c, prev)
        prev = c
    total = d32(int(mma[-1][3]), int(mma[0][3]))
    line = ("%-26s %.1f us | tiles/CTA %d, M x N tiles %d x %d, block_n %d, k-iters %d, stages %d, ring %d, pair %d slab %d "
            "bres %d staged %d groups %d\n    issuer: period %.0f = acc wait %.0f + operand wait %.0f + issue %.0f"
            % (name, prof[idx] * 1e3, n_t, info["m_tiles"], info["n_tiles"], info["block_n"],
               info["taps"] * info["cin_blocks"] // (3 if info["slab3"] else 1), info["stages"], info["ring_slots"],
               info["pair"], info["slab3"], info["bres"], info["staged"], info["epi_groups"], med(period, tiles),
               med(acc_wait, tiles), med(opw, tiles), med(issue, tiles)))
    pr = roles.get("producer")
    if pr is not None:
        w, prevc = {}, None
        for e, t, s, c in pr:
            e, t, c = int(e), int(t), int(c)
            if e == 1 and prevc is not None:
                w[t] = w.get(t, 0) + d32(c, prevc)
            prevc = c
        line += " | producer stage wait %.0f" % med(w, tiles)
    rs = roles.get("residual")
    if rs is not None:
        w, prevc = {}, None
        for e, t, s, c in rs:
            e, t, c = int(e), int(t), int(c)
            if e == 7 and prevc is not None:
                w[t] = w.get(t, 0) + d32(c, prevc)
            prevc = c
        line += " | residual slot wait %.0f" % med(w, tiles)
    for g in range(3):
        ev = roles.get("epi%d" % g)
        if ev is None:
            continue
        full_w, slot_w, ld, comp, st = {}, {}, {}, {}, {}
        prevc = None
        for e, t, s, c in ev:
            e, t, c = int(e), int(t), int(c)
            if prevc is not None:
                tgt = {9: full_w, 10: slot_w, 11: ld, 12: comp, 13: st}.get(e)
                if tgt is not None:
                    tgt[t] = tgt.get(t, 0) + d32(c, prevc)
            prevc = c
        line += "\n    epilogue group %d: wait acc full %.0f, wait slot %.0f, tmem ld %.0f, compute+barrier %.0f, store issue %.0f" % (
            g, med(full_w, tiles), med(slot_w, tiles), med(ld, tiles), med(comp, tiles), med(st, tiles))
    print(line + "\n    CTA total %.0f cycles" % total)
