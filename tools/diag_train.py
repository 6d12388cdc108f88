"""Per-tensor gradient parity table of the native training step vs the CPU oracle (diagnostic, run on the GPU box)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import test_train_gpu as T  # noqa: E402


def main():
    import torch
    from oracle import train as otrain
    from pyrapose_b200 import models, optimizers
    from pyrapose_b200.models import spec
    from pyrapose_b200.utils.anchors import anchor_targets_bbox, anchors_for_shape
    head_init = sys.argv[1] if len(sys.argv) > 1 else "spread"
    emulate = len(sys.argv) > 2 and sys.argv[2] == "bf16"
    C, B, H, W = T.C, T.B, T.H, T.W
    weights = spec.init_weights(C, seed=5, head_init=head_init, randomize_bn=True)
    rng = np.random.default_rng(7)
    img = rng.integers(0, 256, (B, H, W, 3)).astype(np.float32) - np.array([103.939, 116.779, 123.68], np.float32)
    y = T._targets(anchors_for_shape, anchor_targets_bbox)
    ref = otrain.step(img, weights, y, C, lr=T.LR, clipnorm=0.001, emulate_bf16=emulate)
    model = models.backbone("resnet50").retinanet(num_classes=C)
    model.set_weights_dict({k: v.copy() for k, v in weights.items()})
    model.compile(optimizer=optimizers.Adam(lr=T.LR, clipnorm=0.001))
    out = model.train_on_batch(img, y)
    eng = model._train_engine
    print("losses", out, ref["losses"], model.last_step, ref["grad_norm"])
    for name, r in zip(("3Dbox", "cls", "mask"), ref["outputs"]):
        got = eng.buffers[name].cpu().numpy()
        print("output %-6s rel_l2 %.5f max_abs %.5f" % (name, np.linalg.norm(got - r) / np.linalg.norm(r),
                                                         np.abs(got - r).max()))
    for l in reversed(model.layers):
        for bias in (True, False):
            key = l.name + ("/bias" if bias else "/kernel")
            if key not in ref["grads"]:
                continue
            got = eng.train_layer_tensor("grads", l.name, l.ksize, l.cin, l.cout, bias=bias)
            want = ref["grads"][key]
            n = np.linalg.norm(want)
            rel = np.linalg.norm(got - want) / max(n, 1e-30)
            cos = float((got.astype(np.float64) * want).sum() / max(np.linalg.norm(got) * n, 1e-30))
            extra = ""
            if not bias and l.ksize == 3:
                # per-tap relative error (HWIO): localises tap-offset mistakes
                e = [np.linalg.norm(got[r, s] - want[r, s]) / max(np.linalg.norm(want[r, s]), 1e-30)
                     for r in range(3) for s in range(3)]
                extra = " taps " + " ".join("%.3f" % v for v in e)
            print("%-34s norm %.3e rel %.4f cos %.5f%s" % (key, n, rel, cos, extra))


if __name__ == "__main__":
    main()
