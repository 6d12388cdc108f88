"""Forward parity of the engine against the bf16-EMULATING oracle (oracle/train.py rounding hooks), tap by tap.
With identical rounding points the two should agree to ~1e-3; a larger gap localises a forward discrepancy."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch
    from oracle import pfpn, train as otrain
    from pyrapose_b200 import models
    from pyrapose_b200.models import spec
    C, B, H, W = 3, 2, 96, 128
    if len(sys.argv) > 2:
        H, W = int(sys.argv[1]), int(sys.argv[2])
    weights = spec.init_weights(C, seed=5, head_init="spread", randomize_bn=True)
    rng = np.random.default_rng(7)
    img = rng.integers(0, 256, (B, H, W, 3)).astype(np.float32) - np.array([103.939, 116.779, 123.68], np.float32)
    wt = {k: torch.from_numpy(np.ascontiguousarray(v)) for k, v in weights.items()}
    taps = {}
    with torch.no_grad():
        net_w = otrain._emulated_weights(wt)
        x = torch.from_numpy(img).permute(0, 3, 1, 2).contiguous().to(torch.bfloat16).to(torch.float32)
        pfpn._Q[0] = otrain._RoundBF16.apply
        _, C3, C4, C5 = pfpn.resnet50(x, net_w, taps)
        feats = pfpn.sparce_fpn(C3, C4, C5, net_w, taps)
        reg = torch.cat([pfpn.regression_head(f, net_w) for f in feats], dim=1).numpy()
        pfpn._Q[0] = lambda t: t
    ref = pfpn.taps_to_nhwc(taps)
    model = models.backbone("resnet50").retinanet(num_classes=C)
    model.set_weights_dict(weights)
    out = model.predict_on_batch(img)
    names = ["conv1_relu", "pool1"] + [k for k in ref if k.startswith("res")] + \
            ["P3_lateral", "P4_lateral", "P5_lateral", "P4_mid", "P3_mid", "P3_down", "P4_down", "P3", "P4", "P5"]
    for n in names:
        got = model.activation(n, H, W)
        r = ref[n]
        err = got - r
        rel = np.linalg.norm(err) / max(np.linalg.norm(r), 1e-30)
        # error profile over image rows/cols to spot border effects
        py = np.sqrt((err ** 2).sum(axis=(0, 2, 3)) / np.maximum((r ** 2).sum(axis=(0, 2, 3)), 1e-30))
        print("%-14s rel %.5f  rows[0,1,mid,-2,-1] %s" % (n, rel, np.round(py[[0, 1, len(py) // 2, -2, -1]], 4)))
    print("3Dbox rel", np.linalg.norm(out[0] - reg) / np.linalg.norm(reg))


if __name__ == "__main__":
    main()
