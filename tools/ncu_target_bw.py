"""Short target for `ncu --set full` over the bandwidth-class kernels at BASELINE.json sizes: one launch each of
score_mask (B=32, N=56700, C=15), vote filter, the FilterDetections chain (compact / sort / nms / topk), the
materialised compute_overlap (8 images x 100 boxes, 363 MB written) and the fused gt_assign (B=64).
Usage: python tools/ncu_target_bw.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from pyrapose_b200 import _kernels
from pyrapose_b200.utils import anchors as ua

dev = torch.device("cuda:0")
B, N, C = 32, 56700, 15
g = torch.Generator(device=dev).manual_seed(0)
anchors_f32 = ua.anchors_for_shape_device((480, 640, 3)).float()
deltas = torch.randn(B, N, 16, generator=g, device=dev)
sc = torch.sigmoid(torch.randn(B, N, C, generator=g, device=dev) * 1.5 - 6.0)
boxes3d, hull = _kernels.box3d_transform_inv(anchors_f32, deltas, want_hull=True)
_kernels.score_mask(sc, 0.05, want_counts=False)
_kernels.score_mask(sc, 0.5)
_kernels.filter_detections(hull, boxes3d, sc)
anchors64 = ua.anchors_for_shape_device((480, 640, 3))
qs = []
for i in range(64):
    rng = np.random.default_rng(1000 + i)
    x1 = rng.uniform(0, 624, 100); y1 = rng.uniform(0, 464, 100)
    w = rng.uniform(16, 256, 100); h = rng.uniform(16, 256, 100)
    qs.append(np.stack([x1, y1, np.minimum(x1 + w, 640), np.minimum(y1 + h, 480)], axis=1))
q = torch.from_numpy(np.stack(qs)).to(dev)
_kernels.compute_overlap(anchors64, q[:8])
_kernels.gt_assign(anchors64, q)
torch.cuda.synchronize()
print("ok")
