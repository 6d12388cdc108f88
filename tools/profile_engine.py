"""Per-launch device times of one forward pass (CUDA events, eager launches): python tools/profile_engine.py [B]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pyrapose_b200 import models
from pyrapose_b200.models import spec

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
H, W, C = (int(sys.argv[2]), int(sys.argv[3])) + (15,) if len(sys.argv) > 3 else (480, 640, 15)
ALIAS = os.environ.get("PPB_ALIAS", "0") != "0"   # production arena (shared backbone buffers)
m = models.convert_model(models.backbone("resnet50").retinanet(num_classes=C), alias_arena=ALIAS)
m.set_weights_dict(spec.init_weights(C, seed=1, head_init="spread"))
img = torch.randn(B, H, W, 3).mul_(60).pin_memory()
buf = m.forward_device(img, want_hull=True)
eng = m._engines[(H, W, 'alias') if ALIAS else (H, W)]
for _ in range(3):
    eng.run()
torch.cuda.synchronize()
runs = [eng.profile() for _ in range(5)]
ops = eng.op_table()
ms = np.median(np.array([[t for _, t in r] for r in runs]), axis=0)
tot = ms.sum()
print("B=%d %dx%d: %d launches, %.3f ms total (sum of per-launch times); arena %.2f GB%s" %
      (B, H, W, len(ops), tot, eng.device_bytes() / 1e9, " (production arena)" if ALIAS else ""))
for (name, fl), t in zip(ops, ms):
    print("%-28s %8.4f ms %5.1f%% %8.1f TFLOP/s" % (name, t, 100 * t / tot, fl / t / 1e9 if fl else 0.0))
conv = sum(t for (n, fl), t in zip(ops, ms) if fl)
print("conv total %.3f ms, %.1f TFLOP/s; elementwise %.3f ms" % (conv, sum(fl for _, fl in ops) / conv / 1e9, tot - conv))
# graph replay time
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    eng.run()
e1.record(); torch.cuda.synchronize()
print("graph replay: %.3f ms per forward" % (e0.elapsed_time(e1) / 10))
