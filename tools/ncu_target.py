"""Short target for `ncu`: one CUDA-graph forward (78 conv launches) then one eager forward (78 more) at batch B.
Conv launch order inside a forward: 0 conv1 | 1-52 ResNet bottlenecks | 53-55 laterals | 56 P4_mid 57 P3_mid
58 P3_down 59 P3 60 P4_down 61 P4 62 P5 (see engine.cu build_plan) | 63-67 cls head | 68-72 regression head | 73-77 mask head.
Usage: python tools/ncu_target.py [B] [filter]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pyrapose_b200 import _kernels, models
from pyrapose_b200.models import spec

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
H, W, C = 480, 640, 15
m = models.convert_model(models.backbone("resnet50").retinanet(num_classes=C))
m.set_weights_dict(spec.init_weights(C, seed=1, head_init="spread"))
img = torch.randn(B, H, W, 3).mul_(60).pin_memory()
buf = m.forward_device(img, want_hull=True)
eng = m._engines[(H, W)]
torch.cuda.synchronize()
eng.run(eager=True)
torch.cuda.synchronize()
if len(sys.argv) > 2:
    out = _kernels.filter_detections(buf["hull"], buf["boxes3D"], buf["cls"])
    torch.cuda.synchronize()
print("ok", eng.kernels_per_run())
