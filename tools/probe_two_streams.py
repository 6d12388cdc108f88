"""Experiment: one B=32 engine on one stream vs two B=16 engines replayed concurrently on two streams (tail filling).
python tools/probe_two_streams.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pyrapose_b200 import models
from pyrapose_b200.models import spec

H, W, C = 480, 640, 15
w = spec.init_weights(C, seed=1, head_init="spread")


def make(B):
    m = models.convert_model(models.backbone("resnet50").retinanet(num_classes=C))
    m.set_weights_dict(w)
    img = torch.randn(B, H, W, 3).mul_(60).pin_memory()
    m.forward_device(img, want_hull=True)
    return m, m._engines[(H, W)]


def timeit(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


m32, e32 = make(32)
print("one engine  B=32      : %.3f ms per 32 images" % timeit(e32.run))
ma, ea = make(16)
mb, eb = make(16)
print("one engine  B=16 x2 serial: %.3f ms per 32 images" % timeit(lambda: (ea.run(), eb.run())))
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def both():
    cur = torch.cuda.current_stream()
    s1.wait_stream(cur); s2.wait_stream(cur)
    with torch.cuda.stream(s1):
        ea.run()
    with torch.cuda.stream(s2):
        eb.run()
    cur.wait_stream(s1); cur.wait_stream(s2)


print("two engines B=16 on two streams: %.3f ms per 32 images" % timeit(both))
mc, ec = make(32)


def both32():
    cur = torch.cuda.current_stream()
    s1.wait_stream(cur); s2.wait_stream(cur)
    with torch.cuda.stream(s1):
        e32.run()
    with torch.cuda.stream(s2):
        ec.run()
    cur.wait_stream(s1); cur.wait_stream(s2)


print("two engines B=32 on two streams: %.3f ms per 32 images" % (timeit(both32) / 2))
