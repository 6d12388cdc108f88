"""Read-only / write-only / copy HBM bandwidth of this GPU with library kernels (torch fill_, sum, copy_), CUDA events,
best of 10 over 2 GiB buffers. The copy figure is what MEASURED_PEAKS.json records; write-only streams (the materialised
IoU matrix) and read-only streams (score pass) are bounded by the other two."""
import torch
dev = torch.device("cuda:0")
n = 1 << 30
a = torch.empty(n, dtype=torch.int16, device=dev)
b = torch.empty(n, dtype=torch.int16, device=dev)
f = torch.empty(n // 2, dtype=torch.float32, device=dev)


def best(fn, bytes_):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(10):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return bytes_ / (min(ts) * 1e-3) / 1e9


print("write-only  fill_   : %.0f GB/s" % best(lambda: a.fill_(3), 2 * n))
print("write-only  zero_   : %.0f GB/s" % best(lambda: a.zero_(), 2 * n))
print("read-only   sum f32 : %.0f GB/s" % best(lambda: f.sum(), 2 * n))
print("copy (r+w)  copy_   : %.0f GB/s" % best(lambda: b.copy_(a), 4 * n))

# event-to-event floor of one tiny launch, and a library read of the score-pass size (108.9 MB) after an L2 flush
tiny = torch.empty(32, device=dev)
s = torch.rand(32 * 56700 * 15, device=dev)
flush = torch.empty(1 << 30, dtype=torch.uint8, device=dev)


def once(fn):
    ts = []
    for _ in range(10):
        flush.fill_(1)
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


print("tiny launch, event to event : %.2f us" % (1e3 * once(lambda: tiny.fill_(1))))
t = once(lambda: s.sum())
print("torch.sum over 108.9 MB (2 launches): %.2f us = %.0f GB/s" % (1e3 * t, s.numel() * 4 / t / 1e6))
t = once(lambda: s.add_(1.0))
print("torch add_ over 108.9 MB (r+w 217.7 MB): %.2f us = %.0f GB/s" % (1e3 * t, 2 * s.numel() * 4 / t / 1e6))
