"""GPU bring-up probe for the tcgen05 convolution: prints per-case errors (never asserts) and a
first throughput figure. Usage on the GPU box: python tools/probe_conv.py"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))

import torch

from test_conv_igemm_gpu import CASES, _run_case


def main():
    dev = torch.device("cuda:0")
    print(torch.cuda.get_device_name(0))
    for case in CASES:
        try:
            err, rel, border = _run_case(dev, *case)
            print("CASE %s max_abs=%.4g rel_l2=%.4g border_max=%.4g %s" % (case, err, rel, border,
                                                                         "OK" if rel < 1e-2 and border == 0 else "BAD"))
        except Exception as e:  # keep going: this is a diagnostic
            print("CASE %s EXC %r" % (case, e))
            break
    # throughput of the regression-head-shaped conv: 3x3 512->512 on the three pyramid planes
    import torch.nn.functional as F
    from pyrapose_b200 import _native as nat
    lib = nat.load()
    for (B, H, W, Cin, Cout, k) in [(32, 60, 80, 512, 512, 3), (32, 60, 80, 256, 256, 3), (32, 120, 160, 64, 256, 1),
                                    (32, 15, 20, 512, 2048, 1)]:
        x = torch.randn(B, H + 2, W + 2, Cin, device=dev).to(torch.bfloat16)
        w = torch.randn(Cout, k * k * Cin, device=dev).to(torch.bfloat16)
        y = torch.empty(B, H + 2, W + 2, Cout, device=dev, dtype=torch.bfloat16)
        bias = torch.zeros(Cout, device=dev)
        def run():
            nat.check(lib.ppb_conv2d_bf16(nat.ptr(x), nat.ptr(w), nat.ptr(bias), None, nat.ptr(y), B, H, W, Cin, Cout,
                                          k, 1, nat.current_stream_ptr()))
        for _ in range(3):
            run()
        torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        n = 10
        for _ in range(n):
            run()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        flops = 2.0 * B * H * W * Cin * Cout * k * k
        flops_issued = 2.0 * B * (H + 2) * (W + 2) * Cin * Cout * k * k
        print("PERF conv %dx%d %d->%d k%d B%d: %.3f ms  %.1f TFLOP/s useful (%.1f issued)" %
              (H, W, Cin, Cout, k, B, ms, flops / ms / 1e9, flops_issued / ms / 1e9))


if __name__ == "__main__":
    main()
