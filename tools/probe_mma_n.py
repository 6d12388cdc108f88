"""What does a tcgen05.mma with a narrow N cost? 3x3 convolutions with 64 input channels on the stage-2 plane
(B=32, 120x160: 4941 M tiles, slab mode, weights resident) for Cout = 64 / 128 / 256, and the same for 128 input
channels: if the time does not grow with N, the MMA is bound by its A-operand read, not by the math.
Usage on the GPU box: python tools/probe_mma_n.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from pyrapose_b200 import _native as nat


def main():
    dev = torch.device("cuda:0")
    lib = nat.load()
    B, H, W = 32, 120, 160
    for Cin in (64, 128):
        for Cout in (64, 128, 256):
            for k in (3, 1):
                x = torch.randn(B, H + 2, W + 2, Cin, device=dev).to(torch.bfloat16)
                w = torch.randn(Cout, k * k * Cin, device=dev).to(torch.bfloat16)
                y = torch.empty(B, H + 2, W + 2, Cout, device=dev, dtype=torch.bfloat16)
                bias = torch.zeros(Cout, device=dev)

                def run():
                    nat.check(lib.ppb_conv2d_bf16(nat.ptr(x), nat.ptr(w), nat.ptr(bias), None, nat.ptr(y), B, H, W, Cin,
                                                  Cout, k, 1, nat.current_stream_ptr()))
                for _ in range(3):
                    run()
                torch.cuda.synchronize()
                e0 = torch.cuda.Event(enable_timing=True)
                e1 = torch.cuda.Event(enable_timing=True)
                e0.record()
                n = 20
                for _ in range(n):
                    run()
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / n
                m_rows = B * (H + 2) * (W + 2)
                mmas = (m_rows / 128.0) * k * k * (Cin / 16.0) * max(1, Cout // 256)
                hbm = m_rows * (Cin + Cout) * 2
                print("k%d %3d->%3d: %.4f ms  %6.1f TFLOP/s  HBM floor %.4f ms  %.0f clk per MMA per SM at 1.9 GHz (N = %d)"
                      % (k, Cin, Cout, ms, 2.0 * m_rows * Cin * Cout * k * k / ms / 1e9, hbm / 6.5e9,
                         ms * 1e-3 * 1.9e9 / (mmas / 148.0), min(Cout, 256)))


if __name__ == "__main__":
    main()
