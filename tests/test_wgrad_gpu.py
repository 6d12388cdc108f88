"""Parity of the tcgen05 weight-gradient kernel (C-ABI ppb_conv2d_wgrad_bf16) against torch autograd's fp32
convolution weight gradient on the same bf16-rounded operands. The only differences are fp32 accumulation order
(split-K with atomics) -> relative L2 <= 1e-3 (observed ~1e-6)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _run(dev, B, H, W, Cin, Cout, k, g_ld=None, seed=0):
    import torch
    import torch.nn.functional as F
    from pyrapose_b200 import _native as nat
    lib = nat.load()
    gen = torch.Generator(device="cpu").manual_seed(seed)
    x = torch.randn(B, H, W, Cin, generator=gen).to(torch.bfloat16)
    g = torch.randn(B, H, W, Cout, generator=gen).to(torch.bfloat16)
    g_ld = g_ld or Cout
    xp = F.pad(x, (0, 0, 1, 1, 1, 1)).contiguous().to(dev)
    gp = torch.zeros(B, H + 2, W + 2, g_ld, dtype=torch.bfloat16)
    gp[:, 1:-1, 1:-1, :Cout] = g
    gp = gp.to(dev)
    dw = torch.zeros(Cout, k * k * Cin, dtype=torch.float32, device=dev)
    nat.check(lib.ppb_conv2d_wgrad_bf16(nat.ptr(xp), nat.ptr(gp), nat.ptr(dw), B, H, W, Cin, Cout, g_ld, k,
                                        nat.current_stream_ptr()))
    torch.cuda.synchronize()
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    xr = x.to(dev).float().permute(0, 3, 1, 2)
    gr = g.to(dev).float().permute(0, 3, 1, 2)
    ref = torch.nn.grad.conv2d_weight(xr, (Cout, Cin, k, k), gr, padding=k // 2)      # [Cout, Cin, k, k]
    ref = ref.permute(0, 2, 3, 1).reshape(Cout, k * k * Cin)                           # O,HWI like the kernel
    rel = ((dw - ref).norm() / ref.norm()).item()
    return rel, (dw - ref).abs().max().item()


CASES = [
    # B, H, W, Cin, Cout, k, g_ld
    (1, 15, 20, 64, 64, 1, None),
    (2, 15, 20, 64, 128, 3, None),
    (2, 30, 40, 256, 256, 3, None),
    (1, 17, 23, 128, 144, 3, 192),      # head output conv: 9 anchors x 16, gradient matrix padded to 192 columns
    (2, 30, 40, 512, 512, 3, None),
    (1, 15, 20, 2048, 512, 1, None),
    (2, 33, 41, 192, 21, 3, 64),        # mask output conv, 21 classes
]


@pytest.mark.parametrize("case", CASES, ids=lambda c: "x".join(str(v) for v in c))
def test_wgrad_matches_autograd(cuda_device, case):
    rel, mx = _run(cuda_device, *case)
    print("case", case, "rel_l2", rel, "max_abs", mx)
    assert rel < 1e-3, "relative L2 error %g" % rel


def test_wgrad_accumulates_and_rejects_bad_arguments(cuda_device):
    import torch
    from pyrapose_b200 import _native as nat
    lib = nat.load()
    t = torch.zeros(64, device=cuda_device)
    with pytest.raises(ValueError):
        nat.check(lib.ppb_conv2d_wgrad_bf16(nat.ptr(t), nat.ptr(t), nat.ptr(t), 1, 4, 4, 48, 64, 64, 3, None))
    with pytest.raises(ValueError):
        nat.check(lib.ppb_conv2d_wgrad_bf16(nat.ptr(t), nat.ptr(t), nat.ptr(t), 1, 4, 4, 64, 64, 60, 3, None))
