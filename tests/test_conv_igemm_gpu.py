"""Parity of the tcgen05 implicit-GEMM convolution (C-ABI ppb_conv2d_bf16) against an fp32
convolution of the same bf16-rounded operands. Tolerance: the only differences are fp32
accumulation order and the final bf16 rounding of the output -> rel 1e-2 (north_star bf16 bound),
observed ~4e-3."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _run_case(dev, B, H, W, Cin, Cout, k, relu, use_res, use_bias, seed=0):
    import torch
    import torch.nn.functional as F
    from pyrapose_b200 import _native as nat
    lib = nat.load()
    g = torch.Generator(device="cpu").manual_seed(seed)
    x = torch.randn(B, H, W, Cin, generator=g).to(torch.bfloat16)
    w = (torch.randn(Cout, k, k, Cin, generator=g) * (1.0 / np.sqrt(k * k * Cin))).to(torch.bfloat16)
    bias = torch.randn(Cout, generator=g) if use_bias else None
    res = torch.randn(B, H, W, Cout, generator=g).to(torch.bfloat16) if use_res else None

    def pad(t):
        return F.pad(t, (0, 0, 1, 1, 1, 1)).contiguous()

    xp = pad(x).to(dev)
    wd = w.reshape(Cout, k * k * Cin).contiguous().to(dev)
    bd = bias.to(dev) if use_bias else None
    rp = pad(res).to(dev) if use_res else None
    yp = torch.full((B, H + 2, W + 2, Cout), 7.0, dtype=torch.bfloat16, device=dev)
    rc = lib.ppb_conv2d_bf16(nat.ptr(xp), nat.ptr(wd), nat.ptr(bd), nat.ptr(rp), nat.ptr(yp), B, H, W, Cin, Cout, k,
                             int(relu), nat.current_stream_ptr())
    nat.check(rc)
    torch.cuda.synchronize()
    # fp32 reference on the same rounded operands
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    ref = F.conv2d(x.to(dev).float().permute(0, 3, 1, 2), w.to(dev).float().permute(0, 3, 1, 2), bias=bd,
                   padding=k // 2).permute(0, 2, 3, 1)
    if use_res:
        ref = ref + res.to(dev).float()
    if relu:
        ref = ref.clamp_min(0)
    got = yp.float()
    border = torch.ones(H + 2, W + 2, dtype=torch.bool, device=dev)
    border[1:-1, 1:-1] = False
    border_max = got[:, border].abs().max().item()
    inner = got[:, 1:-1, 1:-1]
    err = (inner - ref).abs().max().item()
    rel = ((inner - ref).norm() / ref.norm().clamp_min(1e-12)).item()
    return err, rel, border_max


CASES = [
    # B, H, W, Cin, Cout, k, relu, res, bias
    (1, 15, 20, 64, 64, 1, False, False, False),
    (2, 15, 20, 64, 64, 3, False, False, True),
    (2, 30, 40, 256, 256, 3, True, False, True),
    (1, 15, 20, 512, 2048, 1, True, True, True),
    (2, 60, 80, 128, 128, 3, True, False, True),
    (1, 30, 40, 1024, 256, 1, False, False, True),
    (3, 17, 23, 192, 144, 3, False, False, True),
    (2, 60, 80, 256, 512, 3, True, False, True),
    # >= 2 tiles per SM with a weight tile of at most 72 KB: the resident-weights mode (ConvParams::bres) with one, two
    # and four N tiles, with and without the slab mode / the residual ring
    (4, 120, 160, 64, 64, 3, True, False, True),
    (4, 120, 160, 64, 256, 1, True, True, True),
    (3, 120, 160, 256, 64, 1, True, False, True),
    (6, 60, 80, 128, 512, 1, True, True, True),
]


@pytest.mark.parametrize("case", CASES, ids=lambda c: "x".join(str(int(v)) for v in c))
def test_conv_matches_fp32_reference(cuda_device, case):
    err, rel, border = _run_case(cuda_device, *case)
    print("case", case, "max_abs_err", err, "rel_l2", rel, "border_max", border)
    assert border == 0.0, "border rows must be written as zeros"
    assert rel < 1e-2, "relative L2 error %g exceeds the bf16 tolerance 1e-2" % rel


def test_conv_rejects_bad_arguments(cuda_device):
    import torch
    from pyrapose_b200 import _native as nat
    lib = nat.load()
    t = torch.zeros(16, device=cuda_device)
    rc = lib.ppb_conv2d_bf16(nat.ptr(t), nat.ptr(t), None, None, nat.ptr(t), 1, 4, 4, 48, 64, 3, 0, None)
    with pytest.raises(ValueError):
        nat.check(rc)
