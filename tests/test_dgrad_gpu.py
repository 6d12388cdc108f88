"""Parity of the data-gradient convolution (C-ABI ppb_conv2d_dgrad_bf16: the forward tcgen05 kernel with transposed
weights, negated taps, skip-gradient add and ReLU mask in the epilogue) against torch autograd's fp32 input gradient
on the same bf16-rounded operands. Only the fp32 accumulation order and the final bf16 rounding of dx differ:
relative L2 <= 4e-3 (bf16 output rounding alone is ~2.3e-3)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _run(dev, B, H, W, Cin, Cout, k, residual, mask, seed=0):
    import torch
    import torch.nn.functional as F
    from pyrapose_b200 import _native as nat
    lib = nat.load()
    gen = torch.Generator(device="cpu").manual_seed(seed)
    g_ld = (Cout + 63) // 64 * 64
    wgt = (torch.randn(Cout, Cin, k, k, generator=gen) * 0.05).to(torch.bfloat16)
    g = torch.randn(B, H, W, Cout, generator=gen).to(torch.bfloat16)
    res = torch.randn(B, H, W, Cin, generator=gen).to(torch.bfloat16) if residual else None
    act = torch.randn(B, H, W, Cin, generator=gen).to(torch.bfloat16) if mask else None

    def padded(t, ld=None):
        ld = ld or t.shape[3]
        out = torch.zeros(B, H + 2, W + 2, ld, dtype=torch.bfloat16)
        out[:, 1:-1, 1:-1, :t.shape[3]] = t
        return out.to(dev)

    wt = torch.zeros(Cin, k * k, g_ld, dtype=torch.bfloat16)
    wt[:, :, :Cout] = wgt.permute(1, 2, 3, 0).reshape(Cin, k * k, Cout)    # [ci][tap][co]
    wt = wt.reshape(Cin, k * k * g_ld).contiguous().to(dev)
    gp = padded(g, g_ld)
    rp = padded(res) if residual else None
    mp = padded(act) if mask else None
    dx = torch.full((B, H + 2, W + 2, Cin), 7.0, dtype=torch.bfloat16, device=dev)
    nat.check(lib.ppb_conv2d_dgrad_bf16(nat.ptr(gp), nat.ptr(wt), nat.ptr(rp), nat.ptr(mp), nat.ptr(dx), B, H, W, Cin,
                                        Cout, g_ld, k, nat.current_stream_ptr()))
    torch.cuda.synchronize()
    torch.backends.cudnn.allow_tf32 = False
    ref = torch.nn.grad.conv2d_input((B, Cin, H, W), wgt.float().to(dev), g.float().to(dev).permute(0, 3, 1, 2),
                                     padding=k // 2).permute(0, 2, 3, 1)
    if residual:
        ref = ref + res.float().to(dev)
    if mask:
        ref = ref * (act.float().to(dev) > 0)
    got = dx.float()
    border = torch.ones(B, H + 2, W + 2, dtype=torch.bool, device=dev)
    border[:, 1:-1, 1:-1] = False
    assert float(got[border].abs().max()) == 0.0, "border rows must be written as zeros"
    got = got[:, 1:-1, 1:-1]
    err = (got - ref)
    rel = (err.norm() / ref.norm()).item()
    # where is the error? per-row / per-column relative error profile
    prof_y = (err.pow(2).sum(dim=(0, 2, 3)) / ref.pow(2).sum(dim=(0, 2, 3)).clamp_min(1e-30)).sqrt().cpu().numpy()
    prof_x = (err.pow(2).sum(dim=(0, 1, 3)) / ref.pow(2).sum(dim=(0, 1, 3)).clamp_min(1e-30)).sqrt().cpu().numpy()
    return rel, prof_y, prof_x


CASES = [
    # B, H, W, Cin, Cout, k, residual, mask
    (2, 12, 16, 256, 3, 3, False, True),       # mask_out at the test size of test_train_gpu
    (2, 12, 16, 256, 256, 3, False, True),
    (1, 15, 20, 512, 144, 3, False, True),     # regression output conv: 9 anchors x 16 -> g_ld 192
    (2, 30, 40, 512, 512, 3, False, True),
    (2, 15, 20, 2048, 512, 1, True, True),     # bottleneck 2a with the identity-skip gradient added
    (2, 15, 20, 512, 2048, 1, False, True),    # bottleneck 2c
    (1, 17, 23, 64, 64, 3, True, False),
    (3, 7, 9, 1024, 256, 1, False, False),     # lateral
]


@pytest.mark.parametrize("case", CASES, ids=lambda c: "x".join(str(int(v)) for v in c))
def test_dgrad_matches_autograd(cuda_device, case):
    rel, py, px = _run(cuda_device, *case)
    print("case", case, "rel_l2", rel, "rows", np.round(py, 4), "cols", np.round(px, 4))
    assert rel < 4e-3, "relative L2 error %g" % rel


def test_dgrad_rejects_bad_arguments(cuda_device):
    import torch
    from pyrapose_b200 import _native as nat
    lib = nat.load()
    t = torch.zeros(64, device=cuda_device)
    with pytest.raises(ValueError):
        nat.check(lib.ppb_conv2d_dgrad_bf16(nat.ptr(t), nat.ptr(t), None, None, nat.ptr(t), 1, 4, 4, 48, 64, 64, 3, None))
    with pytest.raises(ValueError):
        nat.check(lib.ppb_conv2d_dgrad_bf16(nat.ptr(t), nat.ptr(t), None, None, nat.ptr(t), 1, 4, 4, 64, 64, 72, 3, None))
