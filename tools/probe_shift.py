"""Experiment (PPB_SHIFT_PROBE=1|2, PPB_PAIR=0): does a K-major SWIZZLE_128B UMMA descriptor that starts one row into a
TMA-written tile read rows 1..128 correctly? The A tile is loaded one row early, so with a working row shift every output
row is right except the last row of each 128-row tile (its operand row lies outside the 16 KB tile)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.nn.functional as F
from pyrapose_b200 import _native as nat
lib = nat.load()
dev = torch.device("cuda:0")
B, H, W, Cin, Cout, k = 2, 30, 40, 128, 128, 3
g = torch.Generator().manual_seed(0)
x = torch.randn(B, H, W, Cin, generator=g).to(torch.bfloat16)
w = (torch.randn(Cout, k, k, Cin, generator=g) / np.sqrt(k * k * Cin)).to(torch.bfloat16)
xp = F.pad(x, (0, 0, 1, 1, 1, 1)).contiguous().to(dev)
wd = w.reshape(Cout, -1).contiguous().to(dev)
yp = torch.zeros(B, H + 2, W + 2, Cout, dtype=torch.bfloat16, device=dev)
nat.check(lib.ppb_conv2d_bf16(nat.ptr(xp), nat.ptr(wd), None, None, nat.ptr(yp), B, H, W, Cin, Cout, k, 0, nat.current_stream_ptr()))
torch.cuda.synchronize()
ref = F.conv2d(x.to(dev).float().permute(0, 3, 1, 2), w.to(dev).float().permute(0, 3, 1, 2), padding=1).permute(0, 2, 3, 1)
refp = F.pad(ref, (0, 0, 1, 1, 1, 1)).reshape(-1, Cout)
got = yp.float().reshape(-1, Cout)
err = (got - refp).abs().max(dim=1).values.cpu().numpy()
rows = np.arange(err.size)
interior = (refp.abs().sum(1) > 0).cpu().numpy()
bad = (err > 0.05) & interior
print("probe=%s rows %d interior %d bad %d; bad rows mod 128 histogram (top): %s" % (
    os.environ.get("PPB_SHIFT_PROBE"), err.size, interior.sum(), bad.sum(),
    sorted(zip(*np.unique(rows[bad] % 128, return_counts=True)), key=lambda t: -t[1])[:6]))
print("bad rows with (row %% 128) != 127: %d ; mod-8 histogram of bad rows: %s" % (
    (bad & (rows % 128 != 127)).sum(), dict(zip(*np.unique(rows[bad] % 8, return_counts=True)))))
