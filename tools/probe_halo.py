#!/usr/bin/env python
"""Per-tap correctness probe of the 3x3 halo kernel: weights with a single non-zero tap isolate each shifted view."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.nn.functional as F
import rdb_b200
from rdb_b200.lib import check, ptr, stream_ptr

lib = rdb_b200.lib.load()
dev = torch.device("cuda:0")
torch.backends.cudnn.allow_tf32 = False
n, h, w, c = 2, 56, 56, 64
g = torch.Generator().manual_seed(1)
x = torch.randn(n, h, w, c, generator=g).to(dev).to(torch.bfloat16)
b = torch.zeros(c, device=dev)
for t in list(range(9)) + [-1]:
    wt = torch.zeros(c, 3, 3, c)
    if t >= 0:
        wt[:, t // 3, t % 3, :] = torch.randn(c, c, generator=g) / 8
    else:
        wt = torch.randn(c, 3, 3, c, generator=g) / 24
    wt = wt.to(dev).to(torch.bfloat16)
    out = torch.full((n, h, w, c), float("nan"), device=dev, dtype=torch.bfloat16)
    check(lib.b200_conv2d_nhwc(ptr(x), n, h, w, c, 0, ptr(wt), ptr(b), None, 0, ptr(out), c, 0, 3, 1, 1, 0, stream_ptr()))
    torch.cuda.synchronize()
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), wt.float().permute(0, 3, 1, 2), b, padding=1).permute(0, 2, 3, 1)
    err = (out.float() - ref).abs()
    bad = err > ref.abs() * 2 ** -8 + 2e-3
    # where are the bad pixels? fraction by x position mod 8
    byx = [bad[:, :, xx::8].float().mean().item() for xx in range(8)]
    print("tap", t, "(dy,dx)=", (t // 3, t % 3) if t >= 0 else "all", "max err %.4f bad %.3f" % (err.max().item(), bad.float().mean().item()),
          "bad by x%8:", ["%.2f" % v for v in byx])
