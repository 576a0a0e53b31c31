#!/usr/bin/env python
"""What a TMA box that is partly outside its tensor map costs on B200 (profiles/r02_tma_box_bounds_probe.txt,
r02_tma_store_bounds_probe.txt): 1x1 convs through b200_conv2d_nhwc on shapes where the 64-channel K box, the 64-row
weight box or the 64-channel store box is clipped by the tensor's extent, against the same convs with every box inside.

"declared 64, rows overlap": the input is given as 64 channels with the true pixel stride (32 or 16 channels), so the
activation box reads into the next pixel and the weights are a zero-padded [cout][64] matrix -- both boxes inside their maps.
The engine reaches the fast case by storing conv weights padded to the box sizes (lowering.weight_k / weight_rows).
Usage: python tools/probe_tma_bounds.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import rdb_b200
from rdb_b200.lib import check, ptr, stream_ptr

lib = rdb_b200.lib.load()
dev = torch.device("cuda:0")


def time_it(run):
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        run()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 20 * 1e3


def load_case(tag, n, h, w, c_phys, c_decl, stride, cout):
    """c_phys valid input channels, declared as c_decl channels with a pixel stride of `stride` channels."""
    P = n * h * w
    x = torch.randn(P + 8, stride, device=dev).to(torch.bfloat16)
    wt = torch.zeros(cout, c_decl, device=dev, dtype=torch.bfloat16)
    wt[:, :c_phys] = (torch.randn(cout, c_phys, device=dev) / c_phys ** 0.5).to(torch.bfloat16)
    b = torch.randn(cout, device=dev)
    out = torch.empty(n, h, w, cout, device=dev, dtype=torch.bfloat16)
    us = time_it(lambda: check(lib.b200_conv2d_nhwc(ptr(x), n, h, w, c_decl, stride, ptr(wt), ptr(b), None, 0, ptr(out), cout,
                                                    0, 1, 1, 0, 1, stream_ptr())))
    ref = torch.relu(x[:P, :c_phys].float() @ wt[:, :c_phys].float().t() + b)
    err = (out.view(P, cout).float() - ref).abs().max().item()
    print("%-58s %8.1f us   max err %.4f" % (tag, us, err), flush=True)


def store_case(tag, n, h, w, cin, cout, ostride, res, wrows):
    P = n * h * w
    x = torch.randn(P, cin, device=dev).to(torch.bfloat16)
    wt = torch.zeros(max(wrows, cout), cin, device=dev, dtype=torch.bfloat16)
    wt[:cout] = (torch.randn(cout, cin, device=dev) / cin ** 0.5).to(torch.bfloat16)
    b = torch.randn(cout, device=dev)
    out = torch.empty(P + 8, ostride, device=dev, dtype=torch.bfloat16)
    r = torch.randn(P + 8, ostride, device=dev).to(torch.bfloat16) if res else None
    us = time_it(lambda: check(lib.b200_conv2d_nhwc(ptr(x), n, h, w, cin, 0, ptr(wt), ptr(b), ptr(r) if res else None,
                                                    ostride if res else 0, ptr(out), cout, ostride, 1, 1, 0, 1, stream_ptr())))
    byt = P * (cin + cout * (2 if res else 1)) * 2
    print("%-58s %8.1f us  %6.0f GB/s" % (tag, us, byt / us / 1e3), flush=True)


for (n, h, w, cout) in ((512, 56, 56, 64), (32, 160, 160, 32)):
    s = "%dx%dx%d " % (n, h, w)
    load_case(s + "cin 32, rows 64 B, boxes clipped", n, h, w, 32, 32, 32, cout)
    load_case(s + "cin 32 of 64-channel pixels, boxes clipped", n, h, w, 32, 32, 64, cout)
    load_case(s + "cin 32 declared 64, rows overlap", n, h, w, 32, 64, 32, cout)
    load_case(s + "cin 64", n, h, w, 64, 64, 64, cout)
    load_case(s + "cin 16 declared 64, rows overlap", n, h, w, 16, 64, 16, cout)
    load_case(s + "cin 16, rows 32 B, boxes clipped", n, h, w, 16, 16, 16, cout)
n, h, w = 32, 160, 160
store_case("64 -> 64", n, h, w, 64, 64, 64, False, 64)
store_case("64 -> 32 (64-byte pixels, store box clipped)", n, h, w, 64, 32, 32, False, 64)
store_case("64 -> 32 into 64-channel pixels (store box clipped)", n, h, w, 64, 32, 64, False, 64)
store_case("64 -> 48 (96-byte pixels, store box clipped)", n, h, w, 64, 48, 48, False, 64)
n, h, w = 128, 14, 14
store_case("960 -> 160 + residual (third box clipped)", n, h, w, 960, 160, 160, True, 256)
store_case("960 -> 192 + residual", n, h, w, 960, 192, 192, True, 256)
store_case("960 -> 128 + residual", n, h, w, 960, 128, 128, True, 128)
