#!/usr/bin/env python
"""Graph-replayed timing of the MBConv project conv: plain, with the SE scale folded in, and the separate scale pass."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, rdb_b200
from rdb_b200.lib import check, ptr, stream_ptr
lib = rdb_b200.lib.load()
dev = torch.device("cuda:0")
s = torch.cuda.Stream()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
def graph_time(f, reps=20):
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(s):
        f(); torch.cuda.synchronize()
        with torch.cuda.graph(g, stream=s):
            for _ in range(reps): f()
    torch.cuda.synchronize()
    g.replay(); torch.cuda.synchronize()
    e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3
for (n, h, w, cin, cout) in [(128, 14, 14, 960, 160), (128, 7, 7, 1536, 256), (128, 14, 14, 512, 128)]:
    x = torch.randn(n, h, w, cin, device=dev).to(torch.bfloat16)
    gate = torch.rand(n, cin, device=dev)
    wt = (torch.randn(cout, 1, 1, cin, device=dev) / cin ** 0.5).to(torch.bfloat16)
    b = torch.randn(cout, device=dev)
    res = torch.randn(n, h, w, cout, device=dev).to(torch.bfloat16)
    out = torch.empty(n, h, w, cout, device=dev, dtype=torch.bfloat16)
    y = torch.empty_like(x)
    tp = graph_time(lambda: check(lib.b200_conv2d_nhwc(ptr(x), n, h, w, cin, 0, ptr(wt), ptr(b), ptr(res), 0, ptr(out), cout, 0, 1, 1, 0, 0, stream_ptr())))
    ts = graph_time(lambda: check(lib.b200_conv1x1_scaled_nhwc(ptr(x), n, h, w, cin, 0, ptr(gate), ptr(wt), ptr(b), ptr(res), 0, ptr(out), cout, 0, 0, stream_ptr())))
    tc = graph_time(lambda: check(lib.b200_channel_scale_nhwc(ptr(x), n, h, w, cin, 0, ptr(gate), ptr(y), 0, stream_ptr())))
    print("%s: plain conv %.1f us, scaled conv %.1f us, scale pass %.1f us" % ((n, h, w, cin, cout), tp, ts, tc))
