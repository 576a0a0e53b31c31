#!/usr/bin/env python
"""Times b200_conv2d_nhwc (or, with a trailing 'g', b200_gather_conv2d_nhwc) on given shapes:
n h w cin cout k stride pad act(0/1/2) res(0/1) [g], comma-separated cases."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import rdb_b200
from rdb_b200.lib import check, ptr, stream_ptr
lib = rdb_b200.lib.load()
dev = torch.device("cuda:0")
for case in sys.argv[1].split(","):
    f = case.split()
    gather = f[-1] == "g"
    n, h, w, cin, cout, k, s, p, act, res = [int(v) for v in f[:10]]
    ho, wo = (h + 2 * p - k) // s + 1, (w + 2 * p - k) // s + 1
    x = torch.randn(n, h, w, cin, device=dev).to(torch.bfloat16)
    wt = (torch.randn(cout, k, k, cin, device=dev) / (k * k * cin) ** 0.5).to(torch.bfloat16)
    b = torch.randn(cout, device=dev)
    r = torch.randn(n, ho, wo, cout, device=dev).to(torch.bfloat16) if res else None
    out = torch.empty(n, ho, wo, cout, device=dev, dtype=torch.bfloat16)
    if gather:
        npad, kpad = lib.b200_gather_conv_npad(cout), lib.b200_gather_conv_kpad(cin, k)
        wg = torch.zeros(npad, kpad, device=dev, dtype=torch.bfloat16)
        wg[:cout, :k * k * cin] = wt.reshape(cout, -1)
        bg = torch.zeros(npad, device=dev)
        bg[:cout] = b
    def run():
        if gather:
            check(lib.b200_gather_conv2d_nhwc(ptr(x), n, h, w, cin, 0, ptr(wg), ptr(bg), ptr(r), 0, ptr(out), cout, 0, k, s, p,
                                              act, 0, stream_ptr()))
            return
        check(lib.b200_conv2d_nhwc(ptr(x), n, h, w, cin, 0, ptr(wt), ptr(b), ptr(r), 0, ptr(out), cout, 0, k, s, p, act, stream_ptr()))
    for _ in range(3): run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): run()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 20 * 1e3
    byt = (n * h * w * cin + n * ho * wo * cout * (2 if res else 1)) * 2 + cout * k * k * cin * 2
    fl = 2.0 * n * ho * wo * cout * cin * k * k
    print("%-40s %8.1f us  %7.0f GB/s  %7.1f TFLOP/s" % (case, us, byt / us / 1e3, fl / us / 1e6))
