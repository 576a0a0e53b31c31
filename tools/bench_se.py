import sys, os
sys.path.insert(0, "/root/repo")
import torch, rdb_b200
from rdb_b200.lib import check, ptr, stream_ptr
lib = rdb_b200.lib.load()
dev = torch.device("cuda:0")
n, c, sq, hw = 128, 960, 40, 196
sums = torch.randn(n, 1, c, device=dev)
w1 = torch.randn(sq, c, device=dev); b1 = torch.randn(sq, device=dev)
w2t = torch.randn(sq, c, device=dev); b2 = torch.randn(c, device=dev)
hid = torch.empty(n, sq, device=dev); gate = torch.empty(n, c, device=dev)
x = torch.randn(n, 14, 14, c, device=dev).to(torch.bfloat16); y = torch.empty_like(x)
def t(f, reps=50):
    for _ in range(5): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3
print("se_hidden %.1f us" % t(lambda: check(lib.b200_se_hidden(ptr(sums), n, 1, c, 1.0 / hw, ptr(w1), ptr(b1), sq, ptr(hid), stream_ptr()))))
print("se_scale  %.1f us" % t(lambda: check(lib.b200_se_scale_nhwc(ptr(x), n, 14, 14, c, 0, ptr(hid), ptr(w2t), ptr(b2), sq, ptr(y), 0, stream_ptr()))))
print("se_gate   %.1f us" % t(lambda: check(lib.b200_se_gate(ptr(sums), n, 1, c, 1.0 / hw, ptr(w1), ptr(b1), ptr(w2t), ptr(b2), sq, ptr(gate), stream_ptr()))))
print("ch_scale  %.1f us" % t(lambda: check(lib.b200_channel_scale_nhwc(ptr(x), n, 14, 14, c, 0, ptr(gate), ptr(y), 0, stream_ptr()))))
def both():
    check(lib.b200_se_hidden(ptr(sums), n, 1, c, 1.0 / hw, ptr(w1), ptr(b1), sq, ptr(hid), stream_ptr()))
    check(lib.b200_se_scale_nhwc(ptr(x), n, 14, 14, c, 0, ptr(hid), ptr(w2t), ptr(b2), sq, ptr(y), 0, stream_ptr()))
print("hidden+scale alternating %.1f us" % t(both))
g = torch.cuda.CUDAGraph()
s = torch.cuda.Stream()
with torch.cuda.stream(s):
    both(); torch.cuda.synchronize()
    with torch.cuda.graph(g, stream=s):
        for _ in range(20): both()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
g.replay(); torch.cuda.synchronize()
e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
print("graph of 20 x (hidden+scale): %.1f us per pair" % (e0.elapsed_time(e1) / 20 * 1e3))
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
print("graph se_hidden alone %.1f us" % graph_time(lambda: check(lib.b200_se_hidden(ptr(sums), n, 1, c, 1.0 / hw, ptr(w1), ptr(b1), sq, ptr(hid), stream_ptr()))))
print("graph se_scale alone  %.1f us" % graph_time(lambda: check(lib.b200_se_scale_nhwc(ptr(x), n, 14, 14, c, 0, ptr(hid), ptr(w2t), ptr(b2), sq, ptr(y), 0, stream_ptr()))))
print("graph se_gate alone   %.1f us" % graph_time(lambda: check(lib.b200_se_gate(ptr(sums), n, 1, c, 1.0 / hw, ptr(w1), ptr(b1), ptr(w2t), ptr(b2), sq, ptr(gate), stream_ptr()))))
print("graph ch_scale alone  %.1f us" % graph_time(lambda: check(lib.b200_channel_scale_nhwc(ptr(x), n, 14, 14, c, 0, ptr(gate), ptr(y), 0, stream_ptr()))))
tiny = torch.zeros(8, device=dev)
print("graph tiny torch op   %.1f us" % graph_time(lambda: tiny.add_(1.0)))
