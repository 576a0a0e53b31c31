import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rdb_b200
from rdb_b200.lib import load, check, ptr, stream_ptr
lib = load()
cuda = torch.device("cuda:0")
n, h, w, cin, c, cp = [int(v) for v in sys.argv[1:7]]
x = torch.randn(n, h, w, cin).to(cuda).to(torch.bfloat16)
parent = torch.randn(n, h, w, 2 * cp).to(cuda).to(torch.bfloat16)
wslots = torch.randn(cp, cin).to(cuda).to(torch.bfloat16)
bslots = torch.randn(cp).to(cuda)
out = torch.full((n, h, w, 2 * cp), float("nan"), device=cuda, dtype=torch.bfloat16)
torch.cuda.synchronize()
t0 = time.time()
rc = lib.b200_conv1x1_shuffle_nhwc(ptr(x), n, h, w, cin, 0, ptr(parent), 2 * cp, c, cp, ptr(wslots), ptr(bslots), ptr(out), 0, stream_ptr())
print("rc", rc, flush=True)
try:
    torch.cuda.synchronize()
    print("ok in %.3f s" % (time.time() - t0), "nan count", torch.isnan(out.float()).sum().item(), flush=True)
except Exception as e:
    print("failed after %.3f s: %s" % (time.time() - t0, str(e).splitlines()[0]), flush=True)
