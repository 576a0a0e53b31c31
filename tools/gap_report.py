#!/usr/bin/env python
"""Per-op gap to the roofline: measured us vs max(bytes / HBM peak, flops / tensor peak), sorted by absolute gap."""
import json, os, re, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import rdb_b200
from bench import plan_traffic, load_peaks
from rdb_b200 import synth
from rdb_b200.engine import B200Model

kind = sys.argv[1] if len(sys.argv) > 1 else "resnet50"
batch = int(sys.argv[2]) if len(sys.argv) > 2 else 256
peaks = load_peaks()
module = synth.make_model(kind) if kind == "yolov8n" else synth.make_model(kind, n_proto=8)
model = B200Model(kind, module, max_batch=batch)
imgs = list(synth.synthetic_images(8, model.in_h, model.in_w)) * (batch // 8 + 1)
model.infer_u8(imgs[:batch]); torch.cuda.synchronize()
ms = np.zeros(model.n_ops)
for r in range(6):
    m, _ = model.profile_ops(batch)
    if r: ms += m
ms /= 5
tr = plan_traffic(model.blob, batch)
rows = []
for i in range(model.n_ops):
    t, b, fl = tr[i]
    bound = max(b / (peaks["hbm"] * 1e9), fl / (peaks["tensor_sustained"] * 1e12)) * 1e6
    rows.append((ms[i] * 1e3 - bound, model.op_names[i], ms[i] * 1e3, bound, "hbm" if b / (peaks["hbm"] * 1e9) >= fl / (peaks["tensor_sustained"] * 1e12) else "tensor"))
rows.sort(reverse=True)
tot, totb = sum(r[2] for r in rows), sum(r[3] for r in rows)
print("%-28s %8s %8s %8s" % ("op", "us", "bound", "gap"))
for g, n, u, b, k in rows[:24]:
    print("%-28s %8.1f %8.1f %8.1f  %s" % (n, u, b, g, k))
if os.environ.get("GAP_ALL"):
    import collections
    agg = collections.defaultdict(lambda: [0, 0.0, 0.0])
    for g, n, u, b, k in rows:
        key = re.sub(r"\d+", "#", n)
        agg[key][0] += 1; agg[key][1] += u; agg[key][2] += b
    print("--- by op family")
    for key, (c, u, b) in sorted(agg.items(), key=lambda kv: -(kv[1][1] - kv[1][2])):
        print("%-28s x%-3d %8.1f %8.1f %8.1f" % (key, c, u, b, u - b))
print("TOTAL measured %.0f us, sum of per-op bounds %.0f us (%.0f%%)" % (tot, totb, 100 * totb / tot))
if os.environ.get("GAP_GREP"):
    for g, n, u, b, k in sorted(rows, key=lambda r: r[1]):
        if re.search(os.environ["GAP_GREP"], n):
            print("%-28s %8.1f %8.1f %8.1f  %s" % (n, u, b, g, k))
