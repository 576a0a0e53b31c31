#!/usr/bin/env python
"""Joins an ncu launch list of ONE step (collate + n_ops kernels, in op order) with the plan's per-op roofline bounds.
Usage: python tools/ncu_gap.py kind batch launches.csv   (CPU only: lowers the model to get names and traffic)"""
import csv, os, re, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import plan_traffic, load_peaks
from rdb_b200 import synth, lowering

kind, batch, path = sys.argv[1], int(sys.argv[2]), sys.argv[3]
module = synth.make_model(kind) if kind == "yolov8n" else synth.make_model(kind, n_proto=8)
pb = lowering.lower(kind, module) if hasattr(lowering, "lower") else None
blob, names = pb.to_blob(), pb.names
peaks = load_peaks()
tr = plan_traffic(blob, batch)
rows = list(csv.reader(open(path)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
hdr = rows[hi]
ik, iv = hdr.index("Kernel Name"), hdr.index("Metric Value")
launches = [(re.sub(r"\(.*", "", r[ik])[:44], float(r[iv].replace(",", "")) / 1e3) for r in rows[hi + 1:] if len(r) > iv]
launches = [l for l in launches if "collate" not in l[0]][:len(names)]
out = []
for (kn, us), nm, (t, b, fl) in zip(launches, names, tr):
    bound = max(b / (peaks["hbm"] * 1e9), fl / (peaks["tensor_sustained"] * 1e12)) * 1e6
    out.append((us - bound, nm, kn, us, bound))
out.sort(reverse=True)
print("%-30s %-44s %8s %8s %8s" % ("op", "kernel", "us", "bound", "gap"))
for g, nm, kn, us, bd in out[:int(os.environ.get("TOP", "30"))]:
    print("%-30s %-44s %8.1f %8.1f %8.1f" % (nm, kn, us, bd, g))
print("TOTAL %.0f us, bounds %.0f us" % (sum(o[3] for o in out), sum(o[4] for o in out)))
