#!/usr/bin/env python
"""Per-kernel table of ONE step from an ncu launch list: launches, time, algorithmic GB/s against the measured HBM peak
and TFLOP/s against the sustained bf16 peak (the figures north_star asks for, per kernel instead of per op).

  python tools/launch_table.py kind batch launches.csv [hw] > profiles/r02_kernels_<kind>.md

The launch list is `ncu --metrics gpu__time_duration.sum --clock-control none --csv` of bench.py or of
tools/eager_forward.py (its per-launch times are cold-cache and serialised: shares and ratios are what to read, not the
absolute step time).  The LAST run of collate +
n_ops kernels in the list is taken as the step and joined, in order, with the plan's ops (bench.plan_traffic)."""
import collections
import csv
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import plan_traffic, load_peaks, make_module  # noqa: E402
from rdb_b200 import lowering  # noqa: E402


def main():
    kind, batch, path = sys.argv[1], int(sys.argv[2]), sys.argv[3]
    hw = int(sys.argv[4]) if len(sys.argv) > 4 else (640 if kind == "yolov8n" else 224)
    module = make_module(kind, hw)
    pb = lowering.lower(kind, module, in_hw=(hw, hw))
    blob, names = pb.to_blob(), pb.names
    peaks = load_peaks()
    tr = plan_traffic(blob, batch)
    rows = list(csv.reader(open(path)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    hdr = rows[hi]
    ik, iv, im = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Name")
    launches = [(re.sub(r"\(.*", "", r[ik]).replace("void ", "").replace("b200::", ""), float(r[iv].replace(",", "")) / 1e3)
                for r in rows[hi + 1:] if len(r) > iv and r[im] == "gpu__time_duration.sum"]      # other metrics may share the file
    # the last "collate + n_ops kernels" block
    idx = [i for i, l in enumerate(launches) if "collate" in l[0] and i + len(names) < len(launches) + 0]
    start = next(i for i in reversed(idx) if i + 1 + len(names) <= len(launches))
    step = launches[start:start + 1 + len(names)]
    coll, ops = step[0], step[1:]
    per = 3 * hw * hw
    agg = collections.OrderedDict()
    agg[coll[0]] = [1, coll[1], batch * (per + hw * hw * 8.0), 0.0, 0.0]
    for (kn, us), nm, (t, b, fl) in zip(ops, names, tr):
        bound = max(b / (peaks["hbm"] * 1e9), fl / (peaks["tensor_sustained"] * 1e12)) * 1e6
        a = agg.setdefault(kn, [0, 0.0, 0.0, 0.0, 0.0])
        a[0] += 1; a[1] += us; a[2] += b; a[3] += fl; a[4] += bound
    total = sum(a[1] for a in agg.values())
    print("| kernel | launches | us | share | algorithmic GB/s | %% of %.0f GB/s | TFLOP/s | %% of %.0f TF/s | bound / measured |"
          % (peaks["hbm"], peaks["tensor_sustained"]))
    print("|---|---|---|---|---|---|---|---|---|")
    for kn, (c, us, b, fl, bd) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        gbs = b / us / 1e3
        tf = fl / us / 1e6
        print("| `%s` | %d | %.1f | %.1f %% | %.0f | %.0f %% | %s | %s | %s |"
              % (kn, c, us, 100 * us / total, gbs, 100 * gbs / peaks["hbm"], ("%.0f" % tf) if fl else "-",
                 ("%.0f %%" % (100 * tf / peaks["tensor_sustained"])) if fl else "-", ("%.2f" % (bd / us)) if bd else "-"))
    bd_all = sum(a[4] for a in agg.values())
    print("\nstep under ncu: %.0f us over %d launches (%s batch %d, input %d); sum of per-op bounds %.0f us = %.2f of it"
          % (total, len(step), kind, batch, hw, bd_all, bd_all / total))


if __name__ == "__main__":
    main()
