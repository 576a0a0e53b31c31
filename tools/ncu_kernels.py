#!/usr/bin/env python
"""Aggregates an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name: count, total us, mean us."""
import collections, csv, re, sys
rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
hdr = rows[hi]
ik, iv = hdr.index("Kernel Name"), hdr.index("Metric Value")
agg = collections.OrderedDict()
tot = 0.0
for r in rows[hi + 1:]:
    if len(r) <= iv:
        continue
    name = re.sub(r"\(.*", "", r[ik])[:64]
    v = float(r[iv].replace(",", "")) / 1e3
    tot += v
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += v
print("total %.1f us over %d launches" % (tot, sum(a[0] for a in agg.values())))
for k, (c, v) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%-66s x%-4d %9.1f us  %7.1f us each" % (k, c, v, v / c))
