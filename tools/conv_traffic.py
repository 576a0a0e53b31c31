#!/usr/bin/env python
"""DRAM bytes of the convolution family per launch, from an ncu capture of one eager forward, for bench.py's
`roofline.traffic`.

On the GPU box (one ncu run, after the same command has exited 0 without ncu):

    python tools/eager_forward.py resnet50 256
    ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
        --log-file gpurun_out/traffic_resnet50.csv python tools/eager_forward.py resnet50 256
    python tools/conv_traffic.py gpurun_out/traffic_resnet50.csv resnet50 256 [commit [output directory]]

writes profiles/r02_traffic_<kind>_b<batch>.json (or into the given directory: only gpurun_out/ travels back from the box) with the hash of the sources the library was built from
(rdb_b200.build.source_stamp()): bench.py reports the figure only while that hash matches the running build.
tools/eager_forward.py runs three forwards; the last one is used (weights and code warm, activations larger than L2)."""
import csv
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

FAMILY = re.compile(r"conv_gemm_persistent_kernel|conv_gemm_2cta_kernel|conv_gemm_kernel|conv3x3_halo_kernel|"
                    r"gather_conv_kernel|stem_pool_kernel|stem_pool_reg_kernel|stem_row_kernel|stem_tc_kernel|stem_conv_kernel")
FORWARDS = 3      # tools/eager_forward.py


def to_bytes(value, unit):
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]
    return float(value.replace(",", "")) * scale


def parse(path):
    """-> list of (kernel name, read bytes, write bytes) in launch order."""
    lines = [ln for ln in open(path) if ln.startswith('"')]
    per_id = {}
    order = []
    for row in csv.DictReader(lines):
        kid = int(row["ID"])
        if kid not in per_id:
            per_id[kid] = [row["Kernel Name"], 0.0, 0.0]
            order.append(kid)
        if row["Metric Name"] == "dram__bytes_read.sum":      # other metrics (durations) may share the file
            per_id[kid][1] = to_bytes(row["Metric Value"], row["Metric Unit"])
        elif row["Metric Name"] == "dram__bytes_write.sum":
            per_id[kid][2] = to_bytes(row["Metric Value"], row["Metric Unit"])
    return [tuple(per_id[k]) for k in order]


def main():
    path, kind, batch = sys.argv[1], sys.argv[2], int(sys.argv[3])
    commit = sys.argv[4] if len(sys.argv) > 4 else os.environ.get("B200_COMMIT")
    launches = [x for x in parse(path) if FAMILY.search(x[0])]
    if not launches or len(launches) % FORWARDS:
        raise SystemExit("%d convolution launches in %s: not a multiple of %d forwards" % (len(launches), path, FORWARDS))
    n = len(launches) // FORWARDS
    last = launches[-n:]
    rd, wr = sum(x[1] for x in last), sum(x[2] for x in last)
    import rdb_b200  # noqa: F401
    from rdb_b200 import build
    out = {
        "kind": kind, "batch": batch, "launches": n, "commit": commit, "source_stamp": build.source_stamp(),
        "traffic_bytes_per_launch": (rd + wr) / n, "dram_read_bytes": rd, "dram_write_bytes": wr,
        "how": "ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none over tools/eager_forward.py "
               "(third forward; per-launch replay, so a layer's output is flushed before the next layer reads it)",
        "per_kernel": {},
    }
    for name, r, w in last:
        short = re.sub(r"^void (b200::)?", "", name).split("(")[0]
        e = out["per_kernel"].setdefault(short, {"launches": 0, "read": 0.0, "write": 0.0})
        e["launches"] += 1
        e["read"] += r
        e["write"] += w
    out_dir = sys.argv[5] if len(sys.argv) > 5 else os.path.join(ROOT, "profiles")
    dst = os.path.join(out_dir, "r02_traffic_%s_b%d.json" % (kind, batch))
    with open(dst, "w") as f:
        json.dump(out, f, indent=1)
    print("%s: %d launches, %.1f MB read + %.1f MB written = %.1f MB per launch" % (dst, n, rd / 1e6, wr / 1e6,
                                                                                  (rd + wr) / n / 1e6))


if __name__ == "__main__":
    main()
