#!/usr/bin/env python
"""Per-op device times of one model at one batch size (eager pass with CUDA events around every op, averaged), with the
algorithmic GB/s and TFLOP/s of each op.  Usage: python tools/profile_ops.py [kind] [batch] [reps]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

import rdb_b200  # noqa: F401
from bench import plan_traffic
from rdb_b200 import synth
from rdb_b200.engine import B200Model


def main():
    kind = sys.argv[1] if len(sys.argv) > 1 else "resnet50"
    batch = int(sys.argv[2]) if len(sys.argv) > 2 else 256
    reps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
    module = synth.make_model(kind) if kind == "yolov8n" else synth.make_model(kind, n_proto=8)
    model = B200Model(kind, module, max_batch=batch)
    imgs = list(synth.synthetic_images(min(batch, 8), model.in_h, model.in_w)) * (batch // min(batch, 8))
    model.infer_u8(imgs[:batch])
    torch.cuda.synchronize()
    ms = np.zeros(model.n_ops)
    for r in range(reps + 1):
        m, _ = model.profile_ops(batch)
        if r:
            ms += m
    ms /= reps
    tr = plan_traffic(model.blob, batch)
    print("%-34s %9s %9s %9s" % ("op", "us", "GB/s", "TFLOP/s"))
    order = range(model.n_ops)
    top = int(os.environ.get("TOP", "0"))
    if top:
        order = sorted(order, key=lambda i: -ms[i])[:top]
        by_type = {}
        for i in range(model.n_ops):
            by_type[tr[i][0]] = by_type.get(tr[i][0], 0.0) + ms[i]
        print("time by op type (ms):", {k: round(v, 3) for k, v in sorted(by_type.items())})
    for i in order:
        t, b, fl = tr[i]
        print("%-34s %9.1f %9.0f %9.1f" % (model.op_names[i], ms[i] * 1e3, b / (ms[i] * 1e-3) / 1e9, fl / (ms[i] * 1e-3) / 1e12))
    tot_b, tot_f = sum(x[1] for x in tr), sum(x[2] for x in tr)
    print("TOTAL %.3f ms  -> %.0f img/s ; %.0f GB/s algorithmic, %.1f TFLOP/s" %
          (ms.sum(), batch / (ms.sum() * 1e-3), tot_b / (ms.sum() * 1e-3) / 1e9, tot_f / (ms.sum() * 1e-3) / 1e12))


if __name__ == "__main__":
    main()
