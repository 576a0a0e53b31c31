#!/usr/bin/env python
"""Device-resident throughput of one model at one batch size through the CUDA-graph path (collate + forward per step).
Usage: python tools/model_throughput.py kind batch [steps]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

import rdb_b200  # noqa: F401
from rdb_b200 import synth
from rdb_b200.engine import B200Model


def main():
    kind, batch = sys.argv[1], int(sys.argv[2])
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 20
    module = synth.make_model(kind) if kind == "yolov8n" else synth.make_model(kind, n_proto=8)
    model = B200Model(kind, module, max_batch=batch)
    h, w = model.in_h, model.in_w
    per = h * w * 3
    dev = model.device
    sets = [torch.randint(0, 256, (batch * per,), dtype=torch.uint8).to(dev) for _ in range(3)]
    offsets = np.arange(batch, dtype=np.uint64) * per
    hw = np.tile(np.array([[h, w]], dtype=np.int32), (batch, 1))
    out = torch.empty(batch, model.out_dim, device=dev)

    def step(i):
        model.collate_u8(sets[i % 3], offsets, hw, batch, batch, 0, batch * per)
        model.infer_resident(batch, out)

    for i in range(4):
        step(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        step(i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    print("%s batch %d: %.3f ms/step -> %.0f img/s (%d ops, CUDA graph)" % (kind, batch, ms, batch / (ms * 1e-3), model.n_ops))


if __name__ == "__main__":
    main()
