#!/usr/bin/env python
"""Three eager (no CUDA graph) forwards of one model at one batch size: the command ncu captures kernels from.
Usage: python tools/eager_forward.py kind batch"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import rdb_b200
from rdb_b200 import synth
from rdb_b200.engine import B200Model
kind, batch = sys.argv[1], int(sys.argv[2])
module = synth.make_model(kind) if kind == "yolov8n" else synth.make_model(kind, n_proto=8)
m = B200Model(kind, module, max_batch=batch, use_graphs=False)
per = m.in_h * m.in_w * 3
src = torch.randint(0, 256, (batch * per,), dtype=torch.uint8).to(m.device)
offsets = np.arange(batch, dtype=np.uint64) * per
hw = np.tile(np.array([[m.in_h, m.in_w]], dtype=np.int32), (batch, 1))
out = torch.empty(batch, m.out_dim, device=m.device)
for i in range(3):
    m.collate_u8(src, offsets, hw, batch, batch, 0, batch * per)
    m.infer_resident(batch, out)
torch.cuda.synchronize()
print("done", float(out.abs().sum()))
