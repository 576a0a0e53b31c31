#!/usr/bin/env python
"""Per-model roofline inputs derived from the lowered plan (CPU only): FLOPs per image (2 x MAC over every conv / FC /
depthwise op), layer-fused bytes per image (every op reads its inputs once and writes its output once, bf16 activations,
residuals counted) and weight bytes per batch -- the SURVEY section 8(d) recipe applied to the plan the engine runs.
Usage: python tools/model_roofline.py [kind ...]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import plan_traffic, load_peaks, MODELS  # noqa: E402
from rdb_b200 import synth, lowering  # noqa: E402

peaks = load_peaks()
for kind in (sys.argv[1:] or sorted(MODELS)):
    hw = MODELS[kind]["hw"]
    module = synth.make_model(kind) if kind == "yolov8n" else synth.make_model(kind, n_proto=8)
    pb = lowering.lower(kind, module)
    blob = pb.to_blob()
    t1 = plan_traffic(blob, 1)
    t2 = plan_traffic(blob, 2)
    flops = sum(f for _, _, f in t2) - sum(f for _, _, f in t1)
    per_img = sum(b for _, b, _ in t2) - sum(b for _, b, _ in t1)          # the batch-proportional part
    weights = sum(b for _, b, _ in t1) - per_img                             # the batch-independent part
    n_params = sum(p.numel() for p in module.parameters())
    B = MODELS[kind]["batch"]
    t_t = flops / (peaks["tensor_sustained"] * 1e12)
    t_h = (per_img + weights / B) / (peaks["hbm"] * 1e9)
    print("%-20s input %d: %.3f GFLOP/img, %.2f MB/img layer-fused (physical channel padding included), %.1f MB weights; "
          "%d parameters; AI @B=%d %.0f FLOP/B; t_tensor %.2f us, t_hbm %.2f us -> %s-bound, roofline %.0f img/s"
          % (kind, hw, flops / 1e9, per_img / 1e6, weights / 1e6, n_params, B, flops / (per_img + weights / B),
             t_t * 1e6, t_h * 1e6, "tensor" if t_t > t_h else "HBM", 1.0 / max(t_t, t_h)))
