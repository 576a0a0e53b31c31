#!/usr/bin/env python
"""Serving simulation of the reference's actor loop on one GPU: a feeder thread submits uint8 requests to a
B200RequestQueue at a given rate (constant, or the reference driver's slope pattern rate = elapsed * slope,
293-project/src/test_scheduler.py:77-79); a B200GPUWorker pulls batches (get_batch -> process_batch) back to back.
Reports batch latency percentiles exactly as RequestQueue.get_stats does (scheduler.py:343-372) against the model's SLO.

  python tools/serve_sim.py --model resnet --rate 20000 --seconds 4 --batch 256 [--slope 0]
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import rdb_b200  # noqa: F401
from rdb_b200 import synth
from rdb_b200.engine import B200Model
from rdb_b200.request_queue import B200RequestQueue, models_config
from rdb_b200.worker import B200GPUWorker, MODEL_KINDS


class Session:
    def __init__(self, model_name, batch_size):
        self.model_name, self.batch_size = model_name, batch_size


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="resnet")
    ap.add_argument("--rate", type=float, default=20000.0)
    ap.add_argument("--slope", type=float, default=0.0)
    ap.add_argument("--seconds", type=float, default=4.0)
    ap.add_argument("--batch", type=int, default=256)
    args = ap.parse_args()
    kind = MODEL_KINDS[args.model]
    module = synth.make_model(kind) if kind == "yolov8n" else synth.make_model(kind, n_proto=8)
    model = B200Model(kind, module, max_batch=2 * args.batch)
    h, w = model.in_h, model.in_w
    pool = list(synth.synthetic_images(32, h, w))
    # batch profile (avg_latency_ms per batch size) measured on this engine, like profiling/ModelProfiler does offline
    profile = {}
    for bs in sorted({1, 8, 32, 64, 128, args.batch, 2 * args.batch}):
        imgs = [pool[i % 32] for i in range(bs)]
        for _ in range(3):
            model.infer_u8(imgs)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(5):
            model.infer_u8(imgs).cpu()
        profile[bs] = {"avg_latency_ms": (time.perf_counter() - t0) / 5 * 1e3}
    slab = 4000 * ((h * w * 3 + 15) // 16 * 16)
    q = B200RequestQueue(args.model, max_size=2000, batch_profile={args.model: profile}, compat=False,
                         pinned_slab_bytes=slab)
    worker = B200GPUWorker("B200_0", 0, [(Session(args.model, args.batch), 1.0)], 1.0, {args.model: model},
                           max_batch=2 * args.batch)
    stop = threading.Event()
    submitted = [0, 0]

    def feeder():
        t0 = time.perf_counter()
        i = 0
        while not stop.is_set():
            el = time.perf_counter() - t0
            if el >= args.seconds:
                break
            target = args.slope * el * el / 2 if args.slope > 0 else args.rate * el
            while i < target:
                ok = q.add_request("r%d" % i, pool[i % 32])
                submitted[0 if ok else 1] += 1
                i += 1
            time.sleep(0.0005)

    th = threading.Thread(target=feeder, daemon=True)
    t_start = time.perf_counter()
    th.start()
    batches = served = 0
    sizes = []
    while th.is_alive() or not q.empty():
        batch, discard = q.get_batch(args.batch)
        if batch is None:
            time.sleep(0.0002)
            continue
        worker.process_batch(batch, {args.model: q}, discard)
        batches += 1
        served += batch.batch_size
        sizes.append(batch.batch_size)
    wall = time.perf_counter() - t_start
    stats = q.get_stats()
    out = {"model": args.model, "kind": kind, "pattern": "slope %.0f/s^2" % args.slope if args.slope else "constant",
           "offered_rate": args.rate if not args.slope else None, "seconds": args.seconds, "batch_target": args.batch,
           "submitted": submitted[0], "dropped_full_queue": submitted[1], "served": served, "batches": batches,
           "mean_batch": sum(sizes) / max(1, len(sizes)), "throughput_img_s": served / wall,
           "slo_ms": models_config[args.model]["SLO"], "slo_violations": stats["slo_violations"],
           "avg_latency_ms": stats["avg_latency"], "p95_latency_ms": stats["p95_latency"],
           "p99_latency_ms": stats["p99_latency"],
           "profile_ms": {k: round(v["avg_latency_ms"], 3) for k, v in profile.items()}}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
