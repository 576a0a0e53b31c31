#!/usr/bin/env python
"""Re-profiles a model on the B200 engine in the reference's `_summary.csv` schema (profiling/ModelProfiler.py).
Usage: python tools/profile_batches.py kind out.csv [max_batch] [--e2e]
--e2e: latency of a batch served alone from page-locked host memory (H2D + collate + forward + D2H; detections for
YOLOv8n) instead of the reference's forward-only figure: the conservative profile for the packing policy.
The CSV has one row per integer batch size 1..max_batch (what nexus.py indexes); the bucket sizes are the timed ones."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rdb_b200  # noqa: F401
from rdb_b200 import profiler, synth
from rdb_b200.engine import B200Model


def main():
    e2e = "--e2e" in sys.argv
    args = [a for a in sys.argv[1:] if a != "--e2e"]
    kind, out = args[0], args[1]
    max_batch = int(args[2]) if len(args) > 2 else 256
    module = synth.make_model(kind) if kind == "yolov8n" else synth.make_model(kind, n_proto=8)
    model = B200Model(kind, module, max_batch=max_batch)
    from rdb_b200.engine import default_buckets
    sizes = default_buckets(max_batch)          # the sizes a worker pads batches to; every integer row is emitted
    if e2e and kind == "yolov8n":
        from rdb_b200.worker import YOLO_NMS
        model.nms = dict(YOLO_NMS)              # what a detection actor returns
    rows = profiler.profile_batches(model, sizes, dense=True, end_to_end=e2e)
    profiler.save_summary_csv(rows, out)
    for r in [r for r in rows if r['batch_size'] in sizes]:
        print("%4d  %8.3f ms  %9.1f img/s  %8.1f MB" % (r['batch_size'], r['avg_latency_ms'], r['throughput'], r['peak_memory_mb']))


if __name__ == "__main__":
    main()
