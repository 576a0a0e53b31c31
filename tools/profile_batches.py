#!/usr/bin/env python
"""Re-profiles a model on the B200 engine in the reference's `_summary.csv` schema (profiling/ModelProfiler.py).
Usage: python tools/profile_batches.py kind out.csv [max_batch]
The CSV has one row per integer batch size 1..max_batch (what nexus.py indexes); the bucket sizes are the timed ones."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rdb_b200  # noqa: F401
from rdb_b200 import profiler, synth
from rdb_b200.engine import B200Model


def main():
    kind, out = sys.argv[1], sys.argv[2]
    max_batch = int(sys.argv[3]) if len(sys.argv) > 3 else 256
    module = synth.make_model(kind) if kind == "yolov8n" else synth.make_model(kind, n_proto=8)
    model = B200Model(kind, module, max_batch=max_batch)
    from rdb_b200.engine import default_buckets
    sizes = default_buckets(max_batch)          # the sizes a worker pads batches to; every integer row is emitted
    rows = profiler.profile_batches(model, sizes, dense=True)
    profiler.save_summary_csv(rows, out)
    for r in [r for r in rows if r['batch_size'] in sizes]:
        print("%4d  %8.3f ms  %9.1f img/s  %8.1f MB" % (r['batch_size'], r['avg_latency_ms'], r['throughput'], r['peak_memory_mb']))


if __name__ == "__main__":
    main()
