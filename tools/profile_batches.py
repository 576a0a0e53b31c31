#!/usr/bin/env python
"""Re-profiles a model on the B200 engine in the reference's `_summary.csv` schema (profiling/ModelProfiler.py).
Usage: python tools/profile_batches.py kind out.csv [max_batch]"""
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
    sizes = [b for b in (1, 2, 4, 8, 16, 32, 64, 96, 128, 192, 256, 384, 512) if b <= max_batch]
    rows = profiler.profile_batches(model, sizes)
    profiler.save_summary_csv(rows, out)
    for r in rows:
        print("%4d  %8.3f ms  %9.1f img/s  %8.1f MB" % (r['batch_size'], r['avg_latency_ms'], r['throughput'], r['peak_memory_mb']))


if __name__ == "__main__":
    main()
