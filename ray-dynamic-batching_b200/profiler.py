"""Batch-size profiler for a B200 engine, in the reference's own CSV schema.

The Nexus packing policy of the reference (293-project/src/nexus.py) consumes, per model, a dict
batch_size -> {'avg_latency_ms', 'peak_memory_mb'} loaded from the `_summary.csv` files that
profiling/ModelProfiler.py writes (columns at ModelProfiler.py:355-357; loader at scheduler.py:95-113).  This module
measures the same quantities on the B200 engine — forward pass from device-resident inputs under CUDA events, warm-up
runs then timed runs, exactly like ModelProfiler._measure_latency (:92-112) — so a maintainer can re-profile and keep
nexus.py unchanged.
"""
import csv
import logging
import statistics
from typing import Dict, List

import torch

from . import lowering as _lowering

FIELDNAMES = ['batch_size', 'status', 'avg_latency_ms', 'std_latency_ms', 'throughput', 'throughput_efficiency',
              'peak_memory_mb', 'memory_per_sample_mb', 'memory_utilization']


def engine_memory_mb(model, batch_size: int) -> float:
    """Device memory an engine created for `batch_size` holds: weights, the activation arena (buffers with disjoint
    lifetimes share addresses), tables -- b200_plan_memory, the same planner b200_engine_create runs.  This is the column
    nexus.py sums to decide which sessions fit on a GPU together (nexus.py:222-226), so it has to be what an actor sized
    for that batch really holds, not a per-image estimate."""
    from .engine import plan_memory_bytes
    return plan_memory_bytes(model.blob, batch_size) / (1024.0 * 1024.0)


def _time_batch(model, bs, out, warmup_runs, num_runs):
    dev = model.device
    for _ in range(warmup_runs + 2):            # +2: the first two calls of a batch size are eager / graph capture
        model.infer_resident(bs, out)
    torch.cuda.synchronize(dev)
    lat = []
    for _ in range(num_runs):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        model.infer_resident(bs, out)
        e1.record()
        e1.synchronize()
        lat.append(e0.elapsed_time(e1))
    return lat


def _time_batch_end_to_end(model, bs, warmup_runs, num_runs):
    """What an actor spends on a batch it serves alone: uint8 requests in page-locked host memory -> H2D -> collate ->
    forward -> D2H of the outputs (detections for a model with device NMS), wall clock, one batch at a time.  The
    reference's profile is forward-only from device-resident inputs (ModelProfiler._measure_latency), and so is the default
    here; a policy that fills a node to the forward-only throughput promises more than the actor delivers, which
    is what this variant is for.  Two batches in flight overlap the copies with the compute, so an actor does better than
    1 / this latency; the figure is the conservative one."""
    import time
    import numpy as np
    per = model.in_h * model.in_w * 3
    host = torch.randint(0, 256, (bs * per,), dtype=torch.uint8).pin_memory()
    offsets = np.arange(bs, dtype=np.uint64) * per
    hw = np.tile(np.array([[model.in_h, model.in_w]], dtype=np.int32), (bs, 1))
    for _ in range(warmup_runs + 2):
        model.infer_packed_async(host, offsets, hw, bs, bs, 0, bs * per).result()
    lat = []
    for _ in range(num_runs):
        t0 = time.perf_counter()
        model.infer_packed_async(host, offsets, hw, bs, bs, 0, bs * per).result()
        lat.append((time.perf_counter() - t0) * 1e3)
    return lat


def profile_batches(model, batch_sizes: List[int], warmup_runs: int = 3, num_runs: int = 10, dense: bool = False,
                    end_to_end: bool = False) -> List[dict]:
    """One row per batch size (ModelProfiler.profile_batch_size, :114-178).

    dense=True emits EVERY integer batch size from 1 to max(batch_sizes), like the reference's sweep
    (profiling/run_profiler.py:188-197, batch_size_step=1): nexus.py indexes the profile by arbitrary integers
    (`profile[ceil(duty * rate / 1000)]`, nexus.py:213-214; `profile[batch_size]`, scheduler.py:281), so a sparse table
    raises KeyError there.  Only the sizes in `batch_sizes` are timed -- they are the buckets a worker pads batches to --
    and a size in between gets the latency of the bucket it would run in; memory is computed per size."""
    dev = model.device
    total_mem_mb = torch.cuda.get_device_properties(dev).total_memory / (1024.0 * 1024.0)
    timed = {}
    for bs in sorted(set(batch_sizes)):
        if bs > model.max_batch:
            continue
        if end_to_end:
            timed[bs] = _time_batch_end_to_end(model, bs, warmup_runs, num_runs)
            continue
        out = torch.empty(bs, model.out_dim, device=dev, dtype=torch.float32)
        timed[bs] = _time_batch(model, bs, out, warmup_runs, num_runs)
    rows = []
    keys = sorted(timed)
    sizes = range(1, keys[-1] + 1) if dense and keys else sorted(set(batch_sizes))
    for bs in sizes:
        if bs > model.max_batch:
            rows.append({'batch_size': bs, 'status': 'exceeds_max_batch'})
            continue
        bucket = next(k for k in keys if k >= bs)
        lat = timed[bucket]
        avg = sum(lat) / len(lat)
        mem = engine_memory_mb(model, bs)
        thr = bs / (avg / 1000.0)
        rows.append({'batch_size': bs, 'status': 'success', 'avg_latency_ms': avg,
                     'std_latency_ms': statistics.pstdev(lat), 'throughput': thr, 'throughput_efficiency': thr / bs,
                     'peak_memory_mb': mem, 'memory_per_sample_mb': mem / bs, 'memory_utilization': mem / total_mem_mb})
    return rows


def save_summary_csv(rows: List[dict], path: str) -> None:
    with open(path, 'w', newline='') as f:
        w = csv.DictWriter(f, fieldnames=FIELDNAMES)
        w.writeheader()
        for r in rows:
            w.writerow({k: r[k] for k in FIELDNAMES if k in r})


def load_csv_to_dict(file_path: str) -> Dict[int, Dict[str, float]]:
    """Mirror of BatchProfiler.load_csv_to_dict (scheduler.py:95-113): batch_size -> {avg_latency_ms, peak_memory_mb}."""
    column_list = ['avg_latency_ms', 'peak_memory_mb']
    result = {}
    try:
        with open(file_path, 'r') as csvfile:
            reader = csv.DictReader(csvfile)
            headers = reader.fieldnames
            key_column = headers[0] if headers else None
            for row in reader:
                key = int(row[key_column])
                result[key] = {h: float(row[h]) for h in headers[1:] if h in column_list and row[h] != ''}
    except (FileNotFoundError, ValueError) as e:
        logging.error("Error loading batch profile from %s: %s", file_path, e)
        return {}
    return result
