"""The B200 re-profile keeps the reference's CSV contract (profiling/ModelProfiler.py:355-357, scheduler.py:95-113) so the
Nexus packing policy can consume it unchanged; checked against one of the reference's own summary files when present."""
import os

import rdb_b200  # noqa: F401
from rdb_b200 import profiler

REF_CSV = "/root/reference/profiling/resnet50_20241117_154052_summary.csv"


def test_schema_and_loader_roundtrip(tmp_path):
    rows = [{'batch_size': b, 'status': 'success', 'avg_latency_ms': 0.1 * b, 'std_latency_ms': 0.01, 'throughput': 1e4,
             'throughput_efficiency': 1e4 / b, 'peak_memory_mb': 100.0 + b, 'memory_per_sample_mb': 1.0,
             'memory_utilization': 0.01} for b in (1, 8, 256)]
    rows.append({'batch_size': 512, 'status': 'exceeds_max_batch'})
    p = tmp_path / "resnet50_b200_summary.csv"
    profiler.save_summary_csv(rows, str(p))
    header = open(p).readline().strip().split(",")
    assert header == profiler.FIELDNAMES
    d = profiler.load_csv_to_dict(str(p))
    assert d[8] == {'avg_latency_ms': 0.8, 'peak_memory_mb': 108.0} and d[512] == {}
    assert profiler.load_csv_to_dict(str(tmp_path / "missing.csv")) == {}


def test_reference_summary_file_has_the_same_columns():
    if not os.path.exists(REF_CSV):
        import pytest
        pytest.skip("reference tree not present on this machine")
    assert open(REF_CSV).readline().strip().split(",") == profiler.FIELDNAMES
    d = profiler.load_csv_to_dict(REF_CSV)
    assert abs(d[256]['avg_latency_ms'] - 107.57) < 0.01          # A6000, 107.57 ms at batch 256
