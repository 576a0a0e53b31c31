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


B200_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "b200_batch_profiles")
B200_FILES = {"resnet": "resnet50_b200_summary.csv", "shufflenet": "shufflenet_b200_summary.csv",
              "efficientnet": "efficientnetv2_b200_summary.csv"}


def test_b200_profiles_have_every_integer_batch_size():
    """nexus.py indexes a profile with arbitrary integers (ceil(duty * rate / 1000), nexus.py:213-214; profile[batch_size],
    scheduler.py:281): the committed B200 tables must have one row per batch size, with memory growing with the batch."""
    for model, f in B200_FILES.items():
        d = profiler.load_csv_to_dict(os.path.join(B200_DIR, f))
        top = max(d)
        assert top >= 256 and sorted(d) == list(range(1, top + 1)), model
        lat = [d[b]['avg_latency_ms'] for b in range(1, top + 1)]
        mem = [d[b]['peak_memory_mb'] for b in range(1, top + 1)]
        assert all(x > 0 for x in lat) and all(b >= a for a, b in zip(mem, mem[1:])), model


def test_packing_policy_consumes_the_b200_profiles():
    from rdb_b200 import packing
    prof = {m: profiler.load_csv_to_dict(os.path.join(B200_DIR, f)) for m, f in B200_FILES.items()}
    for rates in ((300.0, 500.0, 20.0), (26000.0, 90000.0, 7000.0), (140000.0, 15.0, 2.5), (7.0, 3.0, 1.0)):
        sessions = [packing.session("resnet", 2000 / 2.2, rates[0]), packing.session("shufflenet", 1500 / 2.2, rates[1]),
                    packing.session("efficientnet", 40 / 2.2, rates[2])]
        nodes = packing.nexus(prof).squishyBinPacking(sessions)
        assert nodes and all(0 < n.get_occupancy() <= 1.0 + 1e-9 for n in nodes)


def test_reference_nexus_consumes_the_b200_profiles():
    """The reference's own policy module, unmodified, on the B200 tables: load_csv_to_dict, squishyBinPacking and
    mergeNodes (nexus.py:129-133, 204-234) run without a KeyError at arbitrary rates."""
    ref = "/root/reference/293-project/src/nexus.py"
    if not os.path.exists(ref):
        import pytest
        pytest.skip("reference tree not present on this machine")
    import importlib.util
    spec = importlib.util.spec_from_file_location("ref_nexus_b200", ref)
    nx = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(nx)
    prof = {m: nx.load_csv_to_dict(os.path.join(B200_DIR, f)) for m, f in B200_FILES.items()}
    nx.gpu_mem = 160
    pol = nx.nexus(prof)
    for rates in ((300.0, 500.0, 20.0), (26000.0, 90000.0, 7000.0), (140000.0, 15.0, 2.5), (3333.3, 77.7, 911.0)):
        sessions = [nx.session("resnet", 2000 / 2.2, rates[0]), nx.session("shufflenet", 1500 / 2.2, rates[1]),
                    nx.session("efficientnet", 40 / 2.2, rates[2])]
        nodes = pol.squishyBinPacking(sessions)
        assert nodes
        for a in nodes:
            for b in nodes:
                pol.mergeNodes(a, b)               # every pair: indexes profile[ceil(duty * rate / 1000)]


def test_end_to_end_profiles_are_the_conservative_ones():
    """profiles/b200_batch_profiles_e2e/ (tools/profile_batches.py --e2e): same schema and the same integer rows as the
    forward-only profiles, the memory column of the same planner, and a latency that is never below the forward-only one -- so a policy
    that packs with them (nexus.py:129-133) promises at most what the forward-only profile promises."""
    import os
    from rdb_b200 import profiler
    root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles")
    for name in ("resnet50", "shufflenet", "efficientnetv2", "yolov8n"):
        fwd = profiler.load_csv_to_dict(os.path.join(root, "b200_batch_profiles", "%s_b200_summary.csv" % name))
        e2e = profiler.load_csv_to_dict(os.path.join(root, "b200_batch_profiles_e2e", "%s_b200_e2e_summary.csv" % name))
        assert e2e and set(e2e) == set(range(1, max(e2e) + 1))
        common = sorted(set(fwd) & set(e2e))
        assert len(common) >= 128
        for b in common:
            assert e2e[b]['avg_latency_ms'] >= fwd[b]['avg_latency_ms'] * 0.98, (name, b)
            # the same planner; the padded conv weights added a few MB to the engines profiled later
            assert abs(e2e[b]['peak_memory_mb'] - fwd[b]['peak_memory_mb']) <= 0.1 * fwd[b]['peak_memory_mb']
        top = max(common)
        assert top / e2e[top]['avg_latency_ms'] < top / fwd[top]['avg_latency_ms']
