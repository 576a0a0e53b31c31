"""The scheduling policy restatement (rdb_b200.packing) against the reference's own module: every node of every schedule
-- duty cycle, and per session model / rate / batch size / occupancy -- must be identical (exact float equality: the
arithmetic is the same expression by expression).  Golden cases were produced by running 293-project/src/nexus.py
(tests/golden/make_packing_golden.py); where /root/reference exists the two are also run side by side on the reference's
real profile CSVs."""
import importlib.util
import json
import os
import sys

import pytest

import rdb_b200  # noqa: F401
from rdb_b200 import packing

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
from make_packing_golden import describe, synthetic_profile  # noqa: E402

GOLD = json.load(open(os.path.join(HERE, "golden", "packing_golden.json")))
REF = "/root/reference/293-project/src/nexus.py"


@pytest.mark.parametrize("idx", range(len(GOLD["cases"])))
def test_schedule_matches_reference_golden(idx):
    case = GOLD["cases"][idx]
    pol = packing.nexus(synthetic_profile(), gpu_mem=case["gpu_mem"], merge_gpu_mem=11)
    nodes = pol.squishyBinPacking([packing.session(*s) for s in case["sessions"]])
    got = json.loads(json.dumps(describe(nodes)))
    assert got == case["nodes"]


def test_session_validation_like_the_reference():
    for bad in (("", 10, 1), ("m", 0, 1), ("m", 10, -1)):
        with pytest.raises(ValueError):
            packing.session(*bad)
    with pytest.raises(ValueError):
        packing.session("m", 10, 1, batch_size=0)
    assert packing.session("m", 10, 1).batch_size == 0


def test_zero_residual_rate_is_skipped():
    prof = synthetic_profile()
    pol = packing.nexus(prof)
    lat = prof["resnet"][512]["avg_latency_ms"]
    thr = 512 / lat * 1000
    nodes = pol.squishyBinPacking([packing.session("resnet", 2 * lat + 1e-9, 2 * thr)])
    assert len(nodes) == 2 and all(n.node_sessions[0][0].batch_size == 512 for n in nodes)


@pytest.mark.skipif(not os.path.exists(REF), reason="reference tree not present (GPU box)")
def test_side_by_side_on_the_reference_profiles():
    spec = importlib.util.spec_from_file_location("ref_nexus", REF)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    pdir = "/root/reference/293-project/profiling"
    files = {"resnet": "resnet50_20241117_154052_summary.csv", "shufflenet": "shufflenet_20241123_104115_summary.csv",
             "efficientnet": "efficientnetv2_20241123_125206_summary.csv"}
    prof = {m: ref.load_csv_to_dict(os.path.join(pdir, f)) for m, f in files.items()}
    for rates in ((300.0, 500.0, 20.0), (2600.0, 9000.0, 700.0), (40.0, 15000.0, 2.5)):
        sessions = [("resnet", 2000 / 2.2, rates[0]), ("shufflenet", 1500 / 2.2, rates[1]), ("efficientnet", 400.0, rates[2])]
        want = describe(ref.nexus(prof).squishyBinPacking([ref.session(*s) for s in sessions]))
        got = describe(packing.nexus(prof, gpu_mem=11).squishyBinPacking([packing.session(*s) for s in sessions]))
        assert got == want
