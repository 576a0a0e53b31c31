"""The measurement chain around bench.py, without a GPU: the ncu CSV parsers (tools/conv_traffic.py, tools/launch_table.py),
the per-op byte / FLOP accounting (bench.plan_traffic) and the rule that `roofline.traffic` only ever comes from a capture of
the sources that are running."""
import csv
import importlib.util
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import rdb_b200  # noqa: F401,E402
from rdb_b200 import build, lowering, synth  # noqa: E402


def _load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


HEADER = ["ID", "Process ID", "Process Name", "Host Name", "Kernel Name", "Context", "Stream", "Block Size", "Grid Size",
          "Device", "CC", "Section Name", "Metric Name", "Metric Unit", "Metric Value"]


def _write_capture(path, forwards):
    """An ncu --csv log of `forwards` eager forwards of a three-kernel model, three metrics per launch."""
    kernels = ["void b200::collate_u8_flat_kernel(CollateParams)",
               "void b200::stem_pool_kernel<3, 32, 1>(CUtensorMap_st, StemPoolParams)",
               "void b200::conv_gemm_persistent_kernel<64, 0, 1, 2>(b200::ConvMaps, b200::ConvParams, int)",
               "void b200::global_avgpool_kernel(const __nv_bfloat16 *, float *)"]
    with open(path, "w") as f:
        f.write("==PROF== Connected to process 1\n")
        w = csv.writer(f, quoting=csv.QUOTE_ALL)
        w.writerow(HEADER)
        kid = 0
        for fw in range(forwards):
            for k in kernels:
                for name, unit, val in (("gpu__time_duration.sum", "ns", "12,345"), ("dram__bytes_read.sum", "Mbyte", "1.5"),
                                        ("dram__bytes_write.sum", "Kbyte", "%d" % (500 + fw))):
                    w.writerow([kid, 1, "python", "box", k, 1, 7, "(384, 1, 1)", "(148, 1, 1)", 0, "10.0", "Command line profiler metrics",
                                name, unit, val])
                kid += 1


def test_conv_traffic_takes_the_last_forward_of_the_conv_family(tmp_path):
    ct = _load(os.path.join(ROOT, "tools", "conv_traffic.py"), "conv_traffic")
    cap = tmp_path / "cap.csv"
    _write_capture(cap, 3)
    launches = ct.parse(str(cap))
    assert len(launches) == 12 and launches[1][1:] == (1.5e6, 500e3)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "conv_traffic.py"), str(cap), "toy", "7", "abc1234", str(tmp_path)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    out = json.load(open(tmp_path / "r02_traffic_toy_b7.json"))
    assert out["launches"] == 2 and out["commit"] == "abc1234" and out["source_stamp"] == build.source_stamp()
    # stem + conv of the THIRD forward (write = 502 KB each), collate and pooling are not convolutions
    assert out["traffic_bytes_per_launch"] == pytest.approx(1.5e6 + 502e3)
    assert set(out["per_kernel"]) == {"stem_pool_kernel<3, 32, 1>", "conv_gemm_persistent_kernel<64, 0, 1, 2>"}


def test_conv_traffic_refuses_a_partial_capture(tmp_path):
    cap = tmp_path / "cap.csv"
    _write_capture(cap, 3)
    lines = open(cap).read().splitlines()
    open(cap, "w").write("\n".join(lines[:-6]) + "\n")      # the last two launches are missing
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "conv_traffic.py"), str(cap), "toy", "7", "x", str(tmp_path)],
                       capture_output=True, text=True)
    assert r.returncode != 0 and "not a multiple" in (r.stderr + r.stdout)


def test_plan_traffic_counts_the_shuffle_tail_once():
    """bench.plan_traffic: a ShuffleNetV2 unit tail moves its input, the other branch and the unit's output once each; the
    whole model's algorithmic bytes drop by the shuffle passes' traffic when the tail is fused, the FLOPs stay."""
    import bench
    m = synth.make_model("shufflenet_v2_x1_0", n_proto=8)
    fused = lowering.lower("shufflenet_v2_x1_0", m).to_blob()
    os.environ["B200_FUSE_SHUFFLE"] = "0"
    try:
        plain = lowering.lower("shufflenet_v2_x1_0", m).to_blob()
    finally:
        del os.environ["B200_FUSE_SHUFFLE"]
    tf, tp = bench.plan_traffic(fused, 512), bench.plan_traffic(plain, 512)
    assert len(tp) - len(tf) == 16
    assert sum(x[2] for x in tf) == pytest.approx(sum(x[2] for x in tp), rel=1e-9)
    _, bufs, ops, _ = lowering.parse_blob(fused)
    tails = [(op, t) for op, t in zip(ops, tf) if op["type"] == lowering.OP_CONV and op["p"][18] > 0]
    assert len(tails) == 16
    for op, (_, b, fl) in tails:
        p = op["p"]
        px = 512 * bufs[p[3]][0] * bufs[p[3]][1]
        assert b == px * (p[2] + p[18] + p[5]) * 2 + (p[5] // 2) * p[2] * 2 + (p[5] // 2) * 4
        assert fl == 2.0 * px * (p[5] // 2) * p[2]
    shuffle_bytes = sum(b for (t, b, _) in tp if t == lowering.OP_SHUFFLE)
    saved = sum(x[1] for x in tp) - sum(x[1] for x in tf)
    assert 0.45 * shuffle_bytes < saved < 0.8 * shuffle_bytes      # the pass read both branches and wrote the output again


def test_traffic_file_counts_only_for_the_running_sources(tmp_path, monkeypatch):
    """The rule bench.py applies (bench.py, `roofline.traffic`): a capture is reported only while its source hash is the one
    of the library's sources; profiles/ holds captures of HEAD for the four benchmarked configurations."""
    stamp = build.source_stamp()
    for kind, batch in (("resnet50", 256), ("shufflenet_v2_x1_0", 512), ("efficientnet_v2_s", 128), ("yolov8n", 32)):
        path = os.path.join(ROOT, "profiles", "r02_traffic_%s_b%d.json" % (kind, batch))
        assert os.path.exists(path), path
        tj = json.load(open(path))
        assert tj["traffic_bytes_per_launch"] > 0 and tj["launches"] > 0 and tj["commit"]
        assert tj["source_stamp"] == stamp, "%s was captured on other sources: recapture it (tools/conv_traffic.py)" % path
    # any change to a kernel source changes the hash (checked on a copy of the sources)
    import shutil
    copy = tmp_path / "csrc"
    shutil.copytree(build.CSRC, copy)
    monkeypatch.setattr(build, "CSRC", str(copy))
    assert build.source_stamp() == stamp
    with open(copy / "pool.cu", "a") as f:
        f.write("\n// touched\n")
    assert build.source_stamp() != stamp
