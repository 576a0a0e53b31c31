"""The actor loop and the scheduler around it, on the CPU with stand-in engines (the loop is host logic; the engines are
covered by the GPU tests).  Mirrors what the reference's loop does (293-project/src/scheduler.py:483-588, 734-925):

  * execute_schedule: serves the sessions of the current schedule, applies a pushed update (drain to the latest, unload
    the models that left, load the ones that came), survives a failing batch (logs, sleeps 0.1 s, carries on), ends on stop();
  * B200Scheduler (thread mode): submit_request -> rate tracker -> squishy bin packing -> per-worker update queue ->
    batches served from the shared-memory ring -> completions visible in get_stats and in metrics.json (reference schema).
"""
import json
import queue
import threading
import time

import numpy as np
import pytest
import torch

import rdb_b200  # noqa: F401
from rdb_b200 import packing, scheduler as sched_mod, worker as worker_mod
from rdb_b200.request_queue import B200RequestQueue
from rdb_b200.slab_queue import SlabRequestQueue

SIZES = range(1, 65)


class FakeEngine:
    """Stands in for B200Model: same call surface, output row i = [sum of request i, marker]."""
    is_b200_engine = True

    def __init__(self, marker=1.0, max_batch=64, fail_first=0):
        self.marker, self.max_batch, self.fail = marker, max_batch, fail_first
        self.in_h = self.in_w = 8
        self.calls, self.closed = 0, False
        self.out_logical = 2

    def bucket_for(self, n):
        return n

    def __call__(self, x):
        self.calls += 1
        if self.fail > 0:
            self.fail -= 1
            raise RuntimeError("injected engine failure")
        return torch.stack([x.reshape(x.shape[0], -1).sum(1), torch.full((x.shape[0],), self.marker)], 1)

    class _Handle:
        def __init__(self, out):
            self.out = out

        def result(self):
            return self.out

    def infer_runs_async(self, slab, runs, offsets, hw, n, padded_batch=None, mode=0, total_bytes=None):
        self.calls += 1
        staged = torch.cat([slab[src:src + nbytes] for src, nbytes, _ in runs])
        rows = [staged[int(o):int(o) + int(h) * int(w) * 3].float().sum() for o, (h, w) in zip(offsets, hw)]
        return FakeEngine._Handle(torch.stack([torch.stack(rows), torch.full((n,), self.marker)], 1))

    def close(self):
        self.closed = True


class Session:
    def __init__(self, model_name, batch_size):
        self.model_name, self.batch_size = model_name, batch_size


@pytest.fixture
def fake_cuda(monkeypatch):
    monkeypatch.setattr(torch.cuda, "is_available", lambda: True)


def test_execute_schedule_update_error_and_stop(fake_cuda, caplog):
    profile = {m: {b: {"avg_latency_ms": 0.1} for b in SIZES} for m in ("resnet", "shufflenet")}
    queues = {m: B200RequestQueue(m, max_size=1000, batch_profile=profile, compat=False, slo_ms=1e6)
              for m in ("resnet", "shufflenet")}
    a, b = FakeEngine(1.0), FakeEngine(2.0, fail_first=1)
    registry = {"resnet": a, "shufflenet": b}
    w = worker_mod.B200GPUWorker("cpu_0", 0, [(Session("resnet", 4), 1.0)], 1.0, registry, compat_overshoot=False)
    assert list(w.models) == ["resnet"]
    results, updates = [], queue.Queue()
    t = threading.Thread(target=w.execute_schedule, args=(queues, updates), kwargs={"on_result": results.append}, daemon=True)
    t.start()
    reqs = [torch.full((3, 8, 8), float(i)) for i in range(8)]
    for i, r in enumerate(reqs):
        assert queues["resnet"].add_request("a%d" % i, r)
    deadline = time.time() + 10
    while sum(len(r["request_ids"]) for r in results) < 8 and time.time() < deadline:
        time.sleep(0.01)
    got = {rid: float(out[0]) for r in results for rid, out in zip(r["request_ids"], r["outputs"])}
    assert got == {"a%d" % i: float(i) * 192 for i in range(8)}
    assert all(set(r) == {"outputs", "request_ids", "processing_time", "latency"} for r in results)   # scheduler.py:467-472
    assert queues["resnet"].get_stats()["avg_latency"] > 0

    # two updates pushed back to back: only the latest is applied (scheduler.py:489-492); resnet is unloaded, shufflenet loaded
    updates.put(([(Session("resnet", 2), 1.0)], 5.0, time.time()))
    updates.put(([(Session("shufflenet", 4), 0.5)], 2.0, time.time()))
    deadline = time.time() + 10
    while list(w.models) != ["shufflenet"] and time.time() < deadline:
        time.sleep(0.01)
    assert list(w.models) == ["shufflenet"] and w.duty_cycle == 2.0
    assert [s.model_name for s, _ in w.sessions] == ["shufflenet"]

    # the first shufflenet batch hits an engine failure: logged, the loop sleeps 0.1 s and goes on (scheduler.py:586-588)
    n_before = len(results)
    with caplog.at_level("ERROR"):
        for i in range(8):
            assert queues["shufflenet"].add_request("s%d" % i, reqs[i])
        deadline = time.time() + 10
        while len(results) == n_before and time.time() < deadline:
            time.sleep(0.01)
    assert any("injected engine failure" in rec.getMessage() for rec in caplog.records)
    served = [rid for r in results[n_before:] for rid in r["request_ids"]]
    assert served and all(rid.startswith("s") for rid in served)
    assert all(float(r["outputs"][0, 1]) == 2.0 for r in results[n_before:])

    w.stop()
    t.join(timeout=5)
    assert not t.is_alive()
    st = w.get_stats()
    assert st["node_id"] == "cpu_0" and st["processed_batches"] == len(results) and st["avg_processing_time"] >= 0


def test_oversize_batch_is_split_not_refused(fake_cuda):
    profile = {"resnet": {b: {"avg_latency_ms": 0.1} for b in SIZES}}
    q = B200RequestQueue("resnet", max_size=1000, batch_profile=profile, compat=True, slo_ms=1e6)
    eng = FakeEngine(max_batch=8)
    w = worker_mod.B200GPUWorker("cpu_0", 0, [(Session("resnet", 8), 1.0)], 1.0, {"resnet": eng})
    for i in range(20):
        q.add_request("r%d" % i, torch.full((3, 8, 8), float(i)))
    batch, _ = q.get_batch(8)                       # compat overshoot: 16 requests for batch_size 8 (scheduler.py:274)
    assert batch.batch_size == 16
    r = w.process_batch(batch, {"resnet": q})
    assert r["outputs"].shape == (16, 2) and eng.calls == 2
    assert [float(x) for x in r["outputs"][:, 0]] == [i * 192.0 for i in range(16)]


def test_scheduler_thread_mode_end_to_end(fake_cuda, monkeypatch, tmp_path):
    engines = []

    class FakeModel(FakeEngine):                                   # takes B200Model's place inside the worker module
        def __init__(self, kind, module=None, blob=None, max_batch=256, device=0, buckets=None, **kw):
            super().__init__(marker={"resnet50": 1.0, "shufflenet_v2_x1_0": 2.0}[kind], max_batch=max_batch)
            engines.append((kind, self))
    monkeypatch.setattr(worker_mod, "B200Model", FakeModel)

    def fake_pin(self):
        self._pinned_for_test = True
        self.payload_tensor = torch.from_numpy(self.payload)
    monkeypatch.setattr(SlabRequestQueue, "pin", fake_pin)

    profile = {m: {b: {"avg_latency_ms": 0.2 + 0.01 * b, "peak_memory_mb": 50.0 + b} for b in range(1, 513)}
               for m in ("resnet", "shufflenet")}
    registry = {"resnet": {"kind": "resnet50", "blob": b"", "in_hw": (8, 8), "out_logical": 2},
                "shufflenet": {"kind": "shufflenet_v2_x1_0", "blob": b"", "in_hw": (8, 8), "out_logical": 2}}
    mfile = str(tmp_path / "metrics.json")
    s = sched_mod.B200Scheduler(profile, monitoring_interval=0.05, rate_change_threshold=0.05, model_registry=registry,
                                num_gpus=2, worker_mode="thread", metrics_file=mfile, metrics_interval=0.05,
                                max_queue_size=500, in_flight_slots=600, tracker_window=0.5)
    try:
        assert set(s.request_queues) == {"resnet", "shufflenet"} and len(s.workers) == 2
        assert s.submit_request("vit", "x", np.zeros((8, 8, 3), np.uint8)) is False       # no such queue (scheduler.py:740-742)
        s.start_monitoring()
        t_end = time.time() + 1.5
        i = 0
        while time.time() < t_end:
            for _ in range(20):
                img = np.full((8, 8, 3), i % 200, dtype=np.uint8)
                assert s.submit_request("resnet" if i % 3 else "shufflenet", "q%d" % i, img) in (True, False)
                i += 1
            time.sleep(0.01)
        def settled():
            for q in s.request_queues.values():
                st = q.get_stats()
                if st["completed_requests"] + st["discarded_stale"] != st["total_requests"]:
                    return False
            return True
        deadline = time.time() + 20
        while not settled() and time.time() < deadline:
            time.sleep(0.02)
        assert s.metrics["schedule_updates"] >= 1 and len(s.nodes) >= 1
        assert set(s.sessions) == {"resnet", "shufflenet"}
        for model in ("resnet", "shufflenet"):
            st = s.request_queues[model].get_stats()
            assert st["total_requests"] > 0
            assert st["completed_requests"] + st["discarded_stale"] + st["queue_size"] == st["total_requests"]
            assert st["completed_requests"] > 0 and st["p99_latency"] >= st["avg_latency"] * 0 and st["avg_latency"] > 0
        m = json.load(open(mfile))
        for model in ("resnet", "shufflenet"):                      # reference schema: get_stats() keys + request_rate (:969-983)
            for k in ("total_requests", "dropped_requests", "slo_violations", "queue_size", "last_latency", "avg_latency",
                      "p95_latency", "p99_latency", "request_rate"):
                assert k in m[model], k
        assert {k for k, _ in engines} == {"resnet50", "shufflenet_v2_x1_0"}
    finally:
        stats = s.stop(timeout=10)
        s.close()
    assert set(stats) == {"B200_0", "B200_1"}
    assert sum(v["total_requests"] for v in stats.values()) > 0


def test_min_transfer_matching_like_the_reference():
    s = object.__new__(sched_mod.B200Scheduler)

    def nd(*models):
        return packing.node([(packing.session(m, 100.0, 10.0, 4), 0.5) for m in models], duty_cycle=10.0)
    old = [nd("resnet"), nd("shufflenet")]
    new = [nd("shufflenet"), nd("efficientnet"), nd("resnet")]
    assert sched_mod.B200Scheduler.get_transfers(s, old, new[:2]) == 2
    final = sched_mod.B200Scheduler._match(s, old, new)
    assert [[x.model_name for x, _ in n.node_sessions] for n in final] == [["resnet"], ["shufflenet"], ["efficientnet"]]
    fewer = sched_mod.B200Scheduler._match(s, old, [nd("shufflenet")])
    assert len(fewer) == 1


def test_request_tracker_window_resets_only_when_read():
    tr = sched_mod.RequestTracker(window_size=0.05)
    for _ in range(10):
        tr.record_request()
    time.sleep(0.06)
    assert tr.request_count == 10                      # scheduler.py:151-166: the window restarts inside get_request_rate
    rate = tr.get_request_rate()
    assert rate > 0 and tr.request_count == 0 and tr.get_total_requests() == 10
