"""B200Scheduler at deployment shape on a real GPU: one model-actor PROCESS (the `@ray.remote(num_gpus=1)` actor of
293-project/src/scheduler.py:374) fed through the shared-memory request ring, plus the ZMQ ingress in front of it.
What the reference's own driver (src/test_scheduler.py) exercises and never asserts: every submitted request is served,
the driver SEES the completions (they live in the ring, not in a pickled copy), metrics.json has the reference's schema,
and stop() ends the actors."""
import json
import time

import numpy as np
import pytest
import torch

import rdb_b200  # noqa: F401
from rdb_b200 import packing, synth
from rdb_b200.scheduler import B200Scheduler

pytestmark = pytest.mark.gpu


def _wait(cond, timeout):
    t_end = time.time() + timeout
    while time.time() < t_end:
        if cond():
            return True
        time.sleep(0.01)
    return cond()


@pytest.fixture(scope="module")
def deployment(cuda, tmp_path_factory):
    module = synth.make_model("shufflenet_v2_x1_0", n_proto=8)
    profile = {"shufflenet": {b: {"avg_latency_ms": 1.0 + 0.01 * b, "peak_memory_mb": 20.0 + b} for b in range(1, 65)}}
    mfile = str(tmp_path_factory.mktemp("metrics") / "metrics.json")
    s = B200Scheduler(profile, monitoring_interval=0.2, rate_change_threshold=0.05, model_registry={"shufflenet": module},
                      num_gpus=1, worker_mode="process", max_batch=32, preload=("shufflenet",), metrics_file=mfile,
                      metrics_interval=0.2, max_queue_size=256, in_flight_slots=256)
    assert s.wait_ready(300), "the actor process did not come up"
    yield s, mfile, module
    stats = s.stop(timeout=30)
    s.close()
    assert set(stats) == {"B200_0"}
    assert all(not p.is_alive() for p in s.workers)


def test_static_schedule_serves_every_request_and_the_driver_sees_it(deployment):
    s, mfile, _ = deployment
    sess = packing.session("shufflenet", 1500 / 2.2, 1000.0, 16)
    s.set_schedule([packing.node([(sess, 1.0)], duty_cycle=2.0)])
    imgs = synth.synthetic_images(48, seed=6100)
    q = s.request_queues["shufflenet"]
    base = q.get_stats()["completed_requests"]
    for i in range(24):
        assert s.submit_request("shufflenet", "one%d" % i, imgs[i]) is True              # scheduler.py:734-751
    assert s.submit_requests("shufflenet", imgs[24:].contiguous().numpy(), "burst", 0) == 24   # one native call
    assert _wait(lambda: q.get_stats()["completed_requests"] - base == 48, 30)
    st = q.get_stats()
    assert st["queue_size"] == 0 and st["dropped_requests"] == 0 and st["slo_violations"] == 0
    assert 0 < st["avg_latency"] <= st["p99_latency"] <= st["max_latency"] < 1500
    assert st["completed_batches"] >= 3                                                     # batch size 16
    assert _wait(lambda: json.load(open(mfile)).get("shufflenet", {}).get("completed_requests", 0) >= base + 48, 5)
    m = json.load(open(mfile))["shufflenet"]
    for k in ("total_requests", "dropped_requests", "slo_violations", "queue_size", "last_latency", "avg_latency",
              "p95_latency", "p99_latency", "request_rate"):                                # scheduler.py:969-983
        assert k in m, k


def test_rate_monitor_packs_and_pushes_a_schedule(deployment):
    s, _, _ = deployment
    s.set_schedule([])                                       # idle actor, no sessions: the monitor has to build the schedule
    s.sessions = {}
    s.nodes = []
    q = s.request_queues["shufflenet"]
    base = q.get_stats()["completed_requests"]
    updates0 = s.metrics["schedule_updates"]
    s.start_monitoring()
    try:
        imgs = synth.synthetic_images(16, seed=6200).contiguous().numpy()
        sent = 0
        t_end = time.time() + 2.0
        while time.time() < t_end:
            sent += s.submit_requests("shufflenet", imgs, "m", sent)
            time.sleep(0.02)
        assert _wait(lambda: s.metrics["schedule_updates"] > updates0, 5)
        assert "shufflenet" in s.sessions and s.sessions["shufflenet"].request_rate > 0      # what was asked for (scheduler.py:893-896)
        assert s.nodes and s.nodes[0].node_sessions[0][0].batch_size >= 1                    # what the packing made of it
        assert _wait(lambda: q.get_stats()["completed_requests"] + q.get_stats()["discarded_stale"] - base >= sent, 30)
    finally:
        s.stop_monitoring()


def test_zmq_ingress_in_front_of_the_scheduler(deployment):
    zmq = pytest.importorskip("zmq")                                                       # noqa: F841
    from rdb_b200 import ingress
    s, _, _ = deployment
    sess = packing.session("shufflenet", 1500 / 2.2, 1000.0, 8)
    s.set_schedule([packing.node([(sess, 1.0)], duty_cycle=2.0)])
    q = s.request_queues["shufflenet"]
    base = q.get_stats()["completed_requests"]
    ing = ingress.ZmqIngress(s, "tcp://127.0.0.1:*")
    ing.start()
    sender = ingress.RequestSender(ing.endpoint)
    try:
        imgs = synth.synthetic_images(20, seed=6300)
        for i in range(20):
            sender.send("shufflenet", "z%d" % i, imgs[i].numpy(), slo_ms=1500)
        assert _wait(lambda: ing.accepted == 20, 10)
        assert _wait(lambda: q.get_stats()["completed_requests"] - base == 20, 30)
    finally:
        sender.close()
        ing.stop()
