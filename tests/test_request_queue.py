"""Host logic of the product queue (rdb_b200.request_queue) against the reference-generated golden fixtures: in compat mode
the batch composition must be bit-exact with the reference's RequestQueue.get_batch (scheduler.py:258-322)."""
import json
import os
import pickle

import pytest

import rdb_b200  # noqa: F401
from rdb_b200.request_queue import B200RequestQueue, BatchRequest

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "batching_golden.json")))


class Clock:
    def __init__(self):
        self.t, self.step = 1000.0, 0.0

    def __call__(self):
        v = self.t
        self.t += self.step
        return v


def build(case, compat=True):
    clock = Clock()
    profile = {"resnet": {b: {"avg_latency_ms": case["avg_latency_ms"]} for b in (1, 2, 4, 8, 16, 32, 64, 128, 256, 512)}}
    q = B200RequestQueue("resnet", max_size=case["max_size"], batch_profile=profile, compat=compat, clock=clock,
                         slo_ms=case["slo"])
    accepted = []
    for i, t in enumerate(case["arrivals"]):
        clock.t = t
        accepted.append(q.add_request("r%d" % i, i))
    clock.t, clock.step = case["now0"], case["clock_step"]
    return q, clock, accepted


@pytest.mark.parametrize("idx", range(len(GOLD["cases"])))
def test_compat_get_batch_is_bit_exact_with_reference(idx):
    case = GOLD["cases"][idx]
    q, clock, accepted = build(case)
    assert accepted == case["accepted"]
    batch, discard = q.get_batch(case["batch_size"])
    if case["spin"]:
        # the reference busy-waits here; the product returns what it has gathered (documented deviation)
        ids = batch.request_ids if batch else []
        assert ids == case["partial_ids"]
        assert discard == (case["partial_discard"] if batch else 0)
        return
    assert discard == case["discard"]
    if case["batch"] is None:
        assert batch is None
        return
    g = case["batch"]
    assert batch.request_ids == g["request_ids"] and batch.inputs == g["inputs"]
    assert batch.batch_size == g["batch_size"] and batch.arrival_time == g["arrival_time"]
    assert batch.model_name == g["model_name"]
    clock.step = 0.0
    assert q.record_batch_completion(batch, case["completion_time"]) == case["violations"]
    assert q.get_stats() == case["stats"]
    assert [it[0] for it in q.queue.q] == case["remaining"]


@pytest.mark.parametrize("idx", range(0, len(GOLD["cases"]), 3))
def test_strict_mode_is_bounded_fifo_with_same_stale_filter(idx):
    case = GOLD["cases"][idx]
    q, clock, _ = build(case, compat=False)
    n_before = q.queue.qsize()
    batch, discard = q.get_batch(case["batch_size"])
    taken = batch.batch_size if batch else 0
    assert taken <= case["batch_size"]
    if batch:
        kept = [int(r[1:]) for r in batch.request_ids]
        assert kept == sorted(kept), "FIFO order"
        assert taken + discard + q.queue.qsize() == n_before


def test_queue_is_picklable_like_the_reference_requires():
    q = B200RequestQueue("resnet", max_size=10, batch_profile={"resnet": {4: {"avg_latency_ms": 1.0}}})
    q.add_request("a", 1)
    q2 = pickle.loads(pickle.dumps(q))
    assert q2.queue.qsize() == 1 and q2.slo_target == 2000


def test_full_queue_drops_and_counts():
    q = B200RequestQueue("resnet", max_size=2, batch_profile={"resnet": {}})
    assert q.add_request("a", 0) and q.add_request("b", 1) and not q.add_request("c", 2)
    assert q.metrics["dropped_requests"] == 1 and q.metrics["total_requests"] == 2
    assert q.get_batch(0) == (None, 0)
    assert B200RequestQueue("resnet", batch_profile={"resnet": {}}).get_batch(4) == (None, 0)


def test_errors_are_swallowed_like_the_reference():
    q = B200RequestQueue("resnet", max_size=4, batch_profile={"resnet": {}})   # no profile entry for batch 2
    q.add_request("a", 0)
    assert q.get_batch(2) == (None, 0)      # KeyError inside -> logged -> (None, 0)  (scheduler.py:319-322)


def test_batchrequest_fields_match_reference_dataclass():
    b = BatchRequest("resnet", [1], 1, ["a"], 3.0)
    assert (b.model_name, b.inputs, b.batch_size, b.request_ids, b.arrival_time) == ("resnet", [1], 1, ["a"], 3.0)
