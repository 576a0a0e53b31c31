"""ZMQ ingress (rdb_b200.ingress): the reference draft's JSON request (milind-code/request_simulator.py:33-41) with the
pixels riding in a second frame, landing byte for byte in the model's shared-memory request slot."""
import json
import time

import numpy as np
import pytest

import rdb_b200  # noqa: F401
from rdb_b200 import ingress
from rdb_b200.slab_queue import SlabRequestQueue

zmq = pytest.importorskip("zmq")


class StubScheduler:
    """submit_request of NexusScheduler (scheduler.py:734-751) over real request rings, without workers."""

    def __init__(self):
        prof = {m: {b: {"avg_latency_ms": 1.0} for b in range(1, 65)} for m in ("resnet", "yolov8")}
        self.request_queues = {"resnet": SlabRequestQueue("resnet", 16, prof, slot_bytes=24 * 24 * 3, in_flight_slots=16, slo_ms=1e9),
                               "yolov8": SlabRequestQueue("yolov8", 16, prof, slot_bytes=40 * 40 * 3, in_flight_slots=16, slo_ms=1e9)}

    def submit_request(self, model_name, request_id, image):
        if model_name not in self.request_queues:
            return False
        return self.request_queues[model_name].add_request(request_id, image)

    def close(self):
        for q in self.request_queues.values():
            q.close()


def test_header_is_the_reference_draft_plus_shape():
    img = np.arange(4 * 5 * 3, dtype=np.uint8).reshape(4, 5, 3)
    frames = ingress.encode_request("resnet", "resnet7", img, slo_ms=2000)
    header = json.loads(bytes(frames[0]))
    assert {"timestamp", "model_name", "request_id", "SLO"} <= set(header)        # request_simulator.py:35-40
    assert header["shape"] == [4, 5, 3] and header["dtype"] == "uint8" and len(bytes(frames[1])) == 60
    h2, im2 = ingress.decode_request(frames)
    assert h2["request_id"] == "resnet7" and np.array_equal(im2, img)
    gray = ingress.decode_request(ingress.encode_request("resnet", "g", img[:, :, 0]))[1]
    assert gray.shape == (4, 5, 1)
    with pytest.raises(ValueError):
        ingress.decode_request([frames[0], b"short"])
    with pytest.raises(ValueError):
        ingress.decode_request([json.dumps({"model_name": "resnet", "request_id": "x"}).encode()])


def test_requests_travel_from_a_push_socket_into_the_request_slots():
    sched = StubScheduler()
    ing = ingress.ZmqIngress(sched, "tcp://127.0.0.1:*")
    ing.start()
    sender = ingress.RequestSender(ing.endpoint)
    try:
        rng = np.random.default_rng(5)
        sent = {}
        for i in range(10):
            model = "resnet" if i % 2 == 0 else "yolov8"
            s = 24 if model == "resnet" else 40
            img = rng.integers(0, 256, (s, s - (i % 3), 3), dtype=np.uint8)      # ragged widths: any size up to the slot
            sent["%s%d" % (model, i)] = img
            sender.send(model, "%s%d" % (model, i), img, slo_ms=2000)
        sender.send("vit", "nobody-serves-this", np.zeros((2, 2, 3), np.uint8))   # no queue -> rejected (scheduler.py:740-742)
        sender.socket.send(b"not json")                                           # malformed -> logged, dropped
        deadline = time.time() + 10
        while ing.received < 12 and time.time() < deadline:
            time.sleep(0.01)
        assert ing.stats() == {"received": 12, "accepted": 10, "rejected": 1, "malformed": 1}
        for model in ("resnet", "yolov8"):
            batch, _ = sched.request_queues[model].get_batch(8)
            assert batch.batch_size == 5
            for j, rid in enumerate(batch.request_ids):                           # FIFO per model, pixels byte for byte
                assert rid.startswith(model) and np.array_equal(batch.inputs[j].numpy(), sent[rid])
            sched.request_queues[model].release(batch)
            batch = None
    finally:
        sender.close()
        ing.stop()
        sched.close()
