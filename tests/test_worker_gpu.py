"""The actor surface of the reference (GPUWorker.process_batch / duty-cycle loop, scheduler.py:435-588) on top of the
engine: same return dict, same metrics bookkeeping, outputs equal to a direct engine call."""
import queue as pyqueue

import pytest
import torch

import rdb_b200
from rdb_b200 import synth
from rdb_b200.engine import B200Model
from rdb_b200.request_queue import B200RequestQueue
from rdb_b200.worker import B200GPUWorker

pytestmark = pytest.mark.gpu


class Session:
    def __init__(self, model_name, batch_size):
        self.model_name, self.batch_size = model_name, batch_size


class UpdateQueue:
    def __init__(self):
        self.q = pyqueue.Queue()

    def qsize(self):
        return self.q.qsize()

    def get(self):
        return self.q.get()

    def put(self, x):
        self.q.put(x)


@pytest.fixture(scope="module")
def setup(cuda):
    module = synth.make_model("shufflenet_v2_x1_0", n_proto=8)       # the cheapest family to build
    model = B200Model("shufflenet_v2_x1_0", module, max_batch=32)
    profile = {"shufflenet": {b: {"avg_latency_ms": 1.0} for b in (1, 2, 4, 8, 16, 32)}}
    yield module, model, profile
    model.close()


def test_process_batch_contract_uint8_requests_through_pinned_slab(setup):
    module, model, profile = setup
    q = B200RequestQueue("shufflenet", max_size=100, batch_profile=profile, compat=False,
                         pinned_slab_bytes=64 * 150528)
    worker = B200GPUWorker("B200_0", 0, [(Session("shufflenet", 8), 1.0)], 5.0, {"shufflenet": model}, max_batch=32)
    imgs = list(synth.synthetic_images(11))
    for i, im in enumerate(imgs):
        assert q.add_request("r%d" % i, im)
    batch, discard = q.get_batch(8)
    assert batch.batch_size == 8 and discard == 0 and batch.slab_offsets is not None
    res = worker.process_batch(batch, {"shufflenet": q}, discard)
    assert set(res) == {"outputs", "request_ids", "processing_time", "latency"}
    assert res["request_ids"] == ["r%d" % i for i in range(8)]
    assert res["outputs"].device.type == "cpu" and res["outputs"].dtype == torch.float32
    assert res["outputs"].shape == (8, 1000)
    direct = model.infer_u8(imgs[:8]).cpu()
    assert torch.equal(res["outputs"], direct)
    assert res["processing_time"] > 0 and res["latency"] > 0
    assert len(q.metrics["latencies"]) == 8 and worker.get_stats()["processed_batches"] == 1
    assert worker.get_stats()["total_requests"] == 8


def test_process_batch_reference_native_fp32_requests(setup):
    module, model, profile = setup
    q = B200RequestQueue("shufflenet", max_size=100, batch_profile=profile, compat=True)   # reference semantics
    worker = B200GPUWorker("B200_1", 0, [(Session("shufflenet", 4), 1.0)], 5.0, {"shufflenet": model}, max_batch=32)
    reqs = synth.reference_requests(9)
    for i, t in enumerate(reqs):
        q.add_request("x%d" % i, t)
    batch, _ = q.get_batch(4)                       # compat: overshoots to (4, 8]
    assert 4 < batch.batch_size <= 8
    res = worker.process_batch(batch, {"shufflenet": q})
    with torch.no_grad():
        ref = module(torch.stack(batch.inputs))
    assert res["outputs"].shape == ref.shape
    assert (res["outputs"] - ref).abs().max().item() <= 2e-2


def test_duty_cycle_and_schedule_update(setup):
    module, model, profile = setup
    q = B200RequestQueue("shufflenet", max_size=100, batch_profile=profile, compat=False)
    worker = B200GPUWorker("B200_2", 0, [], 5.0, {"shufflenet": model}, max_batch=32)
    assert worker.run_cycle({"shufflenet": q}, sleep=lambda s: None) == []          # no sessions yet
    uq = UpdateQueue()
    uq.put(([(Session("shufflenet", 2), 0.5)], 4.0, 0.0))
    uq.put(([(Session("shufflenet", 4), 1.0)], 8.0, 0.0))                          # only the latest schedule is applied
    assert worker._check_for_updates(uq) is True
    assert worker.duty_cycle == 8.0 and [s.batch_size for s, _ in worker.sessions] == [4]
    for i, im in enumerate(synth.synthetic_images(6)):
        q.add_request("r%d" % i, im)
    slept = []
    results = worker.run_cycle({"shufflenet": q}, sleep=slept.append)
    assert len(results) == 1 and len(results[0]["request_ids"]) == 4
    assert worker._check_for_updates(uq) is False


def test_process_batch_logs_and_reraises(setup):
    module, model, profile = setup
    q = B200RequestQueue("shufflenet", max_size=10, batch_profile=profile, compat=False)
    worker = B200GPUWorker("B200_3", 0, [(Session("shufflenet", 2), 1.0)], 5.0, {"shufflenet": model}, max_batch=32)
    q.add_request("bad", torch.zeros(3, 17, 17))
    batch, _ = q.get_batch(2)
    with pytest.raises(Exception):
        worker.process_batch(batch, {"shufflenet": q})


def test_results_of_earlier_batches_survive_later_ones(setup):
    """The worker hands back outputs it owns: the engine's pinned output slots are reused two batches later, so a result
    that aliased a slot would silently change under the caller (three sessions collected in one cycle would do it)."""
    module, model, profile = setup
    q = B200RequestQueue("shufflenet", max_size=100, batch_profile=profile, compat=False, pinned_slab_bytes=64 * 150528)
    worker = B200GPUWorker("B200_4", 0, [(Session("shufflenet", 4), 1.0)], 5.0, {"shufflenet": model}, max_batch=32)
    imgs = list(synth.synthetic_images(16, seed=7000))
    for i, im in enumerate(imgs):
        assert q.add_request("r%d" % i, im)
    results, snapshots = [], []
    for _ in range(4):
        batch, d = q.get_batch(4)
        assert batch.slab_offsets is not None
        results.append(worker.process_batch(batch, {"shufflenet": q}, d))
        snapshots.append(results[-1]["outputs"].clone())
    for r, snap in zip(results, snapshots):
        assert torch.equal(r["outputs"], snap)
    direct = model.infer_u8(imgs).cpu()
    assert torch.equal(torch.cat([r["outputs"] for r in results]), direct)


def test_pinned_slab_never_overwrites_queued_requests(setup):
    """A slab smaller than the queue: requests that do not fit travel without a slab position (served from their tensors),
    queued payloads are never overwritten, and released space is reused."""
    module, model, profile = setup
    q = B200RequestQueue("shufflenet", max_size=100, batch_profile=profile, compat=False, pinned_slab_bytes=5 * 150528)
    worker = B200GPUWorker("B200_5", 0, [(Session("shufflenet", 4), 1.0)], 5.0, {"shufflenet": model}, max_batch=32)
    imgs = list(synth.synthetic_images(12, seed=7100))
    for i, im in enumerate(imgs):
        assert q.add_request("r%d" % i, im)                   # 12 queued, room for 5 in the slab
    outs = []
    for _ in range(3):
        batch, d = q.get_batch(4)
        outs.append(worker.process_batch(batch, {"shufflenet": q}, d)["outputs"])
    assert torch.equal(torch.cat(outs), model.infer_u8(imgs).cpu())
    assert not q.slab.live                                     # everything released after completion
    assert q.add_request("again", imgs[0]) and q.slab.live     # and the space is used again


def test_shared_memory_queue_through_the_worker(setup):
    """SlabRequestQueue -> get_batch -> process_batch: DMA from the registered segment, bucket padding, pipelined handles."""
    from rdb_b200.slab_queue import SlabRequestQueue
    module, model, profile = setup
    prof = {"shufflenet": {b: {"avg_latency_ms": 1.0} for b in range(1, 65)}}
    q = SlabRequestQueue("shufflenet", max_size=64, batch_profile=prof, slot_bytes=150528, in_flight_slots=64, slo_ms=1e9)
    model.set_buckets([1, 4, 8, 16, 32])
    worker = B200GPUWorker("B200_6", 0, [(Session("shufflenet", 6), 1.0)], 5.0, {"shufflenet": model}, max_batch=32,
                           pipeline_depth=2)
    try:
        imgs = list(synth.synthetic_images(40, seed=7200))
        direct = model.infer_u8(imgs[:32]).cpu()
        direct = torch.cat([direct, model.infer_u8(imgs[32:]).cpu()])
        pend, outs, ids = [], [], []
        for rnd in range(2):                                   # second round wraps the ring (128 slots, 2 x 40 requests ...)
            for i, im in enumerate(imgs):
                assert q.add_request("r%d_%d" % (rnd, i), im)
            while not q.empty():
                batch, d = q.get_batch(6)                      # 6 requests run in the 8 bucket
                pend.append(worker.process_batch_async(batch, {"shufflenet": q}, d))
                if len(pend) == 2:
                    r = worker.finish_batch(pend.pop(0), {"shufflenet": q})
                    outs.append(r["outputs"]); ids += r["request_ids"]
        while pend:
            r = worker.finish_batch(pend.pop(0), {"shufflenet": q})
            outs.append(r["outputs"]); ids += r["request_ids"]
        got = torch.cat(outs)
        assert ids == ["r%d_%d" % (rnd, i) for rnd in range(2) for i in range(40)]
        assert torch.equal(got, torch.cat([direct, direct]))
        st = q.get_stats()
        assert st["completed_requests"] == 80 and st["queue_size"] == 0
        for i in range(64):                                    # every slot was given back: the ring takes a full queue again
            assert q.add_request("z%d" % i, imgs[i % 40])
    finally:
        model.buckets = None
        q.close()


def test_engine_sized_from_sessions_under_the_compat_overshoot(setup):
    module, model, profile = setup
    worker = B200GPUWorker("B200_7", 0, [], 5.0, {"shufflenet": module}, max_batch=8, compat_overshoot=True)
    uq = UpdateQueue()
    uq.put(([(Session("shufflenet", 12), 1.0)], 4.0, 0.0))
    assert worker._check_for_updates(uq)
    assert worker.models["shufflenet"].max_batch == 24         # get_batch may hand over up to 2 x 12 requests
    uq.put(([(Session("shufflenet", 20), 1.0)], 4.0, 0.0))
    assert worker._check_for_updates(uq)
    assert worker.models["shufflenet"].max_batch == 40         # rebuilt for the larger schedule
    worker.models["shufflenet"].close()
