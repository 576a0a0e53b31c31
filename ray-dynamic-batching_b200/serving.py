"""Load generation and measurement for the serving path (queue -> actor loop), shared by bench.py and tools/serve_sim.py.

The reference's driver submits requests in one-second bursts whose size grows with time (`rate = elapsed * slope`,
293-project/src/test_scheduler.py:68-96) and reads latency percentiles from RequestQueue.get_stats
(scheduler.py:343-372); the same two things happen here, at B200 rates: a feeder writes uint8 requests into the
shared-memory ring (one native call per burst), the actor loop serves them, and the percentiles come from get_stats.
"""
import threading
import time

import numpy as np


class Session:
    """The two fields of nexus.session the actor loop reads (nexus.py:17-67)."""

    def __init__(self, model_name, batch_size):
        self.model_name, self.batch_size = model_name, batch_size


def measured_profile(model, batch_sizes, reps=5):
    """batch size -> {'avg_latency_ms'}: end-to-end latency of one batch through the engine (H2D + forward + D2H), what
    the stale-request rule of get_batch needs (scheduler.py:281); like profiling/ModelProfiler.py offline."""
    import torch
    from . import synth
    pool = list(synth.synthetic_images(8, model.in_h, model.in_w))
    prof = {}
    for bs in batch_sizes:
        imgs = [pool[i % 8] for i in range(bs)]
        pb = model.bucket_for(bs)
        for _ in range(2):
            model.infer_u8(imgs, padded_batch=pb)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            model.infer_u8(imgs, padded_batch=pb).cpu()
        prof[bs] = {"avg_latency_ms": (time.perf_counter() - t0) / reps * 1e3}
    return prof


def dense_profile(sparse, max_bs):
    """Every integer batch size 1..max_bs from a profile measured at a few sizes: the next measured size's figures."""
    keys = sorted(sparse)
    out, j = {}, 0
    for b in range(1, max_bs + 1):
        while j < len(keys) - 1 and keys[j] < b:
            j += 1
        out[b] = dict(sparse[keys[j]])
    return out


class Feeder(threading.Thread):
    """Submits requests at `rate` req/s (or the reference's slope pattern) in bursts of `burst` through add_requests."""

    def __init__(self, queue, pool, rate, seconds, slope=0.0, burst=16, prefix="r"):
        super().__init__(daemon=True)
        self.queue, self.pool, self.rate, self.seconds, self.slope = queue, pool, float(rate), float(seconds), float(slope)
        self.burst, self.prefix = int(burst), prefix
        self.submitted = self.dropped = 0
        self.stop_flag = False

    def run(self):
        pool = self.pool                                   # uint8 [P, H, W, 3] contiguous, P a multiple of burst
        p = pool.shape[0]
        t0 = time.perf_counter()
        i = 0
        while not self.stop_flag:
            el = time.perf_counter() - t0
            if el >= self.seconds:
                break
            target = self.slope * el * el / 2 if self.slope > 0 else self.rate * el
            while i + self.burst <= target and not self.stop_flag:
                j = i % p
                if j + self.burst > p:
                    j = 0
                ok = self.queue.add_requests(pool[j:j + self.burst], self.prefix, i)
                self.submitted += ok
                self.dropped += self.burst - ok
                i += self.burst
            time.sleep(0.0002)


def client_process_main(queue, hw, rate, seconds, slope, burst, prefix, seed, barrier, report):
    """One load-generator PROCESS: bursts of `burst` uint8 requests at `rate` req/s (or rate = t * slope, the reference
    driver's pattern, src/test_scheduler.py:77-79) written straight into a model's request ring(s).  Starts on `barrier`."""
    import os
    os.environ.setdefault("OMP_NUM_THREADS", "1")      # a client copies bytes; it must not bring a thread pool per process
    import torch
    torch.set_num_threads(1)
    from . import synth
    if hasattr(queue, "bind"):
        queue.bind(seed // 100)            # clients start their round-robin on different shards
    pool = synth.synthetic_images(4 * burst, hw, hw, seed=8000 + seed).contiguous().numpy()
    p = pool.shape[0]
    barrier.wait()                     # every client (and the parent's clock) starts together, imports done
    t0 = time.perf_counter()
    i = sub = drop = 0
    while True:
        el = time.perf_counter() - t0
        if el >= seconds:
            break
        target = slope * el * el / 2 if slope > 0 else rate * el
        while i + burst <= target and time.perf_counter() - t0 < seconds:
            j = i % p
            if j + burst > p:
                j = 0
            ok = queue.add_requests(pool[j:j + burst], prefix, i)
            sub += ok
            drop += burst - ok
            i += burst
            if ok < burst:
                time.sleep(0.0005)     # full queue: the burst is dropped (counted); do not hammer the ring's counters
        time.sleep(0.0002)
    report.put((prefix, sub, drop, time.perf_counter() - t0))
    queue.close()


def serve_loop(worker, queue, model_name, batch_size, until, idle_sleep=0.0001):
    """The saturated form of GPUWorker.execute_schedule for one session with occupancy 1 (no slack to sleep away):
    get_batch -> process_batch_async, one batch in flight while the next is formed.  Returns the batch sizes served."""
    queues = {model_name: queue}
    sizes, pend = [], []
    serve_loop.last_finish = time.perf_counter()
    while True:
        batch, discard = queue.get_batch(batch_size)
        if batch is None:
            while pend:
                worker.finish_batch(pend.pop(0), queues)
                serve_loop.last_finish = time.perf_counter()
            if until():
                break
            time.sleep(idle_sleep)
            continue
        pend.append(worker.process_batch_async(batch, queues, discard))
        sizes.append(batch.batch_size)
        if len(pend) >= worker.pipeline_depth:
            worker.finish_batch(pend.pop(0), queues)
            serve_loop.last_finish = time.perf_counter()
    return sizes


def bench_serving(model, kind, model_name, batch, capacity_img_s, seconds=3.0, load=0.8, slo_ms=None, max_clients=6):
    """One GPU: client PROCESSES -> SlabRequestQueue -> B200GPUWorker (in this process) at `load` x capacity.  Returns the
    p99 batch latency (RequestQueue.get_stats) against the model's SLO, the served rate and the mean batch size.  A Python
    client moves ~25 k requests of 224 x 224 per second; the offered load is what the clients managed, and is reported."""
    import multiprocessing as mp
    import os
    from .request_queue import models_config
    from .slab_queue import SlabRequestQueue
    from .worker import B200GPUWorker
    h, w = model.in_h, model.in_w
    slo = slo_ms if slo_ms is not None else models_config[model_name]['SLO']
    if model.buckets is None:
        from .engine import default_buckets
        model.set_buckets(default_buckets(model.max_batch))
    prof = dense_profile(measured_profile(model, sorted({1, 8, 32, max(1, batch // 2), batch})), batch)
    q = SlabRequestQueue(model_name, max_size=2000, batch_profile={model_name: prof}, slot_bytes=h * w * 3,
                         in_flight_slots=4 * batch + 64, compat=False, slo_ms=slo)
    worker = B200GPUWorker("B200_0", 0, [(Session(model_name, batch), 1.0)], 1.0, {model_name: model},
                           max_batch=model.max_batch, compat_overshoot=False, pipeline_depth=2)
    rate = load * capacity_img_s
    per_client = 25000.0 * (224 * 224) / float(h * w)
    n_clients = int(max(1, min(max_clients, (os.cpu_count() or 4) // 2, -(-rate // per_client))))
    ctx = mp.get_context("spawn")
    barrier, report = ctx.Barrier(n_clients + 1), ctx.Queue()
    procs = [ctx.Process(target=client_process_main, daemon=True,
                         args=(q, h, rate / n_clients, seconds, 0.0, 16, "c%d_" % c, 100 * c, barrier, report))
             for c in range(n_clients)]
    try:
        q.pin()
        if hasattr(model, "prepare_async"):
            model.prepare_async()
        for p in procs:
            p.start()
        barrier.wait(timeout=300)
        t0 = time.perf_counter()
        alive = lambda: any(p.is_alive() for p in procs)
        sizes = serve_loop(worker, q, model_name, batch, until=lambda: not alive() and q.empty())
        wall = max(1e-6, serve_loop.last_finish - t0)      # to the last completed batch (the clients' teardown is not serving time)
        reports = []
        while len(reports) < n_clients:
            try:
                reports.append(report.get(timeout=5))
            except Exception:  # noqa: BLE001
                break
        st = q.get_stats()
        sub, drop = sum(r[1] for r in reports), sum(r[2] for r in reports)
        feed_s = max([r[3] for r in reports] or [seconds])
        return {"pattern": "constant", "target_rate": rate, "offered_rate": (sub + drop) / feed_s,
                "load_fraction_of_capacity": load, "seconds": seconds, "client_processes": n_clients,
                "submitted": sub, "dropped_full_queue": drop, "served": int(sum(sizes)),
                "batches": len(sizes), "mean_batch": float(np.mean(sizes)) if sizes else 0.0, "batch_target": batch,
                "throughput_img_s": sum(sizes) / wall, "slo_ms": slo, "slo_violations": st["slo_violations"],
                "discarded_stale": st["discarded_stale"], "avg_latency_ms": st["avg_latency"],
                "p95_latency_ms": st["p95_latency"], "p99_latency_ms": st["p99_latency"], "max_latency_ms": st["max_latency"],
                "path": "client processes: add_requests (16 per call) -> shared-memory ring -> get_batch -> "
                        "process_batch_async (2 in flight) -> record_batch_completion; percentiles from get_stats "
                        "(scheduler.py:343-372)"}
    finally:
        for p in procs:
            p.join(timeout=10)
            if p.is_alive():
                p.terminate()
        q.close()
