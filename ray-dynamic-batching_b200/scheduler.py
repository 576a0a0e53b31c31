"""B200Scheduler: the reference's `NexusScheduler` (293-project/src/scheduler.py:602-929) over B200 model actors.

Same surface -- per-model request queues, `submit_request`, one worker per GPU fed through a per-worker update queue, a
rate monitor that re-packs when a model's request rate moves by more than the threshold, min-transfer matching of the new
schedule to the running one -- with the parts that tie the reference to two A6000s and to Ray replaced:

  reference                                                   here
  two hard-coded `'A6000_' + i` Ray actors (:679-690)         one actor PROCESS per visible GPU (`num_gpus`), each pinned
                                                              with CUDA_VISIBLE_DEVICES like `@ray.remote(num_gpus=1)` (:374)
  queues = Ray queue actors, tensors pickled per request      SlabRequestQueue: shared-memory request rings (slab_queue.py)
  metrics dict pickled into each actor, driver never sees     counters live in the ring: `metrics.json` (MetricsWriter, the
  completions                                    headless MetricsDisplay :933-983) reports real completions
  workers cannot be stopped (actor busy in its loop)          `stop()` ends the loops and joins the processes
  model_registry of pretrained torchvision modules pickled    plans are lowered ONCE here (BatchNorm folded, bf16 weights)
  to every actor (:637)                                       and the blobs shipped to the actors

`worker_mode="thread"` runs the same actors as threads of this process (one GPU, tests, bench.py); "process" is the
deployment shape.  Ray is not in this image; with Ray installed `worker.as_ray_actor()` wraps the same class.
"""
import copy
import itertools
import json
import logging
import os
import threading
import time
import traceback
from datetime import datetime
from queue import Queue
from threading import Lock, Thread
from typing import Dict, List, Optional

from . import packing
from .request_queue import models_config
from .slab_queue import ShardedRequestQueue, SlabRequestQueue  # noqa: F401

SLO_hack = 2.2          # scheduler.py:28


class RequestTracker:
    """scheduler.py:115-179: request rate over a window that restarts when it is read after `window_size` seconds."""

    def __init__(self, window_size: float = 60.0):
        self.window_size = window_size
        self.request_count = 0
        self.last_reset = datetime.now()
        self.lock = Lock()
        self._total_requests = 0

    def record_request(self, n: int = 1) -> None:
        with self.lock:
            self.request_count += n
            self._total_requests += n

    def get_request_rate(self) -> float:
        with self.lock:
            now = datetime.now()
            duration = (now - self.last_reset).total_seconds()
            if duration >= self.window_size:
                rate = self.request_count / duration
                self.request_count = 0
                self.last_reset = now
                return rate
            if duration > 0:
                return self.request_count / duration
            return 0.0

    def get_total_requests(self) -> int:
        with self.lock:
            return self._total_requests


class RingRequestTracker(RequestTracker):
    """The same rate, read from the request ring's own counters (accepted + dropped submissions): clients that write into
    the shared-memory ring from other processes never pass through this process's submit_request, and the ring counts
    every submission wherever it came from.  record_request is therefore a no-op here (the ring already counted it)."""

    def __init__(self, queue, window_size: float = 60.0):
        super().__init__(window_size)
        self.queue = queue
        self._seen = self._submitted()
        self._window_base = self._seen

    def _submitted(self) -> int:
        o = self.queue.raw_stats()
        return int(o[0]) + int(o[1])

    def record_request(self, n: int = 1) -> None:
        return

    def get_request_rate(self) -> float:
        with self.lock:
            now = datetime.now()
            total = self._submitted()
            duration = (now - self.last_reset).total_seconds()
            count = total - self._window_base
            if duration >= self.window_size:
                self._window_base = total
                self.last_reset = now
                return count / duration
            return count / duration if duration > 0 else 0.0

    def get_total_requests(self) -> int:
        return self._submitted()


class MetricsWriter:
    """The collector half of MetricsDisplay (scheduler.py:933-983) without the terminal: every `update_interval` seconds
    `metrics.json` = {model: get_stats() + 'request_rate'} -- the schema the reference's viewers read
    (src/metrics_display.py) -- plus, under the extra key '_workers', what each actor reports about itself."""

    def __init__(self, update_interval: float = 1.0, metrics_file: str = "metrics.json"):
        self.update_interval = update_interval
        self.metrics_file = metrics_file
        self.stop_display = False
        self.collector_thread = None
        self.extra = None

    def collect(self, request_queues, request_trackers) -> dict:
        metrics = {}
        for model_name, queue in request_queues.items():
            stats = queue.get_stats()
            stats['request_rate'] = request_trackers[model_name].get_request_rate()
            metrics[model_name] = stats
        if self.extra is not None:
            metrics['_workers'] = self.extra()
        return metrics

    def write_once(self, request_queues, request_trackers) -> dict:
        metrics = self.collect(request_queues, request_trackers)
        tmp = self.metrics_file + ".tmp"
        with open(tmp, 'w') as f:
            json.dump(metrics, f)
        os.replace(tmp, self.metrics_file)          # readers never see a half-written file
        return metrics

    def start(self, request_queues, request_trackers):
        self.stop_display = False
        self.collector_thread = Thread(target=self._collect_metrics, args=(request_queues, request_trackers), daemon=True)
        self.collector_thread.start()

    def stop(self):
        self.stop_display = True
        if self.collector_thread is not None:
            self.collector_thread.join(timeout=self.update_interval + 1.0)
            self.collector_thread = None

    def _collect_metrics(self, request_queues, request_trackers):
        while not self.stop_display:
            try:
                self.write_once(request_queues, request_trackers)
            except Exception as e:  # noqa: BLE001
                logging.error("metrics writer: %s", e)
            t_end = time.time() + self.update_interval
            while not self.stop_display and time.time() < t_end:
                time.sleep(min(0.05, self.update_interval))


MetricsDisplay = MetricsWriter   # the reference's name


def lower_registry(model_registry: Dict) -> Dict:
    """name -> {'kind', 'blob', 'out_logical', 'in_hw'}: plans lowered once, picklable, what every actor loads."""
    from . import lowering
    from .worker import MODEL_KINDS
    out = {}
    for name, entry in model_registry.items():
        if isinstance(entry, dict) and 'blob' in entry:
            out[name] = entry
            continue
        kind = MODEL_KINDS.get(name, name)
        pb = lowering.lower(kind, entry)
        out[name] = {'kind': kind, 'blob': pb.to_blob(), 'out_logical': pb.out_logical, 'in_hw': (pb.in_h, pb.in_w)}
    return out


def _worker_main(node_id, gpu_index, request_queues, update_queue, registry, opts, report_queue, stop_event):
    """Body of one model-actor process: pins its GPU, then runs GPUWorker.execute_schedule until told to stop."""
    # what @ray.remote(num_gpus=1) does (scheduler.py:374): this process sees one GPU -- the gpu_index-th of those the
    # scheduler process itself was allowed to see
    visible = [v for v in os.environ.get("CUDA_VISIBLE_DEVICES", "").split(",") if v.strip()]
    os.environ["CUDA_VISIBLE_DEVICES"] = visible[gpu_index] if gpu_index < len(visible) else str(gpu_index)
    # An actor's host work is a serving loop, not tensor math: one intra-op thread.  Left at the default, every actor
    # starts a pool of as many OpenMP threads as the box has cores for its 1 MB output copies, and those pools spin after
    # each use -- eight actors on a 32-core box then fight over the cores and the loops slow down by an order of magnitude.
    os.environ.setdefault("OMP_NUM_THREADS", str(opts.get("cpu_threads", 1)))
    os.environ.setdefault("MKL_NUM_THREADS", str(opts.get("cpu_threads", 1)))
    logging.basicConfig(level=getattr(logging, opts.get("log_level", "WARNING")),
                        format='%(asctime)s - %(name)s - %(levelname)s - %(message)s')
    from .worker import B200GPUWorker
    import torch
    torch.set_num_threads(int(opts.get("cpu_threads", 1)))
    request_queues = {m: (q.bind(gpu_index) if hasattr(q, "bind") else q) for m, q in request_queues.items()}
    try:
        worker = B200GPUWorker(node_id, 0, [], 1, registry, max_batch=opts.get("max_batch", 256),
                               compat_overshoot=opts.get("compat", False), buckets=opts.get("buckets", "auto"),
                               pipeline_depth=opts.get("pipeline_depth", 2))
        for name in opts.get("preload", ()):                        # engines + graphs before the first schedule arrives
            worker.models[name] = worker._load(name, [(packing.session(name, 1000.0, 1.0, opts.get("max_batch", 256)), 1.0)])
        for q in request_queues.values():
            q.pin()
        report_queue.put(("ready", node_id, sorted(worker.models)))
    except Exception as e:  # noqa: BLE001
        report_queue.put(("failed", node_id, "%s\n%s" % (e, traceback.format_exc())))
        return

    def watch():
        stop_event.wait()
        worker.stop()
    Thread(target=watch, daemon=True).start()
    served = [0]
    keep = opts.get("results", False)

    def on_result(r):
        served[0] += len(r['request_ids']) if keep else 0
    worker.execute_schedule(request_queues, update_queue, on_result=on_result if keep else None)
    report_queue.put(("stopped", node_id, worker.get_stats()))
    for m in list(worker.models.values()):
        try:
            m.close()
        except Exception:  # noqa: BLE001
            pass


class B200Scheduler:
    def __init__(self, batching_profile: Dict[str, Dict[int, Dict[str, float]]], monitoring_interval: float = 5.0,
                 rate_change_threshold: float = 0.05, *, model_registry: Dict, num_gpus: Optional[int] = None,
                 max_queue_size: int = 2000, worker_mode: str = "process", policy=None, compat: bool = False,
                 pipeline_depth: int = 2, buckets="auto", max_batch: int = 256, in_flight_slots: Optional[int] = None,
                 metrics_file: Optional[str] = "metrics.json", metrics_interval: float = 5.0, preload=(),
                 slo_ms: Optional[Dict[str, float]] = None, tracker_window: float = 1.0, log_level: str = "WARNING",
                 max_ring_bytes: int = 4 << 30, shard_queues: bool = True):
        self.batching_profile = batching_profile
        self.nexus_instance = policy if policy is not None else packing.nexus(batching_profile)
        self.sessions: Dict[str, packing.session] = {}
        self.nodes: List[packing.node] = []
        self.request_trackers: Dict[str, RequestTracker] = {}
        self.monitoring_interval = monitoring_interval
        self.rate_change_threshold = rate_change_threshold
        self.lock = Lock()
        self.schedule_changes = Queue()
        self.monitoring_thread: Optional[Thread] = None
        self._stop_monitoring = False
        self.logger = logging.getLogger("NexusScheduler")
        self.worker_mode = worker_mode
        self.slo_ms = dict(slo_ms or {})
        self.max_ring_bytes = int(max_ring_bytes)
        self.shard_queues = bool(shard_queues)
        self.opts = dict(compat=compat, pipeline_depth=pipeline_depth, buckets=buckets, max_batch=max_batch,
                         preload=tuple(preload), log_level=log_level)
        self.model_registry = lower_registry(model_registry)
        if num_gpus is None:
            import torch
            num_gpus = torch.cuda.device_count()
        self.num_gpus = int(num_gpus)

        self.request_queues: Dict[str, ShardedRequestQueue] = {}
        self.update_queues = []
        self._init_queues(max_queue_size, in_flight_slots, tracker_window)
        self.workers = []
        self.futures = []
        self._reports = None
        self._stop_event = None
        self._worker_stats = {}
        self._init_workers()
        self._start_workers()

        self.metrics_display = MetricsWriter(update_interval=metrics_interval, metrics_file=metrics_file or "metrics.json")
        self.metrics_display.extra = lambda: dict(self._worker_stats)
        if metrics_file:
            self.metrics_display.start(self.request_queues, self.request_trackers)
        self.metrics: Dict[str, Dict] = {'schedule_updates': 0, 'total_requests': {}, 'rate_changes': {}, 'node_changes': {}}

    # ------------------------------------------------------------------------------------------ construction
    def _slo(self, model):
        return self.slo_ms.get(model, models_config.get(model, {}).get('SLO', 1000))

    def _init_queues(self, max_queue_size: int, in_flight_slots, tracker_window):
        """scheduler.py:654-661: one queue and one rate tracker per model (of the registry here; the reference creates them
        for every entry of models_config, served or not)."""
        for model, entry in self.model_registry.items():
            h, w = entry.get('in_hw', (224, 224))
            extra = in_flight_slots if in_flight_slots is not None else max(1024, 4 * self.opts["max_batch"] * max(1, self.num_gpus))
            # large requests (a 640 x 640 image is 1.2 MB): keep the segment, which every actor page-locks, within max_ring_bytes
            slot = (h * w * 3 + 255) // 256 * 256
            budget = max(64, int(self.max_ring_bytes // slot))
            if max_queue_size + extra > budget:
                max_queue_size_m = max(32, min(max_queue_size, budget * 3 // 5))
                extra = max(32, budget - max_queue_size_m)
            else:
                max_queue_size_m = max_queue_size
            # one ring per actor (each page-locked by its own actor only); a single ring when the actors are threads
            shards = max(1, self.num_gpus) if self.worker_mode == "process" and self.shard_queues else 1
            self.request_queues[model] = ShardedRequestQueue(model_name=model, max_size=max_queue_size_m,
                                                             batch_profile=self.batching_profile, shards=shards,
                                                             slot_bytes=h * w * 3, in_flight_slots=extra,
                                                             compat=self.opts["compat"], slo_ms=self._slo(model))
            self.request_trackers[model] = RingRequestTracker(self.request_queues[model], tracker_window)

    def _init_workers(self):
        """scheduler.py:670-693, for `num_gpus` actors instead of two."""
        if self.worker_mode == "process":
            import multiprocessing as mp
            ctx = mp.get_context("spawn")
            self._reports = ctx.Queue()
            self._stop_event = ctx.Event()
            for i in range(self.num_gpus):
                uq = ctx.Queue()
                self.update_queues.append(uq)
                p = ctx.Process(target=_worker_main, name="B200_%d" % i,
                                args=("B200_%d" % i, i, self.request_queues, uq, self.model_registry, self.opts,
                                      self._reports, self._stop_event), daemon=True)
                self.workers.append(p)
        elif self.worker_mode == "thread":
            from .worker import B200GPUWorker
            for i in range(max(1, self.num_gpus)):
                w = B200GPUWorker("B200_%d" % i, 0, [], 1, self.model_registry, max_batch=self.opts["max_batch"],
                                  compat_overshoot=self.opts["compat"], buckets=self.opts["buckets"],
                                  pipeline_depth=self.opts["pipeline_depth"])
                for name in self.opts["preload"]:
                    w.models[name] = w._load(name, [(packing.session(name, 1000.0, 1.0, self.opts["max_batch"]), 1.0)])
                self.workers.append(w)
                self.update_queues.append(Queue())
            for q in self.request_queues.values():
                q.pin()
        else:
            raise ValueError("worker_mode must be 'process' or 'thread'")

    def _start_workers(self):
        """scheduler.py:695-714."""
        self.start_time = time.time()
        self.logger.info("Starting schedule execution")
        if self.worker_mode == "process":
            for p in self.workers:
                p.start()
        else:
            for w, uq in zip(self.workers, self.update_queues):
                t = Thread(target=w.execute_schedule, args=(self.request_queues, uq), daemon=True)
                t.start()
                self.futures.append(t)

    def wait_ready(self, timeout: float = 600.0) -> bool:
        """Blocks until every actor process has loaded its preloaded models (process mode; threads load in __init__)."""
        if self.worker_mode != "process":
            return True
        ready, t_end = 0, time.time() + timeout
        while ready < len(self.workers) and time.time() < t_end:
            try:
                kind, node_id, info = self._reports.get(timeout=0.5)
            except Exception:  # noqa: BLE001
                if any(not p.is_alive() for p in self.workers):
                    raise RuntimeError("a model-actor process died while loading")
                continue
            if kind == "failed":
                raise RuntimeError("actor %s failed to start: %s" % (node_id, info))
            if kind == "ready":
                ready += 1
        return ready == len(self.workers)

    # ------------------------------------------------------------------------------------------ request path
    def submit_request(self, model_name: str, request_id: str, input_tensor) -> bool:
        """scheduler.py:734-751."""
        try:
            if model_name not in self.request_queues:
                self.logger.error("No queue found for model %s", model_name)
                return False
            success = self.request_queues[model_name].add_request(request_id, input_tensor)
            self.request_trackers[model_name].record_request()
            return success
        except Exception as e:  # noqa: BLE001
            self.logger.error("Error submitting request: %s", e)
            return False

    def submit_requests(self, model_name: str, images, id_prefix: str = "r", first_index: int = 0) -> int:
        """A burst of equal-size requests in one call (no counterpart in the reference): returns how many were queued."""
        n = self.request_queues[model_name].add_requests(images, id_prefix, first_index)
        self.request_trackers[model_name].record_request(len(images))
        return n

    # ------------------------------------------------------------------------------------------ rate monitor
    def start_monitoring(self) -> None:
        if self.monitoring_thread is not None:
            return
        self._stop_monitoring = False
        self.monitoring_thread = Thread(target=self._monitor_request_rates, daemon=True)
        self.monitoring_thread.start()
        self.logger.info("Request rate monitoring started")

    def stop_monitoring(self) -> None:
        self._stop_monitoring = True
        if self.monitoring_thread:
            self.monitoring_thread.join()
            self.monitoring_thread = None
        self.logger.info("Request rate monitoring stopped")

    def _monitor_request_rates(self) -> None:
        """scheduler.py:763-771."""
        while not self._stop_monitoring:
            try:
                self._check_and_update_schedules()
                t_end = time.time() + self.monitoring_interval
                while not self._stop_monitoring and time.time() < t_end:
                    time.sleep(min(0.05, self.monitoring_interval))
            except Exception as e:  # noqa: BLE001
                self.logger.error("Monitoring error: %s", e)
                self.logger.error("Stack trace: %s", traceback.format_exc())
                time.sleep(0.1)

    def _check_and_update_schedules(self):
        """scheduler.py:773-819: re-pack when a rate appears, or moves by more than the threshold (twice the threshold
        for a decrease)."""
        requires_update = False
        update_info = {}
        with self.lock:
            for model_name, tracker in self.request_trackers.items():
                current_rate = tracker.get_request_rate()
                if current_rate == 0:
                    continue
                if model_name not in self.sessions:
                    requires_update = True
                    update_info[model_name] = current_rate
                    continue
                previous_rate = self.sessions[model_name].request_rate
                rate_diff = current_rate - previous_rate
                factor = 2 if rate_diff < 0 else 1
                rate_diff = abs(rate_diff)
                if (rate_diff / previous_rate) > (self.rate_change_threshold * factor):
                    self.logger.info("Rate change detected for %s: %.2f req/s", model_name, current_rate)
                    requires_update = True
                    update_info[model_name] = current_rate
                    self.metrics['rate_changes'].setdefault(model_name, []).append(
                        {'timestamp': datetime.now(), 'old_rate': previous_rate, 'new_rate': current_rate})
        if requires_update:
            self._update_schedule(update_info)

    def get_transfers(self, old_nodes, new_nodes) -> int:
        """scheduler.py:821-832: models a re-arrangement has to load onto GPUs that do not hold them yet."""
        transfers = 0
        for old_node, new_node in zip(old_nodes, new_nodes):
            old_models = [s.model_name for s, _ in old_node.node_sessions]
            transfers += sum(1 for s, _ in new_node.node_sessions if s.model_name not in old_models)
        return transfers

    def _match(self, old_nodes, new_nodes):
        """scheduler.py:857-891: the assignment of new nodes to running workers with the fewest model transfers (exhaustive
        over permutations, like the reference; ties keep the first one found)."""
        l, n = len(old_nodes), len(new_nodes)
        k = min(l, n)
        best, best_cost = None, None
        for arrangement in itertools.permutations(range(n), k):
            cost = self.get_transfers(old_nodes[:k], [new_nodes[i] for i in arrangement])
            if best_cost is None or best_cost > cost:
                best, best_cost = arrangement, cost
        final = [new_nodes[i] for i in (best or ())]
        if l <= n:
            final += [new_nodes[i] for i in range(n) if i not in (best or ())]
        return final

    def _update_schedule(self, update_info: dict):
        """scheduler.py:834-904."""
        new_sessions = []
        for model, old_session in self.sessions.items():
            new_session = copy.deepcopy(old_session)
            if model in update_info:
                new_session.request_rate = update_info[model]
            if new_session.request_rate > 0:
                new_sessions.append(new_session)
        for model, request_rate in update_info.items():
            if model not in self.sessions:
                new_session = packing.session(model, self._slo(model) / SLO_hack, request_rate)
                if new_session.request_rate > 0:
                    new_sessions.append(new_session)
        new_nodes = self.nexus_instance.squishyBinPacking(new_sessions)
        final_nodes = self._match(self.nodes, new_nodes)
        self.sessions = {s.model_name: s for s in new_sessions}
        self.nodes = final_nodes
        self.metrics['schedule_updates'] += 1
        self.schedule_changes.put({'time': time.time(), 'nodes': [n.to_dict() for n in final_nodes]})
        self._update_workers(final_nodes)

    def set_schedule(self, nodes):
        """Installs a schedule as it is (static deployments and tests; the reference's draft executor does the same with a
        hard-coded schedule, venkat-code/test_script.py:60-81)."""
        self.nodes = list(nodes)
        self.sessions = {s.model_name: s for n in nodes for s, _ in n.node_sessions}
        self._update_workers(self.nodes)

    def _update_workers(self, new_nodes):
        """scheduler.py:906-925: (sessions, duty cycle, time) onto each worker's update queue; workers beyond the schedule
        get an empty one."""
        l, n = len(self.workers), len(new_nodes)
        if n > l:
            self.logger.warning("schedule needs %d GPUs, %d available: the last %d nodes are not served", n, l, n - l)
        for i in range(min(l, n)):
            self.update_queues[i].put((new_nodes[i].node_sessions, new_nodes[i].duty_cycle, time.time()))
        for i in range(n, l):
            self.update_queues[i].put(([], 1, time.time()))

    # ------------------------------------------------------------------------------------------ shutdown
    def stop(self, timeout: float = 30.0):
        """Ends the monitor, the metrics writer and every actor; returns the actors' final stats."""
        if self.monitoring_thread is not None:
            self.stop_monitoring()
        if self.worker_mode == "process":
            self._stop_event.set()
            t_end = time.time() + timeout
            pending = len(self.workers)
            while pending and time.time() < t_end:
                try:
                    kind, node_id, info = self._reports.get(timeout=0.5)
                except Exception:  # noqa: BLE001
                    if not any(p.is_alive() for p in self.workers):
                        break
                    continue
                if kind == "stopped":
                    self._worker_stats[node_id] = info
                    pending -= 1
            for p in self.workers:
                p.join(timeout=5.0)
                if p.is_alive():
                    p.terminate()
        else:
            for w in self.workers:
                w.stop()
            for t in self.futures:
                t.join(timeout=timeout)
            for w in self.workers:
                self._worker_stats[w.node_id] = w.get_stats()
        try:
            if self.metrics_display.collector_thread is not None:
                self.metrics_display.stop()
                self.metrics_display.write_once(self.request_queues, self.request_trackers)
        except Exception as e:  # noqa: BLE001
            self.logger.error("final metrics write failed: %s", e)
        stats = dict(self._worker_stats)
        return stats

    def close(self):
        for q in self.request_queues.values():
            q.close()


NexusScheduler = B200Scheduler   # the reference's name
