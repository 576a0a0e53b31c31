"""Host-side mirror of the reference's request queue and batch record (293-project/src/scheduler.py:181-372).

Same names, argument meaning and return values as `BatchRequest` / `RequestQueue`, so the scheduler and the worker loop
use it unchanged:  add_request -> get_batch -> (worker) -> record_batch_completion -> get_stats.

Differences, all opt-in or forced by the missing Ray runtime:
  * the FIFO is an in-process deque by default; any object with qsize/full/put/get (e.g. ray.util.queue.Queue when Ray
    is installed) can be passed as `backing`;
  * `compat=True` (default) reproduces the live reference `get_batch` bit for bit, including its overshoot to
    (batch_size, 2*batch_size] requests (scheduler.py:274 `<=`); where the reference would busy-wait forever on an empty
    queue (scheduler.py:274-276) this returns what it has collected instead.  `compat=False` is the clean draft
    semantics (venkat-code/schedule_executor.py:85-119: at most batch_size requests) plus the same stale filter;
  * requests may carry uint8 HWC images (engine-native; collated and normalised on the GPU) instead of fp32 CHW tensors;
    with `pinned_slab_bytes > 0` their bytes are copied once, at submit time, into a pinned ring so that a batch goes to
    the GPU with a single H2D copy instead of B pickled 602 KB tensors.
"""
import logging
import time
from collections import deque
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional

models_config = {                      # scheduler.py:30-35 (SLO in ms)
    'vit': {'SLO': 4000, 'base_rate': 0},
    'resnet': {'SLO': 2000, 'base_rate': 0},
    'shufflenet': {'SLO': 1500, 'base_rate': 0},
    'efficientnet': {'SLO': 40, 'base_rate': 0},
    'yolov8': {'SLO': 100, 'base_rate': 0},   # not in the reference (YOLOv8 is workload-only there); ours
}


@dataclass
class BatchRequest:
    """scheduler.py:181-188."""
    model_name: str
    inputs: List[Any]
    batch_size: int
    request_ids: List[str]
    arrival_time: float
    slab_offsets: Optional[List[int]] = field(default=None, compare=False)   # pinned-slab positions (engine-native)
    slab: Any = field(default=None, compare=False, repr=False)               # the PinnedSlab those offsets refer to


class _DequeBacking:
    def __init__(self, maxsize=0):
        self.maxsize = maxsize
        self.q = deque()

    def qsize(self):
        return len(self.q)

    def full(self):
        return self.maxsize > 0 and len(self.q) >= self.maxsize

    def put(self, item):
        self.q.append(item)

    def get(self):
        return self.q.popleft()


class PinnedSlab:
    """Ring of pinned host memory holding uint8 request payloads in arrival order (16-byte aligned slots).

    A payload stays where it is until it is released -- by get_batch for a request it discards, by
    record_batch_completion for the requests of a finished batch.  push() never overwrites an unreleased payload: when
    the ring has no room it returns None and the request is queued without a slab position (its tensor still travels
    in the tuple, so the worker packs it the slow way)."""

    def __init__(self, nbytes):
        import torch
        self.buf = torch.empty(nbytes, dtype=torch.uint8, pin_memory=torch.cuda.is_available())
        self.nbytes = nbytes
        self.head = 0                 # next write position
        self.live = deque()           # [offset, need, released] in push order; live[0] is the oldest payload still held
        self.by_off = {}

    def _place(self, need):
        """Offset for `need` bytes that overlaps no held payload, or None.  Held payloads form one arc of the ring from the
        oldest (live[0]) to head; releases out of order only leave holes inside the arc, which are not reused."""
        if not self.live:
            self.head = 0
            return 0 if need <= self.nbytes else None
        tail = self.live[0][0]
        if self.head > tail:                                   # arc [tail, head): free space is [head, end) and [0, tail)
            if self.head + need <= self.nbytes:
                return self.head
            return 0 if need <= tail else None
        return self.head if self.head + need <= tail else None  # arc wraps the end: free space is [head, tail)

    def push(self, tensor):
        n = tensor.numel()
        need = (n + 15) // 16 * 16
        if need > self.nbytes:
            raise ValueError("request larger than the pinned slab")
        off = self._place(need)
        if off is None:
            return None
        self.buf[off:off + n].copy_(tensor.reshape(-1))
        ent = [off, need, False]
        self.live.append(ent)
        self.by_off[off] = ent
        self.head = off + need
        return off

    def release(self, off):
        ent = self.by_off.pop(off, None)
        if ent is None:
            return
        ent[2] = True
        while self.live and self.live[0][2]:
            self.live.popleft()
        if not self.live:
            self.head = 0


class B200RequestQueue:
    """Drop-in for `RequestQueue(model_name, max_size, batch_profile)` (scheduler.py:190-222)."""

    def __init__(self, model_name: str, max_size: int = 100, batch_profile: dict = None, *, compat: bool = True,
                 clock=time.time, backing=None, slo_ms: Optional[float] = None, pinned_slab_bytes: int = 0):
        batch_profile = batch_profile if batch_profile is not None else {}
        self.model_name = model_name
        self.queue = backing if backing is not None else _DequeBacking(maxsize=max_size)
        self._logger_name = "Queue-%s" % model_name
        self.slo_target = slo_ms if slo_ms is not None else models_config[model_name]['SLO']
        self.profile = batch_profile.get(model_name, {})
        self.compat = compat
        self.clock = clock
        self.slab = PinnedSlab(pinned_slab_bytes) if pinned_slab_bytes > 0 else None
        self.metrics = {'total_requests': 0, 'dropped_requests': 0, 'slo_violations': 0, 'latencies': [],
                        'queue_size': 0, 'last_latency': 0.0}

    @property
    def _logger(self):
        return logging.getLogger(self._logger_name)

    def empty(self) -> bool:
        return self.queue.qsize() == 0

    def add_request(self, request_id: str, input_tensor) -> bool:
        """scheduler.py:238-254: False when the queue is full (request dropped) or on any error."""
        try:
            if self.queue.full():
                self.metrics['dropped_requests'] += 1
                return False
            off = None
            if self.slab is not None and getattr(input_tensor, "dtype", None) is not None \
                    and str(input_tensor.dtype) == "torch.uint8":
                off = self.slab.push(input_tensor)
            self.queue.put((request_id, input_tensor, self.clock()) if off is None
                           else (request_id, input_tensor, self.clock(), off))
            self.metrics['total_requests'] += 1
            self.metrics['queue_size'] = self.queue.qsize()
            return True
        except Exception as e:  # noqa: BLE001 - the reference logs and returns False
            self._logger.error("Error adding request: %s", e)
            return False

    def _stale(self, arrival_time: float, batch_size: int) -> bool:
        """scheduler.py:281: arrival + SLO < now + expected batch latency."""
        eta = self.clock() + (self.profile[batch_size]['avg_latency_ms'] / 1000)
        return (arrival_time + (self.slo_target / 1000)) < eta

    def get_batch(self, batch_size: int):
        """scheduler.py:258-322. Returns (BatchRequest | None, discarded_count)."""
        inputs, request_ids, offs = [], [], []
        earliest_arrival = float('inf')
        try:
            available = min(batch_size, self.queue.qsize())
            if available == 0:
                return None, 0
            discard = 0
            count = 0
            limit = batch_size if self.compat else batch_size - 1       # compat: `while count <= batch_size`
            while count <= limit:
                available = min(batch_size, self.queue.qsize())
                if not self.compat:
                    available = min(available, batch_size - count)
                if available == 0:
                    break                                               # the reference spins here; we return
                for _ in range(available):
                    item = self.queue.get()
                    request_id, input_tensor, arrival_time = item[0], item[1], item[2]
                    if self._stale(arrival_time, batch_size):
                        discard += 1
                        if len(item) > 3 and item[3] is not None and self.slab is not None:
                            self.slab.release(item[3])
                        continue
                    request_ids.append(request_id)
                    inputs.append(input_tensor)
                    offs.append(item[3] if len(item) > 3 else None)
                    earliest_arrival = min(earliest_arrival, arrival_time)
                    count += 1
            if inputs:
                slab_offsets = offs if all(o is not None for o in offs) else None
                if slab_offsets is None and self.slab is not None:
                    for o in offs:                 # mixed batch: it travels by its tensors, the slab copies are not needed
                        if o is not None:
                            self.slab.release(o)
                return BatchRequest(self.model_name, inputs, len(inputs), request_ids, earliest_arrival,
                                    slab_offsets, self.slab if slab_offsets is not None else None), discard
        except Exception as e:  # noqa: BLE001 - reference: log, fall through to (None, 0)
            self._logger.error("Error creating batch: %s", e)
        return None, 0

    def record_batch_completion(self, batch: BatchRequest, completion_time: float) -> int:
        """scheduler.py:324-341."""
        latency = (completion_time - batch.arrival_time) * 1000
        count = 0
        if getattr(batch, "slab", None) is not None and batch.slab_offsets:
            for off in batch.slab_offsets:                     # the batch has left the host: its payloads may be overwritten
                batch.slab.release(off)
            batch.slab_offsets = None
        self.metrics['last_latency'] = latency
        for _ in batch.request_ids:
            if latency > self.slo_target:
                self.metrics['slo_violations'] += 1
                count += 1
            if len(self.metrics['latencies']) >= 1000:
                self.metrics['latencies'] = self.metrics['latencies'][100:]
            self.metrics['latencies'].append(latency)
        self.metrics['queue_size'] = self.queue.qsize()
        return count

    def get_stats(self) -> Dict:
        """scheduler.py:343-372."""
        latencies = self.metrics['latencies']
        stats = {k: self.metrics[k] for k in ('total_requests', 'dropped_requests', 'slo_violations', 'queue_size',
                                              'last_latency')}
        if not latencies:
            stats.update(avg_latency=0, p95_latency=0, p99_latency=0)
            return stats
        s = sorted(latencies)
        stats.update(avg_latency=sum(latencies) / len(latencies), p95_latency=s[int(len(s) * 0.95)],
                     p99_latency=s[int(len(s) * 0.99)])
        return stats


RequestQueue = B200RequestQueue   # the reference's name
