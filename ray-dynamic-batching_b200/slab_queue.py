"""SlabRequestQueue: the reference's per-model `RequestQueue` (293-project/src/scheduler.py:190-372) over the native
request ring of libb200infer (include/b200infer.h, b200_ring_*) in a shared-memory segment.

Same surface as `RequestQueue` -- add_request / get_batch / record_batch_completion / get_stats / empty -- so the
scheduler's submit path and the actor loop use it unchanged; what changes is what a request costs:

  reference                                                        here
  add_request: 3 Ray RPCs + a pickled 602 KB fp32 tensor           one memcpy of the request's uint8 bytes into a slot
  get_batch:   1 RPC per item + a Python loop                      one native call claiming a run of slots
  metrics:     a dict pickled INTO each actor (the driver's copy   counters and the latency window live in the segment:
               never sees completions)                 every process that maps it reads the same numbers

The segment is created by the scheduler process and attached by name in every model-actor process (the object pickles as
its name), where `pin()` registers the payload area with CUDA: a batch then goes to the GPU by DMA from the very bytes the
client wrote -- at most two copies per batch (one per contiguous run of slots) and none of them per request.

`compat=True` keeps the reference's overshoot (rounds of up to batch_size requests while count <= batch_size,
scheduler.py:274-276), with two documented differences: the clock is read once per round, not once per item, and an
empty queue ends the call instead of spinning.  `compat=False` is the clean semantics (at most batch_size requests,
venkat-code/schedule_executor.py:85-119).  The in-process B200RequestQueue remains the bit-exact, golden-pinned mirror
for the reference's own request type (fp32 tensors of any shape).
"""
import ctypes
import logging
import time
from multiprocessing import shared_memory
from typing import Dict, Optional

import numpy as np

from . import lib as _lib
from .request_queue import BatchRequest, models_config

ID_BYTES = 100   # B200_RING_ID_BYTES
_CREATED_HERE = set()   # segments this process created (their tracker entry belongs to the creator)


class SlabInputs:
    """`BatchRequest.inputs` of a slab batch: a sequence of uint8 HWC views into the segment, materialised on demand (the
    hot path never touches them: it copies whole runs of slots)."""

    def __init__(self, queue, slots, hw, channels=3):
        self._q, self._slots, self._hw, self._c = queue, slots, hw, channels

    def __len__(self):
        return len(self._slots)

    def __getitem__(self, i):
        import torch
        h, w = int(self._hw[i, 0]), int(self._hw[i, 1])
        c = self._c if w > 0 else 1                      # negative width: single-channel request
        w = abs(w)
        off = int(self._slots[i]) * self._q.slot_bytes
        return torch.from_numpy(self._q.payload[off:off + h * w * c].reshape(h, w, c))

    def __iter__(self):
        return (self[i] for i in range(len(self)))


class SlabRequestQueue:
    """Drop-in for `RequestQueue(model_name, max_size, batch_profile)` (scheduler.py:190-222) for uint8 HWC requests."""

    def __init__(self, model_name: str, max_size: int = 2000, batch_profile: dict = None, *, slot_bytes: int = 224 * 224 * 3,
                 in_flight_slots: int = 2048, compat: bool = False, clock=time.time, slo_ms: Optional[float] = None,
                 name: Optional[str] = None, create: bool = True):
        self.model_name = model_name
        self.compat = compat
        self.clock = clock
        self._logger_name = "Queue-%s" % model_name
        self.slo_target = slo_ms if slo_ms is not None else models_config[model_name]['SLO']
        batch_profile = batch_profile if batch_profile is not None else {}
        self.profile = batch_profile.get(model_name, {})
        self._lib = _lib.load()
        self._pinned = False
        self._owner = create
        if create:
            nslots = int(max_size) + int(in_flight_slots)
            nbytes = int(self._lib.b200_ring_bytes(nslots, int(slot_bytes)))
            self._shm = shared_memory.SharedMemory(create=True, size=nbytes + 4096, name=name)
            _CREATED_HERE.add(self._shm.name)
        else:
            self._shm = shared_memory.SharedMemory(name=name)
            # An attaching process must not unlink the segment when it exits, but Python < 3.13 registers it with the
            # process's resource tracker anyway.  A multiprocessing child shares its parent's tracker (registering
            # twice is harmless there, unregistering would drop the creator's entry); an unrelated process has a tracker
            # of its own, which would unlink the segment behind the creator's back: take the entry out of that one.
            try:
                import multiprocessing
                from multiprocessing import resource_tracker
                child = multiprocessing.parent_process() is not None or \
                    getattr(multiprocessing.current_process(), "_inheriting", False)   # still unpickling its arguments
                if not child and self._shm.name not in _CREATED_HERE:
                    resource_tracker.unregister(self._shm._name, "shared_memory")
            except Exception:  # noqa: BLE001
                pass
        self.name = self._shm.name
        self._map()
        if create:
            _lib.check(self._lib.b200_ring_init(self._mem, self._nbytes, nslots, int(slot_bytes), int(max_size)),
                       "b200_ring_init")
        self._layout()

    # ---- mapping ----
    def _map(self):
        self._arr = np.frombuffer(self._shm.buf, dtype=np.uint8)
        base = self._arr.ctypes.data
        self._skip = (-base) % 4096                      # the ring wants an aligned start; mmap gives page alignment anyway
        self._mem = ctypes.c_void_p(base + self._skip)
        self._nbytes = self._arr.size - self._skip

    def _layout(self):
        ns, sb, mq = ctypes.c_uint32(), ctypes.c_uint32(), ctypes.c_uint32()
        po, tb = ctypes.c_uint64(), ctypes.c_uint64()
        _lib.check(self._lib.b200_ring_layout(self._mem, ctypes.byref(ns), ctypes.byref(sb), ctypes.byref(mq),
                                              ctypes.byref(po), ctypes.byref(tb)), "b200_ring_layout")
        self.nslots, self.slot_bytes, self.max_size = ns.value, sb.value, mq.value
        start = self._skip + po.value
        self.payload = self._arr[start:start + self.nslots * self.slot_bytes]     # numpy view of the payload area
        self._scratch = {}
        self.payload_tensor = None

    def __getstate__(self):
        return dict(model_name=self.model_name, name=self.name, compat=self.compat, slo=self.slo_target,
                    profile=self.profile)

    def __setstate__(self, st):
        self.__init__(st["model_name"], batch_profile={st["model_name"]: st["profile"]}, compat=st["compat"],
                      slo_ms=st["slo"], name=st["name"], create=False)

    def pin(self):
        """Registers the payload area with CUDA in THIS process (page-locked: async DMA straight from the segment)."""
        if self._pinned:
            return
        import torch
        ptr = self.payload.ctypes.data
        nbytes = self.payload.size
        rc = torch.cuda.cudart().cudaHostRegister(ptr, nbytes, 0)
        if int(rc) != 0:
            raise _lib.B200Error("cudaHostRegister of the request slab failed (%s)" % rc)
        self._pinned = True
        self.payload_tensor = torch.from_numpy(self.payload)

    def close(self, unlink=None):
        try:
            if self._pinned:
                import torch
                torch.cuda.cudart().cudaHostUnregister(self.payload.ctypes.data)
                self._pinned = False
        except Exception:  # noqa: BLE001
            pass
        self.payload = None
        self.payload_tensor = None
        self._arr = None
        self._mem = None
        self._scratch = {}
        try:
            self._shm.close()
        except Exception:  # noqa: BLE001  (a view of the segment is still referenced somewhere: the mapping goes with it)
            pass
        try:
            if self._owner if unlink is None else unlink:
                self._shm.unlink()
        except Exception:  # noqa: BLE001
            pass

    @property
    def _logger(self):
        return logging.getLogger(self._logger_name)

    # ---- RequestQueue surface ----
    def qsize(self) -> int:
        return int(self._lib.b200_ring_qsize(self._mem))

    def empty(self) -> bool:
        return self.qsize() == 0

    def add_request(self, request_id: str, input_tensor) -> bool:
        """scheduler.py:238-254: False when the queue is full (request dropped and counted) or on any error.
        input_tensor: uint8 HWC image (torch tensor or numpy array, contiguous)."""
        try:
            arr = input_tensor.numpy() if hasattr(input_tensor, "numpy") else np.asarray(input_tensor)
            if arr.dtype != np.uint8 or arr.ndim != 3:
                raise TypeError("SlabRequestQueue takes uint8 HWC images (got %s %s)" % (arr.dtype, arr.shape))
            if not arr.flags.c_contiguous:
                arr = np.ascontiguousarray(arr)
            rc = self._lib.b200_ring_push(self._mem, request_id.encode(), arr.ctypes.data, arr.size, arr.shape[0],
                                          arr.shape[1], arr.shape[2], self.clock())
            if rc < 0:
                _lib.check(rc, "b200_ring_push")
            return rc == 1
        except Exception as e:  # noqa: BLE001 - the reference logs and returns False
            self._logger.error("Error adding request: %s", e)
            return False

    def add_requests(self, images, id_prefix="r", first_index=0) -> int:
        """Burst submit: `images` uint8 [n, H, W, C] contiguous; ids are id_prefix + (first_index + i).  One native call;
        returns how many were queued (the rest met a full queue)."""
        arr = images.numpy() if hasattr(images, "numpy") else np.asarray(images)
        n, h, w, c = arr.shape
        rc = self._lib.b200_ring_push_many(self._mem, n, None, id_prefix.encode(), int(first_index), arr.ctypes.data,
                                           arr.strides[0], h * w * c, h, w, c, self.clock())
        if rc < 0:
            _lib.check(rc, "b200_ring_push_many")
        return rc

    def _bufs(self, n):
        b = self._scratch.get(n)
        if b is None:
            b = (np.empty(n, np.uint32), np.empty(n, np.uint64), np.empty(n, np.float64), np.empty((n, 2), np.int32),
                 np.empty(n, np.uint32), np.zeros((n, ID_BYTES), np.uint8))
            self._scratch[n] = b
        return b

    def _pop(self, max_items, batch_size):
        slots, seqs, arr, hw, nb, ids = self._bufs(max_items)
        disc = ctypes.c_int()
        lat_s = self.profile[batch_size]['avg_latency_ms'] / 1000
        n = self._lib.b200_ring_pop(self._mem, max_items, self.clock(), lat_s, self.slo_target / 1000, slots.ctypes.data,
                                    seqs.ctypes.data, arr.ctypes.data, hw.ctypes.data, nb.ctypes.data, ids.ctypes.data,
                                    ctypes.byref(disc))
        if n < 0:
            _lib.check(n, "b200_ring_pop")
        return n, disc.value, (slots[:n].copy(), seqs[:n].copy(), arr[:n].copy(), hw[:n].copy(), ids[:n].copy())

    def get_batch(self, batch_size: int):
        """scheduler.py:258-322. Returns (BatchRequest | None, discarded_count)."""
        try:
            if min(batch_size, self.qsize()) == 0:
                return None, 0
            parts, discard, count = [], 0, 0
            limit = batch_size if self.compat else batch_size - 1          # compat: `while count <= batch_size`
            while count <= limit:
                want = min(batch_size, self.qsize())
                if not self.compat:
                    want = min(want, batch_size - count)
                if want == 0:
                    break                                                   # the reference spins here; we return
                n, d, part = self._pop(want, batch_size)
                discard += d
                if n == 0 and d == 0:
                    break                                                   # front request still being written
                if n:
                    parts.append(part)
                    count += n
            if count:
                slots, seqs, arr, hw, ids = (np.concatenate([p[i] for p in parts]) for i in range(5))
                return SlabBatch(self, slots, seqs, arr, hw, ids), discard
            return None, 0 if self.compat else discard
        except Exception as e:  # noqa: BLE001 - reference: log, fall through to (None, 0)
            self._logger.error("Error creating batch: %s", e)
        return None, 0

    def release(self, batch):
        """Gives the batch's slots back (call once its host-to-device copy has completed)."""
        if getattr(batch, "_released", True):
            return
        batch._released = True
        _lib.check(self._lib.b200_ring_release(self._mem, batch.seqs.ctypes.data, len(batch.seqs)), "b200_ring_release")

    def record_batch_completion(self, batch: BatchRequest, completion_time: float) -> int:
        """scheduler.py:324-341."""
        latency = (completion_time - batch.arrival_time) * 1000
        rc = self._lib.b200_ring_record(self._mem, latency, len(batch.request_ids), float(self.slo_target))
        if rc < 0:
            _lib.check(rc, "b200_ring_record")
        return rc

    def raw_stats(self):
        out = np.zeros(12, np.float64)
        _lib.check(self._lib.b200_ring_stats(self._mem, out.ctypes.data, 12), "b200_ring_stats")
        return out

    def get_stats(self) -> Dict:
        """scheduler.py:343-372 (same keys; integers where the reference has integers), plus three counters of ours."""
        o = self.raw_stats()
        stats = {'total_requests': int(o[0]), 'dropped_requests': int(o[1]), 'slo_violations': int(o[2]),
                 'queue_size': int(o[3]), 'last_latency': float(o[4])}
        if int(o[9]) == 0:
            stats.update(avg_latency=0, p95_latency=0, p99_latency=0)
        else:
            stats.update(avg_latency=float(o[5]), p95_latency=float(o[6]), p99_latency=float(o[7]))
        stats.update(discarded_stale=int(o[8]), completed_requests=int(o[9]), completed_batches=int(o[10]),
                     max_latency=float(o[11]))
        return stats


class SlabBatch(BatchRequest):
    """BatchRequest (scheduler.py:181-188) whose inputs are slots of the queue's segment."""

    def __init__(self, queue, slots, seqs, arrivals, hw, ids):
        self.model_name = queue.model_name
        self.queue = queue
        self.slots, self.seqs, self.arrivals, self.hw, self._ids = slots, seqs, arrivals, hw, ids
        self.batch_size = len(slots)
        self.arrival_time = float(arrivals.min())
        self.inputs = SlabInputs(queue, slots, hw)
        self._request_ids = None
        self._released = False
        self.slab_offsets = None
        self.slab = None

    @property
    def request_ids(self):
        if self._request_ids is None:                 # decoded on first use (the serving loop only needs the count)
            raw = self._ids.tobytes()
            self._request_ids = [raw[i * ID_BYTES:(i + 1) * ID_BYTES].split(b"\0", 1)[0].decode()
                                 for i in range(self.batch_size)]
        return self._request_ids

    def runs(self):
        """Contiguous runs of slots as (first_item, n_items, first_slot): a FIFO batch is one run, two when it wraps
        the ring, a few more when stale requests were dropped from its middle or another actor popped in between."""
        s = self.slots.astype(np.int64)
        brk = np.flatnonzero(np.diff(s) != 1) + 1
        starts = np.concatenate(([0], brk))
        ends = np.concatenate((brk, [len(s)]))
        return [(int(a), int(b - a), int(s[a])) for a, b in zip(starts, ends)]

    def __len__(self):
        return self.batch_size


def _latency_window(q):
    out = np.zeros(1024, np.float64)
    n = q._lib.b200_ring_latencies(q._mem, out.ctypes.data, 1024)
    if n < 0:
        _lib.check(n, "b200_ring_latencies")
    return out[:n]


SlabRequestQueue.latency_window = _latency_window


class ShardedRequestQueue:
    """One model's RequestQueue (scheduler.py:190-372) as one request ring PER MODEL ACTOR.

    The reference has a single queue per model that every actor pulls from (scheduler.py:546-551); with one shared
    segment page-locked by eight actor processes and written by two dozen clients, that single ring is also the single
    point every process meets.  Here each actor owns a ring (and is the only process that page-locks it): a submission goes
    to one shard, round-robin from the submitter's preferred shard, spilling to the next when that one is full; an actor
    serves its own shard and takes from another only when its own is empty (the work sharing the shared FIFO gave).
    Same five methods, so the scheduler and the actor loop use it unchanged; `bind(i)` gives actor or client i its view.
    get_stats sums the counters and takes avg / p95 / p99 over the shards' latency windows together -- with one shard this
    is exactly the single-ring result."""

    def __init__(self, model_name: str, max_size: int = 2000, batch_profile: dict = None, *, shards: int = 1,
                 slot_bytes: int = 224 * 224 * 3, in_flight_slots: int = 2048, compat: bool = False, clock=time.time,
                 slo_ms: Optional[float] = None, names=None):
        self.model_name = model_name
        self.compat, self.clock = compat, clock
        self.profile = (batch_profile or {}).get(model_name, {})
        if names is None:
            per = max(32, -(-int(max_size) // shards))
            self.shards = [SlabRequestQueue(model_name, per, batch_profile, slot_bytes=slot_bytes,
                                            in_flight_slots=max(32, -(-int(in_flight_slots) // shards)), compat=compat,
                                            clock=clock, slo_ms=slo_ms) for _ in range(shards)]
        else:
            self.shards = [SlabRequestQueue(model_name, batch_profile=batch_profile, compat=compat, clock=clock, slo_ms=slo_ms,
                                            name=n, create=False) for n in names]
        self.slo_target = self.shards[0].slo_target
        self.slot_bytes = self.shards[0].slot_bytes
        self.max_size = sum(q.max_size for q in self.shards)
        self.home = 0                 # this process's preferred shard
        self._next = 0

    def bind(self, index: int):
        self.home = int(index) % len(self.shards)
        self._next = self.home
        return self

    def __getstate__(self):
        return dict(model_name=self.model_name, names=[q.name for q in self.shards], compat=self.compat,
                    slo=self.slo_target, profile=self.profile, home=self.home)

    def __setstate__(self, st):
        self.__init__(st["model_name"], batch_profile={st["model_name"]: st["profile"]}, compat=st["compat"],
                      slo_ms=st["slo"], names=st["names"])
        self.bind(st["home"])

    # ---- RequestQueue surface ----
    def qsize(self) -> int:
        return sum(q.qsize() for q in self.shards)

    def empty(self) -> bool:
        return all(q.empty() for q in self.shards)

    def add_request(self, request_id: str, input_tensor) -> bool:
        n = len(self.shards)
        first = self._next
        self._next = (first + 1) % n
        for k in range(n):
            q = self.shards[(first + k) % n]
            if k < n - 1 and q.qsize() >= q.max_size:
                continue                             # full: try the next shard before counting a drop
            if q.add_request(request_id, input_tensor):
                return True
        return False

    def add_requests(self, images, id_prefix="r", first_index=0) -> int:
        n = len(self.shards)
        first = self._next
        self._next = (first + 1) % n
        total, done = len(images), 0
        for k in range(n):
            q = self.shards[(first + k) % n]
            if k < n - 1 and q.qsize() + (total - done) > q.max_size:
                continue
            done += q.add_requests(images[done:], id_prefix, first_index + done)
            if done >= total:
                break
        return done

    def get_batch(self, batch_size: int):
        n = len(self.shards)
        own = self.shards[self.home]
        batch, discard = own.get_batch(batch_size)
        if batch is not None or n == 1:
            return batch, discard
        for k in range(1, n):                        # own shard empty: share another actor's work
            q = self.shards[(self.home + k) % n]
            if q.qsize() == 0:
                continue
            b2, d2 = q.get_batch(batch_size)
            if b2 is not None:
                return b2, discard + d2
            discard += d2
        return None, 0 if self.compat else discard

    def release(self, batch):
        batch.queue.release(batch)

    def record_batch_completion(self, batch, completion_time: float) -> int:
        q = getattr(batch, "queue", None) or self.shards[self.home]
        return q.record_batch_completion(batch, completion_time)

    def pin(self):
        self.shards[self.home].pin()                 # the others are page-locked the first time this actor takes from them

    def raw_stats(self):
        return np.sum([q.raw_stats() for q in self.shards], axis=0)

    def get_stats(self) -> Dict:
        per = [q.raw_stats() for q in self.shards]
        tot = np.sum(per, axis=0)
        stats = {'total_requests': int(tot[0]), 'dropped_requests': int(tot[1]), 'slo_violations': int(tot[2]),
                 'queue_size': int(tot[3]), 'last_latency': float(max(p[4] for p in per))}
        lat = np.concatenate([q.latency_window() for q in self.shards]) if int(tot[9]) else np.zeros(0)
        if lat.size == 0:
            stats.update(avg_latency=0, p95_latency=0, p99_latency=0)
        else:
            s = np.sort(lat)
            stats.update(avg_latency=float(sum(lat.tolist()) / lat.size), p95_latency=float(s[int(lat.size * 0.95)]),
                         p99_latency=float(s[int(lat.size * 0.99)]))
        stats.update(discarded_stale=int(tot[8]), completed_requests=int(tot[9]), completed_batches=int(tot[10]),
                     max_latency=float(max(p[11] for p in per)), shards=len(self.shards))
        return stats

    def close(self, unlink=None):
        for q in self.shards:
            q.close(unlink)
