"""Host-side mirror of the reference's model actor `GPUWorker` (293-project/src/scheduler.py:374-600) for one B200.

Same constructor arguments, methods and return values; the body of `process_batch` -- the reference's
`torch.stack(batch.inputs).cuda()` -> `model(inputs)` -> `outputs.cpu()` (scheduler.py:441-452,468) -- is replaced by
the C-ABI engine: the batch's slots go to the GPU with one DMA per contiguous run, then the fused collate kernel, the
CUDA-graph forward of the batch's bucket, one D2H copy.

Additions the reference does not have (each opt-in through an argument, defaults keep the reference's behaviour):
  * `pipeline_depth=2`: `process_batch_async` returns a handle; the loop forms and uploads batch i+1 while batch i
    computes, which is what keeps a B200 busy (a synchronous loop leaves it idle for the H2D copy and the Python
    between two batches);
  * batch-size buckets (`buckets="auto"`): a batch of n requests runs as the next prepared size with zero images in
    the padding rows, so only graphs captured at load are ever replayed;
  * engines are sized from the sessions (2 x the largest batch size under the compat overshoot) and an oversize batch
    is split, never refused.

The reference makes this class a Ray actor (`@ray.remote(num_gpus=1)`, scheduler.py:374): Ray pins one GPU per actor
process through CUDA_VISIBLE_DEVICES, so the actor always uses cuda:0 (scheduler.py:388,404).  Ray is not part of this
image; `as_ray_actor()` wraps the class when Ray is importable, rdb_b200.scheduler runs it in one process per GPU.
"""
import logging
import time
from collections import deque
from threading import Lock
from typing import Dict, List, Tuple

import numpy as np
import torch

from .engine import B200Model, COLLATE_IDENTITY, COLLATE_LETTERBOX, COLLATE_RESIZE_SHORT_AA
from .request_queue import BatchRequest

YOLO_NMS = {"conf": 0.25, "iou": 0.45, "max_det": 300}      # ultralytics' defaults (utils/ops.py non_max_suppression)
MODEL_KINDS = {'resnet': 'resnet50', 'shufflenet': 'shufflenet_v2_x1_0', 'efficientnet': 'efficientnet_v2_s',
               'yolov8': 'yolov8n'}


class _Pending:
    """One batch in flight: its engine handle(s) and what process_batch needs to finish it."""
    __slots__ = ("batch", "handles", "counts", "start_time", "queue", "copied")

    def __init__(self, batch, start_time):
        self.batch, self.start_time = batch, start_time
        self.handles, self.counts = [], []
        self.queue = None
        self.copied = None


class B200GPUWorker:
    def __init__(self, node_id: str, gpu_id: int, sessions: List[Tuple], duty_cycle: float, model_registry: Dict,
                 max_batch: int = 256, compat_overshoot: bool = True, buckets=None, pipeline_depth: int = 1):
        self.node_id = node_id
        self.gpu_id = 0                       # scheduler.py:380: one visible GPU per actor process
        self.duty_cycle = duty_cycle
        self.sessions = deque(sessions)
        self.models = {}
        self.new_sessions = None
        self.new_duty_cycle = None
        self.lock = Lock()
        self.model_registry = model_registry
        self.device = 'cuda:0'
        self.max_batch = max_batch
        self.compat_overshoot = compat_overshoot
        self.buckets = buckets
        self.pipeline_depth = max(1, int(pipeline_depth))
        self.logger = logging.getLogger("Worker-%s" % node_id)
        if not torch.cuda.is_available():
            raise ValueError("GPU %d not available. Only 0 GPUs found." % self.gpu_id)   # scheduler.py:406-407
        try:
            for session, _ in sessions:
                if session.model_name not in self.models:
                    self.logger.info("Loading %s on GPU %s", session.model_name, gpu_id)
                    self.models[session.model_name] = self._load(session.model_name, sessions)
        except Exception as e:  # noqa: BLE001  (scheduler.py:419-421: log and re-raise)
            self.logger.error("Error initializing models: %s", e)
            raise
        self.active = True
        self.stats = {'processed_batches': 0, 'total_requests': 0, 'processing_times': []}
        self._inflight = deque()

    # ------------------------------------------------------------------------------------------------ model load
    def _engine_batch(self, model_name, sessions):
        """Engine capacity for a model: the largest batch its sessions can hand over -- get_batch in compat mode returns
        up to 2 x batch_size requests (scheduler.py:274-276) -- and never less than the worker's floor."""
        need = max([s.batch_size for s, _ in sessions if s.model_name == model_name and s.batch_size] or [0])
        if self.compat_overshoot:
            need *= 2
        return max(self.max_batch, need)

    def _load(self, model_name, sessions=()):
        """scheduler.py:409-417: registry entry -> device.  A registry entry may be a ready B200Model, an nn.Module (lowered
        here), or a dict {'kind', 'blob'} (a plan lowered once by the scheduler process and shipped to every actor)."""
        entry = self.model_registry[model_name]
        if isinstance(entry, B200Model) or getattr(entry, "is_b200_engine", False):
            return entry
        cap = self._engine_batch(model_name, sessions)
        if isinstance(entry, dict):
            m = B200Model(entry['kind'], blob=entry['blob'], max_batch=cap, device=0, buckets=self.buckets)
            m.out_logical = entry.get('out_logical', m.out_logical)
        else:
            m = B200Model(MODEL_KINDS.get(model_name, model_name), entry, max_batch=cap, device=0, buckets=self.buckets)
        if getattr(m, 'kind', '') == 'yolov8n':
            m.nms = dict(YOLO_NMS)       # a detection actor returns detections, not the 2.8 MB head of every image
        if self.pipeline_depth > 1 and hasattr(m, 'prepare_async'):
            m.prepare_async()            # slots and staging now, not inside the first batch's latency
        return m

    def stop(self):
        self.logger.info("Stopping worker %s", self.node_id)
        self.active = False

    # ------------------------------------------------------------------------------------------------ hot path
    def _submit(self, model: B200Model, batch: BatchRequest, pend: _Pending):
        """Starts the forward of `batch` (asynchronously where the request format allows) and fills `pend`."""
        n = len(batch.inputs)
        queue = getattr(batch, "queue", None)
        if queue is not None and getattr(batch, "slots", None) is not None:      # slots of a shared-memory request ring
            queue.pin()
            pend.queue = queue
            sb = queue.slot_bytes
            ident = bool((batch.hw[:, 0] == model.in_h).all() and (batch.hw[:, 1] == model.in_w).all())
            # requests of another size: the detector letterboxes (YOLOv8's LetterBox), the classifiers get the reference's own
            # preprocessing, transforms.Resize(224) with antialias (profiling/run_profiler.py:63-70) + centre crop
            mode = COLLATE_IDENTITY if ident else \
                (COLLATE_LETTERBOX if getattr(model, "kind", "") == "yolov8n" else COLLATE_RESIZE_SHORT_AA)
            for lo in range(0, n, model.max_batch):                               # an oversize batch is split, not refused
                hi = min(n, lo + model.max_batch)
                runs, rel = [], np.empty(hi - lo, dtype=np.uint64)
                dst = 0
                slots = batch.slots[lo:hi].astype(np.int64)
                brk = np.flatnonzero(np.diff(slots) != 1) + 1
                for a, b in zip(np.concatenate(([0], brk)), np.concatenate((brk, [hi - lo]))):
                    runs.append((int(slots[a]) * sb, int(b - a) * sb, dst))
                    rel[a:b] = dst + np.arange(b - a, dtype=np.uint64) * np.uint64(sb)
                    dst += int(b - a) * sb
                pb = model.bucket_for(hi - lo)
                if pend.handles and not isinstance(pend.handles[-1], torch.Tensor):
                    pend.handles[-1] = pend.handles[-1].result().clone()          # split batch: one chunk in flight at a time
                pend.handles.append(model.infer_runs_async(queue.payload_tensor, runs, rel, batch.hw[lo:hi], hi - lo, pb,
                                                           mode, dst))
                pend.counts.append(hi - lo)
            return
        first = batch.inputs[0]
        if first.dtype == torch.uint8:                                   # engine-native: uint8 HWC requests
            offs = batch.slab_offsets
            if offs is not None and batch.slab is not None and all(b > a for a, b in zip(offs, offs[1:])) \
                    and n <= model.max_batch:
                # payloads already sit, in order, in the queue's pinned ring: one H2D copy of that range, no re-packing
                lo = offs[0]
                hi = offs[-1] + (batch.inputs[-1].numel() + 15) // 16 * 16
                hw = np.asarray([(t.shape[0], t.shape[1]) for t in batch.inputs], dtype=np.int32)
                rel = np.asarray(offs, dtype=np.uint64) - np.uint64(lo)
                mode = 0 if all(tuple(x) == (model.in_h, model.in_w) for x in hw) else 3
                pend.handles.append(model.infer_packed_async(batch.slab.buf[lo:hi], rel, hw, n, model.bucket_for(n), mode,
                                                             hi - lo))
                pend.counts.append(n)
                return
            outs = [model.infer_u8(list(batch.inputs[lo:lo + model.max_batch]),
                                   padded_batch=model.bucket_for(min(n - lo, model.max_batch)))
                    for lo in range(0, n, model.max_batch)]
        else:
            inputs = torch.stack(list(batch.inputs))                     # reference-native fp32 CHW (scheduler.py:443)
            outs = [model(inputs[lo:lo + model.max_batch]) for lo in range(0, n, model.max_batch)]
        pend.handles.append(torch.cat(outs) if len(outs) > 1 else outs[0])
        pend.counts.append(n)

    def _collect(self, pend: _Pending) -> torch.Tensor:
        """Waits for the batch and returns its outputs as a CPU tensor the caller owns (never a view of a staging slot:
        the engine reuses its two pinned output slots two batches later)."""
        outs = []
        for h in pend.handles:
            if isinstance(h, torch.Tensor):
                outs.append(h.cpu() if h.is_cuda else h.clone())         # D2H + the only synchronisation (:452,:468)
            else:
                outs.append(h.result().clone())
        if pend.queue is not None:
            pend.queue.release(pend.batch)                               # H2D long done: the slots go back to the clients
        return torch.cat(outs) if len(outs) > 1 else outs[0]

    def process_batch_async(self, batch: BatchRequest, request_queues: Dict, discard: int = 0) -> _Pending:
        pend = _Pending(batch, time.time())
        try:
            self._submit(self.models[batch.model_name], batch, pend)
        except Exception as e:  # noqa: BLE001
            self.logger.error("Error processing batch: %s", e)
            q = getattr(batch, "queue", None)
            if q is not None:
                q.release(batch)                     # the slots of a failed batch go back to the clients
            if getattr(batch, "slab", None) is not None and batch.slab_offsets:
                for off in batch.slab_offsets:       # same for the in-process pinned slab
                    batch.slab.release(off)
                batch.slab_offsets = None
            raise
        return pend

    def finish_batch(self, pend: _Pending, request_queues: Dict) -> Dict:
        try:
            batch = pend.batch
            outputs_cpu = self._collect(pend)
            completion_time = time.time()
            processing_time = (completion_time - pend.start_time) * 1000
            self.stats['processed_batches'] += 1
            self.stats['total_requests'] += batch.batch_size
            self.stats['processing_times'].append(processing_time)
            if len(self.stats['processing_times']) > 4096:
                del self.stats['processing_times'][:2048]
            request_queues[batch.model_name].record_batch_completion(batch, completion_time)
            return {'outputs': outputs_cpu, 'request_ids': batch.request_ids, 'processing_time': processing_time,
                    'latency': time.time() - batch.arrival_time}
        except Exception as e:  # noqa: BLE001
            self.logger.error("Error processing batch: %s", e)
            raise

    def process_batch(self, batch: BatchRequest, request_queues: Dict, discard: int = 0) -> Dict:
        """scheduler.py:435-475: same keys in the returned dict; logs and re-raises on error."""
        return self.finish_batch(self.process_batch_async(batch, request_queues, discard), request_queues)

    # ------------------------------------------------------------------------------------------------ schedule loop
    def _check_for_updates(self, update_queue) -> bool:
        """scheduler.py:483-523: drain to the latest (sessions, duty_cycle); unload / load models accordingly."""
        if update_queue.qsize() == 0:
            return False
        while update_queue.qsize() > 1:
            update_queue.get()
        latest = update_queue.get()
        self.new_sessions, self.new_duty_cycle = latest[0], latest[1]
        if self.new_sessions is not None:
            new_models = [s.model_name for s, _ in self.new_sessions]
            old_models = [s.model_name for s, _ in self.sessions]
            for name in old_models:
                if name not in new_models and name in self.models:
                    m = self.models.pop(name)
                    if self.model_registry.get(name) is not m:
                        m.close()
            for name in new_models:
                cap = self._engine_batch(name, self.new_sessions)
                if name in self.models and self.models[name].max_batch < cap \
                        and self.model_registry.get(name) is not self.models[name]:
                    self.models.pop(name).close()           # the new schedule asks for larger batches than it was built for
                if name not in self.models:
                    self.models[name] = self._load(name, self.new_sessions)
            self.sessions = deque(list(self.new_sessions))
            self.duty_cycle = self.new_duty_cycle
            self.new_sessions = None
            self.new_duty_cycle = None
            return True
        return False

    @staticmethod
    def _backlog(q) -> int:
        if hasattr(q, "qsize"):
            return q.qsize()
        inner = getattr(q, "queue", None)
        return inner.qsize() if inner is not None else 0

    def _drain(self, request_queues, results, keep=0):
        while len(self._inflight) > keep:
            results.append(self.finish_batch(self._inflight.popleft(), request_queues))

    def run_cycle(self, request_queues: Dict, sleep=time.sleep) -> List[Dict]:
        """One duty cycle of scheduler.py:539-570; returns the results (the reference drops them, :555-556).  With
        pipeline_depth > 1 up to depth-1 batches stay in flight across sessions and cycles; their results come back from
        the cycle in which they finish."""
        results = []
        total_time = self.duty_cycle
        if len(self.sessions) == 0:
            self._drain(request_queues, results)
            sleep(2 / 100)
        for s, occupancy in list(self.sessions):
            time_slice = total_time * occupancy
            t_slice = time.time()
            batch, discard_count = request_queues[s.model_name].get_batch(s.batch_size)
            lt = self.stats.setdefault('loop_s', {'get_batch': 0.0, 'submit': 0.0, 'finish': 0.0, 'sleep': 0.0})
            t1 = time.time()
            lt['get_batch'] += t1 - t_slice
            if batch:
                self._inflight.append(self.process_batch_async(batch, request_queues, discard_count))
                t2 = time.time()
                lt['submit'] += t2 - t1
                before = len(results)
                self._drain(request_queues, results, keep=self.pipeline_depth - 1)
                lt['finish'] += time.time() - t2
                if self.pipeline_depth == 1:
                    spent = results[-1]['processing_time'] if len(results) > before else (time.time() - t_slice) * 1000
                else:
                    # Pipelined: while another full batch is already waiting the loop goes straight on (the wait for the
                    # batch before this one, in _drain above, paces it at the GPU's speed, and the next batch's upload
                    # overlaps this one's forward).  Otherwise there is idle time ahead: a batch is never held in flight
                    # across a sleep, so its latency does not grow by the rest of the slice.
                    if self._backlog(request_queues[s.model_name]) >= s.batch_size:
                        continue
                    self._drain(request_queues, results)
                    spent = (time.time() - t_slice) * 1000
                remaining = time_slice - spent
                if remaining > 0:
                    sleep(remaining / 1000)
            else:
                self._drain(request_queues, results)          # idle slice: nothing to overlap with, hand results back
                sleep(time_slice / 1000)
        return results

    def execute_schedule(self, request_queues: Dict, update_queue, on_result=None):
        """scheduler.py:525-588. Runs until stop(); exceptions are logged and the loop continues after 0.1 s."""
        self.logger.info("Starting schedule execution on %s", self.node_id)
        while self.active:
            try:
                cycle_start = time.time()
                for r in self.run_cycle(request_queues):
                    if on_result is not None:
                        on_result(r)
                if not self._check_for_updates(update_queue):
                    remaining_cycle = (cycle_start + self.duty_cycle / 1000) - time.time()
                    if remaining_cycle > 0:
                        time.sleep(remaining_cycle)
            except Exception as e:  # noqa: BLE001
                self.logger.error("Error in schedule execution: %s", e)
                time.sleep(0.1)
        tail = []
        try:
            self._drain(request_queues, tail)
        except Exception as e:  # noqa: BLE001
            self.logger.error("Error finishing in-flight batches: %s", e)
        for r in tail:
            if on_result is not None:
                on_result(r)

    def get_stats(self) -> Dict:
        """scheduler.py:590-600."""
        pt = self.stats['processing_times'][-100:]
        out = {'node_id': self.node_id, 'gpu_id': self.gpu_id, 'processed_batches': self.stats['processed_batches'],
               'total_requests': self.stats['total_requests'], 'avg_processing_time': sum(pt) / len(pt) if pt else 0}
        # ours, beyond the reference's five keys: where a batch's time goes (device-side upload and compute, host-side loop)
        for name, m in self.models.items():
            t = getattr(m, "timing", None)
            if t and t["batches"]:
                out.setdefault('engines', {})[name] = {'batches': t["batches"], 'h2d_ms': t["h2d_ms"] / t["batches"],
                                                       'gpu_ms': t["gpu_ms"] / t["batches"]}
        lt = self.stats.get('loop_s')
        if lt:
            out['loop_s'] = dict(lt)
        try:
            out['device'] = str(torch.cuda.get_device_properties(0).uuid)
        except Exception:  # noqa: BLE001
            pass
        return out


GPUWorker = B200GPUWorker   # the reference's name


def as_ray_actor(num_gpus=1):
    """`@ray.remote(num_gpus=1) class GPUWorker` (scheduler.py:374) when Ray is installed; raises ImportError otherwise."""
    import ray
    return ray.remote(num_gpus=num_gpus)(B200GPUWorker)
