"""Host-side mirror of the reference's model actor `GPUWorker` (293-project/src/scheduler.py:374-600) for one B200.

Same constructor arguments, methods and return values; the body of `process_batch` — the reference's
`torch.stack(batch.inputs).cuda()` -> `model(inputs)` -> `outputs.cpu()` (scheduler.py:441-452,468) — is replaced by
the C-ABI engine: one H2D copy of the packed requests, the fused collate kernel, the CUDA-graph forward, one D2H copy.

The reference makes this class a Ray actor (`@ray.remote(num_gpus=1)`, scheduler.py:374): Ray pins one GPU per actor
process through CUDA_VISIBLE_DEVICES, so the actor always uses cuda:0 (scheduler.py:388,404).  Ray is not part of this
image; `as_ray_actor()` wraps the class when Ray is importable, and rdb_b200.replicas runs the same class in one
process per GPU otherwise.
"""
import logging
import time
from collections import deque
from threading import Lock
from typing import Dict, List, Tuple

import torch

from .engine import B200Model, COLLATE_IDENTITY
from .request_queue import BatchRequest

MODEL_KINDS = {'resnet': 'resnet50', 'shufflenet': 'shufflenet_v2_x1_0', 'efficientnet': 'efficientnet_v2_s',
               'yolov8': 'yolov8n'}


class B200GPUWorker:
    def __init__(self, node_id: str, gpu_id: int, sessions: List[Tuple], duty_cycle: float, model_registry: Dict,
                 max_batch: int = 256):
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
        self.logger = logging.getLogger("Worker-%s" % node_id)
        if not torch.cuda.is_available():
            raise ValueError("GPU %d not available. Only 0 GPUs found." % self.gpu_id)   # scheduler.py:406-407
        try:
            for session, _ in sessions:
                if session.model_name not in self.models:
                    self.logger.info("Loading %s on GPU %s", session.model_name, gpu_id)
                    self.models[session.model_name] = self._load(session.model_name)
        except Exception as e:  # noqa: BLE001  (scheduler.py:419-421: log and re-raise)
            self.logger.error("Error initializing models: %s", e)
            raise
        self.active = True
        self.stats = {'processed_batches': 0, 'total_requests': 0, 'processing_times': []}

    def _load(self, model_name):
        """scheduler.py:409-417: registry entry -> device. An nn.Module entry is lowered to a B200 engine here."""
        entry = self.model_registry[model_name]
        if isinstance(entry, B200Model):
            return entry
        return B200Model(MODEL_KINDS.get(model_name, model_name), entry, max_batch=self.max_batch, device=0)

    def stop(self):
        self.logger.info("Stopping worker %s", self.node_id)
        self.active = False

    # ------------------------------------------------------------------------------------------------ hot path
    def _forward(self, model: B200Model, batch: BatchRequest) -> torch.Tensor:
        first = batch.inputs[0]
        n = len(batch.inputs)
        if first.dtype == torch.uint8:                                   # engine-native: uint8 HWC requests
            offs = batch.slab_offsets
            if offs is not None and batch.slab is not None and all(b > a for a, b in zip(offs, offs[1:])):
                # payloads already sit, in order, in the queue's pinned ring: one H2D copy of that range, no re-packing
                import numpy as np
                lo = offs[0]
                hi = offs[-1] + (batch.inputs[-1].numel() + 15) // 16 * 16
                hw = np.asarray([(t.shape[0], t.shape[1]) for t in batch.inputs], dtype=np.int32)
                rel = np.asarray(offs, dtype=np.uint64) - np.uint64(lo)
                mode = 0 if all(tuple(x) == (model.in_h, model.in_w) for x in hw) else 3
                return model.infer_packed_async(batch.slab.buf[lo:hi], rel, hw, n, n, mode, hi - lo).result()
            return model.infer_u8(batch.inputs)
        inputs = torch.stack(batch.inputs)                               # reference-native fp32 CHW (scheduler.py:443)
        return model(inputs)[:n]

    def process_batch(self, batch: BatchRequest, request_queues: Dict, discard: int = 0) -> Dict:
        """scheduler.py:435-475: same keys in the returned dict; logs and re-raises on error."""
        try:
            start_time = time.time()
            model = self.models[batch.model_name]
            outputs = self._forward(model, batch)
            outputs_cpu = outputs.cpu()                                   # D2H + the only synchronisation (:452,:468)
            completion_time = time.time()
            processing_time = (completion_time - start_time) * 1000
            self.stats['processed_batches'] += 1
            self.stats['total_requests'] += len(batch.request_ids)
            self.stats['processing_times'].append(processing_time)
            request_queues[batch.model_name].record_batch_completion(batch, completion_time)
            return {'outputs': outputs_cpu, 'request_ids': batch.request_ids, 'processing_time': processing_time,
                    'latency': time.time() - batch.arrival_time}
        except Exception as e:  # noqa: BLE001
            self.logger.error("Error processing batch: %s", e)
            raise

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
                if name not in self.models:
                    self.models[name] = self._load(name)
            self.sessions = deque(list(self.new_sessions))
            self.duty_cycle = self.new_duty_cycle
            self.new_sessions = None
            self.new_duty_cycle = None
            return True
        return False

    def run_cycle(self, request_queues: Dict, sleep=time.sleep) -> List[Dict]:
        """One duty cycle of scheduler.py:539-570; returns the results (the reference drops them, :555-556)."""
        results = []
        total_time = self.duty_cycle
        if len(self.sessions) == 0:
            sleep(2 / 100)
        for s, occupancy in list(self.sessions):
            time_slice = total_time * occupancy
            batch, discard_count = request_queues[s.model_name].get_batch(s.batch_size)
            if batch:
                result = self.process_batch(batch, request_queues, discard_count)
                results.append(result)
                remaining = time_slice - result['processing_time']
                if remaining > 0:
                    sleep(remaining / 1000)
            else:
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

    def get_stats(self) -> Dict:
        """scheduler.py:590-600."""
        pt = self.stats['processing_times'][-100:]
        return {'node_id': self.node_id, 'gpu_id': self.gpu_id, 'processed_batches': self.stats['processed_batches'],
                'total_requests': self.stats['total_requests'], 'avg_processing_time': sum(pt) / len(pt) if pt else 0}


GPUWorker = B200GPUWorker   # the reference's name


def as_ray_actor(num_gpus=1):
    """`@ray.remote(num_gpus=1) class GPUWorker` (scheduler.py:374) when Ray is installed; raises ImportError otherwise."""
    import ray
    return ray.remote(num_gpus=num_gpus)(B200GPUWorker)
