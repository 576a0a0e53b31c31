"""ORACLE (test infrastructure only — never imported by the product path).

CPU restatement of the reference's batch formation and batch-latency bookkeeping, in plain Python with an
injectable clock and an in-process FIFO in place of ``ray.util.queue.Queue``.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference`` legs may import this module.

Pinned: ``tests/golden/batching_golden.json`` holds outputs of the reference's own ``RequestQueue`` class
(293-project/src/scheduler.py:190-372, executed verbatim against a stub queue by
``tests/golden/make_batching_golden.py``); ``tests/test_oracle_batching.py`` checks this restatement against it.
"""
from collections import deque
from dataclasses import dataclass, field
from typing import Any, Callable, Deque, Dict, List, Optional, Tuple


class SpinDetected(BaseException):
    """The reference's ``get_batch`` would busy-wait forever on an empty queue here (scheduler.py:274-276)."""


@dataclass
class OracleBatch:
    """Restates ``BatchRequest`` (293-project/src/scheduler.py:181-188)."""
    model_name: str
    inputs: List[Any]
    batch_size: int
    request_ids: List[str]
    arrival_time: float


@dataclass
class OracleQueue:
    """Restates ``RequestQueue`` (293-project/src/scheduler.py:190-372) minus Ray and logging."""
    model_name: str
    slo_ms: float                       # models_config[model]['SLO'] (scheduler.py:30-35,197)
    profile: Dict[int, Dict[str, float]]  # batch_size -> {'avg_latency_ms': ...} (scheduler.py:198)
    max_size: int = 100
    clock: Callable[[], float] = None
    items: Deque[Tuple[str, Any, float]] = field(default_factory=deque)
    metrics: Dict[str, Any] = None

    def __post_init__(self):
        # scheduler.py:213-220
        self.metrics = {'total_requests': 0, 'dropped_requests': 0, 'slo_violations': 0, 'latencies': [],
                        'queue_size': 0, 'last_latency': 0.0}

    # scheduler.py:223-225
    def empty(self) -> bool:
        return len(self.items) == 0

    # scheduler.py:238-254
    def add_request(self, request_id: str, input_tensor: Any) -> bool:
        if self.max_size > 0 and len(self.items) >= self.max_size:      # RayQueue.full()
            self.metrics['dropped_requests'] += 1
            return False
        self.items.append((request_id, input_tensor, self.clock()))
        self.metrics['total_requests'] += 1
        self.metrics['queue_size'] = len(self.items)
        return True

    # scheduler.py:258-322  (the live, quirky variant: overshoots to (bs, 2*bs] and spins when starved)
    def get_batch(self, batch_size: int, spin_limit: int = 1):
        inputs, request_ids = [], []
        earliest_arrival = float('inf')
        available = min(batch_size, len(self.items))                     # :266
        if available == 0:
            return None, 0                                               # :267-268
        discard = 0
        count = 0
        empty_rounds = 0
        while count <= batch_size:                                       # :274  (note "<=")
            available = min(batch_size, len(self.items))                 # :275
            if available == 0:
                empty_rounds += 1
                if empty_rounds >= spin_limit:
                    exc = SpinDetected()
                    exc.partial = (list(request_ids), discard, earliest_arrival)
                    raise exc
            for _ in range(available):                                   # :276
                request_id, input_tensor, arrival_time = self.items.popleft()   # :277
                slack_deadline = arrival_time + (self.slo_ms / 1000)     # :281 lhs
                eta = self.clock() + (self.profile[batch_size]['avg_latency_ms'] / 1000)  # :281 rhs
                if slack_deadline < eta:
                    discard += 1                                         # :282-283
                    continue
                request_ids.append(request_id)                           # :285-287
                inputs.append(input_tensor)
                earliest_arrival = min(earliest_arrival, arrival_time)   # :288
                count += 1                                               # :289
        if inputs:                                                       # :309-316
            return OracleBatch(self.model_name, inputs, len(inputs), request_ids, earliest_arrival), discard
        return None, 0                                                   # :322

    # 293-project/src/venkat-code/schedule_executor.py:85-119  (the clean draft: <= batch_size, no discard)
    def get_batch_strict(self, batch_size: int) -> Optional[OracleBatch]:
        available = min(batch_size, len(self.items))
        if available == 0:
            return None
        inputs, request_ids = [], []
        earliest_arrival = float('inf')
        for _ in range(available):
            request_id, input_tensor, arrival_time = self.items.popleft()
            request_ids.append(request_id)
            inputs.append(input_tensor)
            earliest_arrival = min(earliest_arrival, arrival_time)
        return OracleBatch(self.model_name, inputs, len(inputs), request_ids, earliest_arrival)

    # scheduler.py:324-341
    def record_batch_completion(self, batch: OracleBatch, completion_time: float) -> int:
        latency = (completion_time - batch.arrival_time) * 1000
        count = 0
        self.metrics['last_latency'] = latency
        for _ in batch.request_ids:
            if latency > self.slo_ms:
                self.metrics['slo_violations'] += 1
                count += 1
            if len(self.metrics['latencies']) >= 1000:
                self.metrics['latencies'] = self.metrics['latencies'][100:]
            self.metrics['latencies'].append(latency)
        self.metrics['queue_size'] = len(self.items)
        return count

    # scheduler.py:343-372
    def get_stats(self) -> Dict:
        latencies = self.metrics['latencies']
        base = {k: self.metrics[k] for k in
                ('total_requests', 'dropped_requests', 'slo_violations', 'queue_size', 'last_latency')}
        if not latencies:
            base.update(avg_latency=0, p95_latency=0, p99_latency=0)
            return base
        s = sorted(latencies)
        base.update(avg_latency=sum(latencies) / len(latencies), p95_latency=s[int(len(s) * 0.95)],
                    p99_latency=s[int(len(s) * 0.99)])
        return base


def stack_batch(inputs):
    """Restates the host gather of ``process_batch`` (scheduler.py:443: ``torch.stack(batch.inputs)``)."""
    import numpy as np
    return np.stack([np.asarray(x) for x in inputs])
