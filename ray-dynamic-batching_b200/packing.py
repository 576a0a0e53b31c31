"""Squishy bin packing: which model runs on which GPU, at what batch size and duty cycle.

Restates the policy the reference scheduler calls (293-project/src/nexus.py, Nexus paper section 6.1) with the same class
and method names -- `session`, `node`, `nexus.squishyBinPacking / scheduleSaturate / scheduleResidue / mergeNodes` -- so
that B200Scheduler is self-contained (the reference module is not on the GPU box) and a maintainer can pass the
reference's own `nexus` object instead (`B200Scheduler(policy=...)`).  Microsecond-scale host code, nothing to accelerate:
it is here because the actor loop needs a schedule.  tests/test_packing.py compares every decision with the reference
module where that is present, and with golden fixtures generated from it elsewhere.

Differences, all at the edges the reference crashes on: a residual rate of exactly 0 is skipped (nexus.py:277 divides by
it), a merged batch size is at least 1 (nexus.py:213 can produce 0, which `session` rejects), the per-GPU memory budget
is an argument (nexus.py:8 hard-codes 11 GB for an A6000; default here: 90 % of a B200's 180 GB).
"""
import copy
import math
from bisect import bisect_right
from datetime import datetime
from typing import Dict, List, Optional, Tuple

GPU_MEM_GB = 162.0       # per-GPU budget of the memory check (nexus.py:8 `gpu_mem`), 90 % of 180 GB


class session:
    """A model deployment request (nexus.py:17-67): model, latency SLO (ms), request rate (req/s), chosen batch size."""

    def __init__(self, model_name: str, latency_SLO: float, request_rate: float, batch_size: Optional[int] = None):
        if not isinstance(model_name, str) or not model_name:
            raise ValueError("Model name must be a non-empty string")
        if not isinstance(latency_SLO, (int, float)) or latency_SLO <= 0:
            raise ValueError("Latency SLO must be a positive number")
        if not isinstance(request_rate, (int, float)) or request_rate < 0:
            raise ValueError("Request rate must be a non-negative number")
        if batch_size is not None and (not isinstance(batch_size, int) or batch_size <= 0):
            raise ValueError("Batch size must be a positive integer")
        self.model_name = model_name
        self.latency_SLO = latency_SLO
        self.request_rate = request_rate
        self.batch_size = batch_size if batch_size is not None else 0
        self.creation_time = datetime.now()

    def to_dict(self) -> dict:
        return {'model_name': self.model_name, 'latency_SLO': self.latency_SLO, 'request_rate': self.request_rate,
                'batch_size': self.batch_size, 'creation_time': self.creation_time.isoformat()}

    def __repr__(self):
        return "session(%s, SLO %.1f ms, %.1f req/s, batch %d)" % (self.model_name, self.latency_SLO, self.request_rate,
                                                                    self.batch_size)


class node:
    """One GPU = one bin (nexus.py:74-109): [(session, occupancy)] served round-robin inside `duty_cycle` ms."""

    def __init__(self, node_sessions: List[Tuple[session, float]] = None, duty_cycle: float = None, gpu_type: str = 'B200',
                 gpu_mem: float = GPU_MEM_GB):
        self.gpu_type = gpu_type
        self.gpu_mem = gpu_mem
        self.duty_cycle = duty_cycle if duty_cycle else float('inf')
        self.node_sessions = node_sessions if node_sessions else []

    def get_occupancy(self) -> float:
        return sum(occ for _, occ in self.node_sessions)

    def to_dict(self) -> dict:
        return {'duty_cycle': self.duty_cycle, 'occupancy': self.get_occupancy(),
                'sessions': [dict(s.to_dict(), occupancy=occ) for s, occ in self.node_sessions]}

    def print_node_pretty(self):
        print("-" * 75)
        print("Node gpu type: %s, Node gpu memory: %sGB" % (self.gpu_type, self.gpu_mem))
        print("Node duty cycle: %sms" % round(self.duty_cycle, 0))
        print("Node sessions: %d" % len(self.node_sessions))
        for i, (s, occ) in enumerate(self.node_sessions):
            print("session number %d has occupancy: %s%%: %r" % (i + 1, round(occ * 100, 0), s))
        print("-" * 75)


class nexus:
    """batching_profile: model -> batch size -> {'avg_latency_ms', 'peak_memory_mb'} (scheduler.py:95-113)."""

    def __init__(self, batching_profile: Dict[str, Dict[int, Dict[str, float]]], gpu_mem: float = GPU_MEM_GB,
                 gpu_type: str = 'B200', merge_gpu_mem: Optional[float] = None):
        self.batching_profile = batching_profile
        self.gpu_mem = gpu_mem            # budget of scheduleSaturate's memory cut (nexus.py:157, the module-level gpu_mem)
        # budget of mergeNodes' memory check: in the reference that is the node's own field, whose default was bound to
        # 11 GB when the class was defined (nexus.py:78), whatever the module variable says later; here: the same budget
        self.merge_gpu_mem = gpu_mem if merge_gpu_mem is None else merge_gpu_mem
        self.gpu_type = gpu_type

    def _columns(self, model):
        prof = self.batching_profile[model]
        sizes = list(prof.keys())                                   # profile order = ascending batch size
        return sizes, [prof[b]['avg_latency_ms'] for b in sizes], [prof[b]['peak_memory_mb'] / 1024 for b in sizes]

    def _node(self, sessions, duty_cycle):
        return node(sessions, duty_cycle=duty_cycle, gpu_type=self.gpu_type, gpu_mem=self.merge_gpu_mem)

    def squishyBinPacking(self, sessions: List[session]) -> List[node]:
        nodes, residual = self.scheduleSaturate(sessions)
        nodes.extend(self.scheduleResidue(residual))
        return nodes

    def scheduleSaturate(self, sessions: List[session]):
        """nexus.py:141-193: whole GPUs at the largest batch whose latency fits SLO / 2 (and whose memory fits the GPU);
        the remainder of the rate goes on as a residual session."""
        nodes, residual = [], []
        for s in sessions:
            sizes, lat, mem = self._columns(s.model_name)
            idx = min(bisect_right(lat, s.latency_SLO / 2), bisect_right(mem, self.gpu_mem))
            idx = max(idx, 1)
            bs, l = sizes[idx - 1], lat[idx - 1]
            throughput = (bs / l) * 1000
            n, r = divmod(s.request_rate, throughput)
            if n > 0:
                full = self._node([(session(s.model_name, s.latency_SLO, throughput, bs), 1.0)], l)
                nodes.extend([full] * int(n))
            residual.append(session(s.model_name, s.latency_SLO, r))
        return nodes, residual

    def mergeNodes(self, node1: node, node2: node) -> Optional[node]:
        """nexus.py:204-234: the sessions of the node with the longer duty cycle are squeezed into the shorter one
        (batch = ceil(duty * rate)); None when they do not fit its time or the GPU's memory."""
        if node1.duty_cycle < node2.duty_cycle:
            node1, node2 = node2, node1
        merged = copy.deepcopy(node2)
        duty = node2.duty_cycle
        for s, _ in node1.node_sessions:
            new_batch = max(1, int(math.ceil((duty * s.request_rate) / 1000)))
            new_latency = self.batching_profile[s.model_name][new_batch]['avg_latency_ms']
            merged.node_sessions.append((session(s.model_name, s.latency_SLO, s.request_rate, batch_size=new_batch),
                                         new_latency / duty))
        if merged.get_occupancy() > 1:
            return None
        total_memory = sum(float(self.batching_profile[s.model_name][s.batch_size]['peak_memory_mb']) / 1024
                           for s, _ in merged.node_sessions)
        if total_memory > merged.gpu_mem:
            return None
        return merged

    def scheduleResidue(self, sessions: List[session]) -> List[node]:
        """nexus.py:245-316: one node per residual session at the largest batch whose (latency + time to gather the batch)
        fits the SLO, then greedy merging in decreasing order of occupancy, each into the node it fills best."""
        singles = []
        for s in sessions:
            if s.request_rate <= 0:
                continue
            sizes, lat, _ = self._columns(s.model_name)
            with_wait = [l + (b / s.request_rate) * 1000 for b, l in zip(sizes, lat)]
            idx = max(bisect_right(with_wait, s.latency_SLO), 1)
            bs, l = sizes[idx - 1], lat[idx - 1]
            duty = (bs / s.request_rate) * 1000
            occ = l / duty
            s.batch_size = bs
            s.occupancy = occ
            singles.append(self._node([(s, occ)], duty))
        nodes: List[node] = []
        for cand in sorted(singles, key=lambda nd: nd.get_occupancy(), reverse=True):
            best, best_i, best_occ = None, None, 0
            for i, nd in enumerate(nodes):
                m = self.mergeNodes(nd, cand)
                if m and m.get_occupancy() > best_occ:
                    best, best_i, best_occ = m, i, m.get_occupancy()
            if best:
                nodes[best_i] = best
            else:
                nodes.append(cand)
        return nodes
