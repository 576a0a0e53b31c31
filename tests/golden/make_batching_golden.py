"""Generates tests/golden/batching_golden.json by EXECUTING THE REFERENCE'S OWN CODE (run in the build container only;
/root/reference does not exist on the GPU box — the committed JSON is what travels).

The reference module cannot be imported (it imports ray and downloads pretrained weights at import time,
293-project/src/scheduler.py:2,14,40-44), so the `BatchRequest` and `RequestQueue` class definitions are cut out of the
source text with `ast` and exec'd, unmodified, in a namespace whose `RayQueue` is an in-process FIFO and whose `time`
is a scripted clock.  Nothing from the reference is copied into this repository: only its outputs are recorded.
"""
import ast
import json
import os
import queue
import random
import sys
from collections import deque
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Tuple

REF = "/root/reference/293-project/src/scheduler.py"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "batching_golden.json")


class Spin(BaseException):
    pass


class StubRayQueue:
    """qsize/full/put/get of ray.util.queue.Queue (python/ray/util/queue.py:20-305) without the actor."""

    def __init__(self, maxsize=0):
        self.maxsize = maxsize
        self.q = deque()
        self.empty_polls = 0
        self.spin_guard = False

    def qsize(self):
        if not self.spin_guard:
            return len(self.q)
        if not self.q:
            self.empty_polls += 1
            if self.empty_polls > 3:      # the reference would spin here forever (scheduler.py:274-276)
                raise Spin()
        else:
            self.empty_polls = 0
        return len(self.q)

    def full(self):
        return self.maxsize > 0 and len(self.q) >= self.maxsize

    def put(self, item):
        self.q.append(item)

    def get(self):
        return self.q.popleft()


class ScriptedTime:
    def __init__(self):
        self.t = 1000.0
        self.step = 0.0

    def time(self):
        v = self.t
        self.t += self.step
        return v


def load_reference_classes(clock, models_config):
    src = open(REF).read()
    tree = ast.parse(src)
    wanted = [n for n in tree.body if isinstance(n, ast.ClassDef) and n.name in ("BatchRequest", "RequestQueue")]
    assert len(wanted) == 2
    code = "\n\n".join(ast.get_source_segment(src, n) for n in wanted)
    # decorators are not part of get_source_segment for ClassDef in all versions: re-attach @dataclass if missing
    if "@dataclass\nclass BatchRequest" not in code:
        code = code.replace("class BatchRequest", "@dataclass\nclass BatchRequest", 1)
    import logging
    import torch
    ns = dict(RayQueue=StubRayQueue, time=clock, models_config=models_config, logging=logging, deque=deque,
              dataclass=dataclass, List=List, Dict=Dict, Optional=Optional, Tuple=Tuple, Any=Any, torch=torch,
              Empty=queue.Empty)
    exec(compile(code, REF, "exec"), ns)
    return ns["BatchRequest"], ns["RequestQueue"]


def main():
    rng_small, rng_large = random.Random(293), random.Random(2930)
    cases = []
    # cases 0..59: batch sizes of configs[0] (1-32); cases 60..83: the sizes configs[1]-[3] are quoted on (64-512).
    # The two groups draw from separate generators so that adding the second group left the first one unchanged.
    for ci in range(84):
        rng = rng_small if ci < 60 else rng_large
        slo = rng.choice([40, 1500, 2000, 4000])
        bs = rng.choice([1, 2, 4, 8, 16, 32]) if ci < 60 else rng.choice([64, 128, 256, 512])
        lat_ms = rng.choice([1.0, 5.0, 14.4, 53.1, 107.57])
        n_items = rng.choice([0, 1, bs, bs + 1, 2 * bs, 2 * bs + 3, 5 * bs])
        max_size = rng.choice([4, 100, 2000]) if ci < 60 else rng.choice([100, 2000, 4000])
        clock = ScriptedTime()
        models_config = {"resnet": {"SLO": slo, "base_rate": 0}}
        BatchRequest, RequestQueue = load_reference_classes(clock, models_config)
        profile = {"resnet": {b: {"avg_latency_ms": lat_ms} for b in (1, 2, 4, 8, 16, 32, 64, 128, 256, 512)}}
        q = RequestQueue("resnet", max_size=max_size, batch_profile=profile)
        arrivals, accepted = [], []
        for i in range(n_items):
            clock.t = 1000.0 + rng.uniform(0, 0.5) * (slo / 1000.0) * 2.5
            arrivals.append(clock.t)
            accepted.append(q.add_request("r%d" % i, i))
        now0 = 1000.0 + rng.uniform(0.0, 1.5) * (slo / 1000.0)
        step = rng.choice([0.0, 1e-4, 5e-3])
        clock.t, clock.step = now0, step
        record = dict(slo=slo, batch_size=bs, avg_latency_ms=lat_ms, max_size=max_size, arrivals=arrivals,
                      accepted=accepted, now0=now0, clock_step=step)
        try:
            q.queue.spin_guard = True
            batch, discard = q.get_batch(bs)
            q.queue.spin_guard = False
            record["spin"] = False
            record["discard"] = discard
            if batch is None:
                record["batch"] = None
            else:
                record["batch"] = dict(request_ids=batch.request_ids, inputs=batch.inputs, batch_size=batch.batch_size,
                                       arrival_time=batch.arrival_time, model_name=batch.model_name)
                clock.step = 0.0
                done = clock.t + 0.0123
                record["completion_time"] = done
                record["violations"] = q.record_batch_completion(batch, done)
        except Spin as e:
            tb = e.__traceback__
            fr = None
            while tb is not None:
                if tb.tb_frame.f_code.co_name == "get_batch":
                    fr = tb.tb_frame
                tb = tb.tb_next
            record["spin"] = True
            q.queue.spin_guard = False
            record["partial_ids"] = list(fr.f_locals["request_ids"])
            record["partial_discard"] = fr.f_locals["discard"]
        record["remaining"] = [it[0] for it in q.queue.q]
        record["metrics"] = {k: v for k, v in q.metrics.items()}
        record["stats"] = q.get_stats() if not record["spin"] else None
        cases.append(record)
    rng = rng_small
    # percentile bookkeeping: one long run of completions through the reference get_stats (scheduler.py:343-372)
    clock = ScriptedTime()
    BatchRequest, RequestQueue = load_reference_classes(clock, {"resnet": {"SLO": 50, "base_rate": 0}})
    q = RequestQueue("resnet", max_size=100, batch_profile={"resnet": {}})
    lat_inputs = []
    for i in range(130):
        n_ids = rng.choice([1, 3, 8, 16])
        arr = 1000.0 + i
        done = arr + rng.uniform(0.001, 0.120)
        lat_inputs.append((n_ids, arr, done))
        q.record_batch_completion(BatchRequest("resnet", [], n_ids, ["x"] * n_ids, arr), done)
    stats_case = dict(slo=50, completions=lat_inputs, stats=q.get_stats(), n_latencies=len(q.metrics["latencies"]))
    json.dump(dict(generated_from=REF, cases=cases, stats_case=stats_case), open(OUT, "w"), separators=(",", ":"))
    print("wrote", OUT, len(cases), "cases;", sum(c["spin"] for c in cases), "spin")


if __name__ == "__main__":
    main()
