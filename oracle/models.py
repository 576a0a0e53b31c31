"""ORACLE (test infrastructure only — never imported by the product path).

The reference's model-actor computation restated on the CPU: GPUWorker.process_batch
(293-project/src/scheduler.py:435-475) is `torch.stack(batch.inputs)` -> `model(inputs)` under `torch.no_grad()` ->
`outputs.cpu()`, with `model` a torchvision module in eval mode, fp32 (scheduler.py:40-44,409-417).  The arithmetic
lives in third-party torch/torchvision (pins: python/requirements/ml/dl-cpu-requirements.txt:18,21 -> torch==2.3.0,
torchvision==0.18.0; installed here: torch 2.11 / torchvision 0.26, same model definitions — parameter counts match the
reference's profiler reports).  The reference has no test that pins a logit ("parity unpinned" at that
boundary); this module *is* the fp32 reference path, executed with the same synthetic weights as the B200 engine.
"""
import time

import torch


def process_batch_cpu(model, inputs_list):
    """scheduler.py:441-450,468 without the device moves: returns {'outputs', 'processing_time' ms}."""
    start = time.time()
    inputs = torch.stack(inputs_list)              # :443
    with torch.no_grad():                          # :449
        outputs = model(inputs)                    # :450
    return {"outputs": outputs, "processing_time": (time.time() - start) * 1000}


def time_cpu_reference(model, inputs_list, budget_s=15.0, min_reps=1, max_reps=20):
    """Times process_batch_cpu on the host cores: 1 warm-up, then repetitions until `budget_s`. Returns (img/s best,
    img/s median, reps, threads)."""
    process_batch_cpu(model, inputs_list)
    times = []
    t_end = time.time() + budget_s
    while len(times) < min_reps or (time.time() < t_end and len(times) < max_reps):
        t0 = time.perf_counter()
        process_batch_cpu(model, inputs_list)
        times.append(time.perf_counter() - t0)
    times.sort()
    n = len(inputs_list)
    return n / times[0], n / times[len(times) // 2], len(times), torch.get_num_threads()
