"""Generates tests/golden/packing_golden.json by RUNNING THE REFERENCE'S OWN POLICY MODULE (build container only):
293-project/src/nexus.py is pure Python and imports as it is.  Only its decisions are recorded (per node: duty cycle and
[(model, request rate, batch size, occupancy)]); the profiles are analytic (formula below, shared with the test) so that
no reference data file has to travel."""
import importlib.util
import json
import os
import random

REF = "/root/reference/293-project/src/nexus.py"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "packing_golden.json")
MODELS = {"resnet": (2000, 0.45, 0.0115, 60.0, 3.8), "shufflenet": (1500, 0.30, 0.0026, 12.0, 1.1),
          "efficientnet": (40, 0.90, 0.0190, 55.0, 2.9), "vit": (4000, 1.2, 0.0800, 340.0, 9.0)}


def synthetic_profile(max_bs=512):
    """model -> batch size -> {avg_latency_ms, peak_memory_mb}: affine in the batch size, B200-like magnitudes."""
    return {m: {b: {"avg_latency_ms": a + k * b, "peak_memory_mb": m0 + m1 * b} for b in range(1, max_bs + 1)}
            for m, (_, a, k, m0, m1) in MODELS.items()}


def describe(nodes):
    return [{"duty_cycle": n.duty_cycle,
             "sessions": [[s.model_name, s.request_rate, s.batch_size, occ] for s, occ in n.node_sessions]} for n in nodes]


def main():
    spec = importlib.util.spec_from_file_location("ref_nexus", REF)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    prof = synthetic_profile()
    rng = random.Random(61)
    cases = []
    for ci in range(80):
        k = rng.choice([1, 2, 3, 4])
        models = rng.sample(sorted(MODELS), k)
        gpu_mem = rng.choice([11, 11, 2.0, 160.0])
        sessions = []
        for m in models:
            slo = MODELS[m][0] / rng.choice([1.0, 2.2, 2.2, 4.0])
            rate = rng.choice([3.0, 40.0, 250.0, 1800.0, 9000.0, 52000.0, 240000.0]) * rng.uniform(0.6, 1.4)
            sessions.append((m, slo, rate))
        ref.gpu_mem = gpu_mem
        pol = ref.nexus(prof)
        try:
            sess = [ref.session(*s) for s in sessions]
            nodes = pol.squishyBinPacking(sess)
            # the reference's node class carries a class-level default gpu_mem of 11 in mergeNodes' memory check
            out = describe(nodes)
            err = None
        except Exception as e:  # noqa: BLE001
            out, err = None, type(e).__name__
        cases.append({"sessions": sessions, "gpu_mem": gpu_mem, "nodes": out, "error": err})
    json.dump({"generated_from": REF, "cases": cases}, open(OUT, "w"), separators=(",", ":"))
    print("wrote", OUT, len(cases), "cases;", sum(c["error"] is not None for c in cases), "errors")


if __name__ == "__main__":
    main()
