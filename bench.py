#!/usr/bin/env python
"""Benchmark of the hot path: batch collate + ResNet-50 forward (bf16 storage, fp32 accumulate) at batch 256.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--impl reference]

One process per GPU (torchrun for N > 1); every rank is an independent model-actor replica fed its own requests
(the path shards by request, SURVEY §8e: no data-path collective) -> "scaling": "weak".
Prints ONE JSON line on rank 0.  A "step" is one pass of the hot path over one batch of synthetic uint8 224x224 requests:
  value : whole-job images/s with the packed uint8 requests already resident in HBM (collate + forward, CUDA events)
  e2e   : same through the public API with HOST buffers: pinned uint8 -> H2D -> collate -> forward -> D2H logits
--impl reference times the reference's own CPU implementation of the path (torch.stack -> torchvision fp32 forward,
oracle/models.py restating scheduler.py:441-450) on the host cores, on a bounded sample per step.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "resnet50_images_per_sec"
UNIT = "images/s"
# algorithmic work per image, every BN/act/residual fused (bf16 activations)
RESNET50_FLOPS_PER_IMG = 8.178e9
RESNET50_FUSED_BYTES_PER_IMG = 56.96e6
RESNET50_WEIGHT_BYTES = 51.0e6


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=float(d["hbm_gbs"]), tensor=float(d["bf16_tflops"]),
                    tensor_sustained=float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), source="measured")
    return dict(hbm=6650.0, tensor=1590.0, tensor_sustained=1400.0, source="fallback")


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], 0, set()
        for ts, line in self.lines:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 8:
                continue
            try:
                mhz = float(f[1])
                mx = max(mx, float(f[2]))
            except ValueError:
                continue
            if t0 - 0.05 <= ts <= t1 + 0.15:
                sm.append(mhz)
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


def plan_traffic(blob, batch):
    """Algorithmic bytes and FLOPs per op of one forward at `batch` (in + out + residual activations once, weights once)."""
    from rdb_b200 import lowering
    hdr, bufs, ops, _ = lowering.parse_blob(blob)
    out = []
    for op in ops:
        t, p = op["type"], op["p"]
        ih, iw = bufs[p[0]][0], bufs[p[0]][1]
        oh, ow = bufs[p[3]][0], bufs[p[3]][1]
        es_in = 4 if bufs[p[0]][3] else 2
        es_out = 4 if bufs[p[3]][3] else 2
        b = batch * (ih * iw * p[2] * es_in + oh * ow * p[5] * es_out)
        fl = 0.0
        if t in (lowering.OP_CONV, lowering.OP_GATHER_CONV):
            k = p[8]
            b += p[5] * k * k * p[2] * 2 + p[5] * 4
            if p[6] >= 0:
                b += batch * oh * ow * p[5] * 2
            fl = 2.0 * batch * oh * ow * p[5] * p[2] * k * k
            if t == lowering.OP_CONV and p[17] > 0:      # folded SE gate [N][Cin] fp32
                b += batch * p[2] * 4
            if t == lowering.OP_CONV and p[13] > 0:      # K-concatenated second input, sampled with stride p[16]
                b += batch * oh * ow * p[15] * 2 + p[5] * p[15] * 2
                fl += 2.0 * batch * oh * ow * p[5] * p[15]
        elif t == lowering.OP_STEM:
            k = p[8]
            b += k * k * 3 * p[5] * 4
            sh, sw_ = (ih + 2 * p[10] - k) // p[9] + 1, (iw + 2 * p[10] - k) // p[9] + 1   # conv output size (pool may follow)
            fl = 2.0 * batch * sh * sw_ * p[5] * 3 * k * k
        elif t == lowering.OP_FC:
            b += p[5] * p[2] * 4
            fl = 2.0 * batch * p[5] * p[2]
        out.append((t, b, fl))
    return out


def run_reference(args):
    """Reference arm: the reference's CPU path (torch.stack -> torchvision fp32 forward) on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    from rdb_b200 import synth
    from oracle.models import process_batch_cpu
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    model = synth.make_model("resnet50")
    sample = 32
    imgs = synth.synthetic_images(sample)
    reqs = list(synth.normalize_u8(imgs, synth.IMAGENET_MEAN, synth.IMAGENET_STD))
    for _ in range(max(1, min(args.warmup, 2))):
        process_batch_cpu(model, reqs)
    steps = max(1, args.steps)
    t0 = time.perf_counter()
    done = 0
    for _ in range(steps):
        process_batch_cpu(model, reqs)
        done += 1
        if time.perf_counter() - t0 > 150:  # keep the whole run within a few minutes on slow hosts
            break
    dt = time.perf_counter() - t0
    value = done * sample / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": done,
        "warmup": args.warmup, "ms_per_step": 1000 * dt / done, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "ResNet-50 dynamic batching, batch 256, synthetic 224x224 (configs[2], one replica)",
                   "note": "reference CPU path timed on a bounded sample of 32 images per step"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
                         "sample": "%d steps x batch %d, torch.stack + torchvision resnet50 fp32 eager" % (done, sample)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import numpy as np
    import torch
    import rdb_b200
    from rdb_b200 import synth
    from rdb_b200.engine import B200Model

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the B200 engine has no CPU fallback")
    torch.cuda.set_device(local)
    from rdb_b200 import replicas
    dist = replicas.init_process_group("nccl") if world > 1 else None
    dev = torch.device("cuda", local)
    W = max(3, args.warmup)
    K = max(1, args.steps)
    B = args.batch
    peaks = load_peaks()

    torch.set_num_threads(max(1, (os.cpu_count() or 8) // max(1, world)))
    module = synth.make_model("resnet50")
    model = B200Model("resnet50", module, max_batch=B, device=local)

    # ---- synthetic requests: NSETS rotating batches so that consecutive steps never reuse L2-resident inputs ----
    NSETS = 4
    per = 224 * 224 * 3
    offsets = (np.arange(B, dtype=np.uint64) * per)
    hw = np.tile(np.array([[224, 224]], dtype=np.int32), (B, 1))
    host_sets, dev_sets = [], []
    g = torch.Generator().manual_seed(1000 + 7919 * rank)
    for s in range(NSETS):
        h = torch.empty(B * per, dtype=torch.uint8, pin_memory=True)
        h.copy_(torch.randint(0, 256, (B * per,), generator=g, dtype=torch.uint8))
        host_sets.append(h)
        dev_sets.append(h.to(dev))
    out_dev = torch.empty(B, model.out_dim, device=dev, dtype=torch.float32)
    out_host = torch.empty(B, model.out_dim, dtype=torch.float32, pin_memory=True)
    stream = torch.cuda.current_stream(dev)

    def step_resident(i):
        model.collate_u8(dev_sets[i % NSETS], offsets, hw, B, B, 0, B * per)
        model.infer_resident(B, out_dev)

    pending = []

    def step_e2e(i):
        # public async API: pinned host requests -> H2D (side stream) -> collate -> forward -> D2H into pinned memory.
        # Two batches in flight: the copy of step i overlaps the forward of step i-1; every step's result is read
        # (handle.result() waits for that batch's D2H) before the step after next is submitted.
        pending.append(model.infer_packed_async(host_sets[i % NSETS], offsets, hw, B, B, 0, B * per))
        if len(pending) > 1:
            pending.pop(0).result()

    def drain_e2e():
        while pending:
            pending.pop(0).result()

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(step_fn, k, drain=None):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record(stream)
        for i in range(k):
            step_fn(i)
        if drain is not None:
            drain()
        e1.record(stream)
        torch.cuda.synchronize(dev)
        ms = replicas.max_over_ranks(e0.elapsed_time(e1), dist, dev)   # the job is as slow as its slowest replica
        barrier()
        return ms

    for i in range(W):
        step_resident(i)
    for i in range(W):
        step_e2e(i)
    drain_e2e()
    torch.cuda.synchronize(dev)

    sampler = ClockSampler(local)
    sampler.start()
    time.sleep(0.3)
    l0 = model.launch_count()
    t0 = time.time()
    ms_res = timed(step_resident, K)
    t1 = time.time()
    launches = model.launch_count() - l0
    ms_e2e = timed(step_e2e, K, drain_e2e)
    clocks = sampler.stop(t0, t1)

    # ---- per-op device times (eager pass with CUDA events on the launching stream) for the roofline ----
    from rdb_b200 import lowering
    traffic = plan_traffic(model.blob, B)
    ms_ops = np.zeros(model.n_ops)
    reps = 3
    for r in range(reps + 1):
        model.collate_u8(dev_sets[r % NSETS], offsets, hw, B, B, 0, B * per)
        ms, _ = model.profile_ops(B)
        if r > 0:
            ms_ops += ms
    ms_ops /= reps
    conv_idx = [i for i, (t, _, _) in enumerate(traffic) if t == lowering.OP_CONV]
    conv_ms = float(ms_ops[conv_idx].sum())
    conv_bytes = float(sum(traffic[i][1] for i in conv_idx))
    conv_flops = float(sum(traffic[i][2] for i in conv_idx))
    n_conv = len(conv_idx)
    achieved_gbs = conv_bytes / (conv_ms * 1e-3) / 1e9
    achieved_tf = conv_flops / (conv_ms * 1e-3) / 1e12
    step_ms = ms_res / K
    per_img_s = step_ms * 1e-3 / B
    model_bound_s = max(RESNET50_FLOPS_PER_IMG / (peaks["tensor_sustained"] * 1e12),
                        (RESNET50_FUSED_BYTES_PER_IMG + RESNET50_WEIGHT_BYTES / B) / (peaks["hbm"] * 1e9))
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "r01d_conv_traffic.json")
    if B == 256 and os.path.exists(tpath):      # dram bytes per launch from the committed ncu capture of this command
        traffic = json.load(open(tpath))["traffic_bytes_per_launch"]
    roofline = {
        "kernel": "conv_gemm_kernel (tcgen05 implicit GEMM), %d launches per step" % n_conv,
        "bound": "hbm", "achieved": achieved_gbs / 1.0, "peak": peaks["hbm"], "unit": "GB/s",
        "frac": achieved_gbs / peaks["hbm"], "traffic": traffic, "peak_source": peaks["source"],
        "bytes_per_launch_avg": conv_bytes / n_conv, "ms_per_launch_avg": conv_ms / n_conv,
        "share_of_step": conv_ms / float(ms_ops.sum()),
        "tensor_achieved_tflops": achieved_tf, "tensor_peak_tflops": peaks["tensor_sustained"],
        "tensor_frac": achieved_tf / peaks["tensor_sustained"],
        "model_frac_of_layer_fused_roofline": model_bound_s / per_img_s,
        "per_op_ms_top": sorted(((float(ms_ops[i]), model.op_names[i]) for i in range(model.n_ops)), reverse=True)[:6],
    }

    value = world * B * K / (ms_res * 1e-3)
    e2e_value = world * B * K / (ms_e2e * 1e-3)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic",
        "config": {"workload": "ResNet-50 dynamic batching, batch %d per replica, synthetic 224x224 uint8 requests "
                               "(configs[2], one request-sharded replica per GPU)" % B,
                   "batch_per_gpu": B, "replicas": world, "weights": "seeded synthetic (rdb_b200.synth)",
                   "l2": "working set >> L2: %d rotating input batches, ~%.1f GB activation traffic per step"
                         % (NSETS, B * RESNET50_FUSED_BYTES_PER_IMG / 1e9),
                   "cuda_graph": True},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(world * (B * per + B * 16)),
                "d2h_bytes_per_step": int(world * B * model.out_dim * 4), "ms_per_step": ms_e2e / K,
                "pipeline": "2 batches in flight (H2D of step i overlaps the forward of step i-1)"},
        "gpu_launches": int(launches),
        "roofline": roofline,
        "clocks": clocks,
    }

    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle.models import time_cpu_reference
        torch.set_num_threads(os.cpu_count() or 1)
        sample = 32
        reqs = list(synth.normalize_u8(synth.synthetic_images(sample), synth.IMAGENET_MEAN, synth.IMAGENET_STD))
        best, med, nrep, threads = time_cpu_reference(module, reqs, budget_s=12.0)
        line["cpu_baseline"] = {"value": med, "unit": UNIT, "cores": threads, "kind": "port", "best": best,
                                "sample": "%d reps x batch %d: torch.stack + torchvision resnet50 fp32 eager "
                                          "(scheduler.py:443-450), same synthetic weights" % (nrep, sample)}
    if rank == 0:
        print(json.dumps(line), flush=True)
    model.close()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
