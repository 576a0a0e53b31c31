#!/usr/bin/env python
"""Benchmark of the hot path: batch collate + forward of one served model (bf16 storage, fp32 accumulate).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--model M] [--batch B] [--hw S] [--impl reference]

--model: resnet50 (default; batch 256, BASELINE.json configs[2] -- the driver's line), shufflenet_v2_x1_0 (batch 512,
configs[1]), efficientnet_v2_s (batch 128, configs[3]; --hw 384 for the size the reference profiled), yolov8n (batch 32,
640x640, configs[4]).  One process per GPU (torchrun for N > 1); every rank is an independent model-actor replica fed its
own requests (the path shards by request, SURVEY section 8e: no data-path collective) -> "scaling": "weak".

Prints ONE JSON line on rank 0.  A "step" is one pass of the hot path over one batch of synthetic uint8 requests:
  value          whole-job images/s with the packed uint8 requests already resident in HBM (collate + forward, CUDA events)
  e2e            the same through the public API with HOST buffers: pinned uint8 -> H2D -> collate -> forward -> D2H
  parity_checked rows of the LAST TIMED step (resident and e2e) compared with the fp32 CPU reference forward of the same
                 requests; the process exits non-zero when they differ by more than the north-star tolerance
  roofline       conv kernel family: algorithmic bytes / measured time against the measured HBM peak, plus
                 bound_frac = sum over launches of max(bytes / HBM peak, flops / sustained bf16 peak) / measured time
  sustained      the resident step repeated for >= 3 s, with the SM clock and power seen meanwhile
  serving        queue -> worker loop at 80 % of the measured capacity: p99 batch latency against the model's SLO
--impl reference times the reference's own CPU implementation of the path (torch.stack -> torchvision fp32 forward,
oracle/models.py restating 293-project/src/scheduler.py:441-450) on the host cores at the same batch size.
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

UNIT = "images/s"
# algorithmic work per image with every BN / activation / residual / SE scale / shuffle fused and
# bf16 activations (flops, layer-fused bytes, weight bytes per batch).  YOLOv8n's row is derived from the lowered plan
# by tools/model_roofline.py.
MODELS = {
    "resnet50": dict(batch=256, hw=224, cfg="configs[2]", name="ResNet-50", flops=8.178e9, bytes=56.96e6, weights=51.0e6,
                     slo_name="resnet"),
    "shufflenet_v2_x1_0": dict(batch=512, hw=224, cfg="configs[1]", name="ShuffleNetV2 x1.0", flops=0.290e9, bytes=8.88e6,
                               weights=4.5e6, slo_name="shufflenet"),
    "efficientnet_v2_s": dict(batch=128, hw=224, cfg="configs[3]", name="EfficientNetV2-S", flops=5.700e9, bytes=59.43e6,
                              weights=42.5e6, slo_name="efficientnet",
                              at={384: dict(flops=16.73e9, bytes=174.2e6)}),
    "yolov8n": dict(batch=32, hw=640, cfg="configs[4]", name="YOLOv8n", flops=8.743e9, bytes=77.6e6, weights=6.3e6,
                    slo_name="yolov8"),
}
LOGIT_TOL = 2e-2        # north star: logits within 2e-2 max-abs
IOU_TOL = 0.99          # north star: YOLO boxes IoU >= 0.99


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=float(d["hbm_gbs"]), tensor=float(d["bf16_tflops"]),
                    tensor_sustained=float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), source="measured")
    return dict(hbm=6650.0, tensor=1590.0, tensor_sustained=1400.0, source="fallback")


class ClockSampler:
    """Samples nvidia-smi clocks / power / throttle reasons during a timed region (B200_PROFILING.md recipe)."""
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

    def window(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm, pw, mx, reasons = [], [], 0, set()
        for ts, line in list(self.lines):
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
                try:
                    pw.append(float(f[3]))
                except ValueError:
                    pass
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(pw) if pw else None}

    def stop(self):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()


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
        if t == lowering.OP_CONV and p[18] > 0:
            # ShuffleNetV2 unit tail: 1x1 conv of p[18] channels written interleaved with the other branch (p[18] logical
            # channels read) into the unit's output of p[5] physical channels; weights hold p[5] / 2 slots
            b = batch * oh * ow * (p[2] * 2 + p[18] * 2 + p[5] * 2) + (p[5] // 2) * p[2] * 2 + (p[5] // 2) * 4
            fl = 2.0 * batch * oh * ow * (p[5] // 2) * p[2]      # physical channels, as for every other conv of the table
        elif t in (lowering.OP_CONV, lowering.OP_GATHER_CONV):
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
        elif t == lowering.OP_DWCONV:
            b += 9 * p[5] * 4
            fl = 2.0 * batch * oh * ow * p[5] * 9
        elif t == lowering.OP_SHUFFLE:                  # two inputs of p[14] logical channels each -> 2 * p[15] physical
            b = batch * ih * iw * (2 * p[14] * 2 + p[5] * 2)
        elif t == lowering.OP_YOLO_DECODE:               # hidden box (64) + class (80 of p[14]) channels in, 84 fp32 out
            b = batch * ih * iw * ((64 + 80) * 2 + 84 * 4) + (64 * 64 + 80 * 80) * 4
            fl = 2.0 * batch * ih * iw * (64 * 64 + 80 * 80)
        elif t in (lowering.OP_SE_GATE, lowering.OP_SE_HIDDEN):
            b = batch * (p[1] * p[2] * 4 + p[5] * 4) + 2 * p[8] * p[2] * 4
        out.append((t, b, fl))
    return out


def cpu_reference_outputs(kind, module, imgs_u8):
    """The reference arm on the host: torchvision fp32 forward of ToTensor + Normalize'd requests (scheduler.py:441-450)."""
    import torch
    from rdb_b200 import synth
    from oracle.models import process_batch_cpu
    if kind == "yolov8n":
        reqs = list(imgs_u8.permute(0, 3, 1, 2).float().div(255))
    else:
        reqs = list(synth.normalize_u8(imgs_u8, synth.IMAGENET_MEAN, synth.IMAGENET_STD))
    return process_batch_cpu(module, reqs)["outputs"]


def compare_rows(kind, got, ref):
    """(ok, figures): logits max-abs for the classifiers; min box IoU and class-probability error for YOLOv8n."""
    import torch
    if kind != "yolov8n":
        err = (got - ref).abs().max().item()
        agree = (got.argmax(1) == ref.argmax(1)).float().mean().item()
        return err <= LOGIT_TOL, {"max_abs": err, "top1_agreement": agree}
    n = got.shape[0]
    g = got.view(n, 84, -1)
    r = ref.view(n, 84, -1)

    def xyxy(t):
        cx, cy, w, h = t[:, 0], t[:, 1], t[:, 2], t[:, 3]
        return torch.stack([cx - w / 2, cy - h / 2, cx + w / 2, cy + h / 2], -1)
    a, b = xyxy(g), xyxy(r)
    lt = torch.max(a[..., :2], b[..., :2])
    rb = torch.min(a[..., 2:], b[..., 2:])
    inter = (rb - lt).clamp(min=0).prod(-1)
    iou = inter / ((a[..., 2] - a[..., 0]) * (a[..., 3] - a[..., 1]) + (b[..., 2] - b[..., 0]) * (b[..., 3] - b[..., 1]) - inter)
    cls_err = (g[:, 4:] - r[:, 4:]).abs().max().item()
    return (iou.min().item() >= IOU_TOL and cls_err <= LOGIT_TOL), {"min_box_iou": iou.min().item(), "max_abs": cls_err}


def make_requests(kind, n, hw, seed):
    from rdb_b200 import synth
    if kind == "yolov8n":
        return synth.synthetic_images(n, hw, hw, seed=seed)
    return synth.heldout_images(n, hw, hw, seed=seed)


def make_module(kind, hw):
    from rdb_b200 import synth
    if kind == "efficientnet_v2_s":
        return synth.make_model(kind, in_hw=(hw, hw))
    return synth.make_model(kind)


def workload_name(kind, B, hw, world):
    m = MODELS[kind]
    return "%s dynamic batching, batch %d per replica, synthetic %dx%d uint8 requests (%s, one request-sharded replica " \
           "per GPU)" % (m["name"], B, hw, hw, m["cfg"])


def run_reference(args):
    """Reference arm: the reference's CPU path (torch.stack -> torchvision fp32 forward) on the host cores, same batch."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    from oracle.models import process_batch_cpu
    from rdb_b200 import synth
    kind = args.model
    B, hw = args.batch, args.hw
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    module = make_module(kind, hw)
    imgs = make_requests(kind, B, hw, seed=910000)
    if kind == "yolov8n":
        reqs = list(imgs.permute(0, 3, 1, 2).float().div(255))
    else:
        reqs = list(synth.normalize_u8(imgs, synth.IMAGENET_MEAN, synth.IMAGENET_STD))
    budget = float(os.environ.get("B200_REF_BUDGET_S", "150"))
    t_w = time.perf_counter()
    for _ in range(max(1, min(args.warmup, 2))):
        process_batch_cpu(module, reqs)
        if time.perf_counter() - t_w > budget / 5:
            break
    steps = max(1, args.steps)
    t0 = time.perf_counter()
    done = 0
    for _ in range(steps):
        process_batch_cpu(module, reqs)
        done += 1
        if time.perf_counter() - t0 > budget:   # keep the whole run within a few minutes on slow hosts
            break
    dt = time.perf_counter() - t0
    value = done * B / dt
    sample = "%d steps x batch %d (the labelled batch), torch.stack + torchvision %s fp32 eager, one host process on %d " \
             "threads" % (done, B, kind, torch.get_num_threads())
    line = {
        "impl": "reference", "metric": "%s_images_per_sec" % kind, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": done, "warmup": args.warmup, "ms_per_step": 1000 * dt / done, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(kind, B, hw, 1), "batch_per_gpu": B,
                   "note": "reference CPU path at the labelled batch size; at N > 1 this is still ONE host process on this "
                           "box's cores (rank 0), not N of them"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--model", default="resnet50", choices=sorted(MODELS))
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--hw", type=int, default=0)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-serving", action="store_true")
    ap.add_argument("--sustain-s", type=float, default=3.0)
    ap.add_argument("--parity-rows", type=int, default=32)
    ap.add_argument("--inflight", type=int, default=2, help="batches in flight on the end-to-end path (engine slots - 1 at most)")
    args = ap.parse_args()
    spec = dict(MODELS[args.model])
    args.batch = args.batch or spec["batch"]
    args.hw = args.hw or spec["hw"]
    if args.hw in spec.get("at", {}):
        spec.update(spec["at"][args.hw])
    if args.impl == "reference":
        return run_reference(args)

    import numpy as np
    import torch
    import rdb_b200
    from rdb_b200 import synth
    from rdb_b200.engine import B200Model

    kind = args.model
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
    B, HW = args.batch, args.hw
    peaks = load_peaks()

    torch.set_num_threads(max(1, (os.cpu_count() or 8) // max(1, world)))
    module = make_module(kind, HW)
    model = B200Model(kind, module, max_batch=B, device=local, in_hw=(HW, HW))
    if kind == "yolov8n":
        # the end-to-end path of a detector returns detections: class-aware NMS on the device (ultralytics' defaults), then
        # D2H of [B, 300, 6] + counts instead of the 2.8 MB [84][8400] head of every image
        model.nms = {"conf": 0.25, "iou": 0.45, "max_det": 300}

    # ---- synthetic requests: NSETS rotating batches so that consecutive steps never reuse L2-resident inputs ----
    NSETS = 4
    per = HW * HW * 3
    offsets = (np.arange(B, dtype=np.uint64) * np.uint64(per))
    hw = np.tile(np.array([[HW, HW]], dtype=np.int32), (B, 1))
    host_sets, dev_sets, img_sets = [], [], []
    for s in range(NSETS):
        imgs = make_requests(kind, B, HW, seed=3000000 + 100000 * rank + 10000 * s)
        h = torch.empty(B * per, dtype=torch.uint8, pin_memory=True)
        h.copy_(imgs.reshape(-1))
        host_sets.append(h)
        dev_sets.append(h.to(dev))
        img_sets.append(imgs)
    out_dev = torch.empty(B, model.out_dim, device=dev, dtype=torch.float32)
    stream = torch.cuda.current_stream(dev)

    def step_resident(i):
        model.collate_u8(dev_sets[i % NSETS], offsets, hw, B, B, 0, B * per)
        model.infer_resident(B, out_dev)

    pending = []
    last_e2e = [None]
    INFLIGHT = max(1, min(args.inflight, B200Model.N_SLOTS))   # pending handles right after a submit

    def step_e2e(i):
        # public async API: pinned host requests -> H2D (side stream) -> collate -> forward -> D2H into pinned memory.
        # Two batches in flight: the copy of step i overlaps the forward of step i-1; every step's result is read
        # (handle.result() waits for that batch's D2H) before the step after next is submitted.
        pending.append(model.infer_packed_async(host_sets[i % NSETS], offsets, hw, B, B, 0, B * per))
        if len(pending) >= INFLIGHT:
            last_e2e[0] = pending.pop(0).result()

    def drain_e2e():
        while pending:
            last_e2e[0] = pending.pop(0).result()

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
    last_resident = out_dev[:, :model.out_logical].cpu()          # the outputs of the last timed resident step
    ms_e2e = timed(step_e2e, K, drain_e2e)
    last_e2e_out = last_e2e[0].clone()                             # ... and of the last timed end-to-end step
    clocks = sampler.window(t0, t1)

    # ---- parity of what was just timed: rows of the last step vs the fp32 CPU reference forward of the same requests ----
    rows = sorted(set(np.linspace(0, B - 1, min(B, max(1, args.parity_rows))).astype(int).tolist()))
    last_set = img_sets[(K - 1) % NSETS]
    ref = cpu_reference_outputs(kind, module, last_set[rows])
    ok_r, fig_r = compare_rows(kind, last_resident[rows], ref)
    if kind == "yolov8n":
        # e2e hands back detections: they must be what the oracle's NMS (oracle/yolov8n.py, ultralytics' rule) makes of the
        # engine's own head for the same requests -- the head itself is the row above, checked against the CPU forward
        from oracle.yolov8n import non_max_suppression
        worst, same_count = 0.0, True
        for r in rows:
            want = non_max_suppression(last_resident[r].view(84, -1))
            got = last_e2e_out[r]
            k = int((got[:, 4] > 0).sum())
            same_count = same_count and k == want.shape[0]
            if k == want.shape[0] and k > 0:
                worst = max(worst, (got[:k] - want).abs().max().item())
        ok_e = same_count and worst <= 1e-3
        fig_e = dict(fig_r, nms_same_counts=bool(same_count), nms_max_abs_vs_oracle=worst)
    else:
        ok_e, fig_e = compare_rows(kind, last_e2e_out[rows], ref)
    parity = {"rows": len(rows), "of_batch": B, "resident": fig_r, "e2e": fig_e,
              "max_abs": max(fig_r["max_abs"], fig_e["max_abs"]), "tolerance": LOGIT_TOL, "ok": bool(ok_r and ok_e),
              "against": "torchvision fp32 CPU forward of the same requests (scheduler.py:441-450), same synthetic weights"}
    if kind == "yolov8n":
        parity["min_box_iou"] = min(fig_r["min_box_iou"], fig_e["min_box_iou"])
        parity["iou_tolerance"] = IOU_TOL

    # ---- sustained: the resident step back to back for >= sustain_s seconds ----
    sustained = None
    if args.sustain_s > 0:
        chunk = max(1, int(0.25 / max(1e-4, ms_res / K * 1e-3)))
        barrier()
        ts0 = time.time()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        n_sus = 0
        while time.time() - ts0 < args.sustain_s:
            for i in range(chunk):
                step_resident(n_sus + i)
            n_sus += chunk
            torch.cuda.current_stream(dev).synchronize()
        e1.record(stream)
        torch.cuda.synchronize(dev)
        ts1 = time.time()
        ms_sus = replicas.max_over_ranks(e0.elapsed_time(e1), dist, dev)
        n_all = replicas.sum_over_ranks(n_sus, dist, dev)
        cs = sampler.window(ts0 + 0.5, ts1)
        sustained = {"value": n_all * B / (ms_sus * 1e-3), "unit": UNIT, "seconds": ms_sus * 1e-3, "steps": int(n_all),
                     "ms_per_step": ms_sus / n_sus, "sm_mhz": cs.get("sm_mhz"), "power_w_max": cs.get("power_w_max"),
                     "reasons": cs.get("reasons")}
    sampler.stop()

    # ---- per-op device times (event nodes inside one captured graph) for the roofline ----
    from rdb_b200 import lowering
    tr = plan_traffic(model.blob, B)
    ms_ops = np.zeros(model.n_ops)
    reps = 3
    for r in range(reps + 1):
        model.collate_u8(dev_sets[r % NSETS], offsets, hw, B, B, 0, B * per)
        ms, _ = model.profile_ops(B)
        if r > 0:
            ms_ops += ms
    ms_ops /= reps
    conv_types = (lowering.OP_CONV, lowering.OP_GATHER_CONV, lowering.OP_STEM)
    conv_idx = [i for i, (t, _, _) in enumerate(tr) if t in conv_types]
    conv_ms = float(ms_ops[conv_idx].sum())
    conv_bytes = float(sum(tr[i][1] for i in conv_idx))
    conv_flops = float(sum(tr[i][2] for i in conv_idx))
    n_conv = len(conv_idx)
    t_hbm = lambda i: tr[i][1] / (peaks["hbm"] * 1e9)
    t_tc = lambda i: tr[i][2] / (peaks["tensor_sustained"] * 1e12)
    conv_bound_ms = 1e3 * sum(max(t_hbm(i), t_tc(i)) for i in conv_idx)
    all_bound_ms = 1e3 * sum(max(t_hbm(i), t_tc(i)) for i in range(len(tr)))
    tensor_bound_launches = sum(1 for i in conv_idx if t_tc(i) > t_hbm(i))
    achieved_gbs = conv_bytes / (conv_ms * 1e-3) / 1e9
    achieved_tf = conv_flops / (conv_ms * 1e-3) / 1e12
    step_ms = ms_res / K
    per_img_s = step_ms * 1e-3 / B
    model_bound_s = max(spec["flops"] / (peaks["tensor_sustained"] * 1e12),
                        (spec["bytes"] + spec["weights"] / B) / (peaks["hbm"] * 1e9))
    # DRAM bytes of the conv family from an `ncu --set full` capture: only a capture of THIS build counts (the file
    # records the hash of the sources it was taken on; tools/conv_traffic.py writes it), otherwise null.
    traffic, traffic_src = None, None
    tpath = os.path.join(ROOT, "profiles", "r02_traffic_%s_b%d.json" % (kind, B))
    if os.path.exists(tpath):
        from rdb_b200 import build as _build
        tj = json.load(open(tpath))
        if tj.get("source_stamp") == _build.source_stamp():
            traffic, traffic_src = tj["traffic_bytes_per_launch"], {"file": os.path.basename(tpath), "commit": tj.get("commit")}
    roofline = {
        "kernel": "convolution family, %d launches per step: tcgen05 implicit GEMM (conv_gemm_persistent_kernel, "
                  "conv_gemm_2cta_kernel, conv3x3_halo_kernel, gather_conv_kernel) and the Toeplitz stem "
                  "(stem_pool_kernel / stem_row_kernel)" % n_conv,
        "bound": "hbm", "achieved": achieved_gbs, "peak": peaks["hbm"], "unit": "GB/s",
        "frac": achieved_gbs / peaks["hbm"], "traffic": traffic, "traffic_source": traffic_src,
        "peak_source": peaks["source"],
        "bytes_per_launch_avg": conv_bytes / n_conv, "ms_per_launch_avg": conv_ms / n_conv,
        "share_of_step": conv_ms / float(ms_ops.sum()),
        "bound_frac": conv_bound_ms / conv_ms,
        "bound_frac_note": "sum over the family's launches of max(bytes / HBM peak, flops / sustained bf16 peak) / measured "
                           "time; %d of %d launches are tensor-bound by that rule" % (tensor_bound_launches, n_conv),
        "tensor_achieved_tflops": achieved_tf, "tensor_peak_tflops": peaks["tensor_sustained"],
        "tensor_frac": achieved_tf / peaks["tensor_sustained"],
        "step_bound_frac": all_bound_ms / float(ms_ops.sum()),
        "model_frac_of_layer_fused_roofline": model_bound_s / per_img_s,
        "per_op_ms_top": sorted(((float(ms_ops[i]), model.op_names[i]) for i in range(model.n_ops)), reverse=True)[:6],
    }

    value = world * B * K / (ms_res * 1e-3)
    e2e_value = world * B * K / (ms_e2e * 1e-3)
    line = {
        "metric": "%s_images_per_sec" % kind, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic",
        "config": {"workload": workload_name(kind, B, HW, world),
                   "batch_per_gpu": B, "replicas": world, "input_hw": HW, "weights": "seeded synthetic (rdb_b200.synth)",
                   "l2": "working set >> L2: %d rotating input batches, ~%.1f GB activation traffic per step"
                         % (NSETS, B * spec["bytes"] / 1e9),
                   "cuda_graph": True},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(world * (B * per + B * 16)),
                "d2h_bytes_per_step": int(world * B * (300 * 6 * 4 + 4 if model.nms else model.out_dim * 4)),
                "ms_per_step": ms_e2e / K,
                "pipeline": "%d batches in flight (H2D of step i overlaps the forward of step i-1)" % INFLIGHT},
        "gpu_launches": int(launches),
        "parity_checked": parity,
        "roofline": roofline,
        "clocks": clocks,
    }
    if sustained is not None:
        line["sustained"] = sustained

    if rank == 0 and world == 1 and not args.no_serving:
        try:
            from rdb_b200 import serving
            line["serving"] = serving.bench_serving(model, kind, spec["slo_name"], B, capacity_img_s=e2e_value,
                                                    seconds=float(os.environ.get("B200_BENCH_SERVE_S", "3")))
            if line["serving"]:
                line["p99_ms"] = line["serving"].get("p99_latency_ms")
                line["slo_violations"] = line["serving"].get("slo_violations")
        except ImportError:
            pass

    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle.models import time_cpu_reference
        torch.set_num_threads(os.cpu_count() or 1)
        imgs = make_requests(kind, B, HW, seed=910000)
        if kind == "yolov8n":
            reqs = list(imgs.permute(0, 3, 1, 2).float().div(255))
        else:
            reqs = list(synth.normalize_u8(imgs, synth.IMAGENET_MEAN, synth.IMAGENET_STD))
        best, med, nrep, threads = time_cpu_reference(module, reqs, budget_s=12.0)
        line["cpu_baseline"] = {"value": med, "unit": UNIT, "cores": threads, "kind": "port", "best": best,
                                "sample": "%d reps x batch %d: torch.stack + torchvision %s fp32 eager "
                                          "(scheduler.py:443-450), same synthetic weights" % (nrep, B, kind)}
    if rank == 0:
        print(json.dumps(line), flush=True)
    model.close()
    if dist is not None:
        dist.destroy_process_group()
    if not parity["ok"]:
        sys.stderr.write("bench.py: outputs of the timed steps differ from the CPU reference: %s\n" % json.dumps(parity))
        sys.exit(3)


if __name__ == "__main__":
    main()
