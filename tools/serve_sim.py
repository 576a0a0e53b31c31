#!/usr/bin/env python
"""Serving run through the drop-in surface at deployment shape: B200Scheduler (one model-actor PROCESS per GPU) fed by
client PROCESSES that write uint8 requests into the shared-memory request rings; every latency figure is read from
RequestQueue.get_stats (293-project/src/scheduler.py:343-372) exactly as the reference's metrics display does.

  python tools/serve_sim.py --models resnet --gpus 1 --rate 70000 --seconds 5 --batch 256
  python tools/serve_sim.py --models resnet --gpus 8 --rate 560000 --seconds 5 --clients 16
  python tools/serve_sim.py --models resnet,efficientnet,yolov8 --gpus 8 --rates 200000,40000,20000 --seconds 6   # configs[4]
  python tools/serve_sim.py --models resnet --gpus 1 --slope 3000 --seconds 10        # reference load pattern, rate = t * slope

Schedules: by default a static one (every GPU serves every listed model, occupancy split evenly, batch sizes from
--batch / --batches) installed with set_schedule; --dynamic lets the rate monitor and the squishy bin packing decide.
Prints one JSON line.
"""
import argparse
import json
import multiprocessing as mp
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

DEFAULT_BATCH = {"resnet": 256, "shufflenet": 512, "efficientnet": 128, "yolov8": 32}
PROFILE_CSV = {"resnet": "resnet50", "shufflenet": "shufflenet", "efficientnet": "efficientnetv2", "yolov8": "yolov8n"}


def client_main(*args):
    """One load-generator process (rdb_b200.serving.client_process_main)."""
    import rdb_b200  # noqa: F401
    from rdb_b200.serving import client_process_main
    return client_process_main(*args)


def load_profiles(models, max_bs=512, kind="forward"):
    """kind "forward": the reference's forward-only profile; "e2e": a batch served alone from host memory
    (tools/profile_batches.py --e2e), the one that does not over-promise a saturated node."""
    from rdb_b200 import profiler, serving
    out = {}
    for m in models:
        path = os.path.join(ROOT, "profiles", "b200_batch_profiles", "%s_b200_summary.csv" % PROFILE_CSV[m])
        if kind == "e2e":
            path = os.path.join(ROOT, "profiles", "b200_batch_profiles_e2e", "%s_b200_e2e_summary.csv" % PROFILE_CSV[m])
        d = profiler.load_csv_to_dict(path) if os.path.exists(path) else {}
        d = {k: v for k, v in d.items() if 'avg_latency_ms' in v}
        if not d:
            raise SystemExit("no batch profile for %s (%s)" % (m, path))
        for v in d.values():
            v.setdefault('peak_memory_mb', 1000.0)
        out[m] = serving.dense_profile(d, max(max_bs, max(d)))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--models", default="resnet")
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--rate", type=float, default=60000.0, help="total offered req/s (single model)")
    ap.add_argument("--rates", default="", help="offered req/s per model, comma separated (mixed runs)")
    ap.add_argument("--slope", type=float, default=0.0)
    ap.add_argument("--seconds", type=float, default=5.0)
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--batches", default="")
    ap.add_argument("--clients", type=int, default=0, help="client processes per model (default: 2 per GPU, at most 16)")
    ap.add_argument("--burst", type=int, default=16)
    ap.add_argument("--dynamic", action="store_true")
    ap.add_argument("--profiles", default="forward", choices=("forward", "e2e"),
                    help="batch profiles the policy and the stale-request rule use")
    ap.add_argument("--pipeline-depth", type=int, default=2)
    ap.add_argument("--no-shards", action="store_true", help="one shared ring per model instead of one per actor")
    ap.add_argument("--max-batch", type=int, default=0, help="engine capacity (default: the largest batch size; 512 with --dynamic)")
    ap.add_argument("--metrics-file", default=os.path.join(ROOT, "gpurun_out", "metrics.json"))
    ap.add_argument("--out", default="")
    args = ap.parse_args()

    import rdb_b200  # noqa: F401
    from rdb_b200 import packing, synth
    from rdb_b200.request_queue import models_config
    from rdb_b200.scheduler import B200Scheduler
    from rdb_b200.worker import MODEL_KINDS

    models = [m for m in args.models.split(",") if m]
    rates = [float(x) for x in args.rates.split(",")] if args.rates else [args.rate / len(models)] * len(models)
    batches = [int(x) for x in args.batches.split(",")] if args.batches else \
        [args.batch or DEFAULT_BATCH[m] for m in models]
    profile = load_profiles(models, kind=args.profiles)
    registry = {}
    for m in models:
        kind = MODEL_KINDS[m]
        registry[m] = synth.make_model(kind) if kind == "yolov8n" else synth.make_model(kind, n_proto=8)
    os.makedirs(os.path.dirname(args.metrics_file), exist_ok=True)
    t_load = time.time()
    sched = B200Scheduler(profile, monitoring_interval=1.0, rate_change_threshold=0.05, model_registry=registry,
                          num_gpus=args.gpus, worker_mode="process",
                          max_batch=args.max_batch or (512 if args.dynamic else max(batches)), preload=models,
                          pipeline_depth=args.pipeline_depth, metrics_file=args.metrics_file, metrics_interval=1.0,
                          max_queue_size=max(2000, 6 * max(batches) * args.gpus),
                          in_flight_slots=4 * max(batches) * args.gpus + 256, shard_queues=not args.no_shards)
    out = {"models": models, "gpus": args.gpus, "seconds": args.seconds, "batches": dict(zip(models, batches))}
    try:
        if not sched.wait_ready(900):
            raise SystemExit("actors did not come up")
        out["actor_load_s"] = time.time() - t_load
        if args.dynamic:
            sched.start_monitoring()
        else:
            nodes = []
            for g in range(args.gpus):
                sessions = []
                duty = 0.0
                for m, b, r in zip(models, batches, rates):
                    lat = profile[m][b]['avg_latency_ms']
                    duty += lat
                    sessions.append((packing.session(m, models_config[m]['SLO'] / 2.2, r / args.gpus, b), lat))
                nodes.append(packing.node([(s, lat / duty) for s, lat in sessions], duty_cycle=duty))
            sched.set_schedule(nodes)
        time.sleep(0.5)
        ctx = mp.get_context("spawn")
        report = ctx.Queue()
        n_clients = args.clients or min(16, 2 * args.gpus)
        barrier = ctx.Barrier(n_clients * len(models) + 1)
        procs = []
        for m, r in zip(models, rates):
            for c in range(n_clients):
                procs.append(ctx.Process(target=client_main, daemon=True,
                                         args=(sched.request_queues[m], sched.model_registry[m]['in_hw'][0], r / n_clients,
                                               args.seconds, args.slope / n_clients, args.burst, "%s_c%d_" % (m, c),
                                               100 * len(procs), barrier, report)))
        for p in procs:
            p.start()
        barrier.wait(timeout=300)
        t0 = time.time()
        reports = [report.get(timeout=args.seconds + 120) for _ in procs]
        t_feed = time.time() - t0
        deadline = time.time() + 30
        while time.time() < deadline:
            if all(q.get_stats()["completed_requests"] + q.get_stats()["discarded_stale"] >= q.get_stats()["total_requests"]
                   for q in sched.request_queues.values()):
                break
            time.sleep(0.005)
        wall = time.time() - t0
        for p in procs:
            p.join(timeout=10)
        out.update(wall_s=wall, feed_s=t_feed, pattern=("slope %.0f/s^2" % args.slope) if args.slope else "constant",
                   clients_per_model=n_clients, burst=args.burst, dynamic=args.dynamic,
                   pipeline_depth=args.pipeline_depth, shards=1 if args.no_shards else args.gpus)
        per = {}
        total_served = 0
        for m, r in zip(models, rates):
            st = sched.request_queues[m].get_stats()
            sub = sum(x[1] for x in reports if x[0].startswith(m + "_c"))
            drop = sum(x[2] for x in reports if x[0].startswith(m + "_c"))
            total_served += st["completed_requests"]
            per[m] = {"offered_rate": r if not args.slope else None, "submitted": sub, "dropped_full_queue": drop,
                      "served": st["completed_requests"], "batches": st["completed_batches"],
                      "mean_batch": st["completed_requests"] / max(1, st["completed_batches"]),
                      "throughput_img_s": st["completed_requests"] / wall, "slo_ms": sched.request_queues[m].slo_target,
                      "slo_violations": st["slo_violations"], "discarded_stale": st["discarded_stale"],
                      "avg_latency_ms": st["avg_latency"], "p95_latency_ms": st["p95_latency"],
                      "p99_latency_ms": st["p99_latency"], "max_latency_ms": st["max_latency"]}
        out["per_model"] = per
        out["throughput_img_s"] = total_served / wall
        if args.dynamic:
            out["schedule_updates"] = sched.metrics["schedule_updates"]
            out["nodes"] = [n.to_dict() for n in sched.nodes]
    finally:
        out["workers"] = sched.stop(timeout=30)
        sched.close()
    line = json.dumps(out, default=str)
    print(line)
    if args.out:
        with open(args.out, "a") as f:
            f.write(line + "\n")


if __name__ == "__main__":
    main()
