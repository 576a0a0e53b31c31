"""Request-sharded replicas: one model-actor process per GPU, no data-path collective.

The reference runs one `@ray.remote(num_gpus=1)` GPUWorker per GPU that pulls from per-model queues
(293-project/src/scheduler.py:374,679-714).  Requests are independent and weights are replicated, so multi-GPU serving
is N independent replicas; torch.distributed is used only to line the replicas up for measurements (barrier, max of the
per-rank elapsed time) — with NCCL on the GPUs, or gloo on CPUs for the tests.
"""
import os
from typing import List, Sequence


def replica_env():
    """(rank, world_size, local_rank) as set by torchrun; (0, 1, 0) when launched as a plain process."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def shard_round_robin(items: Sequence, rank: int, world: int) -> List:
    """Requests are dealt to replicas in arrival order: request i goes to replica i % world (what N actors pulling from a
    shared FIFO converge to under uniform service times).  Disjoint and exhaustive by construction."""
    if not (0 <= rank < world):
        raise ValueError("rank %d outside world of %d" % (rank, world))
    return [x for i, x in enumerate(items) if i % world == rank]


def init_process_group(backend=None):
    """Initialises torch.distributed from the torchrun environment (MASTER_ADDR should be 127.0.0.1 on one node)."""
    import torch
    import torch.distributed as dist
    rank, world, local = replica_env()
    if world == 1:
        return None
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    kw = {}
    if backend == "nccl":
        kw["device_id"] = torch.device("cuda", local)
    dist.init_process_group(backend, rank=rank, world_size=world, **kw)
    return dist


def barrier(dist):
    if dist is not None:
        dist.barrier()


def max_over_ranks(value: float, dist, device=None) -> float:
    """The job-level elapsed time of a measurement is the slowest replica's."""
    if dist is None:
        return float(value)
    import torch
    t = torch.tensor([float(value)], dtype=torch.float64, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(value: float, dist, device=None) -> float:
    if dist is None:
        return float(value)
    import torch
    t = torch.tensor([float(value)], dtype=torch.float64, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def job_throughput(units_this_rank: float, elapsed_ms_this_rank: float, dist, device=None) -> float:
    """Whole-job units/s = units processed by all replicas / max-over-ranks elapsed time."""
    total = sum_over_ranks(units_this_rank, dist, device)
    ms = max_over_ranks(elapsed_ms_this_rank, dist, device)
    return total / (ms * 1e-3)
