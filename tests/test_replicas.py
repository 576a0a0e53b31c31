"""N > 1 host-side logic of the request-sharded replicas, exercised with world_size 2 over gloo on the CPU."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import rdb_b200  # noqa: F401
from rdb_b200 import replicas


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    d = replicas.init_process_group("gloo")
    assert replicas.replica_env() == (rank, world, rank)
    ids = ["req%d" % i for i in range(11)]
    mine = replicas.shard_round_robin(ids, rank, world)
    gathered = [None] * world
    d.all_gather_object(gathered, mine)
    flat = sorted(x for part in gathered for x in part)
    assert flat == sorted(ids) and len(set(flat)) == len(ids)          # disjoint, exhaustive
    replicas.barrier(d)
    elapsed = 10.0 + 5.0 * rank                                        # rank 1 is the slow replica
    assert replicas.max_over_ranks(elapsed, d) == 15.0
    thr = replicas.job_throughput(len(mine), elapsed, d)
    assert abs(thr - 11 / 0.015) < 1e-6                                # all units over the slowest replica's time
    torch.save({"rank": rank, "n": len(mine)}, os.path.join(out_dir, "r%d.pt" % rank))
    d.destroy_process_group()


def test_two_replicas_over_gloo(tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    got = [torch.load(tmp_path / ("r%d.pt" % r)) for r in range(2)]
    assert [g["n"] for g in got] == [6, 5]


def test_single_process_defaults():
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        os.environ.pop(k, None)
    assert replicas.replica_env() == (0, 1, 0)
    assert replicas.init_process_group() is None
    assert replicas.max_over_ranks(3.5, None) == 3.5
    assert replicas.job_throughput(256, 4.0, None) == 64000.0
    with pytest.raises(ValueError):
        replicas.shard_round_robin([1, 2], 2, 2)
