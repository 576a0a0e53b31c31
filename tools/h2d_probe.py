#!/usr/bin/env python
"""Two (or N) processes, one per GPU, each looping: 38.5 MB H2D copy -> 3 ms of GPU work -> 1 MB D2H -> event sync, with
two copies in flight.  Source memory: torch pinned (cudaHostAlloc) or a shared-memory segment page-locked with
cudaHostRegister; optionally with CPU writer processes storing into the segment meanwhile.  Prints ms per iteration."""
import multiprocessing as mp
import os
import sys
import time
from multiprocessing import shared_memory

import numpy as np

NB = 256 * 150528


def writer(name, stop):
    shm = shared_memory.SharedMemory(name=name)
    a = np.frombuffer(shm.buf, dtype=np.uint8)
    src = np.random.randint(0, 255, 16 * 150528, dtype=np.uint8)
    i = 0
    while not stop.is_set():
        o = (i * src.size) % (a.size - src.size)
        a[o:o + src.size] = src
        i += 1
        time.sleep(0.0002)
    del a
    shm.close()


def worker(gpu, kind, name, out, iters, sync):
    os.environ["CUDA_VISIBLE_DEVICES"] = str(gpu)
    import torch
    dev = torch.device("cuda", 0)
    if kind == "pinned":
        host = torch.empty(4 * NB, dtype=torch.uint8, pin_memory=True)
    else:
        shm = shared_memory.SharedMemory(name=name)
        arr = np.frombuffer(shm.buf, dtype=np.uint8)[:4 * NB]
        flags = 1 if kind == "registered_portable" else 0
        rc = torch.cuda.cudart().cudaHostRegister(arr.ctypes.data, arr.size, flags)
        assert int(rc) == 0, rc
        host = torch.from_numpy(arr)
    stage = [torch.empty(NB, dtype=torch.uint8, device=dev) for _ in range(3)]
    outh = [torch.empty(256000, dtype=torch.float32, pin_memory=True) for _ in range(3)]
    a = torch.randn(4096, 4096, device=dev, dtype=torch.bfloat16)
    cs = torch.cuda.Stream(dev)
    ev_c = [torch.cuda.Event() for _ in range(3)]
    ev_d = [torch.cuda.Event(blocking=(sync == "blocking")) for _ in range(3)]
    res = torch.empty(256000, device=dev)
    def submit(i):
        s = i % 3
        with torch.cuda.stream(cs):
            stage[s].copy_(host[(i % 4) * NB:(i % 4 + 1) * NB], non_blocking=True)
            ev_c[s].record(cs)
        torch.cuda.current_stream().wait_event(ev_c[s])
        for _ in range(22):                       # ~3 ms of tensor-core work
            b = a @ a
        outh[s].copy_(res, non_blocking=True)
        ev_d[s].record()
    for i in range(6):
        submit(i); ev_d[i % 3].synchronize()
    t0 = time.perf_counter()
    submit(0)
    for i in range(1, iters):
        submit(i)
        ev_d[(i - 1) % 3].synchronize()
    ev_d[(iters - 1) % 3].synchronize()
    out.put((gpu, kind, (time.perf_counter() - t0) / iters * 1e3))


if __name__ == "__main__":
    ngpu = int(sys.argv[1])
    ctx = mp.get_context("spawn")
    for kind, nwriters, sync in (("pinned", 0, "spin"), ("registered", 0, "spin"), ("registered", 3 * ngpu, "spin"),
                                 ("registered_portable", 0, "spin"), ("registered", 0, "blocking")):
        shms = [shared_memory.SharedMemory(create=True, size=8 * NB) for _ in range(ngpu)]
        out, stop = ctx.Queue(), ctx.Event()
        ps = [ctx.Process(target=worker, args=(g, kind, shms[g].name, out, 200, sync)) for g in range(ngpu)]
        ws = [ctx.Process(target=writer, args=(shms[k % ngpu].name, stop)) for k in range(nwriters)]
        [p.start() for p in ps + ws]
        res = sorted(out.get(timeout=300) for _ in ps)
        stop.set()
        [p.join(timeout=20) for p in ps + ws]
        print("%d GPUs  %-20s writers %2d  %-8s  ms/iter per GPU: %s" % (ngpu, kind, nwriters, sync, ["%.2f" % r[2] for r in res]), flush=True)
        for s in shms:
            s.close(); s.unlink()
