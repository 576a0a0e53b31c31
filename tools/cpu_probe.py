#!/usr/bin/env python
"""How many cores does this box really give us?  N busy processes for 1 s each: aggregate iterations/s vs N."""
import multiprocessing as mp, os, time


def spin(q):
    t0 = time.perf_counter(); n = 0
    while time.perf_counter() - t0 < 1.0:
        for _ in range(1000):
            n += 1
    q.put(n)


if __name__ == "__main__":
    for f in ("/sys/fs/cgroup/cpu.max", "/sys/fs/cgroup/cpu/cpu.cfs_quota_us", "/sys/fs/cgroup/cpu/cpu.cfs_period_us", "/sys/fs/cgroup/cpuset.cpus.effective"):
        try:
            print(f, open(f).read().strip())
        except Exception as e:
            print(f, "n/a")
    print("nproc", os.cpu_count(), "affinity", len(os.sched_getaffinity(0)))
    base = None
    for n in (1, 2, 4, 8, 12, 16, 24, 32):
        q = mp.Queue()
        ps = [mp.Process(target=spin, args=(q,)) for _ in range(n)]
        t0 = time.perf_counter()
        [p.start() for p in ps]
        tot = sum(q.get() for _ in ps)
        [p.join() for p in ps]
        base = base or tot
        print("procs %2d: aggregate %.1fx of one process (wall %.2f s)" % (n, tot / base, time.perf_counter() - t0))
