"""Host copy bandwidth pageable -> pinned (the staging step of device_graph.h2d): torch's own copy against explicit
thread pools; prints the environment that decides torch's parallelism."""
import os, sys, time
from concurrent.futures import ThreadPoolExecutor
import numpy as np
import torch
print("cpu_count", os.cpu_count(), "affinity", len(os.sched_getaffinity(0)), "torch threads", torch.get_num_threads(),
      "OMP_NUM_THREADS", os.environ.get("OMP_NUM_THREADS"), "MKL", os.environ.get("MKL_NUM_THREADS"), flush=True)
n = 366 << 20
src = torch.from_numpy(np.random.RandomState(0).randint(0, 255, n, dtype=np.uint8))
dst = torch.empty(n, dtype=torch.uint8).pin_memory()
def bw(fn, reps=5):
    fn(); ts = []
    for _ in range(reps):
        t0 = time.perf_counter(); fn(); ts.append(time.perf_counter() - t0)
    return n / min(ts) / 1e9, n / max(ts) / 1e9
print("torch copy_: best %.1f GB/s worst %.1f GB/s" % bw(lambda: dst.copy_(src)), flush=True)
for parts in (4, 8, 16):
    pool = ThreadPoolExecutor(max_workers=parts)
    step = -(-n // parts)
    def par():
        fs = [pool.submit(dst[o:o + step].copy_, src[o:o + step]) for o in range(0, n, step)]
        for f in fs: f.result()
    print(f"{parts} python threads x copy_: best %.1f GB/s worst %.1f GB/s" % bw(par), flush=True)
    dn, sn = dst.numpy(), src.numpy()
    def parn():
        fs = [pool.submit(np.copyto, dn[o:o + step], sn[o:o + step]) for o in range(0, n, step)]
        for f in fs: f.result()
    print(f"{parts} python threads x np.copyto: best %.1f GB/s worst %.1f GB/s" % bw(parn), flush=True)
