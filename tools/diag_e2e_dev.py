"""Where does the `e2e_device_graph` step of bench.py (pinned CSR -> DeviceCSR -> sssp_multi_to_host -> pinned D) spend
its time?  Synthetic kNN-like graph of C2's size; prints a per-stage wall-clock breakdown (synchronised)."""
import sys
import time
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from vqvae_b200 import device_graph as dg  # noqa: E402

dev = torch.device("cuda")
n, k = 2_940_000, 16
rng = np.random.RandomState(0)
# ring + random chords, symmetric, sorted rows
i = np.repeat(np.arange(n, dtype=np.int64), k)
j = (i + rng.randint(1, 2000, size=i.size)) % n
from scipy import sparse
W = sparse.coo_matrix((rng.rand(i.size).astype(np.float32) + 0.1, (i, j)), shape=(n, n)).tocsr()
W = W.maximum(W.T).tocsr()
W.sort_indices()
print("graph", W.shape, W.nnz, flush=True)
h_indptr = torch.from_numpy(W.indptr.astype(np.int64)).pin_memory()
h_indices = torch.from_numpy(W.indices.astype(np.int32)).pin_memory()
h_data = torch.from_numpy(W.data.astype(np.float32)).pin_memory()
S = 64
h_D = torch.empty((S, n), dtype=torch.float32).pin_memory()
src = rng.choice(n, S, replace=False)


def t():
    torch.cuda.synchronize()
    return time.perf_counter()


for rep in range(3):
    t0 = t()
    gg = dg.DeviceCSR(h_indptr.to(dev, non_blocking=True), h_indices.to(dev, non_blocking=True), h_data.to(dev, non_blocking=True), n)
    t1 = t()
    gg.adjacency(False)
    t2 = t()
    dg.sssp_multi_to_host(gg, src, h_D)
    t3 = t()
    D = dg.sssp_multi(gg, src)
    t4 = t()
    print(f"rep {rep}: upload {1e3*(t1-t0):.1f} ms  adjacency {1e3*(t2-t1):.1f}  sssp_multi_to_host {1e3*(t3-t2):.1f}  sssp_multi(device) {1e3*(t4-t3):.1f}", flush=True)
    del D
