"""Where does a dijkstra_multi_source(scipy CSR, 64 sources) call spend its time?  Synthetic symmetric graph of C2's
size; wall-clock per stage, synchronised (analysis only)."""
import sys
import time
from pathlib import Path

import numpy as np
import torch
from scipy import sparse

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from vqvae_b200 import device_graph as dg  # noqa: E402
from vqvae_b200.geo.geo_shortest_paths import dijkstra_multi_source, graph_to_device  # noqa: E402

n, k = 2_940_000, 16
rng = np.random.RandomState(0)
i = np.repeat(np.arange(n, dtype=np.int64), k)
j = (i + rng.randint(1, 2000, size=i.size)) % n
W = sparse.coo_matrix((rng.rand(i.size).astype(np.float32) + 0.1, (i, j)), shape=(n, n)).tocsr()
W = W.maximum(W.T).tocsr()
W.sort_indices()
print("graph", W.shape, W.nnz, W.indices.dtype, W.indptr.dtype, W.data.dtype, "torch threads", torch.get_num_threads(), flush=True)
src = rng.choice(n, 64, replace=False)


def t():
    torch.cuda.synchronize()
    return time.perf_counter()


for rep in range(4):
    t0 = t()
    g = graph_to_device(W)
    t1 = t()
    g.adjacency(False)
    t2 = t()
    D = torch.empty((64, n), dtype=torch.float32, pin_memory=True)
    t3 = t()
    dg.sssp_multi_to_host(g, src, D)
    t4 = t()
    del g, D
    t5 = t()
    out = dijkstra_multi_source(W, src)
    t6 = t()
    print(f"rep {rep}: graph_to_device {1e3*(t1-t0):.1f} ms  adjacency {1e3*(t2-t1):.1f}  pinned result alloc {1e3*(t3-t2):.1f}  "
          f"sssp_multi_to_host {1e3*(t4-t3):.1f}  | whole dijkstra_multi_source {1e3*(t6-t5):.1f}", flush=True)
    del out
# upload alone, by array
for rep in range(3):
    t0 = t(); a = dg.h2d(W.indices, torch.device("cuda")); t1 = t(); b = dg.h2d(W.data, torch.device("cuda")); t2 = t()
    print(f"h2d indices {1e3*(t1-t0):.1f} ms ({W.indices.nbytes/1e6/(t1-t0)/1e3:.1f} GB/s)  data {1e3*(t2-t1):.1f} ms", flush=True)
