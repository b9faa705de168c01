"""SSSP-only driver for timing / ncu: kNN graph of synthetic latents with hash weights (no JVP stage).
usage: prof_sssp.py IMAGES SLOTS [LAUNCHES]   (latents = IMAGES*49)"""
import sys, time
from pathlib import Path
import numpy as np, torch
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from vqvae_b200 import device_graph as dg
from vqvae_b200.geo.knn_graph_optimized import knn_graph_device

images, slots = int(sys.argv[1]), int(sys.argv[2])
launches = int(sys.argv[3]) if len(sys.argv) > 3 else 3
dev = torch.device("cuda")
z = torch.randn(images, 16, 7, 7, generator=torch.Generator().manual_seed(42)).permute(0, 2, 3, 1).reshape(-1, 16).contiguous().to(dev)
t0 = time.time()
g, _, _ = knn_graph_device(z, 20, "connectivity", "union")
torch.cuda.synchronize(); t_knn = time.time() - t0
# symmetric pseudo-random weights in [0.25, 2.25): integer hash of (min(i,j), max(i,j))
rows = torch.repeat_interleave(torch.arange(g.n, device=dev), g.indptr[1:] - g.indptr[:-1])
cols = g.indices.long()
a, b = torch.minimum(rows, cols), torch.maximum(rows, cols)
h = (a * 2654435761 + b * 40503) & 0xFFFFFFFF
h = ((h ^ (h >> 15)) * 2246822519) & 0xFFFFFFFF
h = (h ^ (h >> 13)) & 0xFFFFF
g.data = (0.25 + 2.0 * h.double() / (1 << 20)).float()
del rows, cols, a, b, h
g.adjacency(); torch.cuda.synchronize()
rng = np.random.RandomState(0)
for i in range(launches):
    src = rng.choice(g.n, slots, replace=False)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    D, st = dg.sssp_multi(g, src, slots=slots, stats=True)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print(f"n={g.n} nnz={g.nnz} slots={slots} launch {i}: {ms:.3f} ms  {slots*g.nnz/ms/1e6:.2f} G edge-relax/s  "
          f"roofline {slots*16*(g.nnz+g.n)/ms/1e6/6542.1:.3f}  rounds={st['rounds']} scans={st['scans']} "
          f"relax/edge={st['relaxations']/(slots*g.nnz):.3f} knn_s={t_knn:.2f}", flush=True)
