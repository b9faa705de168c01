"""k-means++ chain on a kNN graph of synthetic latents (Euclidean weights): per-iteration time of the Voronoi insert and
of the draw, and the kernels behind them.  usage: prof_kpp_chain.py [IMAGES] [K]"""
import sys, time
from pathlib import Path
import numpy as np, torch
from torch.profiler import profile, ProfilerActivity
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from vqvae_b200 import device_graph as dg
from vqvae_b200.geo.knn_graph_optimized import knn_graph_device
images = int(sys.argv[1]) if len(sys.argv) > 1 else 60000
K = int(sys.argv[2]) if len(sys.argv) > 2 else 256
z = torch.randn(images * 49, 16, generator=torch.Generator().manual_seed(42)).cuda()
g, _, _ = knn_graph_device(z, 20, "distance", "union")
g.adjacency(); torch.cuda.synchronize()
st = dg.VoronoiState(g)
rng = np.random.RandomState(0)
c = int(rng.randint(g.n))
tins, tdraw = [], []
for k in range(K):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    st.insert(c); torch.cuda.synchronize(); t1 = time.perf_counter()
    nxt, exact, _ = st.draw(rng.random_sample()); t2 = time.perf_counter()
    tins.append(t1 - t0); tdraw.append(t2 - t1); c = nxt
print(f"n={g.n} nnz={g.nnz}")
print("insert ms:", " ".join(f"{1e3*t:.2f}" for t in tins[:10]), "... median of the rest", f"{1e3*np.median(tins[10:]):.3f}")
print("draw   ms:", " ".join(f"{1e3*t:.2f}" for t in tdraw[:6]), "... median of the rest", f"{1e3*np.median(tdraw[6:]):.3f}")
t0 = time.perf_counter()
for k in range(64):
    st.insert(c); c, _, _ = st.draw(rng.random_sample())
torch.cuda.synchronize(); print(f"unsynchronised loop: {1e3*(time.perf_counter()-t0)/64:.3f} ms per iteration")
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for k in range(32):
        st.insert(c); c, _, _ = st.draw(rng.random_sample())
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=14, max_name_column_width=60))
