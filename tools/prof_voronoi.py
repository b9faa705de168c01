"""Time incremental Voronoi inserts + k-means++ draws on the cached prof graph (see prof_sssp.py)."""
import sys, time
from pathlib import Path
import numpy as np, torch
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from vqvae_b200 import device_graph as dg
images = int(sys.argv[1]); K = int(sys.argv[2]) if len(sys.argv) > 2 else 64
dev = torch.device("cuda")
ip, ix, dt = torch.load(f"/tmp/prof_graph_{images}.pt")
g = dg.DeviceCSR(ip.to(dev), ix.to(dev), dt.to(dev), ip.numel() - 1)
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
print("insert ms:", " ".join(f"{1e3*t:.2f}" for t in tins[:12]), "... median of rest", f"{1e3*np.median(tins[12:]):.3f}")
print("draw   ms:", " ".join(f"{1e3*t:.2f}" for t in tdraw[:6]), "... median of rest", f"{1e3*np.median(tdraw[6:]):.3f}")
