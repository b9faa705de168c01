import sys, time
sys.path.insert(0, "/root/repo")
import numpy as np, torch
import bench
from vqvae_b200 import device_graph as dg
from vqvae_b200.geo import kmeans_optimized as km
from vqvae_b200.geo.knn_graph_optimized import knn_graph_device
z = torch.randn(60000 * 49, 16, generator=torch.Generator().manual_seed(42)).cuda()
g, _, _ = knn_graph_device(z, 20, "distance", "union")
g.adjacency(); torch.cuda.synchronize()
for rep in range(2):
    st = dg.VoronoiState(g)
    t0 = time.perf_counter()
    centers, exact = km._seed_kpp(st, 512, 42, insert_last=True)
    torch.cuda.synchronize()
    print(f"chain of 512: {time.perf_counter()-t0:.3f} s, exact-path draws {exact}", flush=True)
