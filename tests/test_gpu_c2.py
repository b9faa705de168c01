"""Parity at the HEADLINE scale (BASELINE configs[1]: 2.94M latents d=16, k=20 union) -- the only scale where the
k-means++ chain is numerically delicate (SURVEY 7 hard part 1: 2.94M cdf boundaries per draw).  Slow by design:
builds the C2 graph on the GPU once, then lets the CPU oracle (scipy Dijkstra, NumPy RandomState.choice) run a
prefix of the chain on that very graph.  GEOCB_C2_PREFIX (default 10) centres; GEOCB_C2_IMAGES (default 60000)."""
import os

import numpy as np
import pytest
import torch

import geo_oracle as oracle
from conftest import fashion_decoder

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def c2():
    from vqvae_b200.pipeline import build_codebook_device
    n_img = int(os.environ.get("GEOCB_C2_IMAGES", "60000"))
    z4 = torch.randn(n_img, 16, 7, 7, generator=torch.Generator().manual_seed(42))
    res = build_codebook_device(z4, fashion_decoder().cuda(), k=20, sym="union", K=2, verbose=False)
    return res


def test_c2_knn_slice_vs_sklearn(c2):
    """kNN index rows of a query slice against ALL latents: identical to sklearn's brute search
    (src/geo/knn_graph_optimized.py:39-42), order included."""
    from sklearn.neighbors import NearestNeighbors
    from vqvae_b200 import device_graph as dg
    zf = c2.z_flat.cpu().numpy()
    M = zf.shape[0]
    q = int(os.environ.get("GEOCB_C2_KNN_ROWS", "8192"))
    rows = np.r_[0:q // 2, M - q // 2:M]  # both ends of the index range
    nn_ = NearestNeighbors(n_neighbors=21, algorithm="auto", metric="euclidean").fit(zf)
    _, idx_cpu = nn_.kneighbors(zf[rows])
    lo, _, _ = dg.knn_topk(c2.z_flat, 21, 0, q // 2)
    hi, _, _ = dg.knn_topk(c2.z_flat, 21, M - q // 2, M)
    idx_gpu = torch.cat([lo, hi]).cpu().numpy()
    assert np.array_equal(idx_gpu, idx_cpu)


def test_c2_kpp_prefix_and_sssp_vs_oracle(c2):
    """First centres of the k-means++ chain and the running d_min (float32) on the C2 geodesic graph: bit-equal to
    the oracle's (kmeans_optimized.py:40-61), and full SSSP rows from those centres bit-equal to scipy."""
    from vqvae_b200 import device_graph as dg
    from vqvae_b200.geo import kmeans_optimized as km
    K = int(os.environ.get("GEOCB_C2_PREFIX", "10"))
    g = c2.g_lcc
    W = g.to_scipy()
    trace = []
    exp = oracle.kpp_seed(W, K, seed=42, trace=trace)
    state = dg.VoronoiState(g)
    got, exact_draws = km._seed_kpp(state, K, 42, insert_last=False)
    print(f"C2 chain prefix K={K}: exact-path draws {exact_draws}, centres {got}")
    assert got == exp
    # after K-1 inserts the device d_min equals the oracle's running minimum, bit for bit
    assert np.array_equal(state.dmin32.cpu().numpy(), trace[-1])
    D = dg.sssp_multi(g, np.asarray(exp[:3])).cpu().numpy()
    assert np.array_equal(D, oracle.sssp_multi(W, exp[:3]))
