"""The plain-C restatement (oracle/c/geo_oracle.c) against the reference-generated golden vectors and the
library-backed oracle.  CPU only."""
import numpy as np

import c_oracle
import geo_oracle as oracle
from conftest import s1_latents, csr_from


def test_c_knn_matches_reference_indices(s1):
    _, zf = s1_latents()
    idx, d2 = c_oracle.knn_bruteforce(zf, 11)
    assert np.array_equal(idx[:, 0], np.arange(2000))
    assert np.array_equal(idx[:, 1:], s1["knn_idx"])


def test_c_dijkstra_bit_equal_to_reference(s1):
    W = csr_from(s1["geo_indptr"], s1["geo_indices"], s1["geo_data"], 2000)
    for r, s in enumerate(s1["sssp_src"]):
        assert np.array_equal(c_oracle.dijkstra(W, s).astype(np.float32), s1["sssp_D"][r])


def test_c_pairwise_sum_is_numpy_sum():
    for n in (0, 3, 8, 100, 128, 129, 777, 20000, 2_940_000):
        x = np.random.RandomState(n).rand(n).astype(np.float32) * 3
        assert c_oracle.pairwise_sum_f32(x) == np.sum(x)


def test_c_kpp_chain_reproduces_reference_medoids(s1):
    """Whole k-means++ chain in C arithmetic + NumPy's MT19937 == the reference's medoids."""
    W = csr_from(s1["geo_indptr"], s1["geo_indices"], s1["geo_data"], 2000)
    n = 2000
    rng = np.random.RandomState(42)
    centers = [int(rng.randint(0, n))]
    dmin = np.full(n, np.inf, np.float32)
    is_c = np.zeros(n, np.uint8)
    for _ in range(1, 32):
        is_c[centers[-1]] = 1
        dmin = np.minimum(dmin, c_oracle.dijkstra(W, centers[-1]).astype(np.float32))
        nxt, _ = c_oracle.kpp_draw(dmin, is_c, rng.random_sample())
        centers.append(nxt)
    assert np.array_equal(centers, s1["medoids"])
