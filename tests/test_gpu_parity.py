"""Parity of the CUDA path (through the reference-shaped Python API -> ctypes -> C ABI) against the
golden vectors of the unmodified reference and against the CPU oracle on the same seeded inputs.
Bars: index/structure/medoid/code results bit-exact; SSSP float32 distances bit-equal (<=1e-5 required);
edge lengths <=1e-5 relative except at ReLU kinks."""
import hashlib

import numpy as np
import pytest
import torch
from scipy import sparse

import geo_oracle as oracle
from conftest import s1_latents, c1_latents, fashion_decoder, csr_from
from hashw import reweight

pytestmark = pytest.mark.gpu


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.fixture(scope="module")
def geo():
    assert torch.cuda.is_available()
    from vqvae_b200.geo import knn_graph_optimized, geo_shortest_paths, kmeans_optimized, riemannian_metric
    from vqvae_b200 import _lib
    _lib.load()
    return dict(knn=knn_graph_optimized, sp=geo_shortest_paths, km=kmeans_optimized, rm=riemannian_metric)


# ------------------------------------------------------------------ kNN / CSR
def test_knn_union_connectivity_golden(geo, s1):
    _, zf = s1_latents()
    W, info = geo["knn"].build_knn_graph_auto(zf, k=10, metric="euclidean", mode="connectivity", sym="union")
    assert np.array_equal(info["indices"].astype(np.int32), s1["knn_idx"])
    assert np.array_equal(W.indptr, s1["euc_indptr"]) and np.array_equal(W.indices, s1["euc_indices"])
    assert W.dtype == np.float32 and np.all(W.data == 1.0)


def test_knn_mutual_distance_golden(geo, golden_dir):
    g = np.load(golden_dir / "s1m_knn.npz")
    _, zf = s1_latents()
    W, info = geo["knn"].build_knn_graph(zf, k=6)
    assert np.array_equal(info["indices"].astype(np.int32), g["knn_idx"])
    assert np.array_equal(W.indptr, g["indptr"]) and np.array_equal(W.indices, g["indices"])
    np.testing.assert_allclose(W.data, g["data"], rtol=1e-6)
    np.testing.assert_allclose(info["distances"], g["knn_dist"], rtol=1e-6)
    assert info["distances"].dtype == np.float32 and (W != W.T).nnz == 0 and W.diagonal().sum() == 0


def test_knn_c1_digest(geo, c1):
    _, zf = c1_latents()
    assert sha(zf) == str(c1["z_sha"])
    W, info = geo["knn"].build_knn_graph_auto(zf, k=20, mode="connectivity", sym="union")
    assert sha(info["indices"].astype(np.int32)) == str(c1["knn_idx_sha"])
    assert sha(W.indptr.astype(np.int32)) == str(c1["euc_indptr_sha"])
    assert sha(W.indices.astype(np.int32)) == str(c1["euc_indices_sha"])
    assert W.nnz == int(c1["nnz"])


@pytest.mark.parametrize("d,k", [(4, 5), (6, 3), (8, 10), (32, 20), (64, 7), (100, 4), (16, 40)])
def test_knn_vs_oracle_shapes(geo, d, k):
    z = np.random.RandomState(d * 100 + k).randn(1500, d).astype(np.float32)
    W, info = geo["knn"].build_knn_graph_auto(z, k=k, mode="distance", sym="union")
    Wo, io = oracle.knn_graph(z, k=k, mode="distance", sym="union")
    assert np.array_equal(info["indices"], io["indices"])
    assert np.array_equal(W.indptr, Wo.indptr) and np.array_equal(W.indices, Wo.indices)
    np.testing.assert_allclose(W.data, Wo.data, rtol=1e-6)


def test_knn_cosine_vs_sklearn(geo):
    """metric="cosine" (SURVEY 8f-2): same neighbour lists as sklearn's brute cosine search up to float32 near-ties
    (sklearn ranks by a float32 GEMM, here by the exact order of the unit rows), distances within 2e-6 absolute."""
    rs = np.random.RandomState(5)
    z = rs.randn(3000, 16).astype(np.float32)
    W, info = geo["knn"].build_knn_graph(z, k=10, metric="cosine", mode="distance", sym="union")
    Wo, io = oracle.knn_graph(z, k=10, metric="cosine", mode="distance", sym="union")
    assert info["indices"].shape == (3000, 10) and info["distances"].dtype == np.float32
    same_rows = (info["indices"] == io["indices"]).all(axis=1)
    assert same_rows.mean() > 0.995
    same_sets = np.array([set(a) == set(b) for a, b in zip(info["indices"], io["indices"])])
    assert same_sets.mean() > 0.998
    assert np.abs(info["distances"][same_rows] - io["distances"][same_rows]).max() < 2e-6
    assert W.shape == Wo.shape and abs(W.nnz - Wo.nnz) <= 0.005 * Wo.nnz
    assert (W != W.T).nnz == 0 and W.diagonal().sum() == 0
    with pytest.raises(ValueError):
        geo["knn"].build_knn_graph(z, k=10, metric="manhattan")


def test_knn_edge_cases(geo):
    # reference tests/test_knn_graph.py:11-42
    k = geo["knn"]
    W, info = k.build_knn_graph(np.empty((0, 4), np.float32), k=3)
    assert W.shape == (0, 0) and info["indices"].shape == (0, 0) and info["distances"].shape == (0, 0)
    W, info = k.build_knn_graph(np.zeros((1, 4), np.float32), k=3)
    assert W.shape == (1, 1) and info["indices"].shape == (1, 0)
    z = np.random.RandomState(0).randn(25, 6).astype(np.float32)
    W, info = k.build_knn_graph(z, k=0)
    assert W.nnz == 0 and info["indices"].shape == (25, 0)
    z5 = np.random.RandomState(0).randn(5, 6).astype(np.float32)
    W, info = k.build_knn_graph(z5, k=50)
    assert info["indices"].shape == (5, 4)
    for i in range(5):
        assert i not in info["indices"][i]
    with pytest.raises(ValueError):
        k.build_knn_graph(z5, k=2, sym="bogus")
    with pytest.raises(AssertionError):
        k.build_knn_graph(np.zeros(5, np.float32), k=2)
    W, _ = k.build_knn_graph(z, k=4, mode="connectivity", sym="union")
    assert np.all(W.data == 1.0) and (W != W.T).nnz == 0


def test_knn_duplicates_match_oracle(geo):
    rs = np.random.RandomState(3)
    z = rs.randn(300, 8).astype(np.float32)
    z[10] = z[3]
    z[200] = z[3]
    W, info = geo["knn"].build_knn_graph_auto(z, k=5, mode="connectivity", sym="union")
    Wo, io = oracle.knn_graph(z, k=5, mode="connectivity", sym="union")
    assert np.array_equal(W.indptr, Wo.indptr) and np.array_equal(W.indices, Wo.indices)


def test_lcc(geo):
    A = sparse.block_diag([sparse.csr_matrix(np.ones((3, 3)) - np.eye(3)), sparse.csr_matrix(np.ones((4, 4)) - np.eye(4)),
                           sparse.csr_matrix(np.ones((4, 4)) - np.eye(4))]).tocsr().astype(np.float32)
    assert geo["knn"].largest_connected_component(A).tolist() == [False] * 3 + [True] * 4 + [False] * 4
    _, zf = s1_latents()
    W, _ = geo["knn"].build_knn_graph(zf, k=1, sym="mutual")
    assert np.array_equal(geo["knn"].largest_connected_component(W), oracle.largest_component_mask(W))
    stats = geo["knn"].analyze_graph_connectivity(W)
    assert stats["n_components"] == sparse.csgraph.connected_components(W, directed=False)[0]


# ------------------------------------------------------------------ SSSP
def test_sssp_line_graph_exact(geo):
    n = 5
    r = np.arange(n - 1)
    W = sparse.csr_matrix((np.ones(2 * (n - 1), np.float32), (np.r_[r, r + 1], np.r_[r + 1, r])), shape=(n, n))
    D = geo["sp"].dijkstra_multi_source(W, sources=[0, 2])
    assert D.shape == (2, 5) and D.dtype == np.float32
    assert np.array_equal(D[0], [0, 1, 2, 3, 4]) and np.array_equal(D[1], [2, 1, 0, 1, 2])
    assert np.array_equal(geo["sp"].dijkstra_single_source(W, 3), geo["sp"].dijkstra_multi_source(W, [3])[0])
    assert np.array_equal(geo["sp"].distances_between(W, [0, 1], [2, 4]), [[2, 4], [1, 3]])


def test_sssp_streamed_path_graph(geo):
    """A 300-vertex path (299 hops: hundreds of rounds per source), every vertex a source, through 4 tables: sources
    finish at very different rounds, so the hand-over happens mid-flight of the others; plus an isolated vertex."""
    from vqvae_b200 import device_graph as dg
    n = 301
    r = np.arange(n - 2)
    w = (1 + (r % 7)).astype(np.float32)
    W = sparse.csr_matrix((np.r_[w, w], (np.r_[r, r + 1], np.r_[r + 1, r])), shape=(n, n))
    g = dg.DeviceCSR.from_scipy(W, torch.device("cuda"))
    src = np.r_[np.arange(n), [5, 5, 300]]
    ref = oracle.sssp_multi(W, src)
    for slots, stagger in ((4, 0), (4, 30), (32, 100), (1, 0)):
        D = dg.sssp_multi(g, src, slots=slots, stagger=stagger)
        assert np.array_equal(D.cpu().numpy(), ref), (slots, stagger)
    assert np.isinf(ref[300][:300]).all() and ref[300][300] == 0


def test_sssp_errors_and_variants(geo):
    sp = geo["sp"]
    W = sparse.csr_matrix(np.array([[0, 1, 0], [1, 0, 10], [0, 10, 0]], np.float32))
    assert np.array_equal(sp.dijkstra_single_source(W, 0), [0, 1, 11])
    assert np.array_equal(sp.dijkstra_single_source(W, 0, unweighted=True), [0, 1, 2])
    D, P = sp.dijkstra_multi_source(W, [0], return_predecessors=True)
    assert P.dtype == np.int32 and P.shape == (1, 3) and P[0].tolist() == [-9999, 0, 1]
    two = sparse.block_diag([sparse.csr_matrix(np.array([[0, 1], [1, 0]], np.float32))] * 2).tocsr()
    D = sp.dijkstra_multi_source(two, [0])
    assert np.isinf(D[0, 2:]).all() and D[0, 1] == 1
    with pytest.raises(ValueError):
        sp.dijkstra_multi_source(W, [])
    with pytest.raises(ValueError):
        sp.dijkstra_multi_source(sparse.csr_matrix(np.array([[0, -1.0], [-1.0, 0]])), [0])
    with pytest.raises(ValueError):
        sp.dijkstra_multi_source(sparse.csr_matrix(np.ones((2, 3))), [0])
    with pytest.raises(TypeError):
        sp.dijkstra_multi_source(np.ones((2, 2)), [0])
    with pytest.raises(ValueError):
        sp.distances_between(W, [0], [])
    # directed and asymmetric-undirected semantics follow scipy
    A = sparse.csr_matrix(np.array([[0, 2, 0], [0, 0, 3], [1, 0, 0]], np.float32))
    for directed in (True, False):
        assert np.array_equal(sp.dijkstra_multi_source(A, [0, 1, 2], directed=directed),
                              oracle.sssp_multi(A, [0, 1, 2], directed=directed))


def test_sssp_golden_bit_equal(geo, s1):
    W = csr_from(s1["geo_indptr"], s1["geo_indices"], s1["geo_data"], 2000)
    D = geo["sp"].dijkstra_multi_source(W, s1["sssp_src"])
    assert np.array_equal(D, s1["sssp_D"])
    assert np.array_equal(geo["sp"].dijkstra_multi_source(W, s1["sssp_src"][:2], unweighted=True), s1["sssp_Dunw"])
    D2, P = geo["sp"].dijkstra_multi_source(W, s1["sssp_src"][:2], return_predecessors=True)
    # predecessor trees may differ on ties; they must describe shortest paths
    Wd = W.toarray().astype(np.float64)
    for r in range(2):
        v = np.nonzero(P[r] >= 0)[0]
        D64 = sparse.csgraph.dijkstra(W, indices=[int(s1["sssp_src"][r])])[0]
        assert np.array_equal(D64[P[r][v]] + Wd[P[r][v], v], D64[v])
        assert (P[r] == -9999).sum() == 1


@pytest.mark.parametrize("slots", [1, 3, 8])
def test_sssp_c1_graph_vs_oracle(geo, c1, slots):
    """20k-node union kNN graph with hash weights: every float32 distance bit-equal to scipy."""
    from vqvae_b200 import device_graph as dg
    _, zf = c1_latents()
    We, _ = oracle.knn_graph(zf, k=20, mode="connectivity", sym="union")
    W = reweight(We)
    src = np.random.RandomState(slots).choice(20000, 11, replace=False)
    g = dg.DeviceCSR.from_scipy(W, torch.device("cuda"))
    D, st = dg.sssp_multi(g, src, slots=slots, stats=True)
    assert np.array_equal(D.cpu().numpy(), oracle.sssp_multi(W, src))
    assert st["relaxations"] < 3 * 11 * (W.nnz + 20000)  # near-far keeps redundant work small
    # the 11 sources stream through `slots` tables in one launch; staggered seeding changes the schedule, not the bits
    for stagger in (0, 5, 40):
        Ds, sts = dg.sssp_multi(g, src, slots=slots, stats=True, stagger=stagger)
        assert torch.equal(Ds, D) and sts["stagger"] == stagger and sts["rounds_per_source"] >= 1, stagger
    if slots == 3:
        # sources stream through the slots (11 sources, 3 slots): predecessors are taken at each hand-over,
        # a repeated source is an independent search, and a second call on the reused workspace is clean
        src2 = np.r_[src[:5], src[:2]]
        D2, P2 = dg.sssp_multi(g, src2, slots=3, want_pred=True)
        D2, P2 = D2.cpu().numpy(), P2.cpu().numpy()
        ref = oracle.sssp_multi(W, src2)
        assert np.array_equal(D2, ref)
        D64 = sparse.csgraph.dijkstra(W, directed=False, indices=src2)
        Wc = W.tocsr()
        for r in range(len(src2)):
            v = np.nonzero(P2[r] >= 0)[0]
            assert len(v) == 20000 - 1 - int(np.isinf(ref[r]).sum()) and P2[r][src2[r]] == -9999
            w = np.asarray(Wc[P2[r][v], v]).ravel().astype(np.float64)
            assert np.array_equal(D64[r][P2[r][v]] + w, D64[r][v])


def test_sssp_public_api_asymmetric_and_directed(geo):
    """dijkstra_multi_source on host scipy matrices whose two directions differ (structure and weights): the
    undirected view is merged on the device (geocb_csr_sym_check / geocb_csr_undirected_*) and must equal scipy's
    directed=False result; directed=True with predecessors searches in-edges (transposed adjacency)."""
    rs = np.random.RandomState(11)
    n = 400
    A = sparse.random(n, n, density=0.03, random_state=rs, format="csr", dtype=np.float32)
    A.data = (A.data * 3 + 0.1).astype(np.float32)
    A.setdiag(0)
    A.eliminate_zeros()
    src = [0, 5, 77, 399, -1]  # negative indices wrap like scipy's
    for directed in (False, True):
        D = geo["sp"].dijkstra_multi_source(A, src, directed=directed)
        assert np.array_equal(D, oracle.sssp_multi(A, src, directed=directed)), directed
        D2, P = geo["sp"].dijkstra_multi_source(A, src[:3], directed=directed, return_predecessors=True)
        D64 = sparse.csgraph.dijkstra(A, directed=directed, indices=src[:3])
        Ad = A.toarray().astype(np.float64)
        Wd = Ad if directed else np.where((Ad > 0) & (Ad.T > 0), np.minimum(Ad, Ad.T), np.maximum(Ad, Ad.T))
        for r in range(3):
            v = np.nonzero(P[r] >= 0)[0]
            assert len(v) == np.isfinite(D64[r]).sum() - 1
            assert (Wd[P[r][v], v] > 0).all()
            assert np.array_equal(D64[r][P[r][v]] + Wd[P[r][v], v], D64[r][v])
    # components / LCC see an edge if either direction is stored
    assert np.array_equal(geo["knn"].largest_connected_component(A), oracle.largest_component_mask(A))
    with pytest.raises(ValueError):
        geo["sp"].dijkstra_multi_source(A, [n])


# ------------------------------------------------------------------ K-medoids
def test_kmedoids_golden_s1(geo, s1):
    W = csr_from(s1["geo_indptr"], s1["geo_indices"], s1["geo_data"], 2000)
    med, assign, qe = geo["km"].fit_kmedoids_optimized(W, K=32, init="kpp", seed=42)
    assert med.dtype == np.int64 and assign.dtype == np.int64 and isinstance(qe, float)
    assert np.array_equal(med, s1["medoids"]) and np.array_equal(assign, s1["assign"]) and qe == float(s1["qe"])
    med, assign, qe = geo["km"].fit_kmedoids_optimized(W, K=16, init="random", seed=7)
    assert np.array_equal(med, s1["rand_medoids"]) and np.array_equal(assign, s1["rand_assign"])
    assert qe == float(s1["rand_qe"])
    assert geo["km"].kpp_initialization_graph(W, 8, seed=3) == s1["kpp8_seed3"].tolist()
    a2 = geo["km"].assign_points_to_medoids(W, s1["medoids"])
    assert np.array_equal(a2, s1["assign"])
    assert geo["km"].compute_quantization_error(W, s1["medoids"], s1["assign"]) == float(s1["qe"])
    with pytest.raises(ValueError):
        geo["km"].fit_kmedoids_optimized(W, K=4, init="bogus")


def test_kmedoids_c1_hash_weights_golden(geo, c1):
    _, zf = c1_latents()
    We, _ = geo["knn"].build_knn_graph_auto(zf, k=20, mode="connectivity", sym="union")
    med, assign, qe = geo["km"].fit_kmedoids_optimized(reweight(We), K=64, init="kpp", seed=42)
    assert np.array_equal(med, c1["hashw_medoids"])
    assert sha(assign.astype(np.int32)) == str(c1["hashw_assign_sha"]) and qe == float(c1["hashw_qe"])


def test_kmedoids_toy_graphs_vs_oracle(geo):
    # complete / block-diagonal / disconnected graphs as in reference tests/test_kmeans_optimized.py
    rs = np.random.RandomState(0)
    full = rs.rand(12, 12).astype(np.float32) + 0.1
    full = np.triu(full, 1)
    full = sparse.csr_matrix(full + full.T)
    blocks = sparse.block_diag([full, full[:7][:, :7], sparse.csr_matrix((2, 2), dtype=np.float32)]).tocsr().astype(np.float32)
    for W in (full, blocks):
        for K in (1, 3, 6):
            for init in ("kpp", "random"):
                got = geo["km"].fit_kmedoids_optimized(W, K=K, init=init, seed=5)
                exp = oracle.fit_kmedoids(W, K=K, init=init, seed=5)
                assert np.array_equal(got[0], exp[0]) and np.array_equal(got[1], exp[1]) and got[2] == exp[2]
    res = geo["km"].fit_kmedoids_with_connectivity_check(blocks, K=3, seed=1)
    assert res[3]["n_components"] == 4 and res[3]["method"] == "optimized_kmedoids"


def test_pairwise_sum_matches_numpy(geo):
    from vqvae_b200 import device_graph as dg
    for n in (1, 7, 129, 20000, 2_940_000):
        x = torch.rand(n, generator=torch.Generator().manual_seed(n)).mul_(3.0)
        got, used = dg.pairwise_sum_f32(x.cuda(), square=True)
        assert used == n and got == np.sum(x.numpy() ** 2)
    x = torch.tensor([1.5, float("inf"), 2.0, 0.25] * 100)
    got, used = dg.pairwise_sum_f32(x.cuda(), square=True, finite_only=True)
    assert used == 300 and got == np.sum(x.numpy()[np.isfinite(x.numpy())] ** 2)


def test_kpp_draw_matches_numpy_choice(geo):
    """One draw = RandomState.choice(n, p) bit for bit, incl. the exact sequential fallback."""
    from vqvae_b200 import device_graph as dg
    n = 200_000
    r = np.arange(n - 1)
    W = sparse.csr_matrix((np.ones(2 * (n - 1), np.float32), (np.r_[r, r + 1], np.r_[r + 1, r])), shape=(n, n))
    g = dg.DeviceCSR.from_scipy(W, torch.device("cuda"))
    st = dg.VoronoiState(g)
    d = (np.random.RandomState(1).rand(n).astype(np.float32) * 5 + 0.01)
    st.dmin32.copy_(torch.from_numpy(d))
    centers = [5, 77, 1999]
    st.is_center[centers] = 1
    p = d ** 2
    p[centers] = 0
    p /= p.sum()
    rng = np.random.RandomState(9)
    for _ in range(20):
        state = rng.get_state()
        u = rng.random_sample()
        rng.set_state(state)
        exp = int(rng.choice(n, p=p))
        got, _, psum = st.draw(u)
        assert got == exp
    # force the ambiguous branch: u exactly on a boundary of the fp64 cdf
    cdf = p.astype(np.float64).cumsum()
    cdf /= cdf[-1]
    for i in (10, 1234, n - 2):
        got, exact, _ = st.draw(float(cdf[i]))
        assert got == int(np.searchsorted(cdf, cdf[i], side="right")) and exact == 1
