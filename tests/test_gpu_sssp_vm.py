"""Parity of the source-vectorised SSSP kernels (csrc/sssp_vm.cu through geocb_sssp_multi_vm: 16 sources per launch;
csrc/sssp_vm32.cu through geocb_sssp_multi_vm32: 32 sources, bf16 bound table) with scipy's Dijkstra
(reference call site src/geo/geo_shortest_paths.py:36-49): every float32 distance bit-equal, for every destination
blocking, batch remainder and bucket width."""
import numpy as np
import pytest
import torch
from scipy import sparse

import geo_oracle as oracle
from conftest import c1_latents, csr_from
from hashw import reweight

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dg():
    assert torch.cuda.is_available()
    from vqvae_b200 import device_graph, _lib
    _lib.load()
    return device_graph


def _graph(dg, W):
    return dg.DeviceCSR.from_scipy(W.tocsr().astype(np.float32), torch.device("cuda"))


WIDE = pytest.mark.parametrize("wide", [False, True], ids=["vm16", "vm32"])


@WIDE
def test_vm_toy_graphs(dg, wide):
    n = 70  # a path: deep graph, 69 hops, many threshold raises
    r = np.arange(n - 1)
    W = sparse.csr_matrix((np.ones(2 * (n - 1), np.float32), (np.r_[r, r + 1], np.r_[r + 1, r])), shape=(n, n))
    g = _graph(dg, W)
    for src in ([0], [0, 69, 35], list(range(20)), list(range(69, 20, -1)), [5, 5, 5]):
        for B in (1, 3):
            D = dg.sssp_multi_vm(g, src, blocks=B, blk_min_active=0, wide=wide).cpu().numpy()
            assert np.array_equal(D, oracle.sssp_multi(W, src)), (src, B)
    # two components + an isolated vertex: unreachable = inf
    two = sparse.block_diag([W[:5][:, :5], W[:4][:, :4], sparse.csr_matrix((1, 1), dtype=np.float32)]).tocsr()
    D = dg.sssp_multi_vm(_graph(dg, two), [0, 6, 9], blocks=2, blk_min_active=0, wide=wide).cpu().numpy()
    assert np.array_equal(D, oracle.sssp_multi(two, [0, 6, 9]))
    assert np.isinf(D[0, 5:]).all() and np.isinf(D[2, :9]).all() and D[2, 9] == 0


@WIDE
def test_vm_s1_golden(dg, s1, wide):
    W = csr_from(s1["geo_indptr"], s1["geo_indices"], s1["geo_data"], 2000)
    g = _graph(dg, W)
    for B, mn in ((1, 0), (4, 0), (4, 10**9), (7, 50)):
        D = dg.sssp_multi_vm(g, s1["sssp_src"], blocks=B, blk_min_active=mn, wide=wide).cpu().numpy()
        assert np.array_equal(D, s1["sssp_D"]), (B, mn)
    Du = dg.sssp_multi_vm(g, s1["sssp_src"][:2], unweighted=True, blocks=2, blk_min_active=0, wide=wide).cpu().numpy()
    assert np.array_equal(Du, s1["sssp_Dunw"])


@WIDE
@pytest.mark.parametrize("S,B,scale", [(16, 1, 4.0), (40, 4, 4.0), (17, 8, 1.0), (3, 2, 16.0), (33, 5, 0.25), (70, 3, 4.0)])
def test_vm_c1_hash_graph_vs_oracle(dg, S, B, scale, wide):
    """20k-vertex union kNN graph with hash weights; S not a multiple of 16 exercises the padded last launch."""
    _, zf = c1_latents()
    We, _ = oracle.knn_graph(zf, k=20, mode="connectivity", sym="union")
    W = reweight(We)
    g = _graph(dg, W)
    src = np.random.RandomState(S).choice(20000, S, replace=False)
    _, _, mean_w = g.adjacency(False)
    D, st = dg.sssp_multi_vm(g, src, blocks=B, blk_min_active=0, delta=scale * mean_w, stats=True, wide=wide)
    assert np.array_equal(D.cpu().numpy(), oracle.sssp_multi(W, src))
    assert st["relaxations"] < 4 * S * (W.nnz + 20000)
    # the slot-major kernel gives the same bits
    assert torch.equal(D, dg.sssp_multi(g, src, kernel="slot"))
    # a second call on the reused workspace is clean, and the dispatcher reaches this kernel
    name = "vm32" if wide else "vm"
    D2, st2 = dg.sssp_multi(g, src[::-1].copy(), kernel=name, stats=True)
    assert st2["kernel"] == name and torch.equal(D2, D.flip(0))
