"""tcgen05 kNN filter for every padded dimension (16 / 32 / 64) and the query-vs-candidate entry point
(geocb_knn_query): index lists identical to sklearn's brute search / to an exact float64 argmin."""
import numpy as np
import pytest
import torch

import geo_oracle as oracle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dg():
    assert torch.cuda.is_available()
    from vqvae_b200 import device_graph, _lib
    _lib.load()
    return device_graph


@pytest.mark.parametrize("d,n,k", [(16, 5000, 20), (24, 4000, 20), (32, 6000, 20), (33, 3000, 10), (48, 3000, 25), (64, 5000, 20)])
def test_knn_tc_dims_vs_sklearn(dg, d, n, k):
    """d <= 64 and k+1 <= 26 take the tensor-core filter (no row may need the exact fallback for iid data)."""
    from sklearn.neighbors import NearestNeighbors
    z = np.random.RandomState(d).randn(n, d).astype(np.float32)
    zt = torch.from_numpy(z).cuda()
    idx, dist, nun = dg.knn_topk(zt, k + 1)
    dref, iref = NearestNeighbors(n_neighbors=k + 1, algorithm="brute").fit(z).kneighbors(z)
    assert np.array_equal(idx.cpu().numpy(), iref)
    np.testing.assert_allclose(dist.cpu().numpy(), dref, rtol=2e-6, atol=1e-6)
    assert int(nun.item()) <= n // 100
    # a query slice in the middle of the range (the multi-GPU sharding path)
    idx2, _, _ = dg.knn_topk(zt, k + 1, 1000, 1777)
    assert np.array_equal(idx2.cpu().numpy(), iref[1000:1777])


@pytest.mark.parametrize("env", [{"GEOCB_KNN_TC_FMT": "tf32"}, {"GEOCB_KNN_TC_FMT": "tf32", "GEOCB_KNN_TC_PAIR": "0"},
                                 {"GEOCB_KNN_TC_FMT": "tf32", "GEOCB_KNN_TC_N": "256"},
                                 {"GEOCB_KNN_TC_FMT": "tf32", "GEOCB_KNN_TC_ES": "2"}, {"GEOCB_KNN_TC_N": "64"}],
                         ids=["tf32_pair", "tf32_two_tile_cta", "tf32_n256", "tf32_es2", "bf16_n64"])
def test_knn_tc_d16_other_configurations(dg, env, monkeypatch):
    """d <= 16 runs bf16-split operands, one A tile per CTA and two CTAs per SM by default; the other kernel
    configurations (tf32-split operands with one- or two-tile CTAs, N = 256 MMAs, two epilogue warps per row block; bf16
    with 64-candidate chunks) stay selectable and must give the same lists."""
    from sklearn.neighbors import NearestNeighbors
    z = np.random.RandomState(5).randn(7000, 16).astype(np.float32)
    z[100:140] = z[60:100]  # duplicates: ties broken by index
    zt = torch.from_numpy(z).cuda()
    idx0, dist0, _ = dg.knn_topk(zt, 21)
    for k_, v_ in env.items():
        monkeypatch.setenv(k_, v_)
    idx1, dist1, _ = dg.knn_topk(zt, 21)
    assert torch.equal(idx0, idx1) and torch.equal(dist0, dist1)
    dref, iref = NearestNeighbors(n_neighbors=21, algorithm="brute").fit(z).kneighbors(z)
    np.testing.assert_allclose(dist0.cpu().numpy(), dref, rtol=2e-6, atol=1e-6)
    # (twins tie: sklearn may list them in either order, ours lists the lower index first)
    assert np.array_equal(np.sort(idx0.cpu().numpy()[:, :20], axis=1), np.sort(iref[:, :20], axis=1)) or \
        (np.sort(idx0.cpu().numpy(), axis=1) == np.sort(iref, axis=1)).all(axis=1).mean() > 0.995


def test_knn_tc_tf32_format_other_dims(dg, monkeypatch):
    """d = 32 / 64 with the tf32-split operand format (GEOCB_KNN_TC_FMT=tf32; bf16 is the default everywhere)."""
    for d in (32, 64):
        z = torch.from_numpy(np.random.RandomState(d).randn(4000, d).astype(np.float32)).cuda()
        monkeypatch.delenv("GEOCB_KNN_TC_FMT", raising=False)
        idx0, dist0, _ = dg.knn_topk(z, 21)
        monkeypatch.setenv("GEOCB_KNN_TC_FMT", "tf32")
        idx1, dist1, _ = dg.knn_topk(z, 21)
        assert torch.equal(idx0, idx1) and torch.equal(dist0, dist1)


@pytest.mark.parametrize("d,K", [(16, 512), (32, 512), (64, 1024), (128, 64)])
def test_nearest_medoid_assign(dg, d, K):
    """Downstream assignment of held-out latents (src/eval/evaluate_codebook_health.py:13-26): the code is the exactly
    nearest medoid; equal to a float64 argmin, and to the reference's float32 argmin except at float32 near-ties."""
    from vqvae_b200.eval import nearest_medoid_assign, decode_codes, valid_code_rows
    rs = np.random.RandomState(K + d)
    z = torch.from_numpy(rs.randn(20000, d).astype(np.float32))
    zm = z[rs.choice(20000, K, replace=False)].clone()
    codes = nearest_medoid_assign(z, zm)
    assert codes.dtype == torch.long and codes.shape == (20000,) and codes.device.type == "cpu"
    d2 = ((z.double()[:, None, :] - zm.double()[None, :, :]) ** 2).sum(-1) if K * 20000 * d < 3e8 else None
    if d2 is None:
        d2 = (z.double() ** 2).sum(1, keepdim=True) + (zm.double() ** 2).sum(1)[None] - 2.0 * z.double() @ zm.double().t()
        agree = (codes == d2.argmin(1)).float().mean().item()
        assert agree > 0.9999
    else:
        assert torch.equal(codes, d2.argmin(1))
    # reference formula in float32
    ref = ((z ** 2).sum(1, keepdim=True) + (zm ** 2).sum(1).view(1, -1) - 2.0 * z @ zm.t()).argmin(1)
    assert (codes == ref).float().mean().item() > 0.999
    assert torch.equal(codes[torch.from_numpy(np.isin(np.arange(20000), []))], torch.empty(0, dtype=torch.long))
    # artifact conventions
    grid = codes[:64].reshape(4, 4, 4)
    assert decode_codes(grid, zm).shape == (4, d, 4, 4)
    assert torch.equal(decode_codes(grid, zm)[1, :, 2, 3], zm[grid[1, 2, 3]])
    c = np.zeros((5, 2, 2), np.int32)
    c[3, 1, 0] = -1
    assert valid_code_rows(c).tolist() == [True, True, True, False, True]
