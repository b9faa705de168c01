"""Legacy YAML drivers end to end (BASELINE configs[3] shape) through the reference's module paths
(`src.training.*`, the shim over vqvae_b200.training) against goldens of the UNMODIFIED reference drivers
(tests/golden/make_golden.py --c4): eval-mode BatchNorm, both edge directions re-weighted + maximum, LCC before
re-weighting, 1-D codes.npy; and the Euclidean driver on a unit-weight (hop-count) graph, where argmin ties are
pervasive and the lowest medoid index must win."""
import numpy as np
import pytest
import torch
from scipy import sparse

pytestmark = pytest.mark.gpu

C4_VAE_CFG = dict(in_channels=3, output_image_size=32, latent_dim=32, enc_channels=[16, 32, 64],
                  dec_channels=[64, 32, 16], recon_loss="mse", norm_type="batch", mse_use_sigmoid=False,
                  free_bits_default=0.25, capacity_max_default=25.0, capacity_anneal_steps_default=100000,
                  capacity_mode_default="abs")


@pytest.fixture(scope="module")
def c4(golden_dir):
    return dict(np.load(golden_dir / "c4_legacy.npz"))


@pytest.fixture()
def inputs(tmp_path):
    from src.models.vae import VAE
    z = torch.randn(3000, 32, generator=torch.Generator().manual_seed(42))
    torch.save(z, tmp_path / "z.pt")
    torch.manual_seed(0)
    vae = VAE(**C4_VAE_CFG)
    torch.save({"model_state_dict": vae.state_dict(), "epoch": 0}, tmp_path / "best.pt")
    return tmp_path, z, vae


def test_riemannian_legacy_driver(inputs, c4):
    from src.training.build_riemannian_codebook_legacy import build_and_save
    import kink_margin
    tmp, z, vae = inputs
    cfg = {"data": {"dataset": "cifar10", "latents_path": str(tmp / "z.pt")}, "model": dict(C4_VAE_CFG),
           "checkpoint_path": str(tmp / "best.pt"),
           "graph": {"k": 10, "metric": "euclidean", "sym": "union", "mode": "distance"},
           "riemannian": {"mode": "full", "max_edges": 5000, "batch_size": 512},
           "quantize": {"K": 64, "init": "kpp", "seed": 42}, "out": {"dir": str(tmp / "out")}}
    out = build_and_save(cfg)
    Wr = sparse.load_npz(out / "knn_graph_riemannian.npz").tocsr()
    We = sparse.load_npz(out / "knn_graph_euclidean.npz").tocsr()
    Wr.sort_indices()
    assert np.array_equal(Wr.indptr, c4["riem_indptr"]) and np.array_equal(Wr.indices, c4["riem_indices"])
    assert We.nnz == Wr.nnz and (Wr != Wr.T).nnz == 0
    # weights: 1e-5 relative, ReLU-kink edges excepted (per-edge rule; eval-mode decoder, full edge list in COO order)
    coo = We.tocoo()
    zs, ze = z[coo.row], z[coo.col]
    got_dir = np.asarray(Wr[coo.row, coo.col]).ravel()
    ref_dir = np.asarray(sparse.csr_matrix((c4["riem_data"], c4["riem_indices"], c4["riem_indptr"]), shape=Wr.shape)[coo.row, coo.col]).ravel()
    rel = np.abs(got_dir - ref_dir) / ref_dir
    bad = np.nonzero(rel > 1e-5)[0]
    assert len(bad) <= max(8, 0.005 * len(rel)), (len(bad), rel.max())
    if len(bad):  # the stored weight is max(L_ij, L_ji): either direction's kink explains it
        rev = {(int(r), int(c)): i for i, (r, c) in enumerate(zip(coo.row, coo.col))}
        both = np.unique(np.r_[bad, [rev[(int(coo.col[i]), int(coo.row[i]))] for i in bad]])
        marg = kink_margin.relu_margins(vae.decoder.eval(), zs, ze, 512, both)
        for i in bad:
            j = rev[(int(coo.col[i]), int(coo.row[i]))]
            assert min(marg[int(i)][0], marg[int(j)][0]) < kink_margin.EPS, (i, rel[i], marg[int(i)], marg[int(j)])
    cb = torch.load(out / "codebook.pt", weights_only=False)
    codes = np.load(out / "codes.npy")
    assert codes.ndim == 1 and codes.shape == (3000,) and str(codes.dtype) == str(c4["codes_dtypes"][0])
    assert cb["method"] == str(c4["riem_method"]) and sorted(cb["graph_stats"]) == c4["graph_stats_keys"].tolist()
    assert cb["graph_stats"]["n_components"] == int(c4["n_components"])
    assert cb["medoid_indices"].dtype == np.int32 and cb["z_medoid"].shape == (64, 32)
    # chain parity: exact up to the first k-means++ boundary flip (weights differ in the last bits), see test_cli_artifacts_c1
    got, exp = cb["medoid_indices"].astype(np.int64), c4["riem_medoids"].astype(np.int64)
    diff = np.nonzero(got != exp)[0]
    if len(diff) == 0:
        assert np.array_equal(codes, c4["riem_codes"])
    else:
        print("legacy riemannian chain leaves the reference's at draw", int(diff[0]))
        assert diff[0] >= 1


def test_euclidean_legacy_driver_hop_count_ties(inputs, c4):
    from src.training.build_codebook_legacy import build_and_save
    tmp, z, _ = inputs
    cfg = {"data": {"dataset": "cifar10", "latents_path": str(tmp / "z.pt")},
           "graph": {"k": 10, "metric": "euclidean", "sym": "union", "mode": "connectivity"},
           "quantize": {"K": 64, "init": "kpp", "seed": 42}, "out": {"dir": str(tmp / "out_e")}}
    out = build_and_save(cfg)
    W = sparse.load_npz(out / "knn_graph.npz").tocsr()
    W.sort_indices()
    assert np.array_equal(W.indptr, c4["knn_indptr"]) and np.array_equal(W.indices, c4["knn_indices"]) and (W.data == 1).all()
    cb = torch.load(out / "codebook.pt", weights_only=False)
    codes = np.load(out / "codes.npy")
    # integer hop counts: no rounding anywhere, so medoids and every (tie-broken) code must be identical
    assert np.array_equal(cb["medoid_indices"], c4["euc_medoids"])
    assert np.array_equal(codes, c4["euc_codes"]) and str(codes.dtype) == str(c4["codes_dtypes"][1])
