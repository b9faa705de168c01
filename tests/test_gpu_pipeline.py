"""Edge lengths, the device pipeline and the CLI artifacts against the reference's golden vectors."""
import hashlib
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest
import torch
from scipy import sparse

from conftest import s1_latents, c1_latents, fashion_decoder, FASHION_CFG, ROOT

pytestmark = pytest.mark.gpu


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def rel_err(a, b):
    return np.abs(a - b) / np.maximum(np.abs(b), 1e-30)


def check_edge_lengths(got, ref):
    """Tolerance 1e-5 relative (north star), with the ReLU-kink carve-out measured on the reference itself:
    a pre-activation within float noise of 0 flips a ReLU mask and the JVP jumps.  The reference's own fp32
    result differs from its fp64 evaluation by >1e-5 on 9 of 15641 s1 edges (0.06 %, max 3.5 %; script in
    DESIGN.md), so: median <= 1e-6, at most 0.5 % of edges beyond 1e-5, none beyond 10 %."""
    r = rel_err(got, ref)
    assert np.median(r) <= 1e-6
    assert (r > 1e-5).mean() < 5e-3, f"{(r > 1e-5).sum()} of {len(r)} edges off by >1e-5 (max {r.max():.3g})"
    assert r.max() < 0.1


def test_edge_lengths_s1_golden(s1):
    from vqvae_b200.geo.riemannian_metric import edge_lengths_riemannian
    _, zf = s1_latents()
    dec = fashion_decoder().cuda()
    L = edge_lengths_riemannian(dec, torch.from_numpy(zf[s1["ei"]]), torch.from_numpy(zf[s1["ej"]]), batch_size=512)
    assert L.dtype == torch.float32 and L.is_cuda and L.shape == (len(s1["ei"]),)
    check_edge_lengths(L.cpu().numpy(), s1["edge_len"])


def test_edge_lengths_c1_sample(c1):
    from vqvae_b200.geo.knn_graph_optimized import knn_graph_device
    from vqvae_b200 import device_graph as dg
    from vqvae_b200.geo.riemannian_metric import edge_lengths_device
    _, zf = c1_latents()
    z = torch.from_numpy(zf).cuda()
    g, _, _ = knn_graph_device(z, 20, "connectivity", "union")
    ei, ej, pij, pji = dg.upper_edges(g)
    assert ei.numel() == int(c1["E"])
    L = edge_lengths_device(fashion_decoder().cuda(), z, ei, ej, 512).cpu().numpy()
    check_edge_lengths(L[c1["edge_sample_pos"]], c1["edge_sample_len"])
    np.testing.assert_allclose([L.min(), L.mean(), L.max()], c1["edge_len_min_mean_max"], rtol=1e-4)


def test_edge_lengths_generic_decoder_properties():
    # reference tests/test_riemannian_metric.py with its Linear dummy decoder (no BN), decoder on the CPU
    from vqvae_b200.geo.riemannian_metric import edge_lengths_riemannian
    import geo_oracle as oracle
    torch.manual_seed(0)
    dec = torch.nn.Sequential(torch.nn.Linear(16, 784))
    a, b = torch.randn(127, 16), torch.randn(127, 16)
    L = edge_lengths_riemannian(dec, a, b, batch_size=16)
    assert L.shape == (127,) and L.dtype == torch.float32 and (L >= 0).all() and L.device.type == "cpu"
    torch.testing.assert_close(L, edge_lengths_riemannian(dec, b, a, batch_size=16), rtol=1e-4, atol=0)
    torch.testing.assert_close(L, edge_lengths_riemannian(dec, a, b, batch_size=1024), rtol=1e-5, atol=0)
    torch.testing.assert_close(L, oracle.edge_lengths(dec, a, b, 16), rtol=1e-5, atol=0)
    half = edge_lengths_riemannian(dec, a, a + 0.5 * (b - a), batch_size=64)
    ratio = half / L
    assert ((ratio > 0.3) & (ratio < 0.7)).all()
    with pytest.raises(AssertionError):
        edge_lengths_riemannian(dec, a, b[:5])


@pytest.mark.parametrize("norm", ["group", "none"])
def test_edge_lengths_other_norms_vs_oracle(norm):
    from vqvae_b200.geo.riemannian_metric import edge_lengths_riemannian
    from vqvae_b200.models import SpatialVAE
    import geo_oracle as oracle
    torch.manual_seed(1)
    cfg = dict(FASHION_CFG, norm_type=norm, in_channels=3, output_image_size=32)
    dec = SpatialVAE(**cfg).decoder
    a, b = torch.randn(700, 16), torch.randn(700, 16) * 0.1
    b = a + b
    L = edge_lengths_riemannian(dec.cuda(), a, b, batch_size=512).cpu()
    Lo = oracle.edge_lengths(dec.cpu(), a, b, 512)
    check_edge_lengths(L.numpy(), Lo.numpy())


def test_pipeline_s1_matches_golden(s1):
    """Full device pipeline on the s1 case: structure exact; medoids/codes equal to the reference's."""
    from vqvae_b200.pipeline import build_codebook_device
    z4, _ = s1_latents()
    res = build_codebook_device(z4, fashion_decoder().cuda(), k=10, sym="union", K=32, verbose=False)
    assert np.array_equal(res.ei.cpu().numpy(), s1["ei"]) and np.array_equal(res.ej.cpu().numpy(), s1["ej"])
    check_edge_lengths(res.edge_len.cpu().numpy(), s1["edge_len"])
    W = res.g_lcc.to_scipy()
    assert np.array_equal(W.indptr, s1["geo_indptr"]) and np.array_equal(W.indices, s1["geo_indices"])
    assert (W != W.T).nnz == 0
    assert np.array_equal(res.mask.cpu().numpy().astype(bool), s1["mask"])
    assert np.array_equal(res.medoids, s1["medoids"]), "end-to-end medoids differ (GPU-vs-CPU JVP rounding)"
    assert np.array_equal(res.codes, s1["codes"])
    assert abs(res.qe - float(s1["qe"])) <= 1e-5 * float(s1["qe"])


def test_pipeline_stagewise_with_injected_weights(s1):
    """Oracle edge weights injected -> K-medoids exact (the stage-wise contract of SURVEY 8c)."""
    from vqvae_b200 import device_graph as dg
    from vqvae_b200.geo.kmeans_optimized import fit_kmedoids_optimized
    W = sparse.csr_matrix((s1["geo_data"], s1["geo_indices"], s1["geo_indptr"]), shape=(2000, 2000))
    g = dg.DeviceCSR.from_scipy(W, torch.device("cuda"))
    med, assign, qe = fit_kmedoids_optimized(g, K=32, seed=42)
    assert np.array_equal(med, s1["medoids"]) and np.array_equal(assign, s1["assign"]) and qe == float(s1["qe"])


def test_cli_artifacts_c1(tmp_path, c1):
    """The CLI on BASELINE configs[0]: artifact formats, and (as observed) the reference's medoids/codes."""
    from vqvae_b200.models import SpatialVAE
    z4, _ = c1_latents()
    torch.save(z4, tmp_path / "z.pt")
    torch.manual_seed(0)
    vae = SpatialVAE(**FASHION_CFG)
    torch.save({"model_state_dict": vae.state_dict(), "epoch": 0}, tmp_path / "best.pt")
    out = tmp_path / "out"
    cmd = [sys.executable, str(ROOT / "src" / "scripts" / "build_codebook.py"), "--latents_path", str(tmp_path / "z.pt"),
           "--out_dir", str(out), "--vae_ckpt_path", str(tmp_path / "best.pt"), "--in_channels", "1",
           "--output_image_size", "28", "--latent_dim", "16", "--enc_channels", "64", "128", "256",
           "--dec_channels", "256", "128", "64", "--recon_loss", "mse", "--norm_type", "batch", "--mse_use_sigmoid"]
    p = subprocess.run(cmd, capture_output=True, text=True, cwd=str(ROOT))
    assert p.returncode == 0, p.stderr[-2000:]
    assert "Quantization error:" in p.stdout and "Saved artifacts to:" in p.stdout
    W = sparse.load_npz(out / "knn_graph_geodesic.npz")
    cb = torch.load(out / "codebook.pt", weights_only=False)
    codes = np.load(out / "codes.npy")
    assert W.shape == (20000, 20000) and W.nnz == int(c1["nnz"]) and W.dtype == np.float32
    assert codes.shape == (1250, 4, 4) and codes.dtype == np.int32
    assert cb["medoid_indices"].dtype == np.int32 and cb["medoid_indices"].shape == (512,)
    assert cb["z_medoid"].shape == (512, 16) and cb["z_medoid"].dtype == torch.float32
    assert set(cb["config"]) == {"latents_path", "out_dir", "vae_ckpt_path", "in_channels", "output_image_size",
                                 "latent_dim", "enc_channels", "dec_channels", "recon_loss", "norm_type",
                                 "mse_use_sigmoid", "k", "sym", "K", "init", "seed", "batch_size"}
    _, zf = c1_latents()
    assert torch.equal(cb["z_medoid"], torch.from_numpy(zf[cb["medoid_indices"]]))
    r = rel_err(W.data.min(), c1["edge_len_min_mean_max"][0])
    assert r < 1e-4
    same_medoids = np.array_equal(cb["medoid_indices"], c1["medoids"].astype(np.int32))
    print("c1 end-to-end medoids identical to the reference:", same_medoids)
    if same_medoids:
        assert sha(codes) == str(c1["codes_sha"])
    assert len(np.unique(codes)) == 512
