"""Host-side logic that needs no GPU: decoder unrolling (conv -> dense matrix), shard arithmetic, CLI surface."""
import sys
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))


def _by_identity(m, shape):
    k = shape[0] * shape[1] * shape[2]
    eye = torch.eye(k).view(k, *shape)
    if isinstance(m, nn.ConvTranspose2d):
        out = F.conv_transpose2d(eye, m.weight, None, m.stride, m.padding, m.output_padding, m.groups, m.dilation)
    else:
        out = F.conv2d(eye, m.weight, None, m.stride, m.padding, m.dilation, m.groups)
    return out.reshape(k, -1), tuple(out.shape[1:])


@pytest.mark.parametrize("make,shape", [
    (lambda: nn.Conv2d(16, 32, 3, 1, 1), (16, 1, 1)),                                  # conv_in on a 1x1 latent cell
    (lambda: nn.ConvTranspose2d(32, 16, 3, 2, 1, output_padding=1), (32, 1, 1)),       # reference decoder up-blocks
    (lambda: nn.ConvTranspose2d(8, 4, 3, 2, 1, output_padding=1), (8, 2, 2)),
    (lambda: nn.ConvTranspose2d(8, 4, 4, 2, 1), (8, 2, 2)),
    (lambda: nn.ConvTranspose2d(6, 3, 3, 1, 1), (6, 4, 4)),
    (lambda: nn.Conv2d(4, 6, 3, 2, 1), (4, 5, 5)),
    (lambda: nn.Conv2d(4, 6, 1), (4, 3, 3)),
    (lambda: nn.Conv2d(3, 5, (3, 2), (2, 1), (1, 0)), (3, 6, 4)),
])
def test_conv_matrix_entries_are_the_kernel_weights(make, shape):
    """The dense operator built by index arithmetic equals, bit for bit, the one obtained by pushing an identity
    basis through the layer (what reference-style autograd would differentiate: riemannian_metric.py:12-35)."""
    from vqvae_b200.geo.riemannian_metric import _conv_matrix
    torch.manual_seed(0)
    m = make()
    with torch.no_grad():
        A, osh = _conv_matrix(m, shape)
        R, rsh = _by_identity(m, shape)
    assert osh == rsh and torch.equal(A, R)
    # every non-zero entry is one of the layer's weights
    assert set(A[A != 0].tolist()) <= set(m.weight.detach().reshape(-1).tolist())


def test_conv_matrix_declines_what_it_cannot_express():
    from vqvae_b200.geo.riemannian_metric import _conv_matrix, _unique_taps
    assert _conv_matrix(nn.Conv2d(4, 4, 3, 1, 1, groups=2), (4, 3, 3)) is None
    assert not _unique_taps(nn.Conv2d(4, 4, 3, 1, 2, dilation=2))
    assert _unique_taps(nn.ConvTranspose2d(4, 4, 3, 2, 1))


def test_unrolled_decoder_matches_module_forward():
    """lin/bn/relu program of the FashionMNIST-shape decoder reproduces decoder(z) on a 1x1 grid (eval-mode BN so a
    CPU forward is deterministic): same composition the GPU JVP path runs."""
    import bench
    from vqvae_b200.models import SpatialVAE
    from vqvae_b200.geo import riemannian_metric as rm
    torch.manual_seed(0)
    dec = SpatialVAE(**bench.FASHION_CFG).decoder.eval()
    for mod in dec.modules():
        if isinstance(mod, nn.BatchNorm2d):
            mod.running_mean.normal_(0, 0.1)
            mod.running_var.uniform_(0.5, 1.5)
    z = torch.randn(7, 16)
    with torch.no_grad():
        want = dec(z.view(7, 16, 1, 1)).reshape(7, -1)
        ops = rm._unrolled_ops(rm._flatten_layers(dec), 16)
        x = z
        for op in ops:
            if op[0] == "lin":
                x = x @ op[1] + (0 if op[2] is None else op[2])
            elif op[0] == "bn":
                m, shape = op[1], op[2]
                x = F.batch_norm(x.view(-1, *shape), m.running_mean, m.running_var, m.weight, m.bias, False, 0.0,
                                 m.eps).reshape(x.shape[0], -1)
            elif op[0] == "relu":
                x = torch.relu(x)
    assert x.shape == want.shape
    assert torch.allclose(x, want, rtol=1e-5, atol=1e-6)


def test_decoder_needs_library_gemm_guess():
    """The structural guess behind the pipeline's set-up overlap agrees with the unrolled operators: the reference's
    FashionMNIST decoder (latent 16 -> 4x4 patch = 16 outputs) runs on this repo's kernels only, the CIFAR one
    (3 x 8 x 8 = 192 outputs) and a flat decoder still call a library GEMM."""
    import torch.nn as nn
    from conftest import CIFAR_CFG, FASHION_CFG
    from vqvae_b200.geo import riemannian_metric as rm
    from vqvae_b200.models import SpatialVAE
    for cfg, expect in ((FASHION_CFG, False), (CIFAR_CFG, True)):
        dec = SpatialVAE(**cfg).decoder
        assert rm.decoder_needs_library_gemm(dec, cfg["latent_dim"]) is expect
        ops = rm._unrolled_ops(rm._flatten_layers(dec), cfg["latent_dim"])
        width = [o for o in ops if o[0] == "lin"][-1][1].shape[1]
        assert (width != 16) is expect
    assert rm.decoder_needs_library_gemm(SpatialVAE(**FASHION_CFG).decoder, 24) is True   # first layer not fusable
    assert rm.decoder_needs_library_gemm(nn.Sequential(nn.Linear(16, 8)), 16) is True     # not a conv stack


def test_shard_ranges_keep_whole_bn_batches():
    from vqvae_b200.pipeline import shard_range
    E, bs = 46_400_123, 512
    for world in (1, 2, 3, 4, 8):
        cuts = [shard_range(E, r, world, multiple=bs) for r in range(world)]
        assert cuts[0][0] == 0 and cuts[-1][1] == E
        assert all(a[1] == b[0] for a, b in zip(cuts, cuts[1:]))
        assert all(lo % bs == 0 for lo, _ in cuts)
        sizes = [hi - lo for lo, hi in cuts]
        assert max(sizes) - min(sizes) < 2 * bs  # one unit of imbalance + the ragged tail


def _run_py(code, extra_path=()):
    import os
    import subprocess
    env = dict(os.environ)
    env["PYTHONPATH"] = os.pathsep.join([str(ROOT), *map(str, extra_path)])
    env.pop("VQVAE_REFERENCE_ROOT", None)
    return subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300)


def test_src_shim_standalone():
    """Without a reference checkout the shim exposes exactly this repo's modules."""
    r = _run_py("import src, src.geo.knn_graph_optimized as k, src.scripts.build_codebook as b, src.models as m;"
                "print(src.reference_src(), k.build_knn_graph.__module__, m.SpatialVAE.__module__)")
    assert r.returncode == 0, r.stderr
    assert r.stdout.split() == ["None", "vqvae_b200.geo.knn_graph_optimized", "vqvae_b200.models.spatial_vae"]


@pytest.mark.skipif(not Path("/root/reference/src/training/build_codebook_legacy.py").exists(),
                    reason="needs a checkout of the reference (not present on the GPU box)")
def test_src_shim_overlays_reference_drivers():
    """With the reference behind this repo on PYTHONPATH: the legacy YAML drivers, the flat VAE and the
    geodesic functions are this repo's; modules this repo does not implement come from the reference."""
    code = (
        "import importlib, src\n"
        "e = importlib.import_module('src.training.build_codebook_legacy')\n"
        "r = importlib.import_module('src.training.build_riemannian_codebook_legacy')\n"
        "from src.models import VAE, SpatialVAE\n"
        "import src.scripts.build_codebook as cli, src.utils.spatial_latents as u\n"
        "print(e.build_and_save.__module__, r.build_and_save.__module__, r._reweight_graph_with_riemannian.__module__,\n"
        "      VAE.__module__, SpatialVAE.__module__, cli.__file__, u.__file__)\n")
    r = _run_py(code, ["/root/reference"])
    assert r.returncode == 0, r.stderr
    out = r.stdout.split()
    assert out[0] == "vqvae_b200.training.build_codebook_legacy"
    assert out[1] == out[2] == "vqvae_b200.training.build_riemannian_codebook_legacy"
    assert out[3] == "vqvae_b200.models.vae" and out[4] == "vqvae_b200.models.spatial_vae"
    assert out[5] == str(ROOT / "src" / "scripts" / "build_codebook.py")
    assert out[6] == "/root/reference/src/utils/spatial_latents.py"


def test_legacy_driver_config_plumbing(tmp_path):
    """Config keys of the legacy YAML drivers (build_codebook_legacy.py:29-41, build_riemannian_codebook_legacy.py:
    171-197): explicit paths, dataset-based defaults, the three places a checkpoint path may live."""
    from vqvae_b200.training import _common as C
    assert str(C.latents_path({"data": {"latents_path": "a/z.pt"}})) == "a/z.pt"
    assert str(C.latents_path({"data": {"dataset": "CIFAR10 "}})) == "experiments/vae_cifar10/latents_train/z.pt"
    assert str(C.latents_path({"data": {}})) == "experiments/vae_mnist/latents_train/z.pt"
    assert str(C.checkpoint_path({"checkpoint_path": "c.pt"})) == "c.pt"
    assert str(C.checkpoint_path({"vae": {"ckpt_path": "v.pt"}})) == "v.pt"
    assert str(C.checkpoint_path({"model": {"checkpoint_path": "m.pt"}})) == "m.pt"
    assert str(C.checkpoint_path({"data": {"dataset": "mnist"}})) == "experiments/vae_mnist/checkpoints/best.pt"
    assert C.graph_args({"graph": {"k": "10", "metric": "euclidean", "sym": "union", "mode": "distance"}}) == (10, "euclidean", "union", "distance")
    torch.save({"z": torch.ones(3, 2, dtype=torch.float64)}, tmp_path / "z.pt")
    assert C.load_latents(tmp_path / "z.pt").dtype == torch.float32
    torch.save([1, 2], tmp_path / "bad.pt")
    with pytest.raises(ValueError):
        C.load_latents(tmp_path / "bad.pt")
    mask = np.array([True, False, True])
    assert C.scatter_codes(np.array([4, 5]), mask, 3).tolist() == [4, -1, 5]
    assert C.scatter_codes(np.array([4, 5, 6]), np.ones(3, bool), 3).tolist() == [4, 5, 6]


def test_kink_rule_is_consistent_on_cpu():
    """tools/kink_margin.py: margins are non-negative, the float64 length matches the oracle, and an edge far from any
    kink cannot be excused."""
    import kink_margin
    import geo_oracle as oracle
    from conftest import fashion_decoder
    torch.manual_seed(3)
    dec = fashion_decoder()
    a = torch.randn(40, 16)
    b = a + 0.1 * torch.randn(40, 16)
    marg = kink_margin.relu_margins(dec, a, b, 512, np.arange(40))
    L32 = oracle.edge_lengths(dec, a, b, 512).numpy()
    L64 = np.array([marg[e][2] for e in range(40)])
    assert all(marg[e][0] >= marg[e][1] >= 0 for e in range(40))
    assert np.abs(L32 - L64).max() / L64.max() < 1e-4
    rep = kink_margin.check_edge_lengths(L32, L64, dec, a, b, 512)
    assert rep["median_rel"] <= 1e-6
    bad = L64.copy()
    far = int(np.argmax([marg[e][0] for e in range(40)]))
    bad[far] *= 1.001
    if marg[far][1] >= kink_margin.EPS:
        with pytest.raises(AssertionError):
            kink_margin.check_edge_lengths(bad, L64, dec, a, b, 512)


def test_bench_reference_arm_contract():
    """`bench.py --impl reference` (CPU, SciPy Dijkstra through the oracle) prints exactly one JSON line with the
    contract's keys; the B200 arm refuses to run without a CUDA device instead of falling back."""
    import json
    import os
    import subprocess
    env = dict(os.environ, PYTHONPATH=str(ROOT))
    r = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                        "--images", "80", "--ref-latents", "3000"], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    for key in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better",
                "scaling", "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e"):
        assert key in d, key
    assert d["impl"] == "reference" and d["metric"] == "sssp_edge_relax_per_s" and d["unit"] == "edge-relax/s"
    assert d["value"] > 0 and d["e2e"]["value"] == d["value"] and d["e2e"]["h2d_bytes_per_step"] == 0
    # "reference" when oracle/_ref (the unmodified reference modules, oracle/make_ref.py) travelled with the repo
    expect = "reference" if (ROOT / "oracle" / "_ref" / "src" / "geo" / "geo_shortest_paths.py").exists() else "port"
    assert d["cpu_baseline"]["kind"] == expect and d["cpu_baseline"]["cores"] >= 1 and "workload" in d["config"]
    if not torch.cuda.is_available():
        r2 = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--steps", "1", "--warmup", "0", "--images", "80"],
                            env=env, capture_output=True, text=True, timeout=600)
        assert r2.returncode != 0 and r2.stdout.strip() == ""
