"""Generate the golden vectors in tests/golden/ by running the UNMODIFIED reference.

Run in the build container only (needs /root/reference):
    python tests/golden/make_golden.py [--skip-c1]

Cases
  s1   : z = randn(125,16,4,4; seed 42) -> 2000 latents, k=10 union, K=32 kpp seed 42,
         batch 512, FashionMNIST spatial decoder random-init with torch.manual_seed(0).
         Every stage-boundary array is stored in full (s1_stages.npz).
  s1m  : same latents through `build_knn_graph` defaults (mode=distance, sym=mutual), k=6.
  c1   : BASELINE.json configs[0] (1250x16x4x4 -> 20000 latents, k=20 union, K=512).
         Stored as digests + the small arrays (medoids, qe, degree stats, sampled
         weights/distances) because the full graph is several MB (c1_digest.npz).
  c3   : CIFAR-shape case of BASELINE.json configs[2] at a CPU-runnable size (`--c3`): z = randn(1250,32,4,4;
         seed 42) -> 20000 latents d=32, CIFAR spatial decoder (in_channels 3, output 32, latent_dim 32,
         dec 256-128-64, BatchNorm train mode; configs/cifar10/spatial/geodesic/vae.yaml:36-44 and the CLI flags
         of scripts/run_cifar10_spatial_geodesic_pipeline.py:63-71), JVP output P = 3x8x8 = 192, k=20 union,
         K=512 kpp seed 42.  Stored like c1 (c3_digest.npz).
The decoder weights are not stored: `decoder_sha` pins the state_dict produced by
`torch.manual_seed(0); SpatialVAE(**FASHION_CFG)` so the GPU box can rebuild it.
"""
import argparse
import hashlib
import io
import contextlib
import sys
import time
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, "/root/reference")
from src.geo.knn_graph_optimized import build_knn_graph, build_knn_graph_auto, largest_connected_component  # noqa: E402
from src.geo.riemannian_metric import edge_lengths_riemannian  # noqa: E402
from src.geo.geo_shortest_paths import dijkstra_multi_source  # noqa: E402
from src.geo.kmeans_optimized import fit_kmedoids_optimized, kpp_initialization_graph  # noqa: E402
from src.models import SpatialVAE  # noqa: E402
from scipy import sparse  # noqa: E402

HERE = Path(__file__).resolve().parent
FASHION_CFG = dict(in_channels=1, output_image_size=28, latent_dim=16, enc_channels=[64, 128, 256],
                   dec_channels=[256, 128, 64], recon_loss="mse", norm_type="batch", mse_use_sigmoid=True)


def sha(a):
    a = np.ascontiguousarray(a)
    return hashlib.sha256(a.tobytes()).hexdigest()


def state_sha(sd):
    h = hashlib.sha256()
    for k in sorted(sd):
        h.update(k.encode())
        h.update(np.ascontiguousarray(sd[k].detach().cpu().numpy()).tobytes())
    return h.hexdigest()


CIFAR_CFG = dict(in_channels=3, output_image_size=32, latent_dim=32, enc_channels=[64, 128, 256],
                 dec_channels=[256, 128, 64], recon_loss="mse", norm_type="batch", mse_use_sigmoid=False)


def make_decoder(cfg=None):
    torch.manual_seed(0)
    vae = SpatialVAE(**(cfg or FASHION_CFG))  # train mode, as build_codebook.py:27-29 leaves it
    return vae.decoder


def run_reference(z4, k, sym, K, seed=42, batch=512, cfg=None):
    """build_codebook.py:31-76 verbatim in behaviour, on in-memory inputs."""
    dec = make_decoder(cfg)
    dsha = state_sha(dec.state_dict())
    N, C, H, Wd = z4.shape
    zf = z4.permute(0, 2, 3, 1).reshape(-1, C).numpy()
    tm = {}
    t0 = time.perf_counter()
    We, info = build_knn_graph_auto(zf, k=k, metric="euclidean", mode="connectivity", sym=sym)
    tm["knn"] = time.perf_counter() - t0
    rows, cols = We.nonzero()
    e = np.stack((rows, cols), axis=1)[rows < cols]
    t0 = time.perf_counter()
    L = edge_lengths_riemannian(dec, torch.from_numpy(zf[e[:, 0]]).float(), torch.from_numpy(zf[e[:, 1]]).float(), batch_size=batch)
    tm["edge_len"] = time.perf_counter() - t0
    Wg = sparse.csr_matrix((L.cpu().numpy(), (e[:, 0], e[:, 1])), shape=We.shape)
    Wg = Wg + Wg.T
    mask = largest_connected_component(Wg)
    Wl = Wg[mask][:, mask] if mask.sum() < Wg.shape[0] else Wg
    t0 = time.perf_counter()
    with contextlib.redirect_stdout(io.StringIO()):
        med, assign, qe = fit_kmedoids_optimized(Wl, K=K, init="kpp", seed=seed)
    tm["kmedoids"] = time.perf_counter() - t0
    codes = np.full(zf.shape[0], -1, dtype=np.int32)
    codes[mask] = assign
    return dict(zf=zf, We=We, idx=info["indices"], e=e, L=L.cpu().numpy(), Wg=Wg, mask=mask, Wl=Wl, med=med,
                assign=assign, qe=qe, codes=codes.reshape(N, H, Wd), tm=tm, dsha=dsha)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--skip-c1", action="store_true")
    ap.add_argument("--c3", action="store_true", help="only (re)generate c3_digest.npz")
    ap.add_argument("--c4", action="store_true", help="only (re)generate c4_legacy.npz (legacy YAML drivers)")
    a = ap.parse_args()
    if a.c3:
        return make_c3()
    if a.c4:
        return make_c4()

    # ---------------- s1: full stage-boundary arrays ----------------
    z4 = torch.randn(125, 16, 4, 4, generator=torch.Generator().manual_seed(42))
    r = run_reference(z4, k=10, sym="union", K=32)
    Wl = r["Wl"].tocsr()
    Wl.sort_indices()
    src = np.array([0, 7, 1999, 1234], dtype=np.int64)
    D = dijkstra_multi_source(Wl, src)
    Dp, P = dijkstra_multi_source(Wl, src[:2], return_predecessors=True)
    Du = dijkstra_multi_source(Wl, src[:2], unweighted=True)
    with contextlib.redirect_stdout(io.StringIO()):
        med_r, asg_r, qe_r = fit_kmedoids_optimized(Wl, K=16, init="random", seed=7)
        kpp8 = kpp_initialization_graph(Wl, 8, seed=3)
    np.savez_compressed(
        HERE / "s1_stages.npz",
        knn_idx=r["idx"].astype(np.int32), euc_indptr=r["We"].indptr, euc_indices=r["We"].indices,
        ei=r["e"][:, 0].astype(np.int32), ej=r["e"][:, 1].astype(np.int32), edge_len=r["L"],
        geo_indptr=Wl.indptr, geo_indices=Wl.indices, geo_data=Wl.data, mask=r["mask"],
        medoids=r["med"].astype(np.int64), assign=r["assign"].astype(np.int32), qe=np.float64(r["qe"]),
        codes=r["codes"], sssp_src=src, sssp_D=D, sssp_P=P, sssp_Dunw=Du,
        rand_medoids=med_r.astype(np.int64), rand_assign=asg_r.astype(np.int32), rand_qe=np.float64(qe_r),
        kpp8_seed3=np.array(kpp8, dtype=np.int64), decoder_sha=np.array(r["dsha"]),
        z_sha=np.array(sha(r["zf"])),
    )
    print("s1:", "E", len(r["L"]), "nnz", Wl.nnz, "lcc", int(r["mask"].sum()), "qe", r["qe"], "med[:5]", r["med"][:5], r["tm"])

    # ---------------- s1m: distance / mutual defaults ----------------
    with contextlib.redirect_stdout(io.StringIO()):
        Wm, im = build_knn_graph(r["zf"], k=6)
    Wm.sort_indices()
    np.savez_compressed(HERE / "s1m_knn.npz", indptr=Wm.indptr, indices=Wm.indices, data=Wm.data,
                        knn_idx=im["indices"].astype(np.int32), knn_dist=im["distances"])
    print("s1m: nnz", Wm.nnz)

    if a.skip_c1:
        return
    # ---------------- c1: digests of BASELINE configs[0] ----------------
    z4 = torch.randn(1250, 16, 4, 4, generator=torch.Generator().manual_seed(42))
    t0 = time.perf_counter()
    r = run_reference(z4, k=20, sym="union", K=512)
    total = time.perf_counter() - t0
    Wl = r["Wl"].tocsr()
    Wl.sort_indices()
    samp = np.linspace(0, len(r["L"]) - 1, 4096).astype(np.int64)
    src = np.array([0, 19999, 15795], dtype=np.int64)
    D = dijkstra_multi_source(Wl, src)
    deg = np.diff(Wl.indptr)
    sys.path.insert(0, str(HERE))
    from hashw import reweight
    Wh = reweight(r["We"])
    with contextlib.redirect_stdout(io.StringIO()):
        med_h, asg_h, qe_h = fit_kmedoids_optimized(Wh, K=64, init="kpp", seed=42)
    np.savez_compressed(
        HERE / "c1_digest.npz",
        n=np.int64(Wl.shape[0]), nnz=np.int64(Wl.nnz), E=np.int64(len(r["L"])),
        knn_idx_sha=np.array(sha(r["idx"].astype(np.int32))),
        euc_indptr_sha=np.array(sha(r["We"].indptr.astype(np.int32))),
        euc_indices_sha=np.array(sha(r["We"].indices.astype(np.int32))),
        deg_min_mean_max=np.array([deg.min(), deg.mean(), deg.max()], dtype=np.float64),
        edge_sample_pos=samp, edge_sample_len=r["L"][samp],
        edge_len_min_mean_max=np.array([r["L"].min(), r["L"].mean(), r["L"].max()], dtype=np.float64),
        lcc_size=np.int64(r["mask"].sum()), medoids=r["med"].astype(np.int64), qe=np.float64(r["qe"]),
        codes_sha=np.array(sha(r["codes"].astype(np.int32))), sssp_src=src, sssp_D_sha=np.array(sha(D)),
        sssp_D_sample=D[:, ::97].copy(), cluster_sizes=np.bincount(r["assign"], minlength=512).astype(np.int32),
        decoder_sha=np.array(r["dsha"]), z_sha=np.array(sha(r["zf"])),
        hashw_medoids=med_h.astype(np.int64), hashw_assign_sha=np.array(sha(asg_h.astype(np.int32))), hashw_qe=np.float64(qe_h),
        ref_seconds=np.array([r["tm"]["knn"], r["tm"]["edge_len"], r["tm"]["kmedoids"], total]),
    )
    print("c1:", "E", len(r["L"]), "nnz", Wl.nnz, "qe", r["qe"], "med[:5]", r["med"][:5], r["tm"], "total", total)


def make_c3():
    z4 = torch.randn(1250, 32, 4, 4, generator=torch.Generator().manual_seed(42))
    t0 = time.perf_counter()
    r = run_reference(z4, k=20, sym="union", K=512, cfg=CIFAR_CFG)
    total = time.perf_counter() - t0
    Wl = r["Wl"].tocsr()
    Wl.sort_indices()
    samp = np.linspace(0, len(r["L"]) - 1, 4096).astype(np.int64)
    src = np.array([0, 19999, int(r["med"][0])], dtype=np.int64)
    D = dijkstra_multi_source(Wl, src)
    deg = np.diff(Wl.indptr)
    np.savez_compressed(
        HERE / "c3_digest.npz",
        n=np.int64(Wl.shape[0]), nnz=np.int64(Wl.nnz), E=np.int64(len(r["L"])),
        knn_idx_sha=np.array(sha(r["idx"].astype(np.int32))),
        euc_indptr_sha=np.array(sha(r["We"].indptr.astype(np.int32))),
        euc_indices_sha=np.array(sha(r["We"].indices.astype(np.int32))),
        deg_min_mean_max=np.array([deg.min(), deg.mean(), deg.max()], dtype=np.float64),
        edge_sample_pos=samp, edge_sample_len=r["L"][samp],
        edge_len_min_mean_max=np.array([r["L"].min(), r["L"].mean(), r["L"].max()], dtype=np.float64),
        lcc_size=np.int64(r["mask"].sum()), medoids=r["med"].astype(np.int64), qe=np.float64(r["qe"]),
        codes_sha=np.array(sha(r["codes"].astype(np.int32))), sssp_src=src, sssp_D_sha=np.array(sha(D)),
        sssp_D_sample=D[:, ::97].copy(), cluster_sizes=np.bincount(r["assign"], minlength=512).astype(np.int32),
        decoder_sha=np.array(r["dsha"]), z_sha=np.array(sha(r["zf"])),
        ref_seconds=np.array([r["tm"]["knn"], r["tm"]["edge_len"], r["tm"]["kmedoids"], total]),
    )
    print("c3:", "E", len(r["L"]), "nnz", Wl.nnz, "qe", r["qe"], "med[:5]", r["med"][:5], r["tm"], "total", total)


C4_VAE_CFG = dict(in_channels=3, output_image_size=32, latent_dim=32, enc_channels=[16, 32, 64],
                  dec_channels=[64, 32, 16], recon_loss="mse", norm_type="batch", mse_use_sigmoid=False,
                  free_bits_default=0.25, capacity_max_default=25.0, capacity_anneal_steps_default=100000,
                  capacity_mode_default="abs")


def c4_inputs(tmp):
    """Flat latents + a random-init vanilla VAE checkpoint, written where the legacy drivers read them."""
    from src.models.vae import VAE
    z = torch.randn(3000, 32, generator=torch.Generator().manual_seed(42))
    torch.save(z, tmp / "z.pt")
    torch.manual_seed(0)
    vae = VAE(**C4_VAE_CFG)
    torch.save({"model_state_dict": vae.state_dict(), "epoch": 0}, tmp / "best.pt")
    return z, vae


def c4_configs(tmp):
    """Shape of configs/cifar10/vanilla/{geodesic,euclidean}/codebook.yaml at a CPU-runnable size; `riemannian.mode: full`
    because the shipped `subset` mode samples with NumPy's unseeded global RNG (not reproducible)."""
    riem = {"data": {"dataset": "cifar10", "latents_path": str(tmp / "z.pt")}, "model": dict(C4_VAE_CFG),
            "checkpoint_path": str(tmp / "best.pt"),
            "graph": {"k": 10, "metric": "euclidean", "sym": "union", "mode": "distance"},
            "riemannian": {"mode": "full", "max_edges": 5000, "batch_size": 512},
            "quantize": {"K": 64, "init": "kpp", "seed": 42}, "out": {"dir": str(tmp / "out_riem")}}
    euc = {"data": {"dataset": "cifar10", "latents_path": str(tmp / "z.pt")},
           "graph": {"k": 10, "metric": "euclidean", "sym": "union", "mode": "connectivity"},
           "quantize": {"K": 64, "init": "kpp", "seed": 42}, "out": {"dir": str(tmp / "out_euc")}}
    return riem, euc


def make_c4():
    """Legacy YAML drivers (BASELINE configs[3] shape): src/training/build_riemannian_codebook_legacy.py:169-291 with
    eval-mode BatchNorm and both edge directions re-weighted, and src/training/build_codebook_legacy.py:27-99 on a
    unit-weight (hop count) graph."""
    import tempfile
    from src.training.build_riemannian_codebook_legacy import build_and_save as riem_build
    from src.training.build_codebook_legacy import build_and_save as euc_build
    with tempfile.TemporaryDirectory() as td:
        tmp = Path(td)
        c4_inputs(tmp)
        riem, euc = c4_configs(tmp)
        with contextlib.redirect_stdout(io.StringIO()):
            riem_build(riem)
            euc_build(euc)
        Wr = sparse.load_npz(tmp / "out_riem" / "knn_graph_riemannian.npz").tocsr()
        We = sparse.load_npz(tmp / "out_riem" / "knn_graph_euclidean.npz").tocsr()
        Wk = sparse.load_npz(tmp / "out_euc" / "knn_graph.npz").tocsr()
        for W in (Wr, We, Wk):
            W.sort_indices()
        cb_r = torch.load(tmp / "out_riem" / "codebook.pt", weights_only=False)
        cb_e = torch.load(tmp / "out_euc" / "codebook.pt", weights_only=False)
        codes_r, codes_e = np.load(tmp / "out_riem" / "codes.npy"), np.load(tmp / "out_euc" / "codes.npy")
        np.savez_compressed(
            HERE / "c4_legacy.npz",
            riem_indptr=Wr.indptr, riem_indices=Wr.indices, riem_data=Wr.data, euc_data_sha=np.array(sha(We.data)),
            riem_medoids=cb_r["medoid_indices"], riem_codes=codes_r, riem_method=np.array(cb_r["method"]),
            riem_z_medoid_sha=np.array(sha(cb_r["z_medoid"].numpy())),
            knn_indptr=Wk.indptr, knn_indices=Wk.indices, euc_medoids=cb_e["medoid_indices"], euc_codes=codes_e,
            codes_dtypes=np.array([str(codes_r.dtype), str(codes_e.dtype)]),
            graph_stats_keys=np.array(sorted(cb_r["graph_stats"])), n_components=np.int64(cb_r["graph_stats"]["n_components"]),
        )
        print("c4: riem nnz", Wr.nnz, "medoids[:5]", cb_r["medoid_indices"][:5], "codes", codes_r.shape, codes_r.dtype,
              "| euclid nnz", Wk.nnz, "medoids[:5]", cb_e["medoid_indices"][:5], codes_e.dtype)


if __name__ == "__main__":
    main()
