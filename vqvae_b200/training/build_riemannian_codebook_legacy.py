"""Riemannian legacy builder -- drop-in for the reference's src/training/build_riemannian_codebook_legacy.py:169-291.

Differences from the spatial CLI that this driver keeps: the decoder is the flat-latent `VAE.decoder`
in EVAL mode (:62), the LCC is taken on the Euclidean graph before re-weighting (:233-240), every stored directed
entry is re-weighted (both (i,j) and (j,i), COO order) and the result symmetrised with `maximum(W.T)` (:97-155),
`riemannian.mode: subset` draws its edge sample from NumPy's GLOBAL unseeded RNG (:119, not reproducible by design),
two graphs are saved, codes.npy is 1-D.  YAML keys: data.*, checkpoint_path | vae.ckpt_path | model.checkpoint_path,
vae_config | model | vae, graph.*, riemannian.{mode,max_edges,batch_size}, quantize.*, out.dir."""
from __future__ import annotations

import warnings
from pathlib import Path
from typing import Dict

import numpy as np
import torch
from scipy import sparse

from ..geo.kmeans_optimized import fit_kmedoids_optimized
from ..geo.knn_graph_optimized import analyze_graph_connectivity, build_knn_graph_auto, largest_connected_component
from ..geo.riemannian_metric import edge_lengths_riemannian
from ..models.vae import VAE
from . import _common as C


def _load_vae_model(checkpoint_path: Path, vae_config: Dict, device: torch.device) -> VAE:
    """VAE(**vae_config) with the checkpoint's `model_state_dict` (or a bare state dict), eval mode."""
    ckpt = torch.load(checkpoint_path, map_location=device)
    if not isinstance(ckpt, dict):
        raise ValueError("Expected checkpoint dict with model_state_dict")
    model = VAE(**vae_config)
    model.load_state_dict(ckpt.get("model_state_dict", ckpt))
    return model.to(device).eval()


def _sample_edges(weights: np.ndarray, max_edges: int) -> np.ndarray:
    """Stratified sample over five Euclidean-length quantile bins (:104-122); NumPy global RNG, as the reference."""
    edges = np.linspace(0, 1, 6)
    picked = []
    for lo, hi in zip(edges[:-1], edges[1:]):
        q_lo, q_hi = np.quantile(weights, [lo, hi])
        pool = np.where((weights >= q_lo) & (weights <= q_hi))[0]
        take = min(max_edges // 5, len(pool))
        if take > 0:
            picked.extend(np.random.choice(pool, size=take, replace=False))
    return np.array(picked)


def _reweight_graph_with_riemannian(W: sparse.csr_matrix, z: np.ndarray, decoder: torch.nn.Module, mode: str = "subset",
                                    max_edges: int = 5000, batch_size: int = 512, device: torch.device = None,
                                    ) -> sparse.csr_matrix:
    if device is None:
        device = next(decoder.parameters()).device
    zt = torch.from_numpy(z).float().to(device)
    coo = W.tocoo()
    n_edges = len(coo.row)
    print(f"Graph has {n_edges} edges")
    if mode == "subset" and n_edges > max_edges:
        sel = _sample_edges(coo.data, max_edges)
        print(f"Reweighting {len(sel)} edges (subset mode)")
    else:
        sel = np.arange(n_edges)
        print(f"Reweighting all {len(sel)} edges (full mode)")
    print(f"Computing Riemannian distances for {len(sel)} edges...")
    with torch.no_grad():
        L = edge_lengths_riemannian(decoder, zt[coo.row[sel]], zt[coo.col[sel]], batch_size=batch_size).cpu().numpy()
    new = W.copy().astype(np.float32).tocoo()
    new.data[sel] = L
    Wr = new.tocsr()
    Wr = Wr.maximum(Wr.T)  # undirected: the larger of the two directions' lengths
    bad = ~np.isfinite(Wr.data)
    if bad.any():
        warnings.warn(f"Found {bad.sum()} non-finite Riemannian distances, keeping original Euclidean weights")
        Wr.data[bad] = W.tocsr().data[bad]
    print(f"Riemannian reweighting complete. Edge weight ratio: mean={np.mean(L / coo.data[sel]):.3f}")
    return Wr


def build_and_save(config: Dict) -> Path:
    out_dir = Path(config["out"]["dir"])
    out_dir.mkdir(parents=True, exist_ok=True)
    z = C.load_latents(C.latents_path(config)).numpy()
    N, D = z.shape
    print(f"Loaded latents: N={N}, D={D}")
    device = torch.device("cuda" if torch.cuda.is_available() else "cpu")
    print(f"Using device: {device}")
    vae_config = config.get("vae_config") or config.get("model") or config.get("vae")
    if vae_config is None:
        raise ValueError("VAE configuration not found. Expected 'vae_config', 'model', or 'vae' key in config.")
    vae_config = {k_: v for k_, v in vae_config.items() if k_ not in ("checkpoint_path", "ckpt_path")}
    decoder = _load_vae_model(C.checkpoint_path(config), vae_config, device).decoder
    k, metric, sym, mode = C.graph_args(config)
    print(f"Building k-NN graph: k={k}, metric={metric}, sym={sym}")
    W_euc, _ = build_knn_graph_auto(z, k=k, metric=metric, mode=mode, sym=sym)
    graph_stats = analyze_graph_connectivity(W_euc)
    mask = largest_connected_component(W_euc)
    W_euc, z_lcc = C.restrict_to_lcc(W_euc, z, mask)
    r = config.get("riemannian", {})
    r_mode = r.get("mode", "subset")
    print(f"Re-weighting graph with Riemannian distances (mode={r_mode})")
    W_riem = _reweight_graph_with_riemannian(W_euc, z_lcc, decoder, mode=r_mode, max_edges=int(r.get("max_edges", 5000)),
                                             batch_size=int(r.get("batch_size", 512)), device=device)
    sparse.save_npz(out_dir / "knn_graph_euclidean.npz", W_euc)
    sparse.save_npz(out_dir / "knn_graph_riemannian.npz", W_riem)
    K, init, seed = C.quantize_args(config)
    print(f"Running K-medoids on Riemannian graph: K={K}, init={init}")
    medoids, assign_lcc, qe = fit_kmedoids_optimized(W_riem, K=K, init=init, seed=seed)
    torch.save({"medoid_indices": medoids.astype(np.int32), "z_medoid": torch.from_numpy(z_lcc[medoids]).float(),
                "config": config, "graph_stats": graph_stats, "method": "riemannian_geodesic"}, out_dir / "codebook.pt")
    np.save(out_dir / "codes.npy", C.scatter_codes(assign_lcc, mask, N))
    print(f"Riemannian quantization error: {qe:.3f}")
    print(f"Saved artifacts to: {out_dir}")
    return out_dir


if __name__ == "__main__":
    import argparse
    import yaml
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=str, default="configs/quantize.yaml")
    with open(ap.parse_args().config, "r") as f:
        print(f"Completed: {build_and_save(yaml.safe_load(f))}")
