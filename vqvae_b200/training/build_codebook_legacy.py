"""Euclidean-graph legacy builder -- drop-in for the reference's src/training/build_codebook_legacy.py:27-99.

YAML keys: data.latents_path | data.dataset, graph.{k,metric,sym,mode}, quantize.{K,init,seed}, out.dir.
Artifacts: knn_graph.npz (LCC-restricted CSR), codebook.pt {medoid_indices int32, z_medoid f32, config},
codes.npy (1-D, -1 outside the LCC).  Every geo call runs on the GPU (vqvae_b200.geo); shipped configs use
`mode: connectivity`, i.e. unit weights and hop-count geodesics with pervasive argmin ties (lowest k wins)."""
from __future__ import annotations

from pathlib import Path
from typing import Any, Dict

import numpy as np
import torch
from scipy import sparse

from ..geo.kmeans_optimized import fit_kmedoids_optimized
from ..geo.knn_graph_optimized import analyze_graph_connectivity, build_knn_graph_auto, largest_connected_component
from . import _common as C


def build_and_save(config: Dict[str, Any]) -> Path:
    out_dir = Path(config["out"]["dir"])
    out_dir.mkdir(parents=True, exist_ok=True)
    z = C.load_latents(C.latents_path(config)).numpy()
    N, D = z.shape
    print(f"Loaded latents: N={N}, D={D}")
    k, metric, sym, mode = C.graph_args(config)
    W, _ = build_knn_graph_auto(z, k=k, metric=metric, mode=mode, sym=sym)
    analyze_graph_connectivity(W)
    mask = largest_connected_component(W)
    W, z_lcc = C.restrict_to_lcc(W, z, mask)
    sparse.save_npz(out_dir / "knn_graph.npz", W)
    K, init, seed = C.quantize_args(config)
    medoids, assign_lcc, qe = fit_kmedoids_optimized(W, K=K, init=init, seed=seed)
    torch.save({"medoid_indices": medoids.astype(np.int32), "z_medoid": torch.from_numpy(z_lcc[medoids]).float(),
                "config": config}, out_dir / "codebook.pt")
    np.save(out_dir / "codes.npy", C.scatter_codes(assign_lcc, mask, N))
    print(f"Quantization error: {qe:.3f}")
    print(f"Saved artifacts to: {out_dir}")
    return out_dir


if __name__ == "__main__":
    import argparse
    import yaml
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=str, default="configs/quantize.yaml")
    with open(ap.parse_args().config, "r") as f:
        print(f"Completed: {build_and_save(yaml.safe_load(f))}")
