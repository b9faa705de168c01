"""Shared plumbing of the two legacy drivers: config parsing, latent loading, artifact writing."""
from __future__ import annotations

from pathlib import Path

import numpy as np
import torch

_DEFAULT_LATENTS = {"mnist": "experiments/vae_mnist/latents_train/z.pt", "fashion": "experiments/vae_fashion/latents_train/z.pt",
                    "cifar10": "experiments/vae_cifar10/latents_train/z.pt"}
_DEFAULT_CKPT = {"mnist": "experiments/vae_mnist/checkpoints/best.pt", "fashion": "experiments/vae_fashion/checkpoints/best.pt",
                 "cifar10": "experiments/vae_cifar10/checkpoints/best.pt"}


def load_latents(path) -> torch.Tensor:
    """A bare tensor or a dict with key 'z' (build_codebook_legacy.py:14-24)."""
    obj = torch.load(path, map_location="cpu")
    if isinstance(obj, dict) and "z" in obj:
        obj = obj["z"]
    if not torch.is_tensor(obj):
        raise ValueError("Expected dict with 'z' key or tensor")
    return obj.float()


def _dataset_key(config, default):
    data = config.get("data", {}) if isinstance(config.get("data"), dict) else {}
    return str(data.get("dataset", default)).strip().lower()


def latents_path(config) -> Path:
    data = config["data"] if isinstance(config.get("data"), dict) else {}
    p = data.get("latents_path")
    return Path(p) if p else Path(_DEFAULT_LATENTS.get(_dataset_key(config, "mnist"), _DEFAULT_LATENTS["mnist"]))


def checkpoint_path(config) -> Path:
    model = config.get("model") if isinstance(config.get("model"), dict) else {}
    p = config.get("checkpoint_path") or config.get("vae", {}).get("ckpt_path") or model.get("checkpoint_path")
    return Path(p) if p else Path(_DEFAULT_CKPT.get(_dataset_key(config, "fashion"), _DEFAULT_CKPT["fashion"]))


def graph_args(config):
    g = config["graph"]
    return int(g["k"]), str(g["metric"]), str(g["sym"]), str(g["mode"])


def quantize_args(config):
    q = config["quantize"]
    return int(q["K"]), str(q["init"]), int(q["seed"])


def restrict_to_lcc(W, z, mask):
    """(W[mask][:,mask], z[mask]) when the graph is disconnected, else the inputs."""
    if mask.sum() < W.shape[0]:
        print(f"Using LCC: {mask.sum()}/{W.shape[0]} nodes")
        return W[mask][:, mask], z[mask]
    return W, z


def scatter_codes(assign_lcc, mask, N):
    """1-D int32 codes with -1 outside the LCC (the legacy drivers keep `assign_lcc` as is when nothing is masked)."""
    if mask.sum() < N:
        codes = np.full((N,), -1, dtype=np.int32)
        codes[mask] = assign_lcc
        return codes
    return assign_lcc
