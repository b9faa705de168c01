"""Euclidean nearest-medoid assignment of held-out latents and the two artifact conventions the reference's
consumers rely on -- drop-in for `nearest_medoid_assign` of src/eval/evaluate_codebook_health.py:13-26 and
src/eval/evaluate_quantization_loss.py:37-50 (identical bodies), `CodesDataset`'s -1 filter
(src/data/codes_dataset.py:15-25) and `z_medoid[codes]` of generate_samples.py.

The reference scores a2 + b2 - 2ab in float32 GEMM batches and takes argmin; here the K x M search runs through
the kNN entry point with separate query / candidate sets (`geocb_knn_query`, tcgen05 filter + exact fp64 re-rank),
so the code of a latent is its exactly nearest medoid, lowest index on exact ties.  The two agree except where the
reference's float32 scores of the two nearest medoids tie or cross within rounding.  No CPU fallback."""
from __future__ import annotations

import numpy as np
import torch

from .. import _lib
from .. import device_graph as dg


def nearest_medoid_assign(z: torch.Tensor, z_medoid: torch.Tensor, batch_size: int = 8192) -> torch.Tensor:
    """Assign latent vectors (N,d) to their nearest medoid (K,d) -> int64 codes (N,) on the CPU.
    `batch_size` is accepted for signature compatibility; the device search tiles by itself."""
    assert z.ndim == 2 and z_medoid.ndim == 2 and z.shape[1] == z_medoid.shape[1]
    if z.shape[0] == 0:
        return torch.empty(0, dtype=torch.long)
    dev = _lib.require_cuda()
    q = z.detach().to(dev, torch.float32).contiguous()
    c = z_medoid.detach().to(dev, torch.float32).contiguous()
    idx, _, _ = dg.knn_query(q, c, 1, want_dist=False)
    return idx[:, 0].long().cpu()


def valid_code_rows(codes: np.ndarray) -> np.ndarray:
    """Rows (images) of a codes.npy array (N,H,W) without any -1 (outside-LCC) entry -- the filter CodesDataset
    applies before flattening to sequences (codes_dataset.py:15-18)."""
    return ~(np.asarray(codes) == -1).any(axis=tuple(range(1, np.asarray(codes).ndim)))


def decode_codes(codes: torch.Tensor, z_medoid: torch.Tensor) -> torch.Tensor:
    """codes (..., H, W) int -> latent grid (..., d, H, W) by codebook lookup (generate_samples.py:55-95)."""
    zq = z_medoid[codes.long()]  # (..., H, W, d)
    return zq.movedim(-1, -3).contiguous()
