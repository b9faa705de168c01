"""k-NN graph construction on the GPU -- drop-in for the reference's
`src/geo/knn_graph_optimized.py` (same function names, arguments, defaults, return
types and exceptions; citations below are lines of that file).

The reference picks sklearn (<50k rows) or FAISS; both are exact, so here every
`method` resolves to the same exact CUDA path (`geocb_knn_topk` + `geocb_sym_*`).
There is no CPU fallback.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import numpy as np
import torch
from scipy import sparse

from .. import _lib
from .. import device_graph as dg


def _empty(n):
    return sparse.csr_matrix((n, n), dtype=np.float32)


def _cosine_rows(z_dev: torch.Tensor) -> torch.Tensor:
    """Rows scaled to unit length the way sklearn's cosine_distances does (float32 norms, zero rows left zero).
    For unit rows |a-b|^2 = 2 - 2 a.b, so the exact Euclidean top-k of these rows IS the cosine top-k."""
    nrm = torch.sqrt((z_dev * z_dev).sum(1))
    nrm = torch.where(nrm == 0, torch.ones_like(nrm), nrm)
    return (z_dev / nrm[:, None]).contiguous()


def knn_graph_device(z_dev: torch.Tensor, k: int, mode: str = "distance", sym: str = "mutual",
                     metric: str = "euclidean", after_launch=None):
    """Device-side body shared by every public builder.

    z_dev: (N,D) float32 CUDA tensor, N >= 2, 1 <= k <= N-1.
    Returns (DeviceCSR, nbr_idx int32 (N,k), nbr_dist f32 (N,k)), self already dropped
    following knn_graph_optimized.py:44-52.
    metric "cosine" (reference :86-97 / sklearn metric="cosine"): neighbours are the exact Euclidean
    top-k of the unit-length rows, reported distance = 1 - cosine similarity (fp64 dot, clipped to [0,2], f32).
    A zero row has cosine distance 1 to everything; among all-tied candidates the lowest indices are returned.
    """
    if sym not in ("mutual", "union"):
        raise ValueError(f"Invalid symmetry mode: {sym}")  # :64
    N = z_dev.shape[0]
    if metric == "cosine":
        zn = _cosine_rows(z_dev)
        idx, _, _ = dg.knn_topk(zn, k + 1)
        dist = torch.empty(idx.shape, dtype=torch.float32, device=z_dev.device)
        for lo in range(0, N, 1 << 18):  # bounded temporaries (fp64 gathered neighbours)
            sl = slice(lo, min(N, lo + (1 << 18)))
            dots = torch.einsum("nd,nkd->nk", zn[sl].double(), zn[idx[sl].long()].double())
            dist[sl] = (1.0 - dots).clamp_(0.0, 2.0).float()
    elif metric == "euclidean":
        idx, dist, _ = dg.knn_topk(z_dev, k + 1)  # k+1 query incl. self (:39-42)
        if after_launch is not None:
            after_launch()  # host-side work the caller wants overlapped with the (asynchronous) search
    else:
        raise ValueError(f"metric '{metric}' not supported. Use 'euclidean' or 'cosine'.")  # :99
    me = torch.arange(N, device=z_dev.device, dtype=torch.int32)
    if bool((idx[:, 0] == me).all()):  # :45-46
        idx, dist = idx[:, 1:].contiguous(), dist[:, 1:].contiguous()
    else:  # :47-52 drop each row's first-minimum column (duplicates: self may survive until setdiag)
        drop = dist.argmin(dim=1, keepdim=True)
        keep = torch.ones_like(idx, dtype=torch.bool).scatter_(1, drop, False)
        idx = idx[keep].view(N, -1).contiguous()
        dist = dist[keep].view(N, -1).contiguous()
    g = dg.symmetrize(idx, dist if mode == "distance" else None, sym)  # :54-66
    return g, idx, dist


def _build(z: np.ndarray, k: int, metric: str, mode: str, sym: str):
    assert z.ndim == 2, "z must be (N,D)"  # :28
    N = z.shape[0]
    if N == 0:  # :30-31
        return _empty(0), {"distances": np.empty((0, 0), np.float32), "indices": np.empty((0, 0), dtype=int)}
    k_eff = max(0, min(k, N - 1))  # :34
    if k_eff == 0:  # :35-37
        return _empty(N), {"distances": np.empty((N, 0), np.float32), "indices": np.empty((N, 0), dtype=int)}
    if metric not in ("euclidean", "cosine"):
        raise ValueError(f"metric '{metric}' not supported. Use 'euclidean' or 'cosine'.")  # :99
    if sym not in ("mutual", "union"):
        raise ValueError(f"Invalid symmetry mode: {sym}")
    dev = _lib.require_cuda()
    z_dev = torch.from_numpy(np.ascontiguousarray(z, dtype=np.float32)).to(dev)
    g, idx, dist = knn_graph_device(z_dev, k_eff, mode, sym, metric)
    info = {"distances": dist.cpu().numpy(), "indices": idx.cpu().numpy().astype(int)}
    return g.to_scipy(), info


def build_knn_graph_sklearn(z: np.ndarray, k: int = 10, metric: str = "euclidean", mode: str = "distance",
                            sym: str = "mutual") -> Tuple[sparse.csr_matrix, Dict[str, np.ndarray]]:
    """Same contract as the reference's sklearn builder (:25-67); runs on the GPU."""
    return _build(z, k, metric, mode, sym)


def build_knn_graph_faiss(z: np.ndarray, k: int = 10, metric: str = "euclidean", mode: str = "distance",
                          sym: str = "mutual") -> Tuple[sparse.csr_matrix, Dict[str, np.ndarray]]:
    """Same contract as the reference's FAISS builder (:70-126); runs on the GPU (exact, like IndexFlatL2)."""
    return _build(z, k, metric, mode, sym)


def build_knn_graph_auto(z: np.ndarray, k: int = 10, metric: str = "euclidean", mode: str = "distance",
                         sym: str = "mutual", force_method: Optional[str] = None,
                         size_threshold: int = 50000) -> Tuple[sparse.csr_matrix, Dict[str, np.ndarray]]:
    """:129-170.  `force_method` / `size_threshold` are accepted for compatibility; every value (including the
    reference's "sklearn" / "faiss", and unknown strings, which the reference lets fall through to auto) resolves to
    the one exact CUDA path, so `force_method="faiss"` works without faiss instead of raising (:153-156)."""
    print(f"Building k-NN graph: N={z.shape[0]}, k={k}, method=cuda")  # :164
    return _build(z, k, metric, mode, sym)


def largest_connected_component(W: sparse.csr_matrix) -> np.ndarray:
    """Boolean mask of the largest connected component (:173-181); ties -> lowest component label."""
    n = W.shape[0]
    if n == 0:
        return np.ones(0, dtype=bool)
    g = W if isinstance(W, dg.DeviceCSR) else _undirected_structure(W)
    mask, _, _, ncomp = dg.largest_component(g)
    if ncomp <= 1:
        return np.ones(n, dtype=bool)
    return mask.cpu().numpy().astype(bool)


def _undirected_structure(W) -> dg.DeviceCSR:
    """connected_components(directed=False) sees an edge if either direction is stored: upload, then merge the
    missing directions on the device (geocb_csr_sym_check / geocb_csr_undirected_*)."""
    dev = _lib.require_cuda()
    g = dg.DeviceCSR.from_scipy(W, dev)
    if g.nnz == 0:
        return g
    flags, extra = dg.csr_sym_check(g)
    return dg.csr_undirected(g, extra) if flags & 2 else g


def analyze_graph_connectivity(W: sparse.csr_matrix) -> Dict:
    """Connectivity statistics + the reference's printed lines (:184-219)."""
    N = W.shape[0]
    dev = _lib.require_cuda()
    g = _undirected_structure(W)
    labels = dg.component_labels(g) if g.nnz else torch.arange(N, dtype=torch.int32, device=dev)
    sizes = torch.bincount(labels.long(), minlength=N)
    sizes = sizes[sizes > 0]
    n_components = int(sizes.numel())
    if n_components > 1:
        largest_size = int(sizes.max())
        connectivity_ratio = largest_size / N
    else:
        largest_size, connectivity_ratio = N, 1.0
    Wd = dg.DeviceCSR.from_scipy(W, dev)
    rows = torch.repeat_interleave(torch.arange(N, device=dev), (Wd.indptr[1:] - Wd.indptr[:-1]))
    degrees = torch.zeros(N, dtype=torch.float32, device=dev).index_add_(0, rows, Wd.data).cpu().numpy()  # weight sums, :201
    avg_degree = degrees.mean()
    stats = {"n_nodes": N, "n_edges": W.nnz, "n_components": n_components, "largest_component_size": largest_size,
             "connectivity_ratio": connectivity_ratio, "avg_degree": avg_degree, "min_degree": degrees.min(),
             "max_degree": degrees.max()}
    print("Graph connectivity")
    print(f"nodes={N} edges={W.nnz} avg_deg={avg_degree:.1f}")
    print(f"components={n_components} largest={largest_size} ({100*connectivity_ratio:.1f}%)")
    if n_components > 1:
        print("disconnected -> will use LCC")
    return stats


def build_knn_graph(z: np.ndarray, k: int = 10, metric: str = "euclidean", mode: str = "distance",
                    sym: str = "mutual") -> Tuple[sparse.csr_matrix, Dict[str, np.ndarray]]:
    """Backward-compatible entry point (:223-231)."""
    return build_knn_graph_auto(z, k=k, metric=metric, mode=mode, sym=sym)
