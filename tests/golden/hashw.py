"""Reproducible symmetric pseudo-random edge weights (integer hash -> float32 in [0.25, 2.25))."""
import numpy as np


def hash_weights(i, j):
    """w(i,j)=w(j,i); pure uint64 arithmetic so every machine produces the same float32 bits."""
    a = np.minimum(i, j).astype(np.uint64)
    b = np.maximum(i, j).astype(np.uint64)
    h = (a * np.uint64(0x9E3779B97F4A7C15) + b * np.uint64(0xC2B2AE3D27D4EB4F)) & np.uint64(0xFFFFFFFFFFFFFFFF)
    h ^= h >> np.uint64(29)
    h = (h * np.uint64(0xBF58476D1CE4E5B9)) & np.uint64(0xFFFFFFFFFFFFFFFF)
    h ^= h >> np.uint64(32)
    frac = (h & np.uint64(0xFFFFF)).astype(np.float64) / float(1 << 20)
    return (0.25 + 2.0 * frac).astype(np.float32)


def reweight(W):
    """Copy of CSR W with data replaced by hash_weights(row, col)."""
    W = W.tocsr().copy()
    W.sort_indices()
    rows = np.repeat(np.arange(W.shape[0]), np.diff(W.indptr))
    W.data = hash_weights(rows, W.indices)
    return W
