"""CPU oracle for the geodesic codebook path -- TEST INFRASTRUCTURE ONLY.

This module restates, on the CPU, what the reference's hot path computes
(m4rch1n0/vqvae, `src/geo/*` + `src/scripts/build_codebook.py`).  It is the
checker that `tests/`, `__graft_entry__.smoke()` and `bench.py`'s
`cpu_baseline` / `--impl reference` legs compare the CUDA product against.
Nothing under `vqvae_b200/` may import it.

The reference is a thin Python layer over third-party libraries that are NOT
vendored in /root/reference (pyproject.toml:16-17 pins only lower bounds):
    scikit-learn >=1.3  (installed 1.9.0)   NearestNeighbors / ArgKmin32
    SciPy        >=1.11 (installed 1.18.1)  csgraph.dijkstra, connected_components
    NumPy        unpinned (installed 2.3.5) legacy RandomState, pairwise sum
    PyTorch      unpinned (installed 2.11)  autograd.functional.jvp
The oracle calls the same library entry points with the same arguments the
reference's call sites use, so it reproduces the reference bit for bit on the
same inputs.  `oracle/c/geo_oracle.c` additionally restates the published
algorithms (brute-force fp64 kNN, binary-heap Dijkstra in fp64, NumPy's
float32 pairwise sum, `RandomState.choice` inverse-cdf sampling) in plain C so
that parity does not rest on "same library twice".

Pinning: `tests/golden/make_golden.py` imports the unmodified reference from
/root/reference in the build container and stores its outputs per stage
boundary under `tests/golden/`; `tests/test_oracle_golden.py` checks this file
against those vectors (and against the reference's own toy-graph known answers,
tests/test_geo_shortest_paths.py:37-46).
"""
from __future__ import annotations

import numpy as np
import torch
from scipy import sparse
from scipy.sparse import csgraph
from sklearn.neighbors import NearestNeighbors


# ----------------------------------------------------------------------------
# R1: kNN graph  (reference: src/geo/knn_graph_optimized.py:25-67)
# ----------------------------------------------------------------------------
def knn_graph(z, k=10, metric="euclidean", mode="distance", sym="mutual"):
    """Exact kNN graph, self removed, symmetrised (knn_graph_optimized.py:25-67)."""
    assert z.ndim == 2
    n = z.shape[0]
    if n == 0:  # :30-31
        return (sparse.csr_matrix((0, 0), dtype=np.float32),
                {"distances": np.empty((0, 0), np.float32), "indices": np.empty((0, 0), dtype=int)})
    kk = max(0, min(k, n - 1))  # :34
    if kk == 0:  # :35-37
        return (sparse.csr_matrix((n, n), dtype=np.float32),
                {"distances": np.empty((n, 0), np.float32), "indices": np.empty((n, 0), dtype=int)})
    finder = NearestNeighbors(n_neighbors=min(kk + 1, n), algorithm="auto", metric=metric).fit(z)  # :39-41
    dist, idx = finder.kneighbors(z, return_distance=True)  # :42
    me = np.arange(n)
    if np.array_equal(idx[:, 0], me):  # :45-46
        dist, idx = dist[:, 1:], idx[:, 1:]
    else:  # :47-52  drop each row's first-minimum column
        keep = np.ones(dist.shape, bool)
        keep[me, dist.argmin(axis=1)] = False
        dist = dist[keep].reshape(n, -1)
        idx = idx[keep].reshape(n, -1)
    vals = dist.ravel() if mode == "distance" else np.ones_like(dist).ravel()  # :54
    W = sparse.csr_matrix((vals.astype(np.float32), (np.repeat(me, idx.shape[1]), idx.ravel())), shape=(n, n))
    if sym == "mutual":  # :59-64
        W = W.minimum(W.T)
    elif sym == "union":
        W = W.maximum(W.T)
    else:
        raise ValueError(f"Invalid symmetry mode: {sym}")
    W.setdiag(0.0)  # :65-66
    W.eliminate_zeros()
    return W, {"distances": dist.astype(np.float32, copy=False), "indices": idx}


def largest_component_mask(W):
    """Boolean mask of the largest connected component (knn_graph_optimized.py:173-181)."""
    ncomp, lab = csgraph.connected_components(W, directed=False)
    if ncomp <= 1:
        return np.ones(W.shape[0], dtype=bool)
    return lab == np.argmax(np.bincount(lab))


# ----------------------------------------------------------------------------
# R3: shortest paths  (reference: src/geo/geo_shortest_paths.py:13-76)
# ----------------------------------------------------------------------------
def check_graph(W):
    """geo_shortest_paths.py:13-21."""
    if not sparse.isspmatrix(W):
        raise TypeError("W must be a scipy sparse matrix")
    if W.shape[0] != W.shape[1]:
        raise ValueError("W must be square")
    if W.nnz > 0 and (W.data < 0).any():
        raise ValueError("Negative weights")
    return W.tocsr()


def sssp_multi(W, sources, directed=False, unweighted=False, return_predecessors=False, dtype=np.float32):
    """S independent SSSPs, scipy Dijkstra in fp64 then cast (geo_shortest_paths.py:24-50)."""
    if len(sources) == 0:
        raise ValueError("sources must be a non-empty sequence of node indices")
    W = check_graph(W)
    src = np.asarray(sources, dtype=int)
    if unweighted and W.nnz > 0:
        W = W.copy()
        W.data.fill(1.0)
    out = csgraph.dijkstra(csgraph=W, directed=directed, indices=src, return_predecessors=return_predecessors)
    if return_predecessors:
        return out[0].astype(dtype, copy=False), out[1].astype(np.int32, copy=False)
    return out.astype(dtype, copy=False)


def sssp_single(W, source, **kw):
    """geo_shortest_paths.py:53-63."""
    res = sssp_multi(W, [int(source)], **kw)
    if kw.get("return_predecessors", False):
        return res[0][0], res[1][0]
    return res[0]


# ----------------------------------------------------------------------------
# R4: graph K-medoids  (reference: src/geo/kmeans_optimized.py:14-183)
# ----------------------------------------------------------------------------
def kpp_seed(W, K, seed=42, trace=None):
    """k-means++ over graph distances (kmeans_optimized.py:14-74)."""
    n = W.shape[0]
    rng = np.random.RandomState(seed)  # :30
    chosen = [int(rng.randint(0, n))]  # :33
    nearest = np.full(n, np.inf, dtype=np.float32)  # :36
    for _ in range(1, K):  # :40
        nearest = np.minimum(nearest, sssp_single(W, chosen[-1], dtype=np.float32))  # :42-44
        ok = np.isfinite(nearest)  # :47-53
        if ok.any():
            safe = np.where(ok, nearest, np.max(nearest[ok]) * 2.0)
        else:
            safe = np.ones_like(nearest)
        p = safe ** 2  # :56-57
        p[chosen] = 0.0
        if p.sum() > 0:  # :59-61
            p /= p.sum()
            nxt = int(rng.choice(n, p=p))
        else:  # :62-69
            rest = [i for i in range(n) if i not in chosen]
            if not rest:
                break
            nxt = int(rng.choice(rest))
        chosen.append(nxt)
        if trace is not None:
            trace.append(nearest.copy())
    return chosen


def assign_to_medoids(W, medoids):
    """argmin_k float32(D[k, n]), lowest k on ties (kmeans_optimized.py:77-106)."""
    D = sssp_multi(W, medoids, dtype=np.float32)
    return D.argmin(axis=0).astype(int)


def quantization_error(W, medoids, assign):
    """sum of squared f32 distances to the assigned medoid (kmeans_optimized.py:109-138)."""
    D = sssp_multi(W, medoids, dtype=np.float32)
    d = D[assign, np.arange(len(assign))]
    ok = np.isfinite(d)
    return float(np.sum(d[ok] ** 2)) if ok.any() else float("inf")


def fit_kmedoids(W, K=512, init="kpp", seed=42):
    """kmeans_optimized.py:141-183: seeding -> one assignment -> QE.  No update step."""
    n = W.shape[0]
    if init == "kpp":
        med = np.array(kpp_seed(W, K, seed=seed), dtype=int)
    elif init == "random":
        med = np.random.RandomState(seed).choice(n, size=min(K, n), replace=False).astype(int)
    else:
        raise ValueError("init must be 'kpp' or 'random'")
    assign = assign_to_medoids(W, med)
    return med, assign, quantization_error(W, med, assign)


# ----------------------------------------------------------------------------
# R2: Riemannian edge lengths  (reference: src/geo/riemannian_metric.py:12-66)
# ----------------------------------------------------------------------------
def _jvp_sigmoid_decoder(decoder, z, dz):
    """J(z) dz for z -> sigmoid(decoder(z)).flatten (riemannian_metric.py:12-35)."""
    flat_in = hasattr(next(decoder.children()), "in_features")

    def img(p):
        if not flat_in and p.ndim == 2:
            p = p[:, :, None, None]
        y = torch.sigmoid(decoder(p))
        return y.reshape(y.shape[0], -1)

    return torch.autograd.functional.jvp(img, (z.requires_grad_(True),), (dz,))[1]


@torch.no_grad()
def edge_lengths(decoder, z_start, z_end, batch_size=512):
    """0.5 (|J(zi) d| + |J(zj) d|) per edge in consecutive batches (riemannian_metric.py:37-66).

    The decoder is used in whatever train/eval mode the caller left it in; the
    reference CLI leaves it in TRAIN mode (build_codebook.py:27-29), so BatchNorm
    normalises with the statistics of each 512-edge batch.
    """
    assert z_start.shape == z_end.shape
    dev = next(decoder.parameters()).device
    a, b = z_start.to(dev), z_end.to(dev)
    d = b - a
    out = []
    for lo in range(0, a.shape[0], batch_size):
        hi = min(lo + batch_size, a.shape[0])
        ja = _jvp_sigmoid_decoder(decoder, a[lo:hi], d[lo:hi])
        jb = _jvp_sigmoid_decoder(decoder, b[lo:hi], d[lo:hi])
        out.append(0.5 * (torch.linalg.vector_norm(ja, dim=1) + torch.linalg.vector_norm(jb, dim=1)))
    return torch.cat(out).to(torch.float32)


# ----------------------------------------------------------------------------
# R0: the CLI body up to (not including) artifact I/O  (build_codebook.py:31-76)
# ----------------------------------------------------------------------------
def build_codebook_arrays(z4, decoder, k=20, sym="union", K=512, init="kpp", seed=42, batch_size=512, timings=None):
    """Stage driver on in-memory inputs; returns every stage-boundary array."""
    import time
    t = time.perf_counter
    N, C, H, Wd = z4.shape
    zf = z4.permute(0, 2, 3, 1).reshape(-1, C).numpy()  # :35
    t0 = t()
    We, info = knn_graph(zf, k=k, metric="euclidean", mode="connectivity", sym=sym)  # :38
    t1 = t()
    r, c = We.nonzero()  # :43-45
    up = r < c
    ei, ej = r[up], c[up]
    L = edge_lengths(decoder, torch.from_numpy(zf[ei]).float(), torch.from_numpy(zf[ej]).float(), batch_size)  # :47-51
    t2 = t()
    Wg = sparse.csr_matrix((L.cpu().numpy(), (ei, ej)), shape=We.shape)  # :53-54
    Wg = Wg + Wg.T
    mask = largest_component_mask(Wg)  # :56-63
    Wl = Wg[mask][:, mask] if mask.sum() < Wg.shape[0] else Wg
    t3 = t()
    med, assign, qe = fit_kmedoids(Wl, K=K, init=init, seed=seed)  # :70
    t4 = t()
    codes = np.full(zf.shape[0], -1, dtype=np.int32)  # :74-76
    codes[mask] = assign
    if timings is not None:
        timings.update(knn=t1 - t0, edge_len=t2 - t1, lcc=t3 - t2, kmedoids=t4 - t3)
    return dict(z_flat=zf, knn_idx=info["indices"], W_euc=We, ei=ei, ej=ej, edge_len=L.cpu().numpy(), W_geo=Wg,
                mask=mask, W_lcc=Wl, medoids=med, assign=assign, qe=qe, codes=codes.reshape(N, H, Wd))
