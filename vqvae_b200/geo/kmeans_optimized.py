"""Graph-based geodesic K-medoids on the GPU -- drop-in for the reference's
`src/geo/kmeans_optimized.py` (citations are lines of that file).

The reference runs (K-1) + K + K Dijkstras.  Its outputs only ever use
d_min[v] = min_k float32(D[k,v]), argmin_k and sum d_min^2, so this module keeps a
running Voronoi state on the device (`device_graph.VoronoiState`): inserting a centre
relaxes outwards only where it beats the current distance (exact in fp64), and the
labels follow the reference's "first lowest k wins float32 ties" rule.  The k-means++
draw reproduces NumPy's float32 pairwise sum and fp64 sequential cdf bit for bit; the
MT19937 stream itself is a real `np.random.RandomState` on the host.
There is no medoid-update step in the reference (:166-183) and none here.
"""
from __future__ import annotations

from typing import List, Tuple

import numpy as np
from scipy import sparse

from .. import device_graph as dg
from .geo_shortest_paths import dijkstra_multi_source, graph_to_device


def _seed_kpp(state: dg.VoronoiState, K: int, seed: int, insert_last: bool):
    """k-means++ chain (:30-72).  Returns the centre list; with insert_last the final centre is inserted
    too, leaving `state` equal to the K-source assignment."""
    N = state.g.n
    rng = np.random.RandomState(seed)  # :30
    centers = [int(rng.randint(0, N))]  # :33
    exact_draws = 0
    for _ in range(1, K):  # :40
        state.insert(centers[-1])  # d_min = min(d_min, float32 dijkstra(last centre))  :42-44
        saved = rng.get_state()
        nxt, exact, _psum = state.draw(rng.random_sample())  # :47-61 (one MT19937 double per draw)
        exact_draws += exact
        if nxt < 0:
            # probs.sum() == 0 (:62-69): the reference draws nothing from `choice(p=...)` here, so give the
            # double back before taking its uniform fallback over the remaining candidates.
            rng.set_state(saved)
            taken = set(centers)
            candidates = [i for i in range(N) if i not in taken]
            if not candidates:
                print(f"Warning: Could not find {K} valid centers, stopping at {len(centers)}")
                break
            nxt = int(rng.choice(candidates))
        centers.append(int(nxt))
    if insert_last and len(centers) > state.count:
        state.insert(centers[-1])
    return centers, exact_draws


def kpp_initialization_graph(W: sparse.spmatrix, K: int, seed: int = 42) -> List[int]:
    """K-means++ initialization using graph distances (:14-74)."""
    g = graph_to_device(W)
    N = g.n
    print(f"[kpp] Selecting {K} centers among {N} nodes")  # :38
    state = dg.VoronoiState(g)
    centers, _ = _seed_kpp(state, K, seed, insert_last=False)
    print(f"[kpp] Selected {len(centers)} centers")  # :73
    return centers


def _labels_for(g, medoids):
    """Insert the medoids in order -> (label int32 device, dmin32 device)."""
    state = dg.VoronoiState(g)
    for k, m in enumerate(np.asarray(medoids, dtype=np.int64).tolist()):
        state.insert(int(m), code=k)
    return state


def _finish_assign(state, K):
    state.check()  # a non-converged insert (round limit) must not pass silently
    assign = state.label.cpu().numpy().astype(int)
    assign[assign < 0] = 0  # argmin of an all-inf column is 0 (:100)
    counts = np.bincount(assign, minlength=K)
    print(f"[assign] sizes min={counts.min()}, max={counts.max()}, mean={counts.mean():.1f}")  # :104
    return assign


def assign_points_to_medoids(W: sparse.spmatrix, medoids: np.ndarray) -> np.ndarray:
    """argmin_k float32(D[k,n]) with the lowest k on ties (:77-106)."""
    g = graph_to_device(W)
    K = len(medoids)
    if K == 0:
        raise ValueError("sources must be a non-empty sequence of node indices")
    print(f"[assign] {g.n} points to {K} medoids")  # :94
    return _finish_assign(_labels_for(g, medoids), K)


def _qe_from_dmin(dmin32) -> float:
    total, used = dg.pairwise_sum_f32(dmin32, square=True, finite_only=True)  # :129-132
    return float(total) if used > 0 else float("inf")  # :133-134


def compute_quantization_error(W: sparse.spmatrix, medoids: np.ndarray, assign: np.ndarray) -> float:
    """Sum of squared float32 geodesic distances to the assigned medoid (:109-138).

    `assign` may be any assignment, so D[assign[v], v] is gathered from true multi-source
    distances, a chunk of sources at a time."""
    import torch
    g = graph_to_device(W)
    med = np.asarray(medoids, dtype=np.int64)
    a = torch.as_tensor(np.asarray(assign, dtype=np.int64), device=g.indptr.device)
    picked = torch.full((g.n,), float("inf"), dtype=torch.float32, device=g.indptr.device)
    if g.nnz == 0:
        hit = torch.as_tensor(med, device=a.device)[a] == torch.arange(g.n, device=a.device)
        picked[hit] = 0.0
        return _qe_from_dmin(picked)
    chunk = max(1, min(len(med), (256 << 20) // max(1, 4 * g.n)))
    cols = torch.arange(g.n, device=a.device)
    for lo in range(0, len(med), chunk):
        D = dg.sssp_multi(g, med[lo:lo + chunk])
        sel = (a >= lo) & (a < lo + D.shape[0])
        picked[sel] = D[(a[sel] - lo), cols[sel]]
    return _qe_from_dmin(picked)


def fit_kmedoids_optimized(W: sparse.spmatrix, K: int = 512, init: str = "kpp", seed: int = 42,
                           ) -> Tuple[np.ndarray, np.ndarray, float]:
    """k-means++ (or random) seeding -> one assignment -> quantisation error (:141-183)."""
    g = graph_to_device(W)
    N = g.n
    print(f"[kmedoids] N={N}, K={K}, edges={g.nnz}, avg_deg={g.nnz/max(1,N):.1f}")  # :164
    if init == "kpp":
        print(f"[kpp] Selecting {K} centers among {N} nodes")
        state = dg.VoronoiState(g)
        centers, _ = _seed_kpp(state, K, seed, insert_last=True)
        print(f"[kpp] Selected {len(centers)} centers")
        medoids = np.array(centers, dtype=int)
    elif init == "random":
        rng = np.random.RandomState(seed)
        medoids = rng.choice(N, size=min(K, N), replace=False).astype(int)  # :170-171
        state = _labels_for(g, medoids)
    else:
        raise ValueError("init must be 'kpp' or 'random'")  # :173
    print(f"[assign] {N} points to {len(medoids)} medoids")
    assign = _finish_assign(state, len(medoids))
    qe = _qe_from_dmin(state.dmin32)  # D[assign[v],v] == d_min[v] because assign is the argmin
    print(f"[kmedoids] Done: clusters={len(medoids)}, qe={qe:.3f}")  # :181
    return medoids, assign, qe


def fit_kmedoids_with_connectivity_check(W: sparse.spmatrix, K: int = 512, init: str = "kpp", seed: int = 42,
                                         ) -> Tuple[np.ndarray, np.ndarray, float, dict]:
    """fit_kmedoids_optimized + component metadata (:186-227)."""
    import torch
    g = graph_to_device(W)
    labels = dg.component_labels(g) if g.nnz else torch.arange(g.n, dtype=torch.int32, device=g.indptr.device)
    sizes = torch.bincount(labels.long(), minlength=g.n)
    n_components = int((sizes > 0).sum())
    metadata = {"n_nodes": g.n, "n_edges": g.nnz, "n_components": n_components,
                "largest_component_size": int(sizes.max()) if n_components > 0 else g.n}
    print(f"[graph] components={n_components}, largest={metadata['largest_component_size']}")
    medoids, assign, qe = fit_kmedoids_optimized(g, K=K, init=init, seed=seed)
    metadata.update({"n_medoids": len(medoids), "quantization_error": qe, "method": "optimized_kmedoids"})
    return medoids, assign, qe, metadata
