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
    state.check()  # a non-converged insert anywhere in the chain invalidates every later draw
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


def _shard_group(dist_group):
    """(dist module, group, rank, world) when the K sources are to be split across ranks, else None.
    Sharding is collective: every rank of the group must make the same call.  It is used when a group is passed
    explicitly (the pipeline does) or, for the default group, when GEOCB_SHARD_SOURCES=1."""
    import os
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return None
    if dist_group is None:
        if os.environ.get("GEOCB_SHARD_SOURCES", "0") != "1":
            return None
        dist_group = dist.group.WORLD
    world = dist.get_world_size(dist_group)
    return (dist, dist_group, dist.get_rank(dist_group), world) if world > 1 else None


def _labels_for(g, medoids, dist_group=None):
    """Insert the medoids in order -> Voronoi state (label int32, dmin32 f32 on the device).

    Multi-GPU: rank r inserts only medoids shard_range(K, r, world) -- independent sources, no chain --
    and the per-rank (dmin32, label) pairs are combined with ONE allreduce-MIN over packed
    (float32 bits << 32 | k) int64 keys: non-negative floats order like their bit patterns, so the minimum key is
    the smallest distance and, on ties, the lowest k -- the reference's argmin rule (:100)."""
    import torch
    med = np.asarray(medoids, dtype=np.int64).tolist()
    sh = _shard_group(dist_group)
    state = dg.VoronoiState(g)
    if sh is None:
        for k, m in enumerate(med):
            state.insert(int(m), code=k)
        return state
    dist, group, rank, world = sh
    from ..sharding import shard_range, allreduce_min_assignment
    lo, hi = shard_range(len(med), rank, world)
    for k in range(lo, hi):
        state.insert(int(med[k]), code=k)
    allreduce_min_assignment(state.dmin32, state.label, dist, group)
    state.count = len(med)
    return state


def _finish_assign(state, K):
    state.check()  # a non-converged insert (round limit) must not pass silently
    assign = state.label.cpu().numpy().astype(int)
    assign[assign < 0] = 0  # argmin of an all-inf column is 0 (:100)
    counts = np.bincount(assign, minlength=K)
    print(f"[assign] sizes min={counts.min()}, max={counts.max()}, mean={counts.mean():.1f}")  # :104
    return assign


def assign_points_to_medoids(W: sparse.spmatrix, medoids: np.ndarray, dist_group=None) -> np.ndarray:
    """argmin_k float32(D[k,n]) with the lowest k on ties (:77-106).  `dist_group`: see _labels_for."""
    g = graph_to_device(W)
    K = len(medoids)
    if K == 0:
        raise ValueError("sources must be a non-empty sequence of node indices")
    print(f"[assign] {g.n} points to {K} medoids")  # :94
    return _finish_assign(_labels_for(g, medoids, dist_group), K)


def _qe_from_dmin(dmin32) -> float:
    total, used = dg.pairwise_sum_f32(dmin32, square=True, finite_only=True)  # :129-132
    return float(total) if used > 0 else float("inf")  # :133-134


def compute_quantization_error(W: sparse.spmatrix, medoids: np.ndarray, assign: np.ndarray, dist_group=None) -> float:
    """Sum of squared float32 geodesic distances to the assigned medoid (:109-138).

    `assign` may be any assignment, so D[assign[v], v] is gathered from true multi-source
    distances, a chunk of sources at a time.  Multi-GPU: each rank runs its shard of the sources and the
    picked distances (+inf elsewhere) are combined with an allreduce-MIN."""
    import torch
    g = graph_to_device(W)
    med = np.asarray(medoids, dtype=np.int64)
    sh = _shard_group(dist_group)
    k_lo, k_hi = 0, len(med)
    if sh is not None:
        from ..sharding import shard_range
        k_lo, k_hi = shard_range(len(med), sh[2], sh[3])
    a = torch.as_tensor(np.asarray(assign, dtype=np.int64), device=g.indptr.device)
    picked = torch.full((g.n,), float("inf"), dtype=torch.float32, device=g.indptr.device)
    if g.nnz == 0:
        hit = torch.as_tensor(med, device=a.device)[a] == torch.arange(g.n, device=a.device)
        picked[hit] = 0.0
        return _qe_from_dmin(picked)
    chunk = max(1, min(len(med), (256 << 20) // max(1, 4 * g.n)))
    cols = torch.arange(g.n, device=a.device)
    for lo in range(k_lo, k_hi, chunk):
        D = dg.sssp_multi(g, med[lo:min(lo + chunk, k_hi)])
        sel = (a >= lo) & (a < lo + D.shape[0])
        picked[sel] = D[(a[sel] - lo), cols[sel]]
    if sh is not None:
        sh[0].all_reduce(picked, op=sh[0].ReduceOp.MIN, group=sh[1])
    return _qe_from_dmin(picked)


def fit_kmedoids_optimized(W: sparse.spmatrix, K: int = 512, init: str = "kpp", seed: int = 42, dist_group=None,
                           ) -> Tuple[np.ndarray, np.ndarray, float]:
    """k-means++ (or random) seeding -> one assignment -> quantisation error (:141-183).
    Multi-GPU: the k-means++ chain is sequential and runs replicated; `init="random"` shards its K sources."""
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
        state = _labels_for(g, medoids, dist_group)
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
