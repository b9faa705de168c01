"""Geodesic shortest paths on the GPU -- drop-in for the reference's
`src/geo/geo_shortest_paths.py` (citations are lines of that file).

scipy's heap Dijkstra (fp64 path sums, one source at a time) is replaced by
`geocb_sssp_multi`: near-far frontier relaxation in fp64 whose fixed point is
bit-identical to Dijkstra's result, many sources in flight.  No CPU fallback.
"""
from __future__ import annotations

from typing import Tuple

import numpy as np
import torch
from scipy import sparse

from .. import _lib
from .. import device_graph as dg


def ensure_valid_graph(W: sparse.spmatrix) -> sparse.spmatrix:
    """:13-21 -- same checks, same exception types."""
    if not sparse.isspmatrix(W):
        raise TypeError("W must be a scipy sparse matrix")
    if W.shape[0] != W.shape[1]:
        raise ValueError("W must be square")
    if W.nnz > 0 and (W.data < 0).any():
        raise ValueError("Negative weights")
    return W.tocsr()


def graph_to_device(W, directed: bool = False, dist_group=None) -> dg.DeviceCSR:
    """Validate + upload (or pass through an already device-resident graph).

    Same checks and exception types as `ensure_valid_graph` (:13-21); the value checks run on the device next to
    the symmetry test (`geocb_csr_sym_check`) instead of as host passes over the arrays, and scipy's
    `directed=False` view (edge if either direction is stored, smaller weight wins) is merged on the device.
    `dist_group`: the upload is split across the ranks of that torch.distributed group (DeviceCSR.from_scipy_sharded)."""
    if isinstance(W, dg.DeviceCSR):
        return W
    if not sparse.isspmatrix(W):
        raise TypeError("W must be a scipy sparse matrix")
    if W.shape[0] != W.shape[1]:
        raise ValueError("W must be square")
    dev = _lib.require_cuda()
    if dist_group is not None:  # collective: every rank of the group holds the same W and makes the same call
        import torch.distributed as dist
        g = dg.DeviceCSR.from_scipy_sharded(W, dev, dist, dist_group)
    else:
        g = dg.DeviceCSR.from_scipy(W, dev)
    if g.n == 0 or g.nnz == 0:
        return g
    flags, extra = dg.csr_sym_check(g)
    if flags & 1:
        raise ValueError("Negative weights")
    if not directed and (flags & 6):
        g = dg.csr_undirected(g, extra)
    return g


def dijkstra_multi_source(W: sparse.spmatrix, sources, directed: bool = False, unweighted: bool = False,
                          return_predecessors: bool = False, dtype=np.float32) -> Tuple:
    """Multi-source geodesic distances (:24-50): (S,n) array, inf where unreachable."""
    if len(sources) == 0:
        raise ValueError("sources must be a non-empty sequence of node indices")  # :26-27
    import os
    trace = os.environ.get("GEOCB_TRACE_E2E") == "1"  # per-stage wall clock on stderr (diagnostics; adds synchronisation)
    if trace:
        import sys
        import time
        torch.cuda.synchronize()
        _t = [time.perf_counter()]

        def _mark():
            torch.cuda.synchronize()
            _t.append(time.perf_counter())
    g = graph_to_device(W, directed)
    if trace:
        _mark()
    src = np.asarray(sources, dtype=int)
    if g.n == 0:
        raise ValueError("empty graph")
    src = np.where(src < 0, src + g.n, src)  # scipy wraps negative indices once (csgraph.dijkstra)
    if (src < 0).any() or (src >= g.n).any():
        raise ValueError("source index out of range")
    use_unit = bool(unweighted) and g.nnz > 0  # :31-33
    if g.nnz == 0:
        D = np.full((len(src), g.n), np.inf, dtype=dtype)
        D[np.arange(len(src)), src] = 0
        if return_predecessors:
            return D, np.full((len(src), g.n), -9999, dtype=np.int32)
        return D
    # device->host copies of one batch of sources overlap the next batch's kernel (dg.sssp_multi_to_host); the result
    # lands in pinned memory (torch's caching host allocator recycles the block once the caller drops the array)
    if trace:
        g.adjacency(use_unit)
        _mark()
    big = len(src) * g.n * 4 >= (8 << 20)
    D = torch.empty((len(src), g.n), dtype=torch.float32, pin_memory=big)
    if trace:
        _mark()
    P = torch.empty((len(src), g.n), dtype=torch.int32, pin_memory=big) if return_predecessors else None
    g_in = None
    if return_predecessors and directed and not isinstance(W, dg.DeviceCSR):
        g_in = dg.DeviceCSR.from_scipy(W.T.tocsr(), g.indptr.device)  # predecessors are searched over in-edges
    dg.sssp_multi_to_host(g, src, D, unweighted=use_unit, pred_host=P, g_in=g_in)
    if trace:
        _mark()
        d = [1e3 * (b - a) for a, b in zip(_t[:-1], _t[1:])]
        print(f"[e2e trace] graph_to_device {d[0]:.1f} ms  adjacency {d[1]:.1f}  result alloc {d[2]:.1f}  "
              f"sssp_multi_to_host {d[3]:.1f}", file=sys.stderr, flush=True)
    if return_predecessors:
        return D.numpy().astype(dtype, copy=False), P.numpy()  # :42
    return D.numpy().astype(dtype, copy=False)  # :50


def dijkstra_single_source(W: sparse.spmatrix, source: int, directed: bool = False, unweighted: bool = False,
                           return_predecessors: bool = False, dtype=np.float32) -> Tuple:
    """Single-source wrapper returning 1-D arrays (:53-63)."""
    result = dijkstra_multi_source(W, [int(source)], directed=directed, unweighted=unweighted,
                                   return_predecessors=return_predecessors, dtype=dtype)
    if return_predecessors:
        D, P = result
        return D[0], P[0]
    return result[0]


def distances_between(W: sparse.spmatrix, sources, targets, directed: bool = False, unweighted: bool = False,
                      dtype=np.float32) -> np.ndarray:
    """Compact (S x T) distance matrix (:66-76)."""
    if len(sources) == 0 or len(targets) == 0:
        raise ValueError("sources and targets must be non-empty.")
    sources = np.asarray(sources, dtype=int)
    targets = np.asarray(targets, dtype=int)
    D_full = dijkstra_multi_source(W, sources, directed=directed, unweighted=unweighted, return_predecessors=False,
                                   dtype=dtype)
    return D_full[:, targets]
