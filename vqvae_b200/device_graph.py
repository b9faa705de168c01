"""Device-resident graph objects and thin wrappers over the C ABI (include/geocb.h).

Everything here works on torch CUDA tensors; the reference-facing modules in
`vqvae_b200.geo` convert host arrays <-> these objects at the API boundary.
"""
from __future__ import annotations

import os
import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import check, ptr, stream_ptr


def _i64():
    return C.c_int64(0)


# ------------------------------------------------------------ host <-> device staging
_STAGE_CHUNK = 32 << 20
_stage_bufs = {}


_copy_pool = None


def _host_copy(dst: torch.Tensor, src: torch.Tensor, parts=4):
    """dst.copy_(src) for byte tensors on the host, split over a few Python threads (Tensor.copy_ releases the GIL):
    under torchrun OMP_NUM_THREADS is 1, and one core moves only ~8 GB/s into pinned memory."""
    global _copy_pool
    n = dst.numel()
    if n < (4 << 20) or torch.get_num_threads() >= parts:  # torch's own copy is already parallel
        dst.copy_(src)
        return
    if _copy_pool is None:
        from concurrent.futures import ThreadPoolExecutor
        _copy_pool = ThreadPoolExecutor(max_workers=parts)
    step = -(-n // parts)
    futs = [_copy_pool.submit(dst[o:o + step].copy_, src[o:o + step]) for o in range(0, n, step)]
    for f in futs:
        f.result()


def h2d(arr: np.ndarray, dev) -> torch.Tensor:
    """Pageable NumPy array -> device tensor through two pinned staging buffers: the (multi-threaded) host copy of
    chunk i+1 into pinned memory overlaps the DMA of chunk i, so large arrays move at PCIe speed instead of the
    driver's pageable-copy speed.  Small arrays take the plain path."""
    arr = np.ascontiguousarray(arr)
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")  # read-only NumPy buffers (scipy views): we only read
        src = torch.from_numpy(arr)
    if src.numel() * src.element_size() < (8 << 20):
        return src.to(dev)
    out = torch.empty(src.shape, dtype=src.dtype, device=dev)
    sb, db = src.view(torch.uint8).reshape(-1), out.view(torch.uint8).reshape(-1)
    key = torch.device(dev).index
    if key not in _stage_bufs:
        # (buffers, events): the events outlive the call -- the next upload must not overwrite a staging buffer whose
        # DMA from the previous upload is still in flight
        _stage_bufs[key] = ([torch.empty(_STAGE_CHUNK, dtype=torch.uint8).pin_memory() for _ in range(2)], [None, None])
    bufs, evs = _stage_bufs[key]
    for i, off in enumerate(range(0, sb.numel(), _STAGE_CHUNK)):
        m = min(_STAGE_CHUNK, sb.numel() - off)
        if evs[i & 1] is not None:
            evs[i & 1].synchronize()  # the DMA out of this staging buffer has finished
        _host_copy(bufs[i & 1][:m], sb[off:off + m])
        db[off:off + m].copy_(bufs[i & 1][:m], non_blocking=True)
        evs[i & 1] = torch.cuda.Event()
        evs[i & 1].record()
    return out


class DeviceCSR:
    """CSR matrix in HBM: indptr int64[n+1], indices int32[nnz] (ascending per row), data f32[nnz]."""

    def __init__(self, indptr, indices, data, n):
        self.indptr, self.indices, self.data, self.n = indptr, indices, data, int(n)
        self._adj = {}

    @property
    def nnz(self):
        return int(self.indices.numel())

    @classmethod
    def from_scipy(cls, W, device):
        W = W.tocsr()
        if not W.has_sorted_indices:
            W = W.sorted_indices()
        return cls(h2d(W.indptr.astype(np.int64, copy=False), device), h2d(W.indices.astype(np.int32, copy=False), device),
                   h2d(W.data.astype(np.float32, copy=False), device), W.shape[0])

    @classmethod
    def from_scipy_sharded(cls, W, device, dist, group=None):
        """Collective upload of the SAME host matrix held by every rank of `group`: each rank stages 1/world of the
        indices / data arrays through its own PCIe link and the pieces are all-gathered over NVLink (NCCL), so the
        host->device time of a large graph divides by the number of GPUs instead of being paid by every rank."""
        W = W.tocsr()
        if not W.has_sorted_indices:
            W = W.sorted_indices()
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        nnz = int(W.nnz)
        chunk = max(1, -(-nnz // world))
        lo, hi = min(nnz, rank * chunk), min(nnz, (rank + 1) * chunk)
        parts = []
        for arr, dt, tdt in ((W.indices, np.int32, torch.int32), (W.data, np.float32, torch.float32)):
            mine = torch.zeros(chunk, dtype=tdt, device=device)
            if hi > lo:
                mine[: hi - lo] = h2d(arr[lo:hi].astype(dt, copy=False), device)
            full = torch.empty(world * chunk, dtype=tdt, device=device)
            dist.all_gather_into_tensor(full, mine, group=group)
            parts.append(full[:nnz])
        return cls(h2d(W.indptr.astype(np.int64, copy=False), device), parts[0], parts[1], W.shape[0])

    def to_scipy(self):
        from scipy import sparse
        ip = self.indptr.cpu().numpy()
        ip = ip.astype(np.int32) if self.nnz < 2**31 - 1 else ip
        return sparse.csr_matrix((self.data.cpu().numpy(), self.indices.cpu().numpy(), ip), shape=(self.n, self.n))

    def eliminate_zeros(self):
        """Drop stored entries whose value is exactly 0 (what scipy's `A + A.T` does to the reference's W_geo,
        build_codebook.py:53-54: an edge of Riemannian length 0 -- duplicate latents -- vanishes from the graph).
        Returns self when nothing is stored as zero (the common case: one pass over the data)."""
        keep = self.data != 0
        if bool(keep.all()):
            return self
        rows = torch.repeat_interleave(torch.arange(self.n, device=self.data.device), self.indptr[1:] - self.indptr[:-1])
        counts = torch.zeros(self.n, dtype=torch.int64, device=self.data.device).index_add_(0, rows[keep], torch.ones_like(rows[keep]))
        indptr = torch.zeros(self.n + 1, dtype=torch.int64, device=self.data.device)
        indptr[1:] = torch.cumsum(counts, 0)
        return DeviceCSR(indptr, self.indices[keep].contiguous(), self.data[keep].contiguous(), self.n)

    # -- adjacency-pair form used by the SSSP kernels (rows padded to even length, 16 B aligned) --
    def adjacency(self, unit_weights=False):
        key = bool(unit_weights)
        if key not in self._adj:
            lib = _lib.load()
            dev = self.indptr.device
            row_ptr = torch.empty(self.n + 1, dtype=torch.int64, device=dev)
            npairs = _i64()
            check(lib.geocb_adj_count(ptr(self.indptr), self.n, ptr(row_ptr), C.byref(npairs), stream_ptr()), "adj_count")
            adj = torch.empty(max(npairs.value, 2), dtype=torch.int64, device=dev)  # 8-byte (col,w) pairs
            wsum = torch.zeros(1, dtype=torch.float64, device=dev)
            check(lib.geocb_adj_fill(ptr(self.indptr), ptr(self.indices), ptr(self.data), self.n, int(unit_weights),
                                     ptr(row_ptr), ptr(adj), ptr(wsum), stream_ptr()), "adj_fill")
            mean_w = float(wsum.item()) / max(self.nnz, 1)
            self._adj[key] = (row_ptr, adj, mean_w)
        return self._adj[key]

    def vm_blocks(self, B, unit_weights=False):
        """Destination-block entry offsets blk[(B+1)][n] of the adjacency rows (geocb_sssp_vm_blocks), cached."""
        key = ("blk", bool(unit_weights), int(B))
        if key not in self._adj:
            lib = _lib.load()
            row_ptr, adj, _ = self.adjacency(unit_weights)
            blk = torch.empty((B + 1) * self.n, dtype=torch.int32, device=row_ptr.device)
            check(lib.geocb_sssp_vm_blocks(ptr(row_ptr), ptr(adj), self.n, int(B), ptr(blk), stream_ptr()), "sssp_vm_blocks")
            self._adj[key] = blk
        return self._adj[key]


def csr_sym_check(g: DeviceCSR):
    """(flags, extra): flags 1 = negative weight, 2 = structurally asymmetric, 4 = weights differ across the diagonal."""
    lib = _lib.load()
    extra = torch.empty(g.n, dtype=torch.int32, device=g.indptr.device)
    flags = C.c_int(0)
    check(lib.geocb_csr_sym_check(ptr(g.indptr), ptr(g.indices), ptr(g.data), g.n, ptr(extra), C.byref(flags),
                                  stream_ptr()), "csr_sym_check")
    return flags.value, extra


def csr_undirected(g: DeviceCSR, extra):
    """scipy's directed=False view: edge {i,j} if either direction is stored, weight = the smaller stored value."""
    lib = _lib.load()
    dev = g.indptr.device
    indptr = torch.empty(g.n + 1, dtype=torch.int64, device=dev)
    nnz = _i64()
    check(lib.geocb_csr_undirected_count(ptr(g.indptr), ptr(extra), g.n, ptr(indptr), C.byref(nnz), stream_ptr()),
          "csr_undirected_count")
    indices = torch.empty(nnz.value, dtype=torch.int32, device=dev)
    data = torch.empty(nnz.value, dtype=torch.float32, device=dev)
    check(lib.geocb_csr_undirected_fill(ptr(g.indptr), ptr(g.indices), ptr(g.data), g.n, ptr(indptr), ptr(indices),
                                        ptr(data), stream_ptr()), "csr_undirected_fill")
    return DeviceCSR(indptr, indices, data, g.n)


# --------------------------------------------------------------------------- kNN
def knn_topk(z, kq, q_lo=0, q_hi=None, want_dist=True):
    """kq nearest rows of z (self included) for query rows [q_lo,q_hi): (idx int32, dist f32|None, n_exact_rows)."""
    lib = _lib.load()
    assert z.is_cuda and z.dtype == torch.float32 and z.is_contiguous()
    M, d = z.shape
    q_hi = M if q_hi is None else q_hi
    nq = q_hi - q_lo
    idx = torch.empty((nq, kq), dtype=torch.int32, device=z.device)
    dist = torch.empty((nq, kq), dtype=torch.float32, device=z.device) if want_dist else None
    nun = torch.zeros(1, dtype=torch.int32, device=z.device)
    check(lib.geocb_knn_topk(ptr(z), M, d, q_lo, q_hi, kq, ptr(idx), ptr(dist), ptr(nun), stream_ptr()), "knn_topk")
    return idx, dist, nun


def knn_query(queries, cand, kq=1, want_dist=True):
    """kq nearest rows of cand[M,d] for every row of queries[nq,d] (exact fp64 order, ties -> lowest index):
    (idx int32 [nq,kq], dist f32 | None, n_exact_rows)."""
    lib = _lib.load()
    assert queries.is_cuda and cand.is_cuda and queries.dtype == cand.dtype == torch.float32
    assert queries.is_contiguous() and cand.is_contiguous() and queries.shape[1] == cand.shape[1]
    nq, d = queries.shape
    idx = torch.empty((nq, kq), dtype=torch.int32, device=queries.device)
    dist = torch.empty((nq, kq), dtype=torch.float32, device=queries.device) if want_dist else None
    nun = torch.zeros(1, dtype=torch.int32, device=queries.device)
    check(lib.geocb_knn_query(ptr(queries), nq, ptr(cand), cand.shape[0], d, kq, ptr(idx), ptr(dist), ptr(nun),
                              stream_ptr()), "knn_query")
    return idx, dist, nun


def symmetrize(nbr_idx, nbr_dist, mode):
    """kNN lists (self dropped) -> symmetric DeviceCSR.  mode 'union' | 'mutual'; nbr_dist None => weights 1."""
    lib = _lib.load()
    M, k = nbr_idx.shape
    dev = nbr_idx.device
    m = {"union": 0, "mutual": 1}[mode]
    indptr = torch.empty(M + 1, dtype=torch.int64, device=dev)
    nnz = _i64()
    check(lib.geocb_sym_count(ptr(nbr_idx), ptr(nbr_dist), M, k, m, ptr(indptr), C.byref(nnz), stream_ptr()), "sym_count")
    indices = torch.empty(nnz.value, dtype=torch.int32, device=dev)
    data = torch.empty(nnz.value, dtype=torch.float32, device=dev)
    if nnz.value:
        check(lib.geocb_sym_fill(ptr(nbr_idx), ptr(nbr_dist), M, k, m, ptr(indptr), ptr(indices), ptr(data),
                                 stream_ptr()), "sym_fill")
    return DeviceCSR(indptr, indices, data, M)


def upper_edges(g: DeviceCSR):
    """Upper-triangular entries in row-major order: (ei, ej int32[E], pos_ij, pos_ji int64[E])."""
    lib = _lib.load()
    dev = g.indptr.device
    up = torch.empty(g.n + 1, dtype=torch.int64, device=dev)
    E = _i64()
    check(lib.geocb_upper_count(ptr(g.indptr), ptr(g.indices), g.n, ptr(up), C.byref(E), stream_ptr()), "upper_count")
    ei = torch.empty(E.value, dtype=torch.int32, device=dev)
    ej = torch.empty(E.value, dtype=torch.int32, device=dev)
    pij = torch.empty(E.value, dtype=torch.int64, device=dev)
    pji = torch.empty(E.value, dtype=torch.int64, device=dev)
    if E.value:
        check(lib.geocb_upper_fill(ptr(g.indptr), ptr(g.indices), g.n, ptr(up), ptr(ei), ptr(ej), ptr(pij), ptr(pji),
                                   stream_ptr()), "upper_fill")
    return ei, ej, pij, pji


def edge_len_scatter(ja, jb, e_lo, pos_ij, pos_ji, len_out, data):
    lib = _lib.load()
    E, P = ja.shape
    check(lib.geocb_edge_len_scatter(ptr(ja), ptr(jb), E, P, e_lo, ptr(pos_ij), ptr(pos_ji), ptr(len_out), ptr(data),
                                     stream_ptr()), "edge_len_scatter")


def bn_relu_jvp(x, t, C_, HW, rows_per_group, pre_bias, weight, bias, eps, relu, image=None):
    """Fused BatchNorm(train) JVP (+ReLU) on primal x / tangent t, both [rows, C*HW] contiguous f32: in place, or --
    with image=(aimg, tangent_row_offset) -- written as the operand image of the tcgen05 GEMM that follows.
    Returns False when the shape is outside what the kernel supports (caller falls back to library ops)."""
    lib = _lib.load()
    rows = x.shape[0]
    if not (HW in (1, 2, 4, 8, 16) and (C_ * HW) % 16 == 0 and rows % rows_per_group == 0 and rows_per_group <= 512
            and rows // rows_per_group <= 65535 and x.is_contiguous() and t.is_contiguous()):
        return False
    if image is not None:
        aimg, row_t = image
        check(lib.geocb_bn_relu_jvp_to_image(ptr(x), ptr(t), rows, C_, HW, rows_per_group, ptr(pre_bias), ptr(weight),
                                             ptr(bias), float(eps), int(relu), ptr(aimg), int(row_t), stream_ptr()),
              "bn_relu_jvp_to_image")
        return True
    check(lib.geocb_bn_relu_jvp(ptr(x), ptr(t), rows, C_, HW, rows_per_group, ptr(pre_bias), ptr(weight), ptr(bias),
                                float(eps), int(relu), stream_ptr()), "bn_relu_jvp")
    return True


def lin_bn_relu_jvp_to_image(z, dz, w, C_, HW, rows_per_group, pre_bias, weight, bias, eps, relu, image):
    """z @ w (+pre_bias) -> BatchNorm(train) JVP (+ReLU) -> operand image of the next tcgen05 GEMM, in one kernel
    (geocb_lin_bn_relu_jvp_to_image).  z / dz [rows, kin] primal / tangent inputs, w [kin, C*HW].  Returns False when
    the shape is outside what the kernel supports."""
    lib = _lib.load()
    rows, kin = z.shape
    if not (kin in (16, 32) and HW in (1, 2, 4, 8, 16) and (C_ * HW) % 16 == 0 and rows % rows_per_group == 0
            and rows_per_group <= 512 and rows // rows_per_group <= 65535 and rows % 128 == 0 and z.is_contiguous()
            and dz.is_contiguous() and w.is_contiguous() and tuple(w.shape) == (kin, C_ * HW)):
        return False
    aimg, row_t = image
    check(lib.geocb_lin_bn_relu_jvp_to_image(ptr(z), ptr(dz), kin, ptr(w), rows, C_, HW, rows_per_group, ptr(pre_bias),
                                             ptr(weight), ptr(bias), float(eps), int(relu), ptr(aimg), int(row_t),
                                             stream_ptr()), "lin_bn_relu_jvp_to_image")
    return True


def lin_sigmoid_jvp(x, t, w, bias):
    """s (1 - s) (t @ w) with s = sigmoid(x @ w + bias) in one kernel (geocb_lin_sigmoid_jvp); None when the shape is
    outside what the kernel supports (P = 16 outputs, K a multiple of 128)."""
    K, P = w.shape
    if not (P == 16 and K % 128 == 0 and x.is_contiguous() and t.is_contiguous() and w.is_contiguous()
            and x.shape == t.shape and x.shape[1] == K and x.dtype == torch.float32):
        return None
    out = torch.empty((x.shape[0], P), dtype=torch.float32, device=x.device)
    check(_lib.load().geocb_lin_sigmoid_jvp(ptr(x), ptr(t), x.shape[0], K, P, ptr(w), ptr(bias), ptr(out), stream_ptr()),
          "lin_sigmoid_jvp")
    return out


_gemm3x_weights = {}


def gemm3x_supported(M, K, N):
    return bool(_lib.load().geocb_gemm3x_supported(M, K, N))


def _gemm3x_weight_image(w):
    lib = _lib.load()
    K, N = w.shape
    key = (w.data_ptr(), K, N)
    if key not in _gemm3x_weights:
        bimg = torch.empty(2 * K * N, dtype=torch.float32, device=w.device)
        check(lib.geocb_gemm3x_prep_weights(ptr(w), K, N, ptr(bimg), stream_ptr()), "gemm3x_prep_weights")
        _gemm3x_weights[key] = (bimg, w)  # keep w alive so the pointer stays unique
    return _gemm3x_weights[key][0]


def gemm3x_image_buffer(M, K, device):
    """Scratch operand image (tf32 hi | lo tiles) for an [M,K] activation matrix, cached per device."""
    return _workspace("gemm3x_a", 2 * M * K * 4, device)


def gemm3x(a, w: torch.Tensor, M=None):
    """a[M,K] @ w[K,N] through the tcgen05 3xTF32 GEMM (fp32-grade), or None when the shape is unsupported.
    `a` is the row-major activation matrix, or None when gemm3x_image_buffer(M, K) already holds its operand image
    (written by bn_relu_jvp(..., image=...)).  The weight image is built once per weight tensor."""
    lib = _lib.load()
    K, N = w.shape
    M = a.shape[0] if a is not None else M
    if not (w.is_contiguous() and w.dtype == torch.float32 and lib.geocb_gemm3x_supported(M, K, N)):
        return None
    if a is not None and not (a.is_contiguous() and a.dtype == torch.float32 and a.shape[1] == K):
        return None
    bimg = _gemm3x_weight_image(w)
    aimg = gemm3x_image_buffer(M, K, w.device)
    c = torch.empty((M, N), dtype=torch.float32, device=w.device)
    check(lib.geocb_gemm3x(ptr(a), M, K, N, ptr(bimg), ptr(aimg), ptr(c), stream_ptr()), "gemm3x")
    return c


# ------------------------------------------------------------ components / LCC
def component_labels(g: DeviceCSR):
    lib = _lib.load()
    labels = torch.empty(g.n, dtype=torch.int32, device=g.indptr.device)
    check(lib.geocb_cc_labels(ptr(g.indptr), ptr(g.indices), g.n, ptr(labels), stream_ptr()), "cc_labels")
    return labels


def largest_component(g: DeviceCSR):
    """(mask uint8[n], new_id int32[n], n_keep, n_components)."""
    lib = _lib.load()
    labels = component_labels(g)
    dev = g.indptr.device
    mask = torch.empty(g.n, dtype=torch.uint8, device=dev)
    new_id = torch.empty(g.n, dtype=torch.int32, device=dev)
    nk, nc = _i64(), _i64()
    check(lib.geocb_lcc_mask(ptr(labels), g.n, ptr(mask), ptr(new_id), C.byref(nk), C.byref(nc), stream_ptr()), "lcc_mask")
    return mask, new_id, nk.value, nc.value


def compact(g: DeviceCSR, mask, new_id, n_keep):
    """W[mask][:,mask] where mask is a union of whole components."""
    lib = _lib.load()
    dev = g.indptr.device
    out_ptr = torch.empty(n_keep + 1, dtype=torch.int64, device=dev)
    nnz = _i64()
    check(lib.geocb_csr_compact_count(ptr(g.indptr), ptr(mask), ptr(new_id), g.n, ptr(out_ptr), C.byref(nnz),
                                      stream_ptr()), "csr_compact_count")
    oi = torch.empty(nnz.value, dtype=torch.int32, device=dev)
    od = torch.empty(nnz.value, dtype=torch.float32, device=dev)
    if nnz.value:
        check(lib.geocb_csr_compact_fill(ptr(g.indptr), ptr(g.indices), ptr(g.data), ptr(mask), ptr(new_id), g.n,
                                         ptr(out_ptr), ptr(oi), ptr(od), stream_ptr()), "csr_compact_fill")
    return DeviceCSR(out_ptr, oi, od, n_keep)


# ------------------------------------------------------------------------ SSSP
_ws_cache = {}


def _workspace(tag, nbytes, device, zero=False):
    key = (tag, device)
    buf = _ws_cache.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.zeros(nbytes, dtype=torch.uint8, device=device)
        _ws_cache[key] = buf
    elif zero:
        buf.zero_()
    return buf


def default_slots(n, S=None):
    """Live distance tables per launch.  Measured on B200 at n=2.94M: throughput rises up to 16 tables and is nearly
    flat beyond (profiles/r1_sssp_variant_sweep_delta4.txt); 64 sources streamed through 32 tables take 47.2 ms against
    49.5 ms through 16 (half the rounds, profiles/r2_sssp_streamed_sweep.txt), so calls that say how many sources they
    bring (S >= 32) get 32.  Every table costs 20.4 B per vertex of workspace, so very large graphs are capped by what
    is free on the device (a quarter of it at most)."""
    slots = 32 if (S is not None and S >= 32) else 16
    try:
        free, _ = torch.cuda.mem_get_info()
        slots = int(max(1, min(slots, (free // 4) // max(1, 25 * n))))
    except Exception:
        pass
    return slots


def _delta_for(g, mean_w, unweighted):
    """Bucket width = scale x mean edge weight.  Any width gives the same bits; 4x is best on the C2 geodesic graph
    (1x: 110, 2x: 117, 4x: 120, 6x: 117 G edge-relax/s), but on graphs with a wide weight spread it re-relaxes too much
    (hash-weight proxy: 1.85 relaxations per edge at 4x against 1.13 at 1x).  The kernel returns its relaxation count,
    so the scale adapts per graph: halved while a call spends more than 1.35 relaxations per edge."""
    env = os.environ.get("GEOCB_SSSP_DELTA_SCALE")
    scale = float(env) if env else g._adj.get(("delta_scale", bool(unweighted)), 4.0)
    return (mean_w if mean_w > 0 else 1.0) * scale


def _delta_feedback(g, unweighted, relax, S):
    if os.environ.get("GEOCB_SSSP_DELTA_SCALE") or S <= 0 or g.nnz <= 0:
        return
    key = ("delta_scale", bool(unweighted))
    scale = g._adj.get(key, 4.0)
    if relax / (S * float(g.nnz)) > 1.35 and scale > 0.5:
        g._adj[key] = scale / 2


def _stagger_for(g, unweighted, S, slots):
    """geocb_sssp_multi's stagger_rounds (rounds over which the first seeding of the `slots` tables is spread).
    Default 0: on the C2 graph spreading the sources' phases LOSES (64 sources, 32 tables: 47.2 ms at 0, 48.7 at 8,
    50.0 at 12, 51.2 at 16; profiles/r2_sssp_streamed_sweep.txt) -- a round in which every table is in its wide
    phase runs at 170-180 G relaxations/s, a mixed round at ~120 G/s, so lock-step phases are the better schedule.
    GEOCB_SSSP_STAGGER overrides; `auto` uses the measured rounds per source of the previous call on this graph."""
    env = os.environ.get("GEOCB_SSSP_STAGGER")
    if env is None or S <= slots:
        return 0
    if env == "auto":
        return int(g._adj.get(("rounds_per_source", bool(unweighted)), 12))
    return max(0, int(env))


def vm_default_blocks(n):
    """Destination blocks of the vertex-major kernel: 200 B of randomly accessed state per vertex (dist 128 + ub 64 +
    masks 8), one block's share kept near 40 MB so that it stays L2 resident next to the streamed data."""
    return int(os.environ.get("GEOCB_SSSP_VM_BLOCKS", max(1, min(64, -(-n * 200 // (40 << 20))))))


def _pick_kernel(kernel, S, want_pred):
    """'slot' = csrc/sssp.cu (one table per source), 'vm' = csrc/sssp_vm.cu (16 sources side by side per vertex),
    'vm32' = csrc/sssp_vm32.cu (32 sources per vertex, bf16 bound table).
    'auto' = 'slot': on the C2 graph the vertex-major kernel shares an edge visit among only 3.8 of its 16 sources
    and measures 66 G edge-relax/s against 125 G/s (profiles/r2_sssp_vm_*.txt); it stays selectable for graphs whose
    sources overlap more.  Predecessors need 'slot'."""
    kernel = kernel or os.environ.get("GEOCB_SSSP_KERNEL", "auto")
    if kernel not in ("auto", "slot", "vm", "vm32"):
        raise ValueError(f"unknown SSSP kernel '{kernel}'")
    if want_pred:
        return "slot"
    return "slot" if kernel == "auto" else kernel


def sssp_multi_vm(g: DeviceCSR, sources, unweighted=False, delta=None, out=None, stats=False, blocks=None,
                  blk_min_active=None, wide=False):
    """S independent SSSPs through the vertex-major kernel -> D32[S,n] on the device.
    wide: the 32-source / bf16-bound generation (csrc/sssp_vm32.cu) instead of the 16-source one."""
    lib = _lib.load()
    dev = g.indptr.device
    row_ptr, adj, mean_w = g.adjacency(unweighted)
    B = vm_default_blocks(g.n) if blocks is None else int(blocks)
    blk = g.vm_blocks(B, unweighted)
    src = torch.as_tensor(np.asarray(sources, dtype=np.int32), device=dev)
    S = int(src.numel())
    if delta is None:
        delta = (mean_w if mean_w > 0 else 1.0) * float(os.environ.get("GEOCB_SSSP_VM_DELTA_SCALE", "4"))
    if blk_min_active is None:
        # a blocked round scans the n mask words B times; worth it once the popped rows' gathers would thrash L2
        blk_min_active = int(os.environ.get("GEOCB_SSSP_VM_MIN_ACTIVE", max(1, g.n // 16)))
    if wide:
        ws = _workspace("sssp_vm32", lib.geocb_sssp_vm32_workspace_bytes(g.n), dev)
        fn, name = lib.geocb_sssp_multi_vm32, "sssp_multi_vm32"
    else:
        ws = _workspace("sssp_vm", lib.geocb_sssp_vm_workspace_bytes(g.n), dev)
        fn, name = lib.geocb_sssp_multi_vm, "sssp_multi_vm"
    D = out if out is not None else torch.empty((S, g.n), dtype=torch.float32, device=dev)
    st = (C.c_int64 * 8)()
    check(fn(ptr(adj), ptr(blk), B, g.n, ptr(src), S, float(delta), int(blk_min_active), ptr(ws),
             ws.numel(), ptr(D), st, stream_ptr()), name)
    if stats:
        return D, dict(relaxations=st[0], rounds=st[1], scans=st[2], improvements=st[3], edge_visits=st[4],
                       blocked_rounds=st[5], slots=32 if wide else lib.geocb_sssp_vm_slots(), delta=delta, blocks=B,
                       kernel="vm32" if wide else "vm")
    return D


def sssp_multi(g: DeviceCSR, sources, unweighted=False, want_pred=False, delta=None, slots=None, out=None,
               stats=False, kernel=None, g_in: DeviceCSR | None = None, stagger=None):
    """S independent SSSPs -> D32[S,n] (+pred int32[S,n]) on the device.
    g_in: transpose of a directed (asymmetric) g -- predecessors are searched over in-edges."""
    lib = _lib.load()
    dev = g.indptr.device
    row_ptr, adj, mean_w = g.adjacency(unweighted)
    src = torch.as_tensor(np.asarray(sources, dtype=np.int32), device=dev)
    S = int(src.numel())
    picked = _pick_kernel(kernel, S, want_pred)
    if picked in ("vm", "vm32"):
        return sssp_multi_vm(g, sources, unweighted=unweighted, delta=delta, out=out, stats=stats, wide=picked == "vm32")
    auto_delta = delta is None
    if auto_delta:
        delta = _delta_for(g, mean_w, unweighted)
    slots = default_slots(g.n, None if want_pred else S) if slots is None else slots
    slots = max(1, min(slots, S, 32))
    nbytes = lib.geocb_sssp_workspace_bytes(g.n, slots)
    ws = _workspace("sssp", nbytes, dev)
    D = out if out is not None else torch.empty((S, g.n), dtype=torch.float32, device=dev)
    P = torch.empty((S, g.n), dtype=torch.int32, device=dev) if want_pred else None
    st = (C.c_int64 * 5)()
    in_ptr = in_adj = None
    if want_pred and g_in is not None:
        in_ptr, in_adj, _ = g_in.adjacency(unweighted)
    if stagger is None:
        stagger = 0 if want_pred else _stagger_for(g, unweighted, S, slots)
    check(lib.geocb_sssp_multi(ptr(row_ptr), ptr(adj), g.n, ptr(src), S, float(delta), slots, int(stagger), ptr(ws),
                               ws.numel(), ptr(D), ptr(P), ptr(in_ptr), ptr(in_adj), st, stream_ptr()), "sssp_multi")
    if auto_delta:
        _delta_feedback(g, unweighted, st[0], S)
    if st[4] > 0:
        g._adj[("rounds_per_source", bool(unweighted))] = max(1, min(4096, round(st[4] / S)))
    res = (D, P) if want_pred else D
    if stats:
        return res, dict(relaxations=st[0], rounds=st[1], scans=st[2], improvements=st[3], slots=slots, delta=delta,
                         kernel="slot", stagger=int(stagger), rounds_per_source=st[4] / S)
    return res


def sssp_multi_to_host(g: DeviceCSR, sources, out_host: torch.Tensor, unweighted=False, delta=None, slots=None,
                       pred_host: torch.Tensor | None = None, kernel=None, g_in: DeviceCSR | None = None):
    """S independent SSSPs whose (S,n) float32 result (and int32 predecessors) land in HOST tensors.

    The sources run in lock-step batches of `slots`; the device->host copy of batch i is issued on a side stream and
    overlaps the kernel of batch i+1 (two device staging buffers).  `out_host` should be pinned for the copy to be
    asynchronous; pageable memory works but serialises.  Returns out_host (and pred_host)."""
    dev = g.indptr.device
    src = np.asarray(sources, dtype=np.int32)
    S = int(src.size)
    slots = default_slots(g.n) if slots is None else slots
    slots = max(1, min(slots, S, 32))
    picked = _pick_kernel(kernel, S, pred_host is not None)
    if picked in ("vm", "vm32"):  # one launch per batch: the copy-out overlaps the next launch
        slots = min(S, 32 if picked == "vm32" else _lib.load().geocb_sssp_vm_slots())
    assert out_host.shape == (S, g.n) and out_host.dtype == torch.float32 and out_host.device.type == "cpu"
    main = torch.cuda.current_stream(dev)
    side = _side_stream(dev)
    bufs = [torch.empty((slots, g.n), dtype=torch.float32, device=dev) for _ in range(2)]
    free = [None, None]  # event: the copy out of bufs[i] has finished
    for bi, b0 in enumerate(range(0, S, slots)):
        b1 = min(S, b0 + slots)
        buf = bufs[bi & 1][:b1 - b0]
        if free[bi & 1] is not None:
            main.wait_event(free[bi & 1])
        res = sssp_multi(g, src[b0:b1], unweighted=unweighted, want_pred=pred_host is not None, delta=delta,
                         slots=slots, out=buf, kernel=kernel, g_in=g_in)
        side.wait_stream(main)
        with torch.cuda.stream(side):
            out_host[b0:b1].copy_(buf, non_blocking=True)
            if pred_host is not None:
                pred_host[b0:b1].copy_(res[1], non_blocking=True)
                res[1].record_stream(side)
            ev = torch.cuda.Event()
            ev.record(side)
        free[bi & 1] = ev
    side.synchronize()
    return (out_host, pred_host) if pred_host is not None else out_host


_side_streams = {}


def _side_stream(dev):
    key = torch.device(dev).index if torch.device(dev).index is not None else torch.cuda.current_device()
    if key not in _side_streams:
        _side_streams[key] = torch.cuda.Stream(device=dev)
    return _side_streams[key]


class VoronoiState:
    """Running (dmin64, dmin32, label) over the inserted centres -- see geocb_voronoi_insert."""

    def __init__(self, g: DeviceCSR, unweighted=False):
        lib = _lib.load()
        self.g = g
        dev = g.indptr.device
        self.row_ptr, self.adj, self.mean_w = g.adjacency(unweighted)
        n = g.n
        self.dmin64 = torch.empty(n, dtype=torch.float64, device=dev)
        self.dmin32 = torch.empty(n, dtype=torch.float32, device=dev)
        self.label = torch.empty(n, dtype=torch.int32, device=dev)
        self.is_center = torch.zeros(n, dtype=torch.uint8, device=dev)
        self.ws = torch.empty(lib.geocb_voronoi_workspace_bytes(n), dtype=torch.uint8, device=dev)
        self.kpp_ws = None
        self.plan = None
        self.count = 0
        check(lib.geocb_voronoi_init(n, ptr(self.dmin64), ptr(self.dmin32), ptr(self.label), ptr(self.ws),
                                     self.ws.numel(), stream_ptr()), "voronoi_init")

    def insert(self, center, code=None):
        lib = _lib.load()
        code = self.count if code is None else code
        # near-far only pays off while cells are large; later cells are tiny -> plain frontier relaxation
        delta = self.mean_w if (self.count < 8 and self.mean_w > 0) else float("inf")
        check(lib.geocb_voronoi_insert(ptr(self.row_ptr), ptr(self.adj), self.g.n, int(center), int(code), delta,
                                       ptr(self.dmin64), ptr(self.dmin32), ptr(self.label), ptr(self.ws),
                                       self.ws.numel(), stream_ptr()), "voronoi_insert")
        self.is_center[int(center)] = 1
        self.count += 1

    def check(self):
        """Inserts are asynchronous and their 'hit the round limit' flag is sticky: callers check once, after the
        chain / before the labels are used (kmeans_optimized._seed_kpp, _finish_assign), not per draw."""
        check(_lib.load().geocb_voronoi_check(self.g.n, ptr(self.ws), self.ws.numel(), stream_ptr()), "voronoi_insert")

    def draw(self, u):
        """One k-means++ draw with the host's MT19937 double u -> (index or -1, used_exact_path, probs_sum)."""
        lib = _lib.load()
        n = self.g.n
        dev = self.dmin32.device
        if self.plan is None:
            words = lib.geocb_pairwise_plan_size(n)
            host = (C.c_int64 * words)()
            check(lib.geocb_pairwise_plan_host(n, host), "pairwise_plan_host")
            self.plan = torch.from_numpy(np.frombuffer(host, dtype=np.int64).copy()).to(dev)
            self.kpp_ws = torch.empty(lib.geocb_kpp_workspace_bytes(n), dtype=torch.uint8, device=dev)
        res = (C.c_int64 * 4)()
        check(lib.geocb_kpp_draw(ptr(self.dmin32), ptr(self.is_center), n, float(u), ptr(self.plan), ptr(self.kpp_ws),
                                 self.kpp_ws.numel(), res, stream_ptr()), "kpp_draw")
        psum = np.array([res[2]], dtype=np.uint32).view(np.float32)[0]
        return int(res[0]), int(res[1]), float(psum)


def pairwise_sum_f32(x, square=False, finite_only=False):
    """NumPy-exact float32 np.sum(x) / np.sum(x[isfinite]**2) -> (float32 value, n_used)."""
    lib = _lib.load()
    out = C.c_float(0.0)
    used = _i64()
    check(lib.geocb_pairwise_sum_f32(ptr(x), x.numel(), int(square), int(finite_only), None, 0, C.byref(out),
                                     C.byref(used), stream_ptr()), "pairwise_sum_f32")
    return np.float32(out.value), used.value
