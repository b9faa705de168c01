// graph.cu -- CSR construction around the kNN lists: symmetrise, upper-triangular edge list,
// adjacency-pair packing, connected components, LCC compaction.
// Reference behaviour replaced: scipy CSR algebra at src/geo/knn_graph_optimized.py:54-66,173-181
// and src/scripts/build_codebook.py:43-45,53-63.  All kernels are HBM-bound integer work.
#include "common.cuh"
#include "scan.cuh"

namespace {
constexpr int kT = 256;
inline unsigned blocks_for(int64_t work, int per_block = kT) {
    int64_t b = (work + per_block - 1) / per_block;
    return (unsigned)(b < 1 ? 1 : b);
}

// ---------------------------------------------------------------- symmetrise ----
// entry (i,t) of the kNN list is "valid" when it is not the diagonal (setdiag(0)) and, in
// distance mode, its weight is non-zero (eliminate_zeros).
__device__ __forceinline__ bool nbr_valid(const int32_t* idx, const float* dist, int64_t i, int t, int k) {
    int32_t j = idx[i * k + t];
    return j >= 0 && j != i && (!dist || dist[i * k + t] != 0.0f);
}
// position of i in the valid list of j, or -1
__device__ __forceinline__ int find_recip(const int32_t* idx, const float* dist, int64_t j, int64_t i, int k) {
    for (int t = 0; t < k; ++t)
        if (idx[j * k + t] == i && nbr_valid(idx, dist, j, t, k)) return t;
    return -1;
}

template <bool FILL>
__global__ void sym_kernel(const int32_t* idx, const float* dist, int64_t M, int k, int mutual, int32_t* counter,
                           const int64_t* indptr, int32_t* tmp_col, float* tmp_val) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= M * k) return;
    const int64_t i = e / k;
    const int t = (int)(e - i * k);
    if (!nbr_valid(idx, dist, i, t, k)) return;
    const int64_t j = idx[e];
    const int r = find_recip(idx, dist, j, i, k);
    const float wij = dist ? dist[e] : 1.0f;
    if (mutual) {
        if (r < 0) return;
        const int pos = atomicAdd(counter + i, 1);
        if (FILL) {
            const float wji = dist ? dist[j * k + r] : 1.0f;
            tmp_col[indptr[i] + pos] = (int32_t)j;
            tmp_val[indptr[i] + pos] = fminf(wij, wji);
        }
    } else {
        const int pos = atomicAdd(counter + i, 1);
        float w = wij;
        if (r >= 0 && dist) w = fmaxf(wij, dist[j * k + r]);
        if (FILL) {
            tmp_col[indptr[i] + pos] = (int32_t)j;
            tmp_val[indptr[i] + pos] = w;
        }
        if (r < 0) {  // one-sided edge: the transpose contributes (j,i)
            const int pos2 = atomicAdd(counter + j, 1);
            if (FILL) {
                tmp_col[indptr[j] + pos2] = (int32_t)i;
                tmp_val[indptr[j] + pos2] = w;
            }
        }
    }
}

// rank sort of every row by column (columns are unique within a row); one warp per row
__global__ void row_rank_sort(const int64_t* indptr, int64_t M, const int32_t* tmp_col, const float* tmp_val,
                              int32_t* col, float* val) {
    const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (row >= M) return;
    const int64_t b = indptr[row], e = indptr[row + 1];
    for (int64_t p = b + lane; p < e; p += 32) {
        const int32_t c = tmp_col[p];
        int rank = 0;
        for (int64_t q = b; q < e; ++q) rank += tmp_col[q] < c;
        col[b + rank] = c;
        val[b + rank] = tmp_val[p];
    }
}

// ------------------------------------------------------------ upper triangle ----
__device__ __forceinline__ int64_t first_greater(const int32_t* col, int64_t b, int64_t e, int32_t key) {
    while (b < e) {
        int64_t m = (b + e) >> 1;
        if (col[m] > key) e = m; else b = m + 1;
    }
    return b;
}
__global__ void upper_count_kernel(const int64_t* indptr, const int32_t* col, int64_t M, int32_t* cnt) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= M) return;
    cnt[i] = (int32_t)(indptr[i + 1] - first_greater(col, indptr[i], indptr[i + 1], (int32_t)i));
}
__global__ void upper_fill_kernel(const int64_t* indptr, const int32_t* col, int64_t M, const int64_t* upper_ptr,
                                  int32_t* ei, int32_t* ej, int64_t* pos_ij, int64_t* pos_ji) {
    const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (i >= M) return;
    const int64_t e = indptr[i + 1];
    const int64_t first = first_greater(col, indptr[i], e, (int32_t)i);
    const int64_t out0 = upper_ptr[i];
    for (int64_t p = first + lane; p < e; p += 32) {
        const int32_t j = col[p];
        const int64_t o = out0 + (p - first);
        ei[o] = (int32_t)i;
        ej[o] = j;
        pos_ij[o] = p;
        // locate (j,i): lower_bound of i in row j
        int64_t b = indptr[j], en = indptr[j + 1];
        const int64_t g = first_greater(col, b, en, (int32_t)i - 1);
        pos_ji[o] = (g < en && col[g] == (int32_t)i) ? g : -1;
    }
}

// --------------------------------------------------------- adjacency pairs ----
__global__ void adj_len_kernel(const int64_t* indptr, int64_t n, int32_t* len) {
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    int64_t d = indptr[v + 1] - indptr[v];
    len[v] = (int32_t)(d + (d & 1));
}
__global__ void adj_fill_kernel(const int64_t* indptr, const int32_t* col, const float* val, int64_t n, int unit,
                                const int64_t* row_ptr, uint2* adj, double* wsum) {
    const int64_t v = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    double acc = 0.0;
    if (v < n) {
        const int64_t b = indptr[v], e = indptr[v + 1], o = row_ptr[v];
        for (int64_t p = b + lane; p < e; p += 32) {
            const float w = unit ? 1.0f : val[p];
            adj[o + (p - b)] = make_uint2((uint32_t)col[p], __float_as_uint(w));
            acc += (double)w;
        }
        if (lane == 0 && ((e - b) & 1)) adj[o + (e - b)] = make_uint2((uint32_t)v, F32_INF_BITS);
    }
    if (wsum) {
        for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0 && acc != 0.0) atomicAdd(wsum, acc);
    }
}

// ---------------------------------------------------- connected components ----
__global__ void cc_init_kernel(int32_t* label, int64_t n) {
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v < n) label[v] = (int32_t)v;
}
// hook: every vertex pulls the minimum label of its neighbourhood and pushes it onto its current root
__global__ void cc_hook_kernel(const int64_t* indptr, const int32_t* col, int64_t n, int32_t* label, int* changed) {
    const int64_t u = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3;
    const int lane = threadIdx.x & 7;
    const bool live = u < n;  // no early return: every lane of the warp takes part in the shuffles
    const int32_t lu = live ? ld_cg_s32(label + u) : 0;
    int32_t m = lu;
    if (live)
        for (int64_t p = indptr[u] + lane; p < indptr[u + 1]; p += 8) m = min(m, ld_cg_s32(label + col[p]));
    for (int o = 4; o; o >>= 1) m = min(m, __shfl_xor_sync(0xffffffffu, m, o, 8));
    if (live && lane == 0 && m < lu) {
        atomicMin(label + lu, m);
        atomicMin(label + u, m);
        *changed = 1;
    }
}
__global__ void cc_compress_kernel(int32_t* label, int64_t n) {
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    int32_t l = ld_cg_s32(label + v);
    while (true) {
        int32_t ll = ld_cg_s32(label + l);
        if (ll == l) break;
        l = ll;
    }
    label[v] = l;
}

__global__ void lcc_hist_kernel(const int32_t* label, int64_t n, int32_t* hist) {
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v < n) atomicAdd(hist + label[v], 1);
}
// key = (count << 32) | ~label : max key = largest count, ties -> smallest label
__global__ void lcc_best_kernel(const int32_t* hist, int64_t n, unsigned long long* best, unsigned long long* ncomp) {
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long key = 0;
    if (v < n && hist[v] > 0) key = ((unsigned long long)(uint32_t)hist[v] << 32) | (uint32_t)(0xFFFFFFFFu - (uint32_t)v);
    unsigned long long k2 = key;
    for (int o = 16; o; o >>= 1) {
        unsigned long long t = __shfl_xor_sync(0xffffffffu, k2, o);
        k2 = t > k2 ? t : k2;
    }
    unsigned cnt = __popc(__ballot_sync(0xffffffffu, key != 0));
    if ((threadIdx.x & 31) == 0) {
        if (k2) atomicMax(best, k2);
        if (cnt) atomicAdd(ncomp, (unsigned long long)cnt);
    }
}
__global__ void lcc_mask_kernel(const int32_t* label, int64_t n, const unsigned long long* best, uint8_t* mask) {
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    const int32_t lbl = (int32_t)(0xFFFFFFFFu - (uint32_t)(*best & 0xFFFFFFFFull));
    mask[v] = label[v] == lbl;
}
__global__ void lcc_newid_kernel(const uint8_t* mask, const int64_t* pre, int64_t n, int32_t* new_id) {
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v < n) new_id[v] = mask[v] ? (int32_t)pre[v] : -1;
}

// ------------------------------------------------------------- compaction ----
__global__ void compact_len_kernel(const int64_t* indptr, const uint8_t* mask, const int32_t* new_id, int64_t n,
                                   int32_t* len) {
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v < n && mask[v]) len[new_id[v]] = (int32_t)(indptr[v + 1] - indptr[v]);
}
__global__ void compact_fill_kernel(const int64_t* indptr, const int32_t* col, const float* val, const uint8_t* mask,
                                    const int32_t* new_id, int64_t n, const int64_t* out_ptr, int32_t* out_col,
                                    float* out_val) {
    const int64_t v = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (v >= n || !mask[v]) return;
    const int64_t b = indptr[v], e = indptr[v + 1], o = out_ptr[new_id[v]];
    for (int64_t p = b + lane; p < e; p += 32) {
        out_col[o + (p - b)] = new_id[col[p]];
        out_val[o + (p - b)] = val[p];
    }
}

// ------------------------------------------- undirected view of a CSR graph ----
// scipy's directed=False semantics (csgraph.dijkstra / connected_components): an edge {i,j} exists when either
// (i,j) or (j,i) is stored and its weight is the smaller of the stored values.
__device__ __forceinline__ int64_t find_col(const int32_t* col, int64_t b, int64_t e, int32_t key) {
    const int64_t g = first_greater(col, b, e, key - 1);  // lower_bound(key)
    return (g < e && col[g] == key) ? g : -1;
}
// flags: 1 = negative weight, 2 = some (i,j) without a stored (j,i), 4 = (i,j) and (j,i) stored with different weights;
// extra[j] counts the entries row j gains.  One warp per row.
__global__ void csr_sym_check_kernel(const int64_t* indptr, const int32_t* col, const float* val, int64_t n,
                                     int32_t* extra, int* flags) {
    const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (i >= n) return;
    int f = 0;
    for (int64_t p = indptr[i] + lane; p < indptr[i + 1]; p += 32) {
        const int32_t j = col[p];
        const float w = val[p];
        if (w < 0.0f) f |= 1;
        const int64_t q = find_col(col, indptr[j], indptr[j + 1], (int32_t)i);
        if (q < 0) {
            f |= 2;
            atomicAdd(extra + j, 1);
        } else if (val[q] != w) {
            f |= 4;
        }
    }
    f = __reduce_or_sync(0xffffffffu, f);
    if (lane == 0 && f) atomicOr(flags, f);
}
__global__ void undirected_len_kernel(const int64_t* indptr, const int32_t* extra, int64_t n, int32_t* len) {
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v < n) len[v] = (int32_t)(indptr[v + 1] - indptr[v]) + extra[v];
}
__global__ void undirected_fill_kernel(const int64_t* indptr, const int32_t* col, const float* val, int64_t n,
                                       const int64_t* out_ptr, int32_t* cursor, int32_t* tmp_col, float* tmp_val) {
    const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (i >= n) return;
    const int64_t b = indptr[i], o = out_ptr[i];
    for (int64_t p = b + lane; p < indptr[i + 1]; p += 32) {
        const int32_t j = col[p];
        float w = val[p];
        const int64_t q = find_col(col, indptr[j], indptr[j + 1], (int32_t)i);
        if (q >= 0) {
            w = fminf(w, val[q]);
        } else {  // row j gains (j,i): appended behind its own entries, sorted into place afterwards
            const int64_t slot = out_ptr[j] + (indptr[j + 1] - indptr[j]) + atomicAdd(cursor + j, 1);
            tmp_col[slot] = (int32_t)i;
            tmp_val[slot] = w;
        }
        tmp_col[o + (p - b)] = j;
        tmp_val[o + (p - b)] = w;
    }
}

struct TmpBuf {  // stream-ordered scratch
    void* p = nullptr;
    cudaStream_t st;
    explicit TmpBuf(cudaStream_t s) : st(s) {}
    int alloc(size_t bytes) {
        CU_TRY(cudaMallocAsync(&p, bytes ? bytes : 8, st));
        return 0;
    }
    ~TmpBuf() {
        if (p) cudaFreeAsync(p, st);
    }
};
int read_i64(const int64_t* dev, int64_t* host, cudaStream_t st) {
    CU_TRY(cudaMemcpyAsync(host, dev, 8, cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaStreamSynchronize(st));
    return 0;
}
}  // namespace

GEOCB_API int geocb_sym_count(const int32_t* nbr_idx, const float* nbr_dist, int64_t M, int k, int mode,
                              int64_t* indptr, int64_t* nnz_host, void* stream) {
    ARG_CHECK(nbr_idx && indptr && nnz_host && M > 0 && k > 0 && (mode == 0 || mode == 1));
    cudaStream_t st = (cudaStream_t)stream;
    TmpBuf cnt(st);
    int rc = cnt.alloc((size_t)M * 4);
    if (rc) return rc;
    CU_TRY(cudaMemsetAsync(cnt.p, 0, (size_t)M * 4, st));
    sym_kernel<false><<<blocks_for(M * k), kT, 0, st>>>(nbr_idx, nbr_dist, M, k, mode, (int32_t*)cnt.p, nullptr,
                                                         nullptr, nullptr);
    KERNEL_CHECK();
    rc = scan_exclusive_i32((int32_t*)cnt.p, M, indptr, st);
    if (rc) return rc;
    return read_i64(indptr + M, nnz_host, st);
}

GEOCB_API int geocb_sym_fill(const int32_t* nbr_idx, const float* nbr_dist, int64_t M, int k, int mode,
                             const int64_t* indptr, int32_t* indices, float* data, void* stream) {
    ARG_CHECK(nbr_idx && indptr && indices && data && M > 0 && k > 0 && (mode == 0 || mode == 1));
    cudaStream_t st = (cudaStream_t)stream;
    int64_t nnz = 0;
    int rc = read_i64(indptr + M, &nnz, st);
    if (rc) return rc;
    TmpBuf cur(st), tcol(st), tval(st);
    if ((rc = cur.alloc((size_t)M * 4)) || (rc = tcol.alloc((size_t)nnz * 4)) || (rc = tval.alloc((size_t)nnz * 4)))
        return rc;
    CU_TRY(cudaMemsetAsync(cur.p, 0, (size_t)M * 4, st));
    sym_kernel<true><<<blocks_for(M * k), kT, 0, st>>>(nbr_idx, nbr_dist, M, k, mode, (int32_t*)cur.p, indptr,
                                                        (int32_t*)tcol.p, (float*)tval.p);
    row_rank_sort<<<blocks_for(M * 32), kT, 0, st>>>(indptr, M, (int32_t*)tcol.p, (float*)tval.p, indices, data);
    KERNEL_CHECK();
    return 0;
}

GEOCB_API int geocb_upper_count(const int64_t* indptr, const int32_t* indices, int64_t M, int64_t* upper_ptr,
                                int64_t* E_host, void* stream) {
    ARG_CHECK(indptr && indices && upper_ptr && E_host && M > 0);
    cudaStream_t st = (cudaStream_t)stream;
    TmpBuf cnt(st);
    int rc = cnt.alloc((size_t)M * 4);
    if (rc) return rc;
    upper_count_kernel<<<blocks_for(M), kT, 0, st>>>(indptr, indices, M, (int32_t*)cnt.p);
    KERNEL_CHECK();
    if ((rc = scan_exclusive_i32((int32_t*)cnt.p, M, upper_ptr, st))) return rc;
    return read_i64(upper_ptr + M, E_host, st);
}

GEOCB_API int geocb_upper_fill(const int64_t* indptr, const int32_t* indices, int64_t M, const int64_t* upper_ptr,
                               int32_t* ei, int32_t* ej, int64_t* pos_ij, int64_t* pos_ji, void* stream) {
    ARG_CHECK(indptr && indices && upper_ptr && ei && ej && pos_ij && pos_ji && M > 0);
    upper_fill_kernel<<<blocks_for(M * 32), kT, 0, (cudaStream_t)stream>>>(indptr, indices, M, upper_ptr, ei, ej,
                                                                            pos_ij, pos_ji);
    KERNEL_CHECK();
    return 0;
}

GEOCB_API int geocb_adj_count(const int64_t* indptr, int64_t n, int64_t* row_ptr, int64_t* n_pairs_host,
                              void* stream) {
    ARG_CHECK(indptr && row_ptr && n_pairs_host && n > 0);
    cudaStream_t st = (cudaStream_t)stream;
    TmpBuf len(st);
    int rc = len.alloc((size_t)n * 4);
    if (rc) return rc;
    adj_len_kernel<<<blocks_for(n), kT, 0, st>>>(indptr, n, (int32_t*)len.p);
    KERNEL_CHECK();
    if ((rc = scan_exclusive_i32((int32_t*)len.p, n, row_ptr, st))) return rc;
    if ((rc = read_i64(row_ptr + n, n_pairs_host, st))) return rc;
    if (*n_pairs_host >= (1LL << 32)) {  // the SSSP kernels address (col, weight) pairs with 32-bit offsets
        geocb_set_error("graph has %lld adjacency pairs (limit 2^32 - 1)", (long long)*n_pairs_host);
        return GEOCB_ERR_UNSUPPORTED;
    }
    return 0;
}

GEOCB_API int geocb_adj_fill(const int64_t* indptr, const int32_t* indices, const float* data, int64_t n,
                             int unit_weights, const int64_t* row_ptr, void* adj, double* weight_sum, void* stream) {
    ARG_CHECK(indptr && indices && (data || unit_weights) && row_ptr && adj && n > 0);
    cudaStream_t st = (cudaStream_t)stream;
    if (weight_sum) CU_TRY(cudaMemsetAsync(weight_sum, 0, 8, st));
    adj_fill_kernel<<<blocks_for(n * 32), kT, 0, st>>>(indptr, indices, data, n, unit_weights, row_ptr, (uint2*)adj,
                                                        weight_sum);
    KERNEL_CHECK();
    return 0;
}

GEOCB_API int geocb_cc_labels(const int64_t* indptr, const int32_t* indices, int64_t n, int32_t* labels,
                              void* stream) {
    ARG_CHECK(indptr && indices && labels && n > 0);
    cudaStream_t st = (cudaStream_t)stream;
    TmpBuf flag(st);
    int rc = flag.alloc(4);
    if (rc) return rc;
    cc_init_kernel<<<blocks_for(n), kT, 0, st>>>(labels, n);
    for (int it = 0; it < 10000; ++it) {
        int changed = 0;
        CU_TRY(cudaMemsetAsync(flag.p, 0, 4, st));
        cc_hook_kernel<<<blocks_for(n * 8), kT, 0, st>>>(indptr, indices, n, labels, (int*)flag.p);
        cc_compress_kernel<<<blocks_for(n), kT, 0, st>>>(labels, n);
        KERNEL_CHECK();
        CU_TRY(cudaMemcpyAsync(&changed, flag.p, 4, cudaMemcpyDeviceToHost, st));
        CU_TRY(cudaStreamSynchronize(st));
        if (!changed) return 0;
    }
    geocb_set_error("connected components did not converge");
    return GEOCB_ERR_OVERFLOW;
}

GEOCB_API int geocb_lcc_mask(const int32_t* labels, int64_t n, uint8_t* mask, int32_t* new_id, int64_t* n_keep_host,
                             int64_t* n_comp_host, void* stream) {
    ARG_CHECK(labels && mask && new_id && n_keep_host && n_comp_host && n > 0);
    cudaStream_t st = (cudaStream_t)stream;
    TmpBuf hist(st), best(st), pre(st);
    int rc;
    if ((rc = hist.alloc((size_t)n * 4)) || (rc = best.alloc(16)) || (rc = pre.alloc((size_t)(n + 1) * 8))) return rc;
    CU_TRY(cudaMemsetAsync(hist.p, 0, (size_t)n * 4, st));
    CU_TRY(cudaMemsetAsync(best.p, 0, 16, st));
    unsigned long long* b = (unsigned long long*)best.p;
    lcc_hist_kernel<<<blocks_for(n), kT, 0, st>>>(labels, n, (int32_t*)hist.p);
    lcc_best_kernel<<<blocks_for(n), kT, 0, st>>>((int32_t*)hist.p, n, b, b + 1);
    lcc_mask_kernel<<<blocks_for(n), kT, 0, st>>>(labels, n, b, mask);
    KERNEL_CHECK();
    if ((rc = scan_exclusive_u8(mask, n, (int64_t*)pre.p, st))) return rc;
    lcc_newid_kernel<<<blocks_for(n), kT, 0, st>>>(mask, (int64_t*)pre.p, n, new_id);
    KERNEL_CHECK();
    unsigned long long hb[2];
    CU_TRY(cudaMemcpyAsync(hb, b, 16, cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaStreamSynchronize(st));
    *n_keep_host = (int64_t)(hb[0] >> 32);
    *n_comp_host = (int64_t)hb[1];
    return 0;
}

GEOCB_API int geocb_csr_compact_count(const int64_t* indptr, const uint8_t* mask, const int32_t* new_id, int64_t n,
                                      int64_t* out_indptr, int64_t* nnz_host, void* stream) {
    ARG_CHECK(indptr && mask && new_id && out_indptr && nnz_host && n > 0);
    cudaStream_t st = (cudaStream_t)stream;
    TmpBuf len(st), pre(st);
    int rc;
    if ((rc = len.alloc((size_t)n * 4)) || (rc = pre.alloc((size_t)(n + 1) * 8))) return rc;
    // n_keep = popcount(mask)
    if ((rc = scan_exclusive_u8(mask, n, (int64_t*)pre.p, st))) return rc;
    int64_t n_keep = 0;
    if ((rc = read_i64((int64_t*)pre.p + n, &n_keep, st))) return rc;
    ARG_CHECK(n_keep > 0);
    compact_len_kernel<<<blocks_for(n), kT, 0, st>>>(indptr, mask, new_id, n, (int32_t*)len.p);
    KERNEL_CHECK();
    if ((rc = scan_exclusive_i32((int32_t*)len.p, n_keep, out_indptr, st))) return rc;
    return read_i64(out_indptr + n_keep, nnz_host, st);
}

GEOCB_API int geocb_csr_compact_fill(const int64_t* indptr, const int32_t* indices, const float* data,
                                     const uint8_t* mask, const int32_t* new_id, int64_t n, const int64_t* out_indptr,
                                     int32_t* out_indices, float* out_data, void* stream) {
    ARG_CHECK(indptr && indices && data && mask && new_id && out_indptr && out_indices && out_data && n > 0);
    compact_fill_kernel<<<blocks_for(n * 32), kT, 0, (cudaStream_t)stream>>>(indptr, indices, data, mask, new_id, n,
                                                                              out_indptr, out_indices, out_data);
    KERNEL_CHECK();
    return 0;
}

GEOCB_API int geocb_csr_sym_check(const int64_t* indptr, const int32_t* indices, const float* data, int64_t n,
                                  int32_t* extra, int* flags_host, void* stream) {
    ARG_CHECK(indptr && indices && data && extra && flags_host && n > 0);
    cudaStream_t st = (cudaStream_t)stream;
    TmpBuf flags(st);
    int rc = flags.alloc(4);
    if (rc) return rc;
    CU_TRY(cudaMemsetAsync(flags.p, 0, 4, st));
    CU_TRY(cudaMemsetAsync(extra, 0, (size_t)n * 4, st));
    csr_sym_check_kernel<<<blocks_for(n * 32), kT, 0, st>>>(indptr, indices, data, n, extra, (int*)flags.p);
    KERNEL_CHECK();
    CU_TRY(cudaMemcpyAsync(flags_host, flags.p, 4, cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaStreamSynchronize(st));
    return 0;
}

GEOCB_API int geocb_csr_undirected_count(const int64_t* indptr, const int32_t* extra, int64_t n, int64_t* out_indptr,
                                         int64_t* nnz_host, void* stream) {
    ARG_CHECK(indptr && extra && out_indptr && nnz_host && n > 0);
    cudaStream_t st = (cudaStream_t)stream;
    TmpBuf len(st);
    int rc = len.alloc((size_t)n * 4);
    if (rc) return rc;
    undirected_len_kernel<<<blocks_for(n), kT, 0, st>>>(indptr, extra, n, (int32_t*)len.p);
    KERNEL_CHECK();
    if ((rc = scan_exclusive_i32((int32_t*)len.p, n, out_indptr, st))) return rc;
    return read_i64(out_indptr + n, nnz_host, st);
}

GEOCB_API int geocb_csr_undirected_fill(const int64_t* indptr, const int32_t* indices, const float* data, int64_t n,
                                        const int64_t* out_indptr, int32_t* out_indices, float* out_data,
                                        void* stream) {
    ARG_CHECK(indptr && indices && data && out_indptr && out_indices && out_data && n > 0);
    cudaStream_t st = (cudaStream_t)stream;
    int64_t nnz = 0;
    int rc = read_i64(out_indptr + n, &nnz, st);
    if (rc) return rc;
    TmpBuf cur(st), tcol(st), tval(st);
    if ((rc = cur.alloc((size_t)n * 4)) || (rc = tcol.alloc((size_t)nnz * 4)) || (rc = tval.alloc((size_t)nnz * 4)))
        return rc;
    CU_TRY(cudaMemsetAsync(cur.p, 0, (size_t)n * 4, st));
    undirected_fill_kernel<<<blocks_for(n * 32), kT, 0, st>>>(indptr, indices, data, n, out_indptr, (int32_t*)cur.p,
                                                               (int32_t*)tcol.p, (float*)tval.p);
    row_rank_sort<<<blocks_for(n * 32), kT, 0, st>>>(out_indptr, n, (int32_t*)tcol.p, (float*)tval.p, out_indices,
                                                      out_data);
    KERNEL_CHECK();
    return 0;
}
