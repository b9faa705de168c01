/* geocb.h -- C ABI of libgeocb.so: the sm_100a kernels behind the geodesic
 * codebook builder (drop-in for m4rch1n0/vqvae `src/geo/*` + `build_codebook.py`).
 *
 * The reference has no FFI: its boundary is Python functions over host arrays
 * (SURVEY.md §8b).  Each entry point below replaces the third-party CPU call a
 * reference function makes; the reference call site is cited per function.
 * INTEGRATION.md shows the ctypes stub a reference maintainer would add.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer unless its name ends in `_host`;
 *  - `stream` is a cudaStream_t passed as void*; calls are stream-ordered and
 *    do not synchronise unless stated ("syncs");
 *  - return value: 0 = ok, otherwise a negative geocb code or a positive
 *    cudaError_t; geocb_last_error() gives the text (thread-local);
 *  - no CPU fallback exists: every function fails if no sm_100 device is usable.
 */
#ifndef GEOCB_H
#define GEOCB_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define GEOCB_ABI_VERSION 1
#define GEOCB_ERR_ARG (-1)      /* bad argument */
#define GEOCB_ERR_UNSUPPORTED (-2)
#define GEOCB_ERR_WORKSPACE (-3) /* workspace too small */
#define GEOCB_ERR_OVERFLOW (-4)

int geocb_abi_version(void);
const char* geocb_last_error(void);
/* number of SMs, compute capability major*10+minor of the current device (syncs) */
int geocb_device_info(int* sm_count, int* cc);

/* ---------------------------------------------------------------- kNN ----
 * replaces sklearn NearestNeighbors.fit/kneighbors (knn_graph_optimized.py:39-42).
 * For query rows [q_lo,q_hi) of z[M,d] (row-major f32) find the kq nearest rows
 * of z (self included), ordered by (exact fp64 squared distance, index).
 *   pass 1: filter  |b|^2 - 2ab  over all candidates, keeping the best kq+slack per row:
 *           d <= 64 and kq <= 26: tcgen05 tensor cores, operands split into bf16 (default) or tf32
 *           hi + lo pieces (fp32-grade scores, csrc/knn_tc.cu; env GEOCB_KNN_TC_FMT=tf32 selects the tf32
 *           pieces, GEOCB_KNN_TC=0 disables the tensor-core filter); otherwise fused fp32 FFMA2 tiles (d <= 64, kq <= 56)
 *           or, beyond that, the exact fp64 kernel for every row.
 *   pass 2: exact fp64 re-rank of the candidates; rows whose filter margin is
 *           not provably safe are redone by an exact fp64 brute-force kernel.
 *   The result never depends on which filter ran.
 * out_idx[(q_hi-q_lo),kq] int32, out_dist same shape f32 = sqrt(fp64 d2) (nullable).
 * n_unsafe (device int32[1], nullable): rows that took the exact brute-force path. */
int geocb_knn_topk(const float* z, int64_t M, int d, int64_t q_lo, int64_t q_hi, int kq,
                   int32_t* out_idx, float* out_dist, int32_t* n_unsafe, void* stream);

/* Same search with SEPARATE query and candidate sets: queries[nq,d] against cand[M,d] (no self to drop).
 * kq = 1 is the downstream "nearest medoid" assignment of held-out latents to the K codebook entries
 * (src/eval/evaluate_codebook_health.py:13-26, evaluate_quantization_loss.py:37-50: d2.argmin(dim=1)), which the
 * reference ranks in float32 (a2 + b2 - 2ab); here the order is by exact fp64 squared distance, ties -> lowest index,
 * so results differ from the reference only where its float32 scores tie or cross within rounding. */
int geocb_knn_query(const float* queries, int64_t nq, const float* cand, int64_t M, int d, int kq,
                    int32_t* out_idx, float* out_dist, int32_t* n_unsafe, void* stream);

/* ------------------------------------------------- symmetrise -> CSR -----
 * replaces scipy csr_matrix(...) ; W.maximum/minimum(W.T) ; setdiag(0) ;
 * eliminate_zeros  (knn_graph_optimized.py:54-66).
 * nbr_idx[M,k] (self already dropped), nbr_dist[M,k] or NULL (connectivity: data=1).
 * mode: 0 = union (maximum), 1 = mutual (minimum).
 * Two calls: _count fills indptr[M+1] (int64) and returns nnz in *nnz_host (syncs);
 * _fill writes indices[nnz] (int32, ascending per row) and data[nnz] (f32). */
int geocb_sym_count(const int32_t* nbr_idx, const float* nbr_dist, int64_t M, int k, int mode,
                    int64_t* indptr, int64_t* nnz_host, void* stream);
int geocb_sym_fill(const int32_t* nbr_idx, const float* nbr_dist, int64_t M, int k, int mode,
                   const int64_t* indptr, int32_t* indices, float* data, void* stream);

/* -------------------------------------------------- edge enumeration -----
 * replaces W.nonzero(); keep rows<cols (build_codebook.py:43-45): upper-triangular
 * entries in row-major order.  _count fills upper_ptr[M+1] (exclusive scan of the
 * per-row count of cols>row) and returns E (syncs); _fill writes ei,ej[E] and the
 * CSR positions of (i,j) and (j,i). */
int geocb_upper_count(const int64_t* indptr, const int32_t* indices, int64_t M, int64_t* upper_ptr,
                      int64_t* E_host, void* stream);
int geocb_upper_fill(const int64_t* indptr, const int32_t* indices, int64_t M, const int64_t* upper_ptr,
                     int32_t* ei, int32_t* ej, int64_t* pos_ij, int64_t* pos_ji, void* stream);

/* ------------------------------------------- Riemannian edge reduce ------
 * replaces vector_norm(dim=1) x2, 0.5*(a+b) (riemannian_metric.py:60-63) and the
 * csr_matrix((L,(i,j))) + transpose of build_codebook.py:53-54.
 * ja, jb [E,P] f32 = J(z_i)dz, J(z_j)dz for edges [e_lo, e_lo+E) ; writes
 * len_out[e] (nullable) and data[pos_ij[e]] = data[pos_ji[e]] = L (pos nullable). */
int geocb_edge_len_scatter(const float* ja, const float* jb, int64_t E, int P, int64_t e_lo,
                           const int64_t* pos_ij, const int64_t* pos_ji, float* len_out, float* data,
                           void* stream);

/* BatchNorm2d(train) forward-mode JVP fused with the following ReLU, in place (part of the JVP stack behind
 * riemannian_metric.py:12-35; the reference CLI keeps the decoder in train mode, build_codebook.py:27-29).
 * x, t [rows][C*HW] f32 primal / tangent; statistics per group of rows_per_group consecutive rows (one reference
 * batch of edges) and channel; pre_bias[C*HW] (nullable) is added to x first (bias of the preceding linear map);
 * weight/bias [C] nullable (no affine); relu != 0 applies max(y,0) and masks the tangent.
 * Supported: HW in {1,2,4,8,16}, (C*HW) % 16 == 0, rows % rows_per_group == 0, rows_per_group <= 512;
 * otherwise GEOCB_ERR_UNSUPPORTED (the caller then composes the same formulas from library ops). */
int geocb_bn_relu_jvp(float* x, float* t, int64_t rows, int C, int HW, int rows_per_group, const float* pre_bias,
                      const float* weight, const float* bias, float eps, int relu, void* stream);

/* fp32-grade GEMM on the tcgen05 tensor cores for the decoder-JVP layers (csrc/jvp_gemm.cu; replaces the cuBLAS SGEMM
 * behind the 512 -> 1024 layer of the unrolled decoder): C[M,N] = A[M,K] @ W[K,N], every operand split into tf32
 * hi + lo parts, three MMAs per K-step (hi.hi + lo.hi + hi.lo), fp32 accumulation in TMEM.
 * geocb_gemm3x_prep_weights writes the weight image bimg[2*K*N] once per matrix; geocb_gemm3x needs a scratch image
 * aimg[2*M*K] and shapes with M % 256 == 0, K % 16 == 0, N % 256 == 0 (geocb_gemm3x_supported), else
 * GEOCB_ERR_UNSUPPORTED and the caller keeps the library GEMM. */
/* geocb_bn_relu_jvp_to_image: same arithmetic as geocb_bn_relu_jvp, but the result is written as the tf32 hi | lo
 * operand image of the GEMM that follows (primal rows at [0, rows), tangent rows at [img_row_t, ...) of its A matrix;
 * rows and img_row_t multiples of 128, K = C*HW) instead of in place; geocb_gemm3x then takes A = NULL. */
int geocb_bn_relu_jvp_to_image(const float* x, const float* t, int64_t rows, int C, int HW, int rows_per_group,
                               const float* pre_bias, const float* weight, const float* bias, float eps, int relu,
                               float* aimg, int64_t img_row_t, void* stream);
/* geocb_lin_bn_relu_jvp_to_image: the linear map in FRONT of that BatchNorm fused in as well (the decoder's first
 * layer on the 1x1 latent grid, riemannian_metric.py:12-35): zin / tin [rows][kin] are the primal / tangent latents,
 * w_in [kin][C*HW] row-major the dense operator, kin = 16 or 32 (else GEOCB_ERR_UNSUPPORTED).  Pre-activations are kin
 * sequential FMAs (k ascending) + pre_bias: the [rows][C*HW] matrix is never written or read. */
int geocb_lin_bn_relu_jvp_to_image(const float* zin, const float* tin, int kin, const float* w_in, int64_t rows, int C,
                                   int HW, int rows_per_group, const float* pre_bias, const float* weight,
                                   const float* bias, float eps, int relu, float* aimg, int64_t img_row_t, void* stream);
/* geocb_lin_sigmoid_jvp: the decoder's LAST linear map fused with the sigmoid and its tangent
 * (riemannian_metric.py:29-35): out[r][p] = s (1 - s) (t[r,:] . w[:,p]), s = sigmoid(x[r,:] . w[:,p] + bias[p]).
 * x / t [rows][K] primal / tangent activations, w [K][P] row-major, bias [P] or NULL, out [rows][P].
 * P = 16 and K a multiple of 128 (else GEOCB_ERR_UNSUPPORTED: the caller keeps the library GEMM). */
int geocb_lin_sigmoid_jvp(const float* x, const float* t, int64_t rows, int K, int P, const float* w, const float* bias,
                          float* out, void* stream);
int geocb_gemm3x_supported(int64_t M, int K, int N);
int geocb_gemm3x_prep_weights(const float* W, int K, int N, float* bimg, void* stream);
int geocb_gemm3x(const float* A, int64_t M, int K, int N, const float* bimg, float* aimg, float* C, void* stream);

/* --------------------------------------------- connected components ------
 * replaces scipy.sparse.csgraph.connected_components (knn_graph_optimized.py:175).
 * labels[n] = smallest vertex id of the component (undirected: the CSR must be
 * structurally symmetric).  Syncs (iterates to a fixed point). */
int geocb_cc_labels(const int64_t* indptr, const int32_t* indices, int64_t n, int32_t* labels, void* stream);
/* largest component (ties -> smallest label, = np.argmax(np.bincount)); mask[n] u8,
 * new_id[n] int32 (-1 outside); *n_keep_host, *n_comp_host returned (syncs). */
int geocb_lcc_mask(const int32_t* labels, int64_t n, uint8_t* mask, int32_t* new_id, int64_t* n_keep_host,
                   int64_t* n_comp_host, void* stream);
/* W[mask][:,mask] for a union of whole components: rows copied, columns relabelled.
 * out_indptr[n_keep+1] must be pre-sized; returns nnz_out in *nnz_host (syncs) from _count. */
int geocb_csr_compact_count(const int64_t* indptr, const uint8_t* mask, const int32_t* new_id, int64_t n,
                            int64_t* out_indptr, int64_t* nnz_host, void* stream);
int geocb_csr_compact_fill(const int64_t* indptr, const int32_t* indices, const float* data, const uint8_t* mask,
                           const int32_t* new_id, int64_t n, const int64_t* out_indptr, int32_t* out_indices,
                           float* out_data, void* stream);

/* ------------------------------------------ undirected view of a CSR ------
 * replaces the host-side W.T / maximum / minimum passes that scipy's `directed=False` implies
 * (csgraph.dijkstra at geo_shortest_paths.py:36-49, connected_components at knn_graph_optimized.py:175):
 * an edge {i,j} exists when (i,j) or (j,i) is stored; its weight is the smaller stored value.
 * geocb_csr_sym_check (rows sorted by column, no duplicates) returns in *flags_host (syncs):
 *   1 = a negative weight (the reference raises ValueError, geo_shortest_paths.py:18-19),
 *   2 = some (i,j) has no stored (j,i),  4 = (i,j) and (j,i) differ in weight,
 * and fills extra[n] (int32) with the number of entries each row gains.  When flags & 6,
 * _count (syncs) sizes and _fill writes the merged, column-sorted CSR. */
int geocb_csr_sym_check(const int64_t* indptr, const int32_t* indices, const float* data, int64_t n, int32_t* extra,
                        int* flags_host, void* stream);
int geocb_csr_undirected_count(const int64_t* indptr, const int32_t* extra, int64_t n, int64_t* out_indptr,
                               int64_t* nnz_host, void* stream);
int geocb_csr_undirected_fill(const int64_t* indptr, const int32_t* indices, const float* data, int64_t n,
                              const int64_t* out_indptr, int32_t* out_indices, float* out_data, void* stream);

/* ----------------------------------------------------------- SSSP --------
 * Graph in "adjacency-pair" form: adj[row_ptr[v] .. row_ptr[v+1]) are 8-byte
 * (int32 col, f32 weight) pairs; every row is padded to an even count so rows
 * start 16-byte aligned (pad entries: col = v, weight = +inf).
 * geocb_adj_count / _fill build it from CSR (unit_weights != 0 -> all weights 1,
 * the reference's `unweighted=True`, geo_shortest_paths.py:31-33). */
/* (GEOCB_ERR_UNSUPPORTED if the graph has >= 2^32 pairs: the kernels use 32-bit pair offsets) */
int geocb_adj_count(const int64_t* indptr, int64_t n, int64_t* row_ptr, int64_t* n_pairs_host, void* stream);
int geocb_adj_fill(const int64_t* indptr, const int32_t* indices, const float* data, int64_t n, int unit_weights,
                   const int64_t* row_ptr, void* adj, double* weight_sum /* device f64[1], nullable */, void* stream);

/* workspace bytes for geocb_sssp_multi with `slots` concurrent sources */
int64_t geocb_sssp_workspace_bytes(int64_t n, int slots);

/* replaces scipy.sparse.csgraph.dijkstra(indices=sources) + astype(float32)
 * (geo_shortest_paths.py:36-49).  S independent single-source problems solved by a
 * near-far (delta-stepping) frontier relaxation that accumulates path sums in fp64
 * exactly like scipy (monotone rounded addition => same fixed point as Dijkstra),
 * through `slots` (1..32) live distance tables; results rounded to f32 at the end.
 * Without predecessors all S sources run in ONE cooperative launch: a table whose source has converged writes its
 * row out and takes the next source at once, and table i is first seeded in round i*stagger_rounds/slots (0 = all
 * at once), so the narrow first and last rounds of one source overlap the wide middle rounds of the others
 * (callers pass the measured rounds per source, stats[4]/S of an earlier call).  With predecessors: batches of
 * `slots` sources, one launch each.
 * Per relaxation the kernel gathers only a float32 upper bound of dist[v] (exact filter),
 * frontiers are bitmaps swept in vertex order (see csrc/sssp.cu).
 *   out_d32[S,n] f32 (+inf unreachable); out_pred[S,n] int32 or NULL
 *   (predecessor = lowest-index u with an edge u->v and fl(d[u]+w)==d[v]; -9999 for the
 *   source and unreachable vertices, scipy's convention).  Predecessors are found over v's
 *   IN-edges: in_row_ptr / in_adj = adjacency pairs of the transposed graph for a directed
 *   (asymmetric) graph, both NULL for a symmetric one.
 * delta > 0 is the bucket width (callers pass 4x the mean edge weight; any value gives the same result).  Syncs.  stats_host (nullable):
 * int64[5] = {edge relaxations, rounds, threshold scans, improvements, sum over sources of the rounds each was live}.
 * Returns GEOCB_ERR_OVERFLOW if the kernel hit its round limit (8n+4096) without converging. */
int geocb_sssp_multi(const int64_t* row_ptr, const void* adj, int64_t n, const int32_t* sources, int S,
                     double delta, int slots, int stagger_rounds, void* workspace, int64_t workspace_bytes,
                     float* out_d32, int32_t* out_pred, const int64_t* in_row_ptr, const void* in_adj,
                     int64_t* stats_host, void* stream);

/* Source-vectorised variant of geocb_sssp_multi (csrc/sssp_vm.cu): same results bit for bit, different layout.
 * The geocb_sssp_vm_slots() (=16) sources of a launch sit side by side per vertex (dist[v][16], ub[v][16], one
 * frontier mask word per vertex); lane&15 = source, a half-warp = one edge, so an edge pair is loaded once and one
 * coalesced 64-byte gather serves all 16 sources.  The vertex range is cut into B destination blocks so that the
 * randomly accessed state of one block stays L2 resident: geocb_sssp_vm_blocks fills blk[(B+1)][n] (uint32 pair offset
 * where row v enters block b; needs < 2^32 adjacency pairs, else GEOCB_ERR_UNSUPPORTED; syncs) once per graph.
 * Rounds that pop fewer than blk_min_active vertices run unblocked.  No predecessors (use geocb_sssp_multi).
 * stats_host (nullable): int64[8] = {relaxations, rounds, scans, improvements, edge visits, blocked rounds, -, -}. */
int geocb_sssp_vm_slots(void);
int64_t geocb_sssp_vm_workspace_bytes(int64_t n);
int geocb_sssp_vm_blocks(const int64_t* row_ptr, const void* adj, int64_t n, int B, uint32_t* blk, void* stream);
int geocb_sssp_multi_vm(const void* adj, const uint32_t* blk, int B, int64_t n, const int32_t* sources, int S,
                        double delta, int64_t blk_min_active, void* workspace, int64_t workspace_bytes,
                        float* out_d32, int64_t* stats_host, void* stream);
/* Second generation (csrc/sssp_vm32.cu): 32 sources per launch, the gathered bound table is bf16 (ub[v][32] = 64 B:
 * an exact filter, ub = bf16 round-up of dist, a candidate is dropped iff its bf16 round-down >= ub), four lanes per
 * edge with 8 sources each.  Same blk table, same results bit for bit, same stats layout; n <= 2^26. */
int64_t geocb_sssp_vm32_workspace_bytes(int64_t n);
int geocb_sssp_multi_vm32(const void* adj, const uint32_t* blk, int B, int64_t n, const int32_t* sources, int S,
                          double delta, int64_t blk_min_active, void* workspace, int64_t workspace_bytes,
                          float* out_d32, int64_t* stats_host, void* stream);

/* ------------------------------------------- incremental Voronoi SSSP ----
 * The K-medoids path (kmeans_optimized.py:40-44, 97-100, 125-134) only ever uses
 * d_min[v] = min_k D[k,v], argmin_k and sum d_min^2, so instead of K full SSSPs the
 * state {dmin64, dmin32, label} is grown one centre at a time: inserting centre c
 * with code k relaxes outwards from c and stops wherever the new distance is not
 * smaller (exact in fp64 by monotonicity of rounded addition), then
 *   dmin32[v] = f32(dmin64[v]);  label[v] = k  iff f32 value strictly decreased
 * which is the reference's  argmin over k of float32(D[k,v]) with lowest-k ties.
 * geocb_voronoi_init sets dmin64=+inf, dmin32=+inf, label=-1 and initialises the workspace, which must then be
 * passed unchanged to every geocb_voronoi_insert of that state (it holds the kernel's per-vertex bookkeeping). */
int64_t geocb_voronoi_workspace_bytes(int64_t n);
int geocb_voronoi_init(int64_t n, double* dmin64, float* dmin32, int32_t* label, void* workspace,
                       int64_t workspace_bytes, void* stream);
int geocb_voronoi_insert(const int64_t* row_ptr, const void* adj, int64_t n, int32_t center, int32_t code,
                         double delta, double* dmin64, float* dmin32, int32_t* label, void* workspace,
                         int64_t workspace_bytes, void* stream);
/* inserts are asynchronous; this reads back whether any of them hit the round limit (8n+4096) without
 * converging (GEOCB_ERR_OVERFLOW).  Syncs. */
int geocb_voronoi_check(int64_t n, void* workspace, int64_t workspace_bytes, void* stream);

/* ------------------------------------------------ k-means++ sampling -----
 * replaces kmeans_optimized.py:47-61: inf handling, probs = d_min_safe**2,
 * probs[centers]=0, probs.sum() (NumPy float32 pairwise tree, reproduced exactly),
 * probs/=sum, and RandomState.choice's  cdf=cumsum(float64(p)); cdf/=cdf[-1];
 * searchsorted(u,'right').  `u` is the MT19937 double drawn by the host.
 * is_center[n] u8.  plan = device copy of the pairwise-sum plan built by
 * geocb_pairwise_plan_host (host, int64 triples).  Returns in result_host[4] (syncs):
 *   [0] chosen index (or -1 if probs.sum()==0 -> caller takes the reference's uniform fallback)
 *   [1] 1 if the exact sequential-cumsum kernel had to be used, else 0
 *   [2] float bits of probs.sum() before normalisation   [3] reserved */
int64_t geocb_pairwise_plan_size(int64_t n);               /* number of int64 words */
int geocb_pairwise_plan_host(int64_t n, int64_t* plan_host); /* host-side, no GPU needed */
int64_t geocb_kpp_workspace_bytes(int64_t n);
int geocb_kpp_draw(const float* dmin32, const uint8_t* is_center, int64_t n, double u, const int64_t* plan,
                   void* workspace, int64_t workspace_bytes, int64_t* result_host, void* stream);
/* float32 NumPy-pairwise sum of x[n] (exposed for tests and for QE) -> *out_host (syncs).
 * square != 0: sums x*x (f32 multiply); only finite x are included when finite_only != 0,
 * with the survivors compacted first exactly like d[finite_mask] (kmeans_optimized.py:129-132). */
int geocb_pairwise_sum_f32(const float* x, int64_t n, int square, int finite_only, void* workspace,
                           int64_t workspace_bytes, float* out_host, int64_t* n_used_host, void* stream);

/* Allocation-free variant of geocb_pairwise_sum_f32 (finite_only = 0).  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Calls no cudaMalloc / cudaFree / allocating library routine: its scratch (summation
 * plan + partial sums) is a 3 MiB file-scope `static __device__` array that every GPU loading the library holds until
 * exit.  n <= geocb_pairwise_sum_f32_static_max_n() = 4 194 304 elements, else GEOCB_ERR_WORKSPACE with a message.
 * All calls on a device share the array: calls that overlap on the GPU race; serialise them.  Same bits as
 * geocb_pairwise_sum_f32 -> *out_host (syncs the stream). */
int64_t geocb_pairwise_sum_f32_static_max_n(void);
int geocb_pairwise_sum_f32_static(const float* x, int64_t n, int square, float* out_host, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GEOCB_H */
