// knn_tc.cuh -- internal interface of the tcgen05 kNN filter (knn_tc.cu), used by knn.cu.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

// d <= 64.  Candidates are padded to whole 256-row tiles, queries to 256 rows (every configuration's CTA divides it).
int knn_tc_dp(int d);  // padded dimension of the operand images: 16, 32 or 64
int64_t knn_tc_padded_candidates(int64_t M);
int64_t knn_tc_padded_queries(int64_t nq, int d);
int64_t knn_tc_image_floats(int64_t rows_padded, int d);  // floats of an operand image with that many rows
// |filter score - exact score| <= knn_tc_eps_unit() * (|b|^2 + 2|a||b|) for the operand format in use (tf32 / bf16 split)
double knn_tc_eps_unit(void);
// filter pass: queries zq[nq][d] against candidates z[M][d]; writes cand_idx[nq][32] (candidate ids, -1 = empty,
// ids >= M = pad rows) and cand_thr[nq] (every candidate not in the list has tensor-core score >= cand_thr).
// bimg/aimg are scratch images.
int knn_filter_tc(const float* z, const float* zq, int64_t M, int d, int64_t nq, float* bimg, float* aimg,
                  unsigned int* max_nrm_bits, int32_t* cand_idx, float* cand_thr, cudaStream_t st);
