// knn_tc.cuh -- internal interface of the tcgen05 kNN filter (knn_tc.cu), used by knn.cu.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

// d <= 64.  Candidates are padded to whole 128-row tiles, queries to whole CTAs (256 rows, 128 for d > 32).
int knn_tc_dp(int d);  // padded dimension of the operand images: 16, 32 or 64
int64_t knn_tc_padded_candidates(int64_t M);
int64_t knn_tc_padded_queries(int64_t nq, int d);
int64_t knn_tc_image_floats(int64_t rows_padded, int d);  // floats of an operand image with that many rows
// filter pass: queries zq[nq][d] against candidates z[M][d]; writes cand_idx[nq][32] (candidate ids, -1 = empty,
// ids >= M = pad rows) and cand_thr[nq] (every candidate not in the list has tensor-core score >= cand_thr).
// bimg/aimg are scratch images.
int knn_filter_tc(const float* z, const float* zq, int64_t M, int d, int64_t nq, float* bimg, float* aimg,
                  unsigned int* max_nrm_bits, int32_t* cand_idx, float* cand_thr, cudaStream_t st);
