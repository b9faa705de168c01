// knn_tc.cuh -- internal interface of the tcgen05 kNN filter (knn_tc.cu), used by knn.cu.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

// candidates are padded to whole 128-row tiles, queries to whole 256-row CTAs
int64_t knn_tc_padded_candidates(int64_t M);
int64_t knn_tc_padded_queries(int64_t nq);
int64_t knn_tc_image_floats(int64_t rows_padded);  // floats of an operand image with that many rows
// filter pass: writes cand_idx[nq][32] (candidate ids, -1 = empty, ids >= M = pad rows) and cand_thr[nq]
// (every candidate not in the list has tensor-core score >= cand_thr).  bimg/aimg are scratch images.
int knn_filter_tc(const float* z, int64_t M, int d, int64_t q_lo, int64_t nq, float* bimg, float* aimg,
                  unsigned int* max_nrm_bits, int32_t* cand_idx, float* cand_thr, cudaStream_t st);
