// scan.cuh -- exclusive prefix sums into int64 (out has n+1 entries; out[n] = total).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
int scan_exclusive_i32(const int32_t* in, int64_t n, int64_t* out, cudaStream_t st);
int scan_exclusive_u8(const uint8_t* in, int64_t n, int64_t* out, cudaStream_t st);
