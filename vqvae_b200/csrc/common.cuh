// common.cuh -- shared helpers for libgeocb (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cooperative_groups.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include "../../include/geocb.h"

namespace cg = cooperative_groups;

#define GEOCB_API extern "C" __attribute__((visibility("default")))

void geocb_set_error(const char* fmt, ...);

#define CU_TRY(expr)                                                                   \
    do {                                                                               \
        cudaError_t _e = (expr);                                                       \
        if (_e != cudaSuccess) {                                                       \
            geocb_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return (int)_e;                                                            \
        }                                                                              \
    } while (0)

#define ARG_CHECK(cond)                                                    \
    do {                                                                   \
        if (!(cond)) {                                                     \
            geocb_set_error("%s:%d bad argument: %s", __FILE__, __LINE__, #cond); \
            return GEOCB_ERR_ARG;                                          \
        }                                                                  \
    } while (0)

#define KERNEL_CHECK() CU_TRY(cudaGetLastError())

static inline int64_t align_up(int64_t x, int64_t a) { return (x + a - 1) / a * a; }

// cached device properties
struct DevInfo {
    int sm_count;
    int cc;
};
int geocb_devinfo(DevInfo* out);

// +inf as raw bits; non-negative IEEE doubles order like their uint64 bit patterns,
// which is what lets atomicMin(unsigned long long) implement an fp64 min.
#define D64_INF_BITS 0x7FF0000000000000ULL
#define F32_INF_BITS 0x7F800000u

__device__ __forceinline__ unsigned long long ld_cg_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.global.cg.u64 %0, [%1];" : "=l"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ int ld_cg_s32(const int* p) {
    int v;
    asm volatile("ld.global.cg.s32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
// streaming 128-bit read-only load that does not pollute L1 (edge pairs are read once per relaxation)
__device__ __forceinline__ uint4 ld_stream_v4(const void* p) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p));
    return v;
}
