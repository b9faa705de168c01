// Microbenchmark: random 8-byte gathers out of an L2-resident array -- LSU path (divergent LDG) vs the
// TMA bulk-copy path (cp.async.bulk 16 B global->shared) vs both at once.  Decides whether the SSSP kernel's
// L1TEX wavefront wall (one 128-byte line per ~2 cycles per SM for divergent LDGs) can be side-stepped.
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o gather gather.cu ; run: ./gather [n]
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
typedef unsigned long long u64;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); exit(1);} } while (0)

__device__ __forceinline__ uint32_t hash32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x;
}
__global__ void fill_idx(uint32_t* idx, int64_t m, uint32_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (int64_t)gridDim.x * blockDim.x)
        idx[i] = hash32((uint32_t)i * 2654435761u + 12345u) % n;
}
__global__ void fill_arr(u64* a, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) a[i] = i;
}

// LSU path: U independent divergent 8-byte gathers in flight per lane
template <int U, typename T>
__global__ void __launch_bounds__(256) k_ldg(const T* __restrict__ arr, const uint32_t* __restrict__ idx, int64_t m, u64* out) {
    const int64_t gt = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nt = (int64_t)gridDim.x * blockDim.x;
    u64 acc = 0;
    for (int64_t i = gt; i + (U - 1) * nt < m; i += U * nt) {
        uint32_t ix[U];
        T v[U];
#pragma unroll
        for (int k = 0; k < U; ++k) ix[k] = idx[i + k * nt];
#pragma unroll
        for (int k = 0; k < U; ++k) asm volatile("ld.global.cg.b%1 %0, [%2];" : "=l"(*(u64*)&v[k]) : "n"(sizeof(T) * 8), "l"(arr + ix[k]));
#pragma unroll
        for (int k = 0; k < U; ++k) acc += (u64)v[k];
    }
    if (acc == 0x123456789ULL) out[0] = acc;
}
template <int U>
__global__ void __launch_bounds__(256) k_ldg32(const uint32_t* __restrict__ arr, const uint32_t* __restrict__ idx, int64_t m, u64* out) {
    const int64_t gt = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nt = (int64_t)gridDim.x * blockDim.x;
    u64 acc = 0;
    for (int64_t i = gt; i + (U - 1) * nt < m; i += U * nt) {
        uint32_t ix[U], v[U];
#pragma unroll
        for (int k = 0; k < U; ++k) ix[k] = idx[i + k * nt];
#pragma unroll
        for (int k = 0; k < U; ++k) asm volatile("ld.global.cg.u32 %0, [%1];" : "=r"(v[k]) : "l"(arr + ix[k]));
#pragma unroll
        for (int k = 0; k < U; ++k) acc += v[k];
    }
    if (acc == 0x123456789ULL) out[0] = acc;
}

// TMA path: every lane issues U 16-byte bulk copies into its warp's shared buffer, one mbarrier per warp.
// frac_ldg (0..U): that many of the U gathers per lane go through the LSU instead (hybrid).
template <int U>
__global__ void __launch_bounds__(256) k_bulk(const u64* __restrict__ arr, const uint32_t* __restrict__ idx, int64_t m, u64* out, int n_ldg) {
    __shared__ __align__(16) u64 buf[8][U][32][2];
    __shared__ __align__(8) u64 bar[8];
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    const int64_t gt = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nt = (int64_t)gridDim.x * blockDim.x;
    const uint32_t bar_a = (uint32_t)__cvta_generic_to_shared(&bar[w]);
    if (l == 0) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_a));
    __syncwarp();
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    u64 acc = 0;
    uint32_t phase = 0;
    const int n_bulk = U - n_ldg;
    for (int64_t i = gt; i + (U - 1) * nt < m; i += U * nt) {
        uint32_t ix[U];
#pragma unroll
        for (int k = 0; k < U; ++k) ix[k] = idx[i + k * nt];
        if (l == 0 && n_bulk > 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_a), "r"(n_bulk * 32 * 16) : "memory");
        __syncwarp();
        u64 v[U];
#pragma unroll
        for (int k = 0; k < U; ++k) {
            if (k < n_bulk) {
                const uint32_t dst = (uint32_t)__cvta_generic_to_shared(&buf[w][k][l][0]);
                asm volatile("cp.async.bulk.shared::cta.global.mbarrier::complete_tx::bytes [%0], [%1], 16, [%2];"
                             ::"r"(dst), "l"(arr + (ix[k] & ~1u)), "r"(bar_a) : "memory");
            } else {
                asm volatile("ld.global.cg.u64 %0, [%1];" : "=l"(v[k]) : "l"(arr + ix[k]));
            }
        }
        if (n_bulk > 0) {
            uint32_t done = 0;
            while (!done)
                asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                             : "=r"(done) : "r"(bar_a), "r"(phase) : "memory");
            phase ^= 1;
        }
#pragma unroll
        for (int k = 0; k < U; ++k) acc += k < n_bulk ? buf[w][k][l][ix[k] & 1] : v[k];
        __syncwarp();
    }
    if (acc == 0x123456789ULL) out[0] = acc;
}

template <typename F>
float timeit(F f, int reps = 5) {
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    f();
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < reps; ++r) {
        cudaEventRecord(a); f(); cudaEventRecord(b);
        CK(cudaEventSynchronize(b));
        float ms; cudaEventElapsedTime(&ms, a, b);
        best = ms < best ? ms : best;
    }
    return best;
}

int main(int argc, char** argv) {
    const int64_t n = argc > 1 ? atoll(argv[1]) : 2940000;
    const int64_t m = 1LL << 28;  // gathers per launch
    u64 *arr, *out; uint32_t* idx;
    CK(cudaMalloc(&arr, n * 8)); CK(cudaMalloc(&out, 64)); CK(cudaMalloc(&idx, m * 4));
    fill_arr<<<1024, 256>>>(arr, n); fill_idx<<<4096, 256>>>(idx, m, (uint32_t)n);
    CK(cudaDeviceSynchronize());
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    printf("n=%lld (%.1f MB of u64), %lld gathers/launch, %d SMs\n", (long long)n, n * 8 / 1e6, (long long)m, sms);
    for (int bps : {4, 8}) {
        const int grid = sms * bps;
        float t;
        t = timeit([&] { k_ldg<4, u64><<<grid, 256>>>(arr, idx, m, out); });
        printf("LDG.64  U=4  %d blk/SM: %.2f ms  %.1f G gathers/s\n", bps, t, m / t / 1e6);
        t = timeit([&] { k_ldg<8, u64><<<grid, 256>>>(arr, idx, m, out); });
        printf("LDG.64  U=8  %d blk/SM: %.2f ms  %.1f G gathers/s\n", bps, t, m / t / 1e6);
        t = timeit([&] { k_ldg32<8><<<grid, 256>>>((const uint32_t*)arr, idx, m, out); });
        printf("LDG.32  U=8  %d blk/SM: %.2f ms  %.1f G gathers/s\n", bps, t, m / t / 1e6);
    }
    for (int bps : {2, 4}) {
        const int grid = sms * bps;
        for (int nl : {0, 2, 4, 6}) {
            float t = timeit([&] { k_bulk<8><<<grid, 256>>>(arr, idx, m, out, nl); });
            printf("bulk16 U=8 (%d of 8 via LDG) %d blk/SM: %.2f ms  %.1f G gathers/s\n", nl, bps, t, m / t / 1e6);
        }
    }
    CK(cudaDeviceSynchronize());
    return 0;
}
