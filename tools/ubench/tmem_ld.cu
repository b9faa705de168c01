// Microbenchmark: how fast can the epilogue warps of knn_tc.cu read scores out of TMEM?
// Every CTA (one per SM) allocates all 512 TMEM columns and its warps loop over tcgen05.ld.32x32b.x32 reads of their
// lane quarter -- no MMA, no shared memory, the accumulators hold whatever is there -- in three flavours:
//   ld        loads only (one value kept alive)
//   ld+min    loads + the 31-deep FMNMX tree + threshold test of the kNN filter's epilogue
//   ld+min2   the same with two loads in flight before the first wait (x32 + x32)
// Prints scores per clock per SM and chip-wide T scores/s: the ceiling DESIGN.md 4.2 compares the filter against.
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tmem_ld tmem_ld.cu ; run: ./tmem_ld
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); exit(1);} } while (0)

__device__ __forceinline__ void ld32_nowait(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
}
__device__ __forceinline__ void ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

template <int MODE>
__global__ void __launch_bounds__(512, 1) k_tmem(int iters, float thr, unsigned long long* clocks, float* sink) {
    __shared__ uint32_t tmem_base_s;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(
                         (uint32_t)__cvta_generic_to_shared(&tmem_base_s))
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t base = tmem_base_s;
    // warp w reads lane quarter w%4; warps sharing a quarter read different 128-column accumulators
    const uint32_t taddr = base + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)(((warp >> 2) & 3) * 128);
    float acc = 0.f;
    int hits = 0;
    __syncthreads();
    const unsigned long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        if (MODE == 2) {
#pragma unroll 1
            for (int part = 0; part < 4; part += 2) {
                uint32_t a[32], b[32];
                ld32_nowait(taddr + part * 32, a);
                ld32_nowait(taddr + part * 32 + 32, b);
                ld_wait();
                float m = __uint_as_float(a[0]);
#pragma unroll
                for (int i = 1; i < 32; ++i) m = fminf(m, __uint_as_float(a[i]));
#pragma unroll
                for (int i = 0; i < 32; ++i) m = fminf(m, __uint_as_float(b[i]));
                if (m < thr) ++hits;
            }
        } else {
#pragma unroll 1
            for (int part = 0; part < 4; ++part) {
                uint32_t a[32];
                ld32_nowait(taddr + part * 32, a);
                ld_wait();
                if (MODE == 0) {
                    acc += __uint_as_float(a[it & 31]);
                } else {
                    float m = __uint_as_float(a[0]);
#pragma unroll
                    for (int i = 1; i < 32; ++i) m = fminf(m, __uint_as_float(a[i]));
                    if (m < thr) ++hits;
                }
            }
        }
    }
    const unsigned long long t1 = clock64();
    if (threadIdx.x == 0) clocks[blockIdx.x] = t1 - t0;
    if (acc == 12345.f || hits == -1) sink[0] = acc + hits;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(base) : "memory");
}

template <int MODE>
void run(const char* name, int warps, int sms, double ghz) {
    unsigned long long* clk;
    float* sink;
    CK(cudaMalloc(&clk, sms * 8));
    CK(cudaMalloc(&sink, 4));
    const int iters = 20000;
    k_tmem<MODE><<<sms, warps * 32>>>(100, -1e30f, clk, sink);
    k_tmem<MODE><<<sms, warps * 32>>>(iters, -1e30f, clk, sink);
    CK(cudaDeviceSynchronize());
    unsigned long long* h = (unsigned long long*)malloc(sms * 8);
    CK(cudaMemcpy(h, clk, sms * 8, cudaMemcpyDeviceToHost));
    double mx = 0;
    for (int i = 0; i < sms; ++i) mx = h[i] > mx ? (double)h[i] : mx;
    const double scores = (double)warps * 32 * 128 * iters;
    printf("%-8s %2d warps/SM: %.2f scores/clk/SM  -> %.2f T scores/s chip-wide at %.2f GHz\n", name, warps, scores / mx,
           scores / mx * sms * ghz * 1e9 / 1e12, ghz);
    free(h);
    CK(cudaFree(clk));
    CK(cudaFree(sink));
}

int main() {
    cudaDeviceProp p;
    CK(cudaGetDeviceProperties(&p, 0));
    const double ghz = p.clockRate / 1e6;
    printf("%s, %d SMs, %.3f GHz max\n", p.name, p.multiProcessorCount, ghz);
    for (int warps : {4, 8, 16}) {
        run<0>("ld", warps, p.multiProcessorCount, ghz);
        run<1>("ld+min", warps, p.multiProcessorCount, ghz);
        run<2>("ld+min2", warps, p.multiProcessorCount, ghz);
    }
    return 0;
}
