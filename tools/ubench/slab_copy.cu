// slab_copy.cu -- how fast can a [rows][1024] float matrix be read and rewritten when every CTA owns a column slab of W
// floats x 512 rows (the access pattern of bn_relu_jvp_kernel: W = 16 -> 64-byte pieces at a 4 KB stride)?
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o slab_copy slab_copy.cu ; run: ./slab_copy
#include <cstdio>
#include <cuda_runtime.h>

template <int W>  // floats per row and CTA
__global__ void __launch_bounds__(256) slab(float* __restrict__ x, float* __restrict__ t, int ld) {
    constexpr int V = W / 4;  // float4 per row
    const int col0 = blockIdx.x * W;
    const long long row0 = (long long)blockIdx.y * 512;
    float4 a[2][V], b[2][V];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const long long r = row0 + threadIdx.x + k * 256;
#pragma unroll
        for (int q = 0; q < V; ++q) {
            a[k][q] = ((const float4*)(x + r * ld + col0))[q];
            b[k][q] = ((const float4*)(t + r * ld + col0))[q];
        }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const long long r = row0 + threadIdx.x + k * 256;
#pragma unroll
        for (int q = 0; q < V; ++q) {
            float4 u = a[k][q], v = b[k][q];
            u.x += v.y; v.x += u.y;
            ((float4*)(x + r * ld + col0))[q] = u;
            ((float4*)(t + r * ld + col0))[q] = v;
        }
    }
}

template <int W>
void run(float* x, float* t, int rows, int ld) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    dim3 grid(ld / W, rows / 512);
    for (int i = 0; i < 3; ++i) slab<W><<<grid, 256>>>(x, t, ld);
    cudaEventRecord(e0);
    for (int i = 0; i < 10; ++i) slab<W><<<grid, 256>>>(x, t, ld);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double bytes = 4.0 * rows * (double)ld * 4.0;  // x and t, read and written
    printf("slab of %2d floats (%3d B per row): %.1f us per pass, %.2f TB/s  (%s)\n", W, W * 4, ms * 100, bytes / (ms * 1e-4) / 1e12,
           cudaGetErrorString(cudaGetLastError()));
}

int main() {
    const int rows = 65536, ld = 1024;
    float *x, *t;
    cudaMalloc(&x, (size_t)rows * ld * 4); cudaMalloc(&t, (size_t)rows * ld * 4);
    cudaMemset(x, 0, (size_t)rows * ld * 4); cudaMemset(t, 0, (size_t)rows * ld * 4);
    run<8>(x, t, rows, ld);
    run<16>(x, t, rows, ld);
    run<32>(x, t, rows, ld);
    run<64>(x, t, rows, ld);
    return 0;
}
