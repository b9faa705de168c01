// jvp_gemm.cu -- fp32-grade GEMM on the 5th-generation tensor cores (tcgen05 + TMEM) for the decoder JVP.
//
// The Riemannian edge lengths (reference: src/geo/riemannian_metric.py:12-35, 50-63) are built from JVPs of the
// decoder; with the decoder unrolled on its 1x1 latent grid the work is three GEMMs per stacked (primal | tangent)
// row, and the 512 -> 1024 one is 90 % of the flops (vqvae_b200/geo/riemannian_metric.py).  The 1e-5 parity bar
// rules out plain TF32 (2.7e-4), so every operand is split into tf32 hi + lo parts and
//     C = A_hi B_hi + A_lo B_hi + A_hi B_lo          (the dropped lo.lo term is ~2^-22 relative)
// accumulates in fp32 in TMEM: three tcgen05.mma.kind::tf32 per K-step instead of cuBLAS' SIMT SGEMM.
//
//   operand images   global memory already holds UMMA canonical K-major no-swizzle tiles (8x16-byte core matrices,
//                    SBO = 128 B between 8-row groups, LBO = rows*16 B between 16-byte K-groups), hi and lo halves
//                    back to back per (tile, 16-float K-chunk), so ONE 1-D cp.async.bulk stages an operand chunk:
//                      A: [M/128][K/16][hi 8 KB | lo 8 KB]   written per call by gemm3x_split_a_kernel
//                      B: [N/256][K/16][hi 16 KB | lo 16 KB] written once per weight matrix
//   CTA tile         256 (two 128-row A tiles) x 256 columns, K-chunks of 16 floats: 64 KB per stage, 3 stages;
//                    12 MMAs (128x256x8) per chunk; the two accumulators fill all 512 TMEM columns.
//   persistent CTAs  one per SM, tiles in N-fastest order so the four column tiles of an A row block run side by side
//                    (A chunks hit L2 three times out of four).
//   warps            8 = TMA producer (one lane), 9 = MMA issuer (one lane), 0-7 = epilogue (thread = row:
//                    tcgen05.ld 32 columns -> 128 B of the fp32 row-major C).
#include <stdlib.h>
#include "common.cuh"

namespace {

constexpr int kBM = 256, kBN = 256, kBK = 16;
constexpr int kATileBytes = 2 * 128 * kBK * 4;  // hi | lo of one 128-row A tile chunk: 16 KB
constexpr int kBTileBytes = 2 * kBN * kBK * 4;  // hi | lo of one 256-row B tile chunk: 32 KB
constexpr int kStageBytes = 2 * kATileBytes + kBTileBytes;  // 64 KB
constexpr int kStages = 3;
constexpr int kThreads = 320;
constexpr int kLboA = 128 * 16, kLboB = kBN * 16;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// bounded wait: a protocol bug must trap (error returned to the caller), never hang the device
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok = 0;
    for (uint32_t spin = 0; !ok; ++spin) {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(ok)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
        if (spin > (1u << 26)) __trap();
    }
}
__device__ __forceinline__ void tma_load_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ float to_tf32(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}
// UMMA shared-memory descriptor, K-major, no swizzle: start>>4 [0,14), LBO>>4 [16,30), SBO>>4 [32,46), version 1
__device__ __forceinline__ uint64_t umma_desc(uint32_t smem_addr, uint32_t lbo) {
    return (uint64_t)((smem_addr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(128 >> 4) << 32) | (1ULL << 46);
}
// instruction descriptor: D=f32, A=B=tf32, K-major both, M=128, N=256
constexpr uint32_t kIdesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(kBN >> 3) << 17) | ((128u >> 4) << 24);

__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(kIdesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// ---------------------------------------------------------------- operand images ----
// one thread per (row, 16-byte K-group): hi = tf32(x), lo = x - hi.  rows_per_tile = 128 (A) or 256 (B).
// TRANS = false: src is row-major [rows][K] (activations); true: src is [K][rows] (weight matrix W[k][n], row = n).
template <bool TRANS>
__global__ void gemm3x_split_kernel(const float* __restrict__ src, int64_t rows, int K, int rows_per_tile, float* img,
                                    int round_lo) {
    const int kgroups = K / 4;
    const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t row = gid % rows;  // consecutive threads = consecutive rows: the image writes coalesce
    const int g = (int)(gid / rows);
    if (g >= kgroups) return;
    float v[4];
    if (TRANS) {
#pragma unroll
        for (int i = 0; i < 4; ++i) v[i] = src[(int64_t)(4 * g + i) * rows + row];
    } else {
        const float4 x = *reinterpret_cast<const float4*>(src + row * K + 4 * g);
        v[0] = x.x, v[1] = x.y, v[2] = x.z, v[3] = x.w;
    }
    float hi[4], lo[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) hi[i] = to_tf32(v[i]), lo[i] = round_lo ? to_tf32(v[i] - hi[i]) : v[i] - hi[i];
    const int nchunks = K / kBK;
    const int64_t tile = row / rows_per_tile;
    const int r = (int)(row % rows_per_tile);
    const int chunk = g / 4, kgi = g % 4;
    const int64_t half = (int64_t)rows_per_tile * kBK;  // floats of one hi (or lo) chunk
    float* base = img + ((tile * nchunks + chunk) * 2) * half + (int64_t)kgi * rows_per_tile * 4 + (int64_t)r * 4;
    *reinterpret_cast<float4*>(base) = make_float4(hi[0], hi[1], hi[2], hi[3]);
    *reinterpret_cast<float4*>(base + half) = make_float4(lo[0], lo[1], lo[2], lo[3]);
}

// ---------------------------------------------------------------- GEMM ----
struct alignas(8) GemmBars {
    uint64_t full[kStages], empty[kStages], acc_full, acc_empty;
    uint32_t tmem_base;
};

__global__ void __launch_bounds__(kThreads, 1) gemm3x_kernel(const float* __restrict__ aimg, const float* __restrict__ bimg,
                                                              int64_t M, int K, int N, float* __restrict__ C, int lolo) {
    extern __shared__ __align__(1024) unsigned char smem[];
    GemmBars* bars = (GemmBars*)(smem + kStages * kStageBytes);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int nchunks = K / kBK;
    const int64_t mtiles = M / kBM, ntiles = N / kBN, ntot = mtiles * ntiles;

    if (tid == 0) {
        for (int s = 0; s < kStages; ++s) mbar_init(&bars->full[s], 1), mbar_init(&bars->empty[s], 1);
        mbar_init(&bars->acc_full, 1);
        mbar_init(&bars->acc_empty, 8);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 9) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&bars->tmem_base)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *(volatile uint32_t*)&bars->tmem_base;

    if (warp == 8) {
        // ===== TMA producer =====
        if (lane == 0) {
            int64_t it = 0;
            for (int64_t t = blockIdx.x; t < ntot; t += gridDim.x) {
                const int64_t mt = t / ntiles, nt = t % ntiles;
                for (int c = 0; c < nchunks; ++c, ++it) {
                    const int st = (int)(it % kStages);
                    mbar_wait(&bars->empty[st], (uint32_t)(((it / kStages) & 1) ^ 1));
                    mbar_expect_tx(&bars->full[st], kStageBytes);
                    unsigned char* s = smem + st * kStageBytes;
                    tma_load_1d(s, aimg + ((2 * mt) * nchunks + c) * (int64_t)(kATileBytes / 4), kATileBytes, &bars->full[st]);
                    tma_load_1d(s + kATileBytes, aimg + ((2 * mt + 1) * nchunks + c) * (int64_t)(kATileBytes / 4), kATileBytes,
                                &bars->full[st]);
                    tma_load_1d(s + 2 * kATileBytes, bimg + (nt * nchunks + c) * (int64_t)(kBTileBytes / 4), kBTileBytes,
                                &bars->full[st]);
                }
            }
        }
    } else if (warp == 9) {
        // ===== MMA issuer =====
        if (lane == 0) {
            int64_t it = 0, tile_n = 0;
            for (int64_t t = blockIdx.x; t < ntot; t += gridDim.x, ++tile_n) {
                mbar_wait(&bars->acc_empty, (uint32_t)((tile_n & 1) ^ 1));  // epilogue has drained the accumulators
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                for (int c = 0; c < nchunks; ++c, ++it) {
                    const int st = (int)(it % kStages);
                    mbar_wait(&bars->full[st], (uint32_t)((it / kStages) & 1));
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t s = smem_u32(smem + st * kStageBytes);
                    const uint32_t b_hi = s + 2 * kATileBytes, b_lo = b_hi + kBTileBytes / 2;
#pragma unroll
                    for (int a = 0; a < 2; ++a) {
                        const uint32_t a_hi = s + a * kATileBytes, a_lo = a_hi + kATileBytes / 2;
                        const uint32_t d_tmem = tmem_base + (uint32_t)(a * kBN);
#pragma unroll
                        for (int ks = 0; ks < kBK / 8; ++ks) {  // K-step of 8 floats = 2 K-groups
                            const uint32_t ao = ks * 2 * kLboA, bo = ks * 2 * kLboB;
                            umma_tf32(d_tmem, umma_desc(a_hi + ao, kLboA), umma_desc(b_hi + bo, kLboB), (c | ks) ? 1u : 0u);
                            umma_tf32(d_tmem, umma_desc(a_lo + ao, kLboA), umma_desc(b_hi + bo, kLboB), 1u);
                            umma_tf32(d_tmem, umma_desc(a_hi + ao, kLboA), umma_desc(b_lo + bo, kLboB), 1u);
                            if (lolo) umma_tf32(d_tmem, umma_desc(a_lo + ao, kLboA), umma_desc(b_lo + bo, kLboB), 1u);
                        }
                    }
                    umma_commit(&bars->empty[st]);  // stage reusable once these MMAs have read it
                }
                umma_commit(&bars->acc_full);  // both accumulators of this tile complete
            }
        }
    } else {
        // ===== epilogue: thread = output row, warps 0-3 -> A tile 0, warps 4-7 -> A tile 1 =====
        const int a = warp >> 2;
        const int row = a * 128 + (warp & 3) * 32 + lane;
        int64_t tile_n = 0;
        for (int64_t t = blockIdx.x; t < ntot; t += gridDim.x, ++tile_n) {
            const int64_t mt = t / ntiles, nt = t % ntiles;
            mbar_wait(&bars->acc_full, (uint32_t)(tile_n & 1));
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t taddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)(a * kBN);
            float* crow = C + (mt * kBM + row) * (int64_t)N + nt * kBN;
#pragma unroll 1
            for (int part = 0; part < kBN / 32; ++part) {
                uint32_t v[32];
                tmem_ld32(taddr + part * 32, v);
#pragma unroll
                for (int i = 0; i < 8; ++i)
                    *reinterpret_cast<uint4*>(crow + part * 32 + 4 * i) = make_uint4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(&bars->acc_empty);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 9) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
    }
}

}  // namespace

static int gemm_mode() {
    const char* e = getenv("GEOCB_GEMM3X_MODE");  // experiment switch: bit 0 = round the lo parts to tf32, bit 1 = add lo.lo
    return e ? atoi(e) : 0;
}

GEOCB_API int geocb_gemm3x_supported(int64_t M, int K, int N) {
    return M > 0 && M % kBM == 0 && K > 0 && K % kBK == 0 && N > 0 && N % kBN == 0;
}

GEOCB_API int geocb_gemm3x_prep_weights(const float* W, int K, int N, float* bimg, void* stream) {
    ARG_CHECK(W && bimg && K > 0 && K % kBK == 0 && N > 0 && N % kBN == 0);
    const int64_t threads = (int64_t)N * (K / 4);
    gemm3x_split_kernel<true><<<(unsigned)((threads + 255) / 256), 256, 0, (cudaStream_t)stream>>>(W, N, K, kBN, bimg, gemm_mode() & 1);
    KERNEL_CHECK();
    return 0;
}

GEOCB_API int geocb_gemm3x(const float* A, int64_t M, int K, int N, const float* bimg, float* aimg, float* C, void* stream) {
    ARG_CHECK(bimg && aimg && C);  // A == NULL: aimg already holds the operand image (geocb_bn_relu_jvp_to_image)
    if (!geocb_gemm3x_supported(M, K, N)) {
        geocb_set_error("gemm3x needs M %% 256 == 0, K %% 16 == 0, N %% 256 == 0 (got %lld, %d, %d)", (long long)M, K, N);
        return GEOCB_ERR_UNSUPPORTED;
    }
    cudaStream_t st = (cudaStream_t)stream;
    DevInfo di;
    int rc = geocb_devinfo(&di);
    if (rc) return rc;
    if (A) {
        const int64_t threads = M * (K / 4);
        gemm3x_split_kernel<false><<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(A, M, K, 128, aimg, gemm_mode() & 1);
        KERNEL_CHECK();
    }
    const size_t smem = (size_t)kStages * kStageBytes + sizeof(GemmBars) + 64;
    CU_TRY(cudaFuncSetAttribute(gemm3x_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t ntot = (M / kBM) * (N / kBN);
    const unsigned grid = (unsigned)(ntot < di.sm_count ? ntot : di.sm_count);
    gemm3x_kernel<<<grid, kThreads, smem, st>>>(aimg, bimg, M, K, N, C, (gemm_mode() >> 1) & 1);
    KERNEL_CHECK();
    return 0;
}
