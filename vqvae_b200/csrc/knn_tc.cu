// knn_tc.cu -- kNN filter pass on the 5th-generation tensor cores (tcgen05 + TMEM), d <= 64.
//
// Replaces pass 1 of knn.cu (the fp32 FFMA2 filter) for the reference's kNN query
// (src/geo/knn_graph_optimized.py:39-42).  Only the contraction runs on tensor cores:
//     score(q,c) = |c|^2 - 2 q.c      (|q|^2 is constant per row and dropped)
// is one tensor-core GEMM with fp32-grade accuracy by splitting every operand into hi + lo pieces
// (hi*hi + lo*hi + hi*lo; the lo*lo term is dropped) and folding |c|^2 in as one more K-step.  Two operand formats:
//   bf16 pieces (default; kind::f16 MMAs of K = 16, twice the tf32 rate), DP = 16:
//     A row (query)     = [ -2a_hi (16) | -2a_lo (16) | 1 1 1 0 ... ]            48 bf16 = 6 K-groups of 16 bytes
//     B row (candidate) = [   b_hi (16) |   b_lo (16) | n_0 n_1 n_2 0 ... ]      (|c|^2 in three bf16 pieces: 2^-24)
//     4 MMAs per tile and chunk:  hi.hi  lo.hi  hi.lo  ones.norm         error <= 2^-14 (|b|^2 + 2|a||b|)
//   tf32 pieces (GEOCB_KNN_TC_FMT=tf32; kind::tf32 MMAs of K = 8), DP = 16:
//     A row = [ -2a_hi (16) | -2a_lo (16) | 1 1 0 0 0 0 0 0 ],  B row = [ b_hi | b_lo | n_hi n_lo 0 ... ]   40 floats
//     7 MMAs per tile and chunk:  hi.hi (2)  lo.hi (2)  hi.lo (2)  ones.norm (1)     error <= 2^-17 (...)
// The error bound only sets the margin of the exact re-rank's certificate (knn.cu): results never depend on the format.
// Operand tiles are stored in global memory already in the UMMA canonical K-major no-swizzle image
// (8x16-byte core matrices; LBO = rows * 16 B between K-groups, SBO = 128 B between 8-row groups),
// so one 1-D TMA bulk copy (cp.async.bulk + mbarrier) stages a 128-candidate tile.
// CTA shapes (Cfg): one 128-row A tile per CTA and TWO CTAs per SM where shared memory and 256 TMEM columns allow it
// (bf16 DP <= 32, tf32 DP = 16: two independent pipelines fill each other's stalls), else 256 rows = two A tiles that
// share every staged B tile.  Accumulators are double-buffered in TMEM.
//   warp E   lane 0 : TMA producer (A tiles once, then the B tile ring)          (E = number of epilogue warps: 4 or 8)
//   warp E+1 lane 0 : tcgen05.mma issuer; tcgen05.commit releases smem stages / publishes accumulators
//   warps 0..E-1    : epilogue -- thread = query row: tcgen05.ld 32 scores at a time, a min-tree (FMNMX3) screens them
//                     against the row's current threshold, the rare survivors enter the row's L-entry list.
// Output is the same (cand_idx, cand_thr) pair knn.cu's exact fp64 re-rank + certificate consume.
// Measured phase by phase with `make knntrace` + tools/prof_knn_trace.py (profiles/r2_knn_phase_clocks.txt).
#include <stdlib.h>
#include "common.cuh"
#include "knn_tc.cuh"


#ifndef KNN_TC_DEFAULT_BF16
#define KNN_TC_DEFAULT_BF16 1
#endif

namespace {

constexpr int kTileRows = 128;
constexpr int kLBO = kTileRows * 16;        // bytes between K-groups
constexpr int kL = 32;                      // list length (kq <= 26)

// BF = 0: operands split into tf32 hi + lo, 4 values per 16-byte K-group, MMAs of K = 8 (kind::tf32);
// BF = 1: operands split into bf16 hi + lo, 8 values per K-group, MMAs of K = 16 (kind::f16) at twice the tensor rate.
template <int DP_, int ATILES_, int STAGES_, int NSUB_, int NB_, int ES_, int BF_ = 0>
struct Cfg {
    static constexpr int DP = DP_, ATILES = ATILES_, STAGES = STAGES_, NSUB = NSUB_, NB = NB_, ES = ES_, BF = BF_;
    static constexpr int kHiGroups = BF ? DP / 8 : DP / 4;  // K-groups of the hi part (= of the lo part)
    // ES epilogue warps share each 32-row block (each takes NB/ES of a chunk's columns)
    static constexpr int kEpiWarps = 4 * ATILES * ES, kThreads = (kEpiWarps + 2) * 32;
    static constexpr int kKGroups = 2 * kHiGroups + 2;   // 16-byte K-groups per row: hi | lo | norm / ones (one MMA step)
    static constexpr int kTileBytes = kKGroups * kLBO;   // A tile (128 rows): tf32 20480 / 36864 / 69632, bf16 12288 / 20480 / 36864
    static constexpr int kLboB = NB * 16;                // candidate (B) tiles hold NB = 128 or 256 rows: N of the MMA
    static constexpr int kTileBytesB = kKGroups * kLboB;
    static constexpr int kRowsPerCta = ATILES * kTileRows;
    // B sub-tiles (K-group ranges): one, or [hi) and [lo + norm)
    static constexpr int kSplit = NSUB == 1 ? kKGroups : kHiGroups;
    static constexpr int kStageBytes = (NSUB == 1 ? kKGroups : (kKGroups - kHiGroups)) * kLboB;
    static constexpr int kChunks = kHiGroups / 2;        // MMA-K sized pieces of the hi part (an MMA reads two K-groups)
    static constexpr int kSteps = 3 * kChunks + 1;       // hi.hi, lo.hi, hi.lo per piece + ones.norm
    static constexpr int NBUF = 512 / (ATILES * NB);     // accumulator buffers per A tile that fit the 512 TMEM columns
    static_assert(NBUF == 1 || NBUF == 2 || NBUF == 4 || NBUF == 8, "TMEM budget");
    static constexpr int kBufs = NBUF > 2 ? 2 : NBUF;
    static constexpr int kTmemCols = kBufs * ATILES * NB;  // 512, or 256 for the one-tile configurations
    // one-tile CTAs whose shared memory fits twice (and 256 TMEM columns each) run two per SM
    static constexpr int kSmemEstimate = ATILES * kTileBytes + STAGES * kStageBytes + kL * kRowsPerCta * 8 + 2048;
    static constexpr int kBySmem = (225 * 1024) / kSmemEstimate, kByTmem = 512 / kTmemCols;
    static constexpr int kMinCtas = ATILES != 1 ? 1 : (kBySmem < kByTmem ? (kBySmem < 1 ? 1 : kBySmem) : kByTmem);
    // D=f32, A=B=tf32 (2) or bf16 (1), K-major both, M=128, N=NB  (cute::UMMA::InstrDescriptor)
    static constexpr uint32_t kFmt = BF ? 1u : 2u;
    static constexpr uint32_t kIdesc = (1u << 4) | (kFmt << 7) | (kFmt << 10) | ((uint32_t)(NB >> 3) << 17) | ((128u >> 4) << 24);
};

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
// UMMA shared-memory descriptor, K-major, no swizzle (cute::UMMA::SmemDescriptor): start>>4 [0,14),
// LBO>>4 [16,30), SBO>>4 [32,46), version=1 [46,48), layout_type=0 [61,64)
__device__ __forceinline__ uint64_t umma_desc(uint32_t smem_addr, uint32_t lbo) {
    return (uint64_t)((smem_addr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(128 >> 4) << 32) |
           (1ULL << 46);
}
// (instruction descriptor, cute::UMMA::InstrDescriptor layout: Cfg::kIdesc)
template <int BF>
__device__ __forceinline__ void umma_issue(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    if constexpr (BF) {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "setp.ne.b32 p, %4, 0;\n"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
            "}\n" ::"r"(tmem_d),
            "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
            : "memory");
    } else {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "setp.ne.b32 p, %4, 0;\n"
            "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
            "}\n" ::"r"(tmem_d),
            "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
            : "memory");
    }
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}
// (no wait inside: the epilogue puts all four 32-column loads of an accumulator in flight before one
//  tcgen05.wait::ld -- tools/ubench/tmem_ld.cu: 23.6 scores/clk/SM with a wait per load, 68 with two in flight)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ float fmin3(float a, float b, float c) {  // FMNMX3 on sm_100a
    float r;
    asm("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// byte offset of element (row r, float k) inside a 128-row operand tile image
__device__ __forceinline__ int img_off(int r, int k, int lbo) { return (k >> 2) * lbo + (r >> 3) * 128 + (r & 7) * 16 + (k & 3) * 4; }
// K-group pairs (A kg, B kg) of MMA step t for padded dimension DP: steps are grouped by the B sub-tile they read:
// first every step on B's hi part (hi.hi and lo.hi per 8-float chunk), then hi.lo per chunk, then ones.norm.
template <int H>  // H = K-groups of the hi part
__device__ __forceinline__ void mma_step(int t, int& akg, int& bkg) {
    constexpr int C = H / 2;
    if (t < 2 * C) {
        const int j = t >> 1;
        akg = (t & 1) ? H + 2 * j : 2 * j;
        bkg = 2 * j;
    } else if (t < 3 * C) {
        const int j = t - 2 * C;
        akg = 2 * j;
        bkg = H + 2 * j;
    } else {
        akg = bkg = 2 * H;
    }
}

// ------------------------------------------------------------------ prep ----
// one thread per (row, K-group): writes the candidate image (all M rows, padded to whole tiles) and the query image
// (nq rows of zq, padded to whole CTAs).  Pad candidates get a huge finite norm so they are never kept.
__global__ void knn_tc_prep_kernel(const float* __restrict__ z, const float* __restrict__ zq, int64_t M, int d, int dp,
                                   int nb, int64_t Mp, int64_t nq, int64_t nq_pad, float* bimg, float* aimg,
                                   unsigned int* max_nrm_bits) {
    const int kgroups = (2 * dp + 8) / 4, hi_groups = dp / 4;
    const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t row = gid / kgroups;
    const int kg = (int)(gid % kgroups);
    const bool is_b = row < Mp;
    const int64_t qrow = row - Mp;  // rows >= Mp enumerate the query image
    if (!is_b && qrow >= nq_pad) return;
    const bool live = is_b ? row < M : qrow < nq;
    const float* src = is_b ? z + row * d : zq + qrow * d;
    float out[4] = {0.f, 0.f, 0.f, 0.f};
    if (kg < 2 * hi_groups) {  // hi (first dp/4 groups) / lo (next dp/4 groups) parts of 4 coordinates
        const int c0 = (kg % hi_groups) * 4;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float v = (live && c0 + i < d) ? src[c0 + i] : 0.f;
            if (!is_b) v *= -2.0f;
            const float hi = to_tf32(v);
            out[i] = kg < hi_groups ? hi : v - hi;
        }
    } else if (kg == 2 * hi_groups) {
        if (is_b) {
            float s = 0.f;
            if (live) {
                for (int i = 0; i < d; ++i) s = __fmaf_rn(src[i], src[i], s);
                atomicMax(max_nrm_bits, __float_as_uint(s));
            } else {
                s = 3.0e38f;
            }
            const float hi = to_tf32(s);
            out[0] = hi;
            out[1] = live ? s - hi : 0.f;
        } else {
            out[0] = 1.f;
            out[1] = 1.f;
        }
    }
    const int tile_rows = is_b ? nb : kTileRows;  // candidate tiles hold nb rows (the MMA's N), query tiles 128
    const int64_t tile = (is_b ? row : qrow) / tile_rows;
    const int r = (int)((is_b ? row : qrow) % tile_rows);
    float* base = (is_b ? bimg : aimg) + tile * ((int64_t)kgroups * tile_rows * 4);
    *(float4*)((char*)base + img_off(r, kg * 4, tile_rows * 16)) = make_float4(out[0], out[1], out[2], out[3]);
}

// bf16 operand images (Cfg<..., BF = 1>): K-groups of 8 bf16: hi | lo | norm.  x = hi + lo + r with |r| <= 2^-16 |x|
// (two round-to-nearest 8-bit pieces); |c|^2 goes in as three bf16 pieces (2^-24) against three ones.
__device__ __forceinline__ uint32_t bf16_bits(float x) {
    uint16_t r;
    asm("cvt.rn.bf16.f32 %0, %1;" : "=h"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float bf16_val(uint32_t b) { return __uint_as_float(b << 16); }
__global__ void knn_tc_prep_bf16_kernel(const float* __restrict__ z, const float* __restrict__ zq, int64_t M, int d, int dp,
                                        int nb, int64_t Mp, int64_t nq, int64_t nq_pad, float* bimg, float* aimg,
                                        unsigned int* max_nrm_bits) {
    const int hi_groups = dp / 8, kgroups = 2 * hi_groups + 2;
    const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t row = gid / kgroups;
    const int kg = (int)(gid % kgroups);
    const bool is_b = row < Mp;
    const int64_t qrow = row - Mp;
    if (!is_b && qrow >= nq_pad) return;
    const bool live = is_b ? row < M : qrow < nq;
    const float* src = is_b ? z + row * d : zq + qrow * d;
    uint32_t h[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (kg < 2 * hi_groups) {
        const int c0 = (kg % hi_groups) * 8;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            float v = (live && c0 + i < d) ? src[c0 + i] : 0.f;
            if (!is_b) v *= -2.0f;
            const uint32_t hi = bf16_bits(v);
            h[i] = kg < hi_groups ? hi : bf16_bits(v - bf16_val(hi));
        }
    } else if (kg == 2 * hi_groups) {
        if (is_b) {
            float s = 0.f;
            if (live) {
                for (int i = 0; i < d; ++i) s = __fmaf_rn(src[i], src[i], s);
                atomicMax(max_nrm_bits, __float_as_uint(s));
                h[0] = bf16_bits(s);
                const float r1 = s - bf16_val(h[0]);
                h[1] = bf16_bits(r1);
                h[2] = bf16_bits(r1 - bf16_val(h[1]));
            } else {
                h[0] = bf16_bits(3.0e38f);  // pad candidates: never kept
            }
        } else {
            h[0] = h[1] = h[2] = 0x3F80u;  // 1.0
        }
    }
    const int tile_rows = is_b ? nb : kTileRows;
    const int64_t tile = (is_b ? row : qrow) / tile_rows;
    const int r = (int)((is_b ? row : qrow) % tile_rows);
    float* base = (is_b ? bimg : aimg) + tile * ((int64_t)kgroups * tile_rows * 4);
    *(uint4*)((char*)base + img_off(r, kg * 4, tile_rows * 16)) =
        make_uint4(h[0] | (h[1] << 16), h[2] | (h[3] << 16), h[4] | (h[5] << 16), h[6] | (h[7] << 16));
}

// ------------------------------------------------------------------ filter ----
// order-preserving float <-> int map (an involution): lets __reduce_max_sync find the largest list entry
__device__ __forceinline__ int okey(float f) {
    const int b = __float_as_int(f);
    return b ^ ((b >> 31) & 0x7fffffff);
}
__device__ __forceinline__ float unkey(int k) { return __int_as_float(k ^ ((k >> 31) & 0x7fffffff)); }

// Warp-cooperative list insert (lists row-major, sc[row][kL]): for every lane with a pending survivor (s < its row
// threshold), the 32 lanes load the 32 entries of THAT row's list, one __reduce_max_sync finds the entry to replace, a
// second one the new threshold -- ~20 warp-wide instructions per survivor instead of a 31-step dependent shared-memory
// scan by one lane while 31 wait.  Kept small and out of line on purpose: it is called from 128 unrolled sites with ~130
// live registers, and a callee that needs many registers makes every call spill them (measured: 2x slower).
// The caller holds the row block's lock when two warps share the rows.  Returns the caller's (lowered) threshold.
__device__ __noinline__ float coop_insert(float* sc, int32_t* ix, int row0w, bool pend, float s, int32_t cidx, float thr) {
    const int lane = threadIdx.x & 31;
    unsigned todo = __ballot_sync(0xffffffffu, pend);
    while (todo) {
        const int j = __ffs(todo) - 1;
        todo &= todo - 1;
        const float sj = __shfl_sync(0xffffffffu, s, j);
        const int32_t cj = __shfl_sync(0xffffffffu, cidx, j);
        float* lst = sc + (row0w + j) * kL;
        int ke = okey(lst[lane]);
        const int mx = __reduce_max_sync(0xffffffffu, ke);
        const int pos = __ffs(__ballot_sync(0xffffffffu, ke == mx)) - 1;
        if (lane == pos) {
            lst[pos] = sj;
            ix[(row0w + j) * kL + pos] = cj;
            ke = okey(sj);
        }
        const int mx2 = __reduce_max_sync(0xffffffffu, ke);
        if (lane == j) thr = unkey(mx2);
        __syncwarp();
    }
    return thr;
}

#ifdef GEOCB_KNN_TRACE  // profiling builds only (`make knntrace`, tools/prof_knn_trace.py): clocks per phase, summed per warp
__device__ unsigned long long g_knn_trace[16];
#define KT_NOW() clock64()
#define KT_DECL() unsigned long long kt_acc[16] = {0}
#define KT_ADD(slot, dt) kt_acc[slot] += (unsigned long long)(dt)
#define KT_FLUSH() do { if ((threadIdx.x & 31) == 0) for (int kt_i = 0; kt_i < 16; ++kt_i) if (kt_acc[kt_i]) atomicAdd(&g_knn_trace[kt_i], kt_acc[kt_i]); } while (0)
#else
#define KT_NOW() 0LL
#define KT_DECL() do { } while (0)
#define KT_ADD(slot, dt) do { } while (0)
#define KT_FLUSH() do { } while (0)
#endif

template <int STAGES>
struct alignas(8) TcBars {
    uint64_t a_full, full[STAGES], empty[STAGES], tmem_full[4], tmem_empty[4];  // accumulator barriers: [buffer][A tile]
    uint32_t tmem_base;
    int lock[8];  // one per 32-row block of the CTA (epilogue warps sharing the block's lists)
};

template <class C>
__global__ void __launch_bounds__(C::kThreads, C::kMinCtas) knn_filter_tc_kernel(const float* __restrict__ aimg,
                                                                        const float* __restrict__ bimg, int64_t Mp,
                                                                        int64_t nq, int32_t* cand_idx, float* cand_thr) {
    constexpr int kTileBytes = C::kTileBytes, kStages = C::STAGES, kRowsPerCta = C::kRowsPerCta, NSUB = C::NSUB;
    constexpr int NB = C::NB, kBufs = C::kBufs, AT = C::ATILES, ES = C::ES;
    constexpr int kProducer = C::kEpiWarps, kIssuer = C::kEpiWarps + 1;
    extern __shared__ __align__(1024) unsigned char smem[];
    unsigned char* a_s = smem;                                   // ATILES A tiles
    unsigned char* b_s = smem + AT * kTileBytes;                 // kStages B (sub-)tiles
    float* sc = (float*)(b_s + kStages * C::kStageBytes);        // [rows][kL] list scores
    int32_t* ix = (int32_t*)(sc + kL * kRowsPerCta);             // [rows][kL] list indices
    float* thr_s = (float*)(ix + kL * kRowsPerCta);              // [rows] current list maximum (the row's threshold)
    TcBars<kStages>* bars = (TcBars<kStages>*)(thr_s + kRowsPerCta);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int64_t nchunks = Mp / NB;
    const int64_t tile0 = (int64_t)blockIdx.x * AT;  // first of this CTA's A tiles

    if (tid == 0) {
        mbar_init(&bars->a_full, 1);
        for (int s = 0; s < kStages; ++s) mbar_init(&bars->full[s], 1), mbar_init(&bars->empty[s], 1);
        for (int b = 0; b < 4; ++b) mbar_init(&bars->tmem_full[b], 1), mbar_init(&bars->tmem_empty[b], 4 * ES);
        for (int b = 0; b < 8; ++b) bars->lock[b] = 0;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = tid; i < kL * kRowsPerCta; i += C::kThreads) {
        sc[i] = __uint_as_float(F32_INF_BITS);
        ix[i] = -1;
    }
    for (int i = tid; i < kRowsPerCta; i += C::kThreads) thr_s[i] = __uint_as_float(F32_INF_BITS);
    if (warp == kIssuer) {  // one warp allocates the CTA's TMEM columns (all 512, or 256 when two CTAs share the SM)
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&bars->tmem_base)),
                     "n"(C::kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *(volatile uint32_t*)&bars->tmem_base;
    KT_DECL();  // [0] epilogue: wait for the accumulator [2] TMEM loads [3] screen + insert [5] warp-chunks;
                // [8] issuer: wait B stage [9] wait accumulators drained [10] issue + [9] [11] chunks

    if (warp == kProducer) {
        // ===== TMA producer: A tiles once, then the ring of B (sub-)tiles =====
        if (lane == 0) {
            mbar_expect_tx(&bars->a_full, AT * kTileBytes);
            tma_load_1d(a_s, aimg + tile0 * (kTileBytes / 4), AT * kTileBytes, &bars->a_full);
            for (int64_t it = 0; it < nchunks * NSUB; ++it) {
                const int64_t c = it / NSUB;
                const int sub = (int)(it % NSUB);
                const int st = (int)(it % kStages);
                const int kg0 = sub == 0 ? 0 : C::kSplit, kg1 = (NSUB == 1 || sub == 1) ? C::kKGroups : C::kSplit;
                const uint32_t bytes = (uint32_t)((kg1 - kg0) * C::kLboB);
                mbar_wait(&bars->empty[st], (uint32_t)(((it / kStages) & 1) ^ 1));
                mbar_expect_tx(&bars->full[st], bytes);
                tma_load_1d(b_s + st * C::kStageBytes, bimg + c * (C::kTileBytesB / 4) + kg0 * (C::kLboB / 4), bytes,
                            &bars->full[st]);
            }
        }
    } else if (warp == kIssuer) {
        // ===== MMA issuer.  Accumulators are tracked per (buffer, A tile): with one buffer per A tile (N = 256) the
        //       epilogue of A tile 0 overlaps the MMAs of A tile 1 and vice versa. =====
        if (lane == 0) {
            mbar_wait(&bars->a_full, 0);
            const uint32_t a_addr = smem_u32(a_s), b_addr = smem_u32(b_s);
            constexpr int kHiSteps = NSUB == 1 ? C::kSteps : 2 * C::kChunks;  // steps that read B sub-tile 0
            for (int64_t c = 0; c < nchunks; ++c) {
                const int buf = (int)(c % kBufs);
                const uint32_t use = (uint32_t)(c / kBufs);  // how often this buffer has been used before
#pragma unroll
                for (int sub = 0; sub < NSUB; ++sub) {
                    const int64_t it = c * NSUB + sub;
                    const int st = (int)(it % kStages);
                    const long long ki0 = KT_NOW();
                    mbar_wait(&bars->full[st], (uint32_t)((it / kStages) & 1));
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const long long ki1 = KT_NOW();
                    KT_ADD(8, ki1 - ki0);
                    const int t0 = sub == 0 ? 0 : kHiSteps, t1 = (NSUB == 1 || sub == 1) ? C::kSteps : kHiSteps;
                    const int kg0 = sub == 0 ? 0 : C::kSplit;
#pragma unroll
                    for (int a = 0; a < AT; ++a) {
                        if (sub == 0) {  // the epilogue must have drained this accumulator
                            const long long kw = KT_NOW();
                            mbar_wait(&bars->tmem_empty[buf * AT + a], (use & 1) ^ 1);
                            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                            KT_ADD(9, KT_NOW() - kw);
                        }
                        const uint32_t d_tmem = tmem_base + (uint32_t)((buf * AT + a) * NB);
#pragma unroll
                        for (int t = t0; t < t1; ++t) {
                            int akg, bkg;
                            mma_step<C::kHiGroups>(t, akg, bkg);
                            umma_issue<C::BF>(d_tmem, umma_desc(a_addr + a * kTileBytes + akg * kLBO, kLBO),
                                      umma_desc(b_addr + st * C::kStageBytes + (bkg - kg0) * C::kLboB, C::kLboB), C::kIdesc,
                                      t > 0 ? 1u : 0u);
                        }
                        if (NSUB == 1 || sub == 1) umma_commit(&bars->tmem_full[buf * AT + a]);  // this accumulator is complete
                    }
                    umma_commit(&bars->empty[st]);  // smem stage reusable once these MMAs have read it
                    KT_ADD(10, KT_NOW() - ki1);
                    KT_ADD(11, 1);
                }
            }
        }
    } else {
        // ===== epilogue: thread = query row; ES warps share a 32-row block, each screens NB/ES columns of a chunk =====
        const int a = warp / (4 * ES);      // which A tile; warp % 4 selects the TMEM lane quarter this warp may read
        const int h = (warp >> 2) % ES;     // which slice of the chunk's columns
        const int row0w = a * 128 + (warp & 3) * 32;  // first row of this warp
        const int row = row0w + lane;
        int* lock = ES > 1 ? &bars->lock[row0w >> 5] : nullptr;
        constexpr int kCols = NB / ES;      // columns per warp and chunk (64, 128 or 256)
        constexpr int kGroup = kCols >= 128 ? 4 : kCols / 32;  // 32-column loads in flight per wait
        float thr = __uint_as_float(F32_INF_BITS);
        for (int64_t c = 0; c < nchunks; ++c) {
            const int buf = (int)(c % kBufs);
            const long long kt0 = KT_NOW();
            mbar_wait(&bars->tmem_full[buf * AT + a], (uint32_t)((c / kBufs) & 1));
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            KT_ADD(0, KT_NOW() - kt0);
            KT_ADD(5, 1);
            const uint32_t taddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)((buf * AT + a) * NB + h * kCols);
            const int32_t c0 = (int32_t)(c * NB + h * kCols);
            if (ES > 1) thr = fminf(thr, *(volatile float*)&thr_s[row]);  // the partner warp may have tightened it
#pragma unroll 1
            for (int q4 = 0; q4 < kCols / (32 * kGroup); ++q4) {
                // screen with the values kept in registers: kGroup 32-column loads in flight, one wait, min-trees; a hit
                // part is searched with one warp vote per column and every survivor is inserted by the whole warp
                float v[kGroup][32];
                const long long kt1 = KT_NOW();
#pragma unroll
                for (int part = 0; part < kGroup; ++part) tmem_ld32(taddr + (q4 * kGroup + part) * 32, v[part]);
                tmem_ld_wait();
                const long long kt2 = KT_NOW();
                KT_ADD(2, kt2 - kt1);
#pragma unroll
                for (int part = 0; part < kGroup; ++part) {
                    // minimum of each 8-column octet with three-input minima (FMNMX3: 4 per octet), then of the part
                    float g8[4];
#pragma unroll
                    for (int o = 0; o < 4; ++o) {
                        const float* u = &v[part][8 * o];
                        g8[o] = fminf(fmin3(u[0], u[1], u[2]), fmin3(fmin3(u[3], u[4], u[5]), u[6], u[7]));
                    }
                    const float tmin = fminf(fmin3(g8[0], g8[1], g8[2]), g8[3]);
                    if (__any_sync(0xffffffffu, tmin < thr)) {
                        if (ES > 1) {  // the partner warp shares these rows' lists: take the block's lock, refresh thr
                            if (lane == 0)
                                while (atomicCAS(lock, 0, 1) != 0) {
                                }
                            __syncwarp();
                            __threadfence_block();
                            thr = *(volatile float*)&thr_s[row];
                        }
                        const float thr_before = thr;
#pragma unroll
                        for (int o = 0; o < 4; ++o) {
                            if (!__any_sync(0xffffffffu, g8[o] < thr)) continue;  // (thr only ever drops: still exact)
#pragma unroll
                            for (int i = 8 * o; i < 8 * o + 8; ++i) {
                                const bool pend = v[part][i] < thr;
                                if (__any_sync(0xffffffffu, pend))
                                    thr = coop_insert(sc, ix, row0w, pend, v[part][i], c0 + (q4 * kGroup + part) * 32 + i, thr);
                            }
                        }
                        if (thr != thr_before) thr_s[row] = thr;  // publish: final write-out and (ES > 1) the partner
                        if (ES > 1) {
                            __threadfence_block();
                            __syncwarp();
                            if (lane == 0) atomicExch(lock, 0);
                        }
                    }
                }
                KT_ADD(3, KT_NOW() - kt2);
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(&bars->tmem_empty[buf * AT + a]);
        }
    }
    KT_FLUSH();
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (tid < kRowsPerCta) {  // every epilogue warp is done: thread = row writes its list out
        const int64_t q = (int64_t)blockIdx.x * kRowsPerCta + tid;
        if (q < nq) {
            for (int l = 0; l < kL; ++l) cand_idx[q * kL + l] = ix[tid * kL + l];
            cand_thr[q] = thr_s[tid];
        }
    }
    if (warp == kIssuer) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(C::kTmemCols) : "memory");
    }
}

template <class C>
int launch_tc(const float* z, const float* zq, int64_t M, int d, int64_t nq, float* bimg, float* aimg,
              unsigned int* max_nrm_bits, int32_t* cand_idx, float* cand_thr, cudaStream_t st) {
    const int64_t Mp = align_up(M, 256), nqp = align_up(nq, C::kRowsPerCta);  // (256: every config's B tile divides it)
    const int64_t threads = (Mp + nqp) * C::kKGroups;
    if (C::BF)
        knn_tc_prep_bf16_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(z, zq, M, d, C::DP, C::NB, Mp, nq, nqp,
                                                                                    bimg, aimg, max_nrm_bits);
    else
        knn_tc_prep_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(z, zq, M, d, C::DP, C::NB, Mp, nq, nqp, bimg,
                                                                               aimg, max_nrm_bits);
    KERNEL_CHECK();
    const size_t smem = (size_t)C::ATILES * C::kTileBytes + (size_t)C::STAGES * C::kStageBytes +
                        (size_t)kL * C::kRowsPerCta * 8 + (size_t)C::kRowsPerCta * 4 + sizeof(TcBars<C::STAGES>) + 64;
    auto kern = knn_filter_tc_kernel<C>;
    CU_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(unsigned)(nqp / C::kRowsPerCta), C::kThreads, smem, st>>>(aimg, bimg, Mp, nq, cand_idx, cand_thr);
    KERNEL_CHECK();
    return 0;
}

typedef Cfg<16, 2, 4, 1, 128, 1> Cfg16;    // N = 128 MMAs, two accumulator buffers per A tile, 8 epilogue warps
typedef Cfg<16, 1, 3, 1, 128, 1, 1> Cfg16B;  // bf16 operands: 4 MMAs of K = 16 per tile and chunk instead of 7 of K = 8; 2 CTAs/SM
typedef Cfg<16, 1, 2, 1, 64, 1, 1> Cfg16B64;  // the same with 64-candidate chunks: 128 TMEM columns, 3 CTAs/SM (experiment)
typedef Cfg<32, 1, 2, 1, 128, 1, 1> Cfg32B;  // bf16: 7 MMAs instead of 13, one-tile CTA (95 KB): 2 CTAs/SM
typedef Cfg<64, 2, 2, 1, 128, 1, 1> Cfg64B;  // bf16: 13 MMAs instead of 25, two 36 KB A tiles per CTA
typedef Cfg<16, 1, 2, 1, 128, 1> Cfg16P;   // one A tile per CTA, TWO CTAs per SM: each CTA's MMAs fill the other's epilogue stalls
typedef Cfg<16, 2, 4, 1, 128, 2> Cfg16E2;  // the same with 16 epilogue warps: two per 32-row block, 64 columns each
typedef Cfg<16, 2, 3, 1, 256, 1> Cfg16W;   // N = 256 MMAs (half as many instructions), one buffer per A tile
typedef Cfg<32, 2, 2, 1, 128, 1> Cfg32;
typedef Cfg<64, 1, 3, 2, 128, 1> Cfg64;
static_assert(2 * Cfg16::kTileBytes + 4 * Cfg16::kStageBytes + kL * 256 * 8 < 227 * 1024, "smem");
static_assert(2 * Cfg16W::kTileBytes + 3 * Cfg16W::kStageBytes + kL * 256 * 8 + 512 < 227 * 1024, "smem");
static_assert(2 * Cfg32::kTileBytes + 2 * Cfg32::kStageBytes + kL * 256 * 8 < 227 * 1024, "smem");
static_assert(1 * Cfg64::kTileBytes + 3 * Cfg64::kStageBytes + kL * 128 * 8 < 227 * 1024, "smem");

}  // namespace

int knn_tc_dp(int d) { return d <= 16 ? 16 : d <= 32 ? 32 : 64; }
int64_t knn_tc_padded_candidates(int64_t M) { return align_up(M, 256); }
int64_t knn_tc_padded_queries(int64_t nq, int d) {
    (void)d;
    return align_up(nq, 256);  // every configuration's CTA (128 or 256 query rows) divides it
}
int64_t knn_tc_image_floats(int64_t rows_padded, int d) { return rows_padded * (2 * knn_tc_dp(d) + 8); }

// operand format of the filter: GEOCB_KNN_TC_FMT = bf16 | tf32
static bool knn_tc_bf16() {
    const char* e = getenv("GEOCB_KNN_TC_FMT");
    return KNN_TC_DEFAULT_BF16 ? !(e && e[0] == 't') : (e && e[0] == 'b');
}
// bound on |filter score - exact score| / (|b|^2 + 2|a||b|), used by the certificate of the re-rank pass (knn.cu):
//   tf32 hi+lo: dropped lo.lo term, tf32 truncation of lo, fp32 accumulation of <= 136 terms: 2^-17;
//   bf16 hi+lo: x = hi + lo + r, |r| <= 2^-16|x| for both operands and the dropped lo.lo term: 3 * 2^-16 on the dot
//   product, plus the fp32 accumulation and the 2^-24 norm split: 2^-14.
double knn_tc_eps_unit(void) { return knn_tc_bf16() ? 6.103515625e-05 : 7.62939453125e-06; }

int knn_filter_tc(const float* z, const float* zq, int64_t M, int d, int64_t nq, float* bimg, float* aimg,
                  unsigned int* max_nrm_bits, int32_t* cand_idx, float* cand_thr, cudaStream_t st) {
    if (d > 64) {
        geocb_set_error("knn_filter_tc supports d <= 64");
        return GEOCB_ERR_UNSUPPORTED;
    }
    if (knn_tc_bf16()) {
        switch (knn_tc_dp(d)) {
            case 16: {
                const char* e = getenv("GEOCB_KNN_TC_N");
                if (e && atoi(e) == 64) return launch_tc<Cfg16B64>(z, zq, M, d, nq, bimg, aimg, max_nrm_bits, cand_idx, cand_thr, st);
                return launch_tc<Cfg16B>(z, zq, M, d, nq, bimg, aimg, max_nrm_bits, cand_idx, cand_thr, st);
            }
            case 32: return launch_tc<Cfg32B>(z, zq, M, d, nq, bimg, aimg, max_nrm_bits, cand_idx, cand_thr, st);
            default: return launch_tc<Cfg64B>(z, zq, M, d, nq, bimg, aimg, max_nrm_bits, cand_idx, cand_thr, st);
        }
    }
    switch (knn_tc_dp(d)) {
        case 16: {
            // N of the tcgen05.mma instructions at d <= 16: 128 (default) | 256.  Measured equal at C2 (2.44 vs 2.50 s): the
            // filter is bound by the tensor pipe's sustained tf32 rate, not by instruction count
            // Default: one A tile per CTA and two CTAs per SM.  The phase clocks (tools/prof_knn_trace.py,
            // profiles/r2_knn_phase_clocks.txt) show the two-tile CTA fully serialised: its issuer spends 1058 clocks per
            // chunk issuing 14 MMAs at the tensor pipe's tf32 rate, then waits for the slowest of 8 epilogue warps, which
            // in turn wait for the MMAs (56 % of their time).  Two independent CTAs fill each other's gaps: tensor pipe
            // 91 % busy, 1.75 s against 1.89 s at C2.  GEOCB_KNN_TC_PAIR=0 selects the two-tile CTA.
            const char* pp = getenv("GEOCB_KNN_TC_PAIR");
            if (!(pp && atoi(pp) == 0) && !getenv("GEOCB_KNN_TC_N") && !getenv("GEOCB_KNN_TC_ES"))
                return launch_tc<Cfg16P>(z, zq, M, d, nq, bimg, aimg, max_nrm_bits, cand_idx, cand_thr, st);
            const char* e = getenv("GEOCB_KNN_TC_N");
            if (e && atoi(e) == 256) return launch_tc<Cfg16W>(z, zq, M, d, nq, bimg, aimg, max_nrm_bits, cand_idx, cand_thr, st);
            // epilogue warps per 32-row block: GEOCB_KNN_TC_ES=2 splits each block's columns over two warps that share the
            // lists under a lock -- measured SLOWER at C2 (4.6-5.7 s against 2.2 s: lock contention), kept for experiments
            const char* es = getenv("GEOCB_KNN_TC_ES");
            if (es && atoi(es) == 2) return launch_tc<Cfg16E2>(z, zq, M, d, nq, bimg, aimg, max_nrm_bits, cand_idx, cand_thr, st);
            return launch_tc<Cfg16>(z, zq, M, d, nq, bimg, aimg, max_nrm_bits, cand_idx, cand_thr, st);
        }
        case 32: return launch_tc<Cfg32>(z, zq, M, d, nq, bimg, aimg, max_nrm_bits, cand_idx, cand_thr, st);
        default: return launch_tc<Cfg64>(z, zq, M, d, nq, bimg, aimg, max_nrm_bits, cand_idx, cand_thr, st);
    }
}

#ifdef GEOCB_KNN_TRACE
extern "C" __attribute__((visibility("default"))) int geocb_knn_trace_read(unsigned long long* out16, int reset) {
    if (cudaMemcpyFromSymbol(out16, g_knn_trace, sizeof(g_knn_trace)) != cudaSuccess) return -1;
    if (reset) {
        unsigned long long z[16] = {0};
        if (cudaMemcpyToSymbol(g_knn_trace, z, sizeof(z)) != cudaSuccess) return -1;
    }
    return 0;
}
#endif
