// knn.cu -- exact Euclidean kNN: fused fp32 tiled score + per-thread top-L filter, exact fp64 re-rank.
//
// Replaces sklearn NearestNeighbors(n_neighbors=k+1).fit(z).kneighbors(z)
// (reference: src/geo/knn_graph_optimized.py:39-42; sklearn brute ArgKmin32 ranks by the fp64
// value of |x|^2+|y|^2-2xy with the earlier index winning ties).
//
// pass 0  prep: candidates re-laid out pair-interleaved  zt[pair][dim][2]  (+ fp32 |b|^2), padded to
//         whole tiles with |b|^2 = +inf so pad rows can never be selected.
// pass 1  filter: each thread owns RQ query rows (registers, pre-scaled by -2, duplicated into f32x2
//         lanes); candidate tiles are staged into shared memory with 1-D TMA bulk copies
//         (cp.async.bulk + mbarrier, double buffered) and scored two candidates at a time with
//         packed FFMA2 (`fma.rn.f32x2`): score = |b|^2 - 2 a.b.  Scores below the thread's current
//         threshold enter an L-entry list in shared memory (L = kq + slack).
// pass 2  re-rank: one warp per query recomputes the L candidates' squared distances exactly in fp64,
//         orders them by (d2, index) and certifies the row: every excluded candidate has fp32 score
//         >= T, hence exact score >= T - eps; if the kq-th exact score is not < T - eps the row is
//         redone by pass 3.
// pass 3  exact fp64 brute force for uncertified rows (one CTA per row; rare).
#include <stdlib.h>
#include "common.cuh"
#include "knn_tc.cuh"

namespace {

constexpr int kFilterThreads = 128;
constexpr int kTileBytes = 16384;  // candidate coordinates per stage

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// bounded wait: a protocol bug must trap (error returned to the caller), never hang the device
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok = 0;
    for (uint32_t spin = 0; !ok; ++spin) {
        if (spin > (1u << 26)) __trap();
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(ok)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
    }
}
// 1-D TMA bulk copy global -> shared, completion counted on the mbarrier
__device__ __forceinline__ void tma_load_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ unsigned long long pack2(float lo, float hi) {
    unsigned long long r;
    asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(unsigned long long v, float& lo, float& hi) {
    asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ unsigned long long ffma2(unsigned long long a, unsigned long long b, unsigned long long c) {
    unsigned long long d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}

// ------------------------------------------------------------------ pass 0 ----
__global__ void knn_prep_kernel(const float* z, int64_t M, int d, int dpad, int64_t Mp, float* zt, float* nrm,
                                unsigned int* max_nrm_bits) {
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= Mp) return;
    float s = 0.f;
    float* dst = zt + (c >> 1) * (int64_t)dpad * 2 + (c & 1);
    for (int i = 0; i < dpad; ++i) {
        float v = (c < M && i < d) ? z[c * d + i] : 0.f;
        dst[i * 2] = v;
        s = __fmaf_rn(v, v, s);
    }
    nrm[c] = c < M ? s : __uint_as_float(F32_INF_BITS);
    if (c < M) atomicMax(max_nrm_bits, __float_as_uint(s));
}

// ------------------------------------------------------------------ pass 1 ----
template <int DPAD, int RQ, int L>
__global__ void __launch_bounds__(kFilterThreads) knn_filter_kernel(const float* __restrict__ zq, int d,
                                                                    const float* __restrict__ zt,
                                                                    const float* __restrict__ nrm, int64_t Mp,
                                                                    int64_t nq, int32_t* cand_idx,
                                                                    float* cand_thr) {
    constexpr int TILE_C = kTileBytes / (DPAD * 4);
    constexpr int NT = kFilterThreads * RQ;
    extern __shared__ __align__(128) unsigned char smem[];
    float* tile[2] = {(float*)smem, (float*)(smem + kTileBytes)};
    float* tnrm[2] = {(float*)(smem + 2 * kTileBytes), (float*)(smem + 2 * kTileBytes + TILE_C * 4)};
    float* sc = (float*)(smem + 2 * kTileBytes + 2 * TILE_C * 4);
    int32_t* ix = (int32_t*)(sc + L * NT);
    uint64_t* bars = (uint64_t*)(ix + L * NT);

    const int tid = threadIdx.x;
    const int64_t ntiles = Mp / TILE_C;
    constexpr uint32_t kStageBytes = kTileBytes + TILE_C * 4;
    if (tid == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (tid == 0) {
        for (int s = 0; s < 2 && s < ntiles; ++s) {
            mbar_expect_tx(&bars[s], kStageBytes);
            tma_load_1d(tile[s], zt + (int64_t)s * TILE_C * DPAD, kTileBytes, &bars[s]);
            tma_load_1d(tnrm[s], nrm + (int64_t)s * TILE_C, TILE_C * 4, &bars[s]);
        }
    }
    // query rows -> registers, scaled by -2 and duplicated into both f32x2 lanes
    unsigned long long qd[RQ][DPAD];
    float thr[RQ];
    int maxpos[RQ];
#pragma unroll
    for (int r = 0; r < RQ; ++r) {
        const int64_t q = (int64_t)blockIdx.x * NT + r * kFilterThreads + tid;
        const float* src = zq + (q < nq ? q : 0) * d;
#pragma unroll
        for (int i = 0; i < DPAD; ++i) {
            float v = i < d ? -2.0f * src[i] : 0.f;
            qd[r][i] = pack2(v, v);
        }
        thr[r] = __uint_as_float(F32_INF_BITS);
        maxpos[r] = 0;
        for (int l = 0; l < L; ++l) {
            sc[l * NT + r * kFilterThreads + tid] = __uint_as_float(F32_INF_BITS);
            ix[l * NT + r * kFilterThreads + tid] = -1;
        }
    }

    auto insert = [&](int r, float s, int32_t c) {
        const int t = r * kFilterThreads + tid;
        sc[maxpos[r] * NT + t] = s;
        ix[maxpos[r] * NT + t] = c;
        float m = sc[t];
        int mp = 0;
#pragma unroll 8
        for (int l = 1; l < L; ++l) {
            float v = sc[l * NT + t];
            if (v > m) {
                m = v;
                mp = l;
            }
        }
        thr[r] = m;
        maxpos[r] = mp;
    };

    for (int64_t t = 0; t < ntiles; ++t) {
        const int st = (int)(t & 1);
        mbar_wait(&bars[st], (uint32_t)((t >> 1) & 1));
        const float4* tp = (const float4*)tile[st];
        const float2* np = (const float2*)tnrm[st];
        const int32_t c0 = (int32_t)(t * TILE_C);
#pragma unroll 2
        for (int p = 0; p < TILE_C / 2; ++p) {
            const float2 nb = np[p];
            unsigned long long acc[RQ];
#pragma unroll
            for (int r = 0; r < RQ; ++r) acc[r] = pack2(nb.x, nb.y);
#pragma unroll
            for (int i = 0; i < DPAD; i += 2) {
                const float4 v = tp[p * (DPAD / 2) + (i >> 1)];
                const unsigned long long b0 = pack2(v.x, v.y), b1 = pack2(v.z, v.w);
#pragma unroll
                for (int r = 0; r < RQ; ++r) {
                    acc[r] = ffma2(qd[r][i], b0, acc[r]);
                    acc[r] = ffma2(qd[r][i + 1], b1, acc[r]);
                }
            }
#pragma unroll
            for (int r = 0; r < RQ; ++r) {
                float s0, s1;
                unpack2(acc[r], s0, s1);
                if (fminf(s0, s1) < thr[r]) {
                    if (s0 < thr[r]) insert(r, s0, c0 + 2 * p);
                    if (s1 < thr[r]) insert(r, s1, c0 + 2 * p + 1);
                }
            }
        }
        __syncthreads();  // everyone is done with stage st
        if (tid == 0 && t + 2 < ntiles) {
            mbar_expect_tx(&bars[st], kStageBytes);
            tma_load_1d(tile[st], zt + (t + 2) * TILE_C * DPAD, kTileBytes, &bars[st]);
            tma_load_1d(tnrm[st], nrm + (t + 2) * TILE_C, TILE_C * 4, &bars[st]);
        }
    }
#pragma unroll
    for (int r = 0; r < RQ; ++r) {
        const int64_t q = (int64_t)blockIdx.x * NT + r * kFilterThreads + tid;
        if (q < nq) {
            for (int l = 0; l < L; ++l) cand_idx[q * L + l] = ix[l * NT + r * kFilterThreads + tid];
            cand_thr[q] = thr[r];
        }
    }
}

// ------------------------------------------------------------------ pass 2 ----
__device__ __forceinline__ double exact_d2(const float* a, const float* b, int d) {
    double s = 0.0;
    for (int i = 0; i < d; ++i) {
        double t = (double)a[i] - (double)b[i];
        s = fma(t, t, s);
    }
    return s;
}
__device__ __forceinline__ bool key_less(double d0, int i0, double d1, int i1) {
    return d0 < d1 || (d0 == d1 && i0 < i1);
}

template <int L>
__global__ void knn_rerank_kernel(const float* __restrict__ z, const float* __restrict__ zq, int64_t M, int d, int64_t nq, int kq,
                                  const int32_t* cand_idx, const float* cand_thr, const unsigned int* max_nrm_bits,
                                  double eps_unit, int32_t* out_idx, float* out_dist, int32_t* unsafe_rows,
                                  int32_t* n_unsafe) {
    constexpr int PER = L / 32;
    const int64_t q = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (q >= nq) return;
    const float* a = zq + q * d;
    double d2[PER];
    int id[PER];
    int nvalid = 0;
#pragma unroll
    for (int e = 0; e < PER; ++e) {
        id[e] = cand_idx[q * L + e * 32 + lane];
        if (id[e] >= M) id[e] = -1;  // pad rows of the tensor-core filter
        d2[e] = id[e] >= 0 ? exact_d2(a, z + (int64_t)id[e] * d, d) : __longlong_as_double(D64_INF_BITS);
        if (id[e] < 0) id[e] = 0x7fffffff;
        nvalid += __popc(__ballot_sync(0xffffffffu, id[e] != 0x7fffffff));
    }
    // rank of every candidate among the L (keys are distinct because indices are)
    int rank[PER];
#pragma unroll
    for (int e = 0; e < PER; ++e) rank[e] = 0;
#pragma unroll
    for (int f = 0; f < PER; ++f)
        for (int src = 0; src < 32; ++src) {
            const double od = __shfl_sync(0xffffffffu, d2[f], src);
            const int oi = __shfl_sync(0xffffffffu, id[f], src);
#pragma unroll
            for (int e = 0; e < PER; ++e) rank[e] += key_less(od, oi, d2[e], id[e]);
        }
    double kth = 0.0;
    bool have_kth = false;
#pragma unroll
    for (int e = 0; e < PER; ++e) {
        if (rank[e] < kq && id[e] != 0x7fffffff) {
            out_idx[q * kq + rank[e]] = id[e];
            if (out_dist) out_dist[q * kq + rank[e]] = (float)sqrt(d2[e]);
        }
        if (rank[e] == kq - 1) {
            kth = d2[e];
            have_kth = true;
        }
    }
    // certificate
    const unsigned long long kb = __double_as_longlong(have_kth ? kth : -1.0);
    unsigned long long kall = 0;
    for (int src = 0; src < 32; ++src) {
        unsigned long long t = __shfl_sync(0xffffffffu, kb, src);
        if ((long long)t >= 0) kall = t;
    }
    if (lane == 0) {
        bool safe;
        if (nvalid < L) {
            safe = nvalid >= kq;  // every candidate entered the list
        } else {
            double a2 = 0.0;
            for (int i = 0; i < d; ++i) a2 = fma((double)a[i], (double)a[i], a2);
            const double b2 = (double)__uint_as_float(*max_nrm_bits) * (1.0 + 1e-6);
            // filter score error bound: eps_unit = (2d+8)*2^-24*1.01 for the fp32 FFMA2 filter, 2^-17 for the
            // 3xTF32 tensor-core filter (dropped lo*lo term, tf32 truncation of lo, fp32 accumulation of 56 terms)
            const double eps = eps_unit * (b2 + 2.0 * sqrt(a2 * b2));
            const double T = (double)cand_thr[q];
            safe = (__longlong_as_double((long long)kall) - a2) < (T - eps);
        }
        if (!safe) unsafe_rows[atomicAdd(n_unsafe, 1)] = (int32_t)q;
    }
}

// ------------------------------------------------------------------ pass 3 ----
// exact fp64 selection for one query per CTA: kq rounds of "smallest key greater than the last one".
__global__ void knn_exact_kernel(const float* __restrict__ z, const float* __restrict__ zq, int64_t M, int d, int kq,
                                 const int32_t* rows, const int32_t* n_rows, int all_rows, int32_t* out_idx,
                                 float* out_dist) {
    const int nr = all_rows ? all_rows : *n_rows;
    __shared__ double s_d[256];
    __shared__ int s_i[256];
    __shared__ double last_d;
    __shared__ int last_i;
    for (int rr = blockIdx.x; rr < nr; rr += gridDim.x) {
        const int64_t q = all_rows ? rr : rows[rr];
        const float* a = zq + q * d;
        if (threadIdx.x == 0) {
            last_d = -1.0;
            last_i = -1;
        }
        __syncthreads();
        for (int sel = 0; sel < kq; ++sel) {
            const double ld = last_d;
            const int li = last_i;
            double bd = __longlong_as_double(D64_INF_BITS);
            int bi = 0x7fffffff;
            for (int64_t c = threadIdx.x; c < M; c += blockDim.x) {
                const double dd = exact_d2(a, z + c * d, d);
                if (key_less(ld, li, dd, (int)c) && key_less(dd, (int)c, bd, bi)) {
                    bd = dd;
                    bi = (int)c;
                }
            }
            s_d[threadIdx.x] = bd;
            s_i[threadIdx.x] = bi;
            __syncthreads();
            for (int o = blockDim.x / 2; o; o >>= 1) {
                if (threadIdx.x < o && key_less(s_d[threadIdx.x + o], s_i[threadIdx.x + o], s_d[threadIdx.x], s_i[threadIdx.x])) {
                    s_d[threadIdx.x] = s_d[threadIdx.x + o];
                    s_i[threadIdx.x] = s_i[threadIdx.x + o];
                }
                __syncthreads();
            }
            if (threadIdx.x == 0) {
                last_d = s_d[0];
                last_i = s_i[0];
                if (s_i[0] != 0x7fffffff) {
                    out_idx[q * kq + sel] = s_i[0];
                    if (out_dist) out_dist[q * kq + sel] = (float)sqrt(s_d[0]);
                }
            }
            __syncthreads();
        }
    }
}

template <int DPAD, int RQ, int L>
int launch_filter(const float* zq, int d, const float* zt, const float* nrm, int64_t Mp, int64_t nq,
                  int32_t* cand_idx, float* cand_thr, cudaStream_t st) {
    constexpr int TILE_C = kTileBytes / (DPAD * 4);
    constexpr int NT = kFilterThreads * RQ;
    const size_t smem = 2 * kTileBytes + 2 * TILE_C * 4 + (size_t)L * NT * 8 + 64;
    auto kern = knn_filter_kernel<DPAD, RQ, L>;
    CU_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const unsigned grid = (unsigned)((nq + NT - 1) / NT);
    kern<<<grid, kFilterThreads, smem, st>>>(zq, d, zt, nrm, Mp, nq, cand_idx, cand_thr);
    KERNEL_CHECK();
    return 0;
}

template <int L>
int dispatch_filter(int dpad, const float* zq, int d, const float* zt, const float* nrm, int64_t Mp,
                    int64_t nq, int32_t* cand_idx, float* cand_thr, cudaStream_t st) {
    switch (dpad) {
        case 4: return launch_filter<4, 2, L>(zq, d, zt, nrm, Mp, nq, cand_idx, cand_thr, st);
        case 8: return launch_filter<8, 2, L>(zq, d, zt, nrm, Mp, nq, cand_idx, cand_thr, st);
        case 16: return launch_filter<16, 2, L>(zq, d, zt, nrm, Mp, nq, cand_idx, cand_thr, st);
        case 32: return launch_filter<32, 1, L>(zq, d, zt, nrm, Mp, nq, cand_idx, cand_thr, st);
        case 64: return launch_filter<64, 1, L>(zq, d, zt, nrm, Mp, nq, cand_idx, cand_thr, st);
    }
    return GEOCB_ERR_UNSUPPORTED;
}

struct Scratch {
    cudaStream_t st;
    void* ptrs[8];
    int n = 0;
    explicit Scratch(cudaStream_t s) : st(s) {}
    int alloc(void** p, size_t bytes) {
        CU_TRY(cudaMallocAsync(p, bytes ? bytes : 8, st));
        ptrs[n++] = *p;
        return 0;
    }
    ~Scratch() {
        for (int i = 0; i < n; ++i) cudaFreeAsync(ptrs[i], st);
    }
};
}  // namespace

// queries zq[nq][d] against candidates z[M][d] (zq may point into z)
static int knn_query_impl(const float* z, const float* zq, int64_t M, int d, int64_t nq, int kq, int32_t* out_idx,
                          float* out_dist, int32_t* n_unsafe, cudaStream_t st) {
    DevInfo di;
    int rc = geocb_devinfo(&di);
    if (rc) return rc;
    CU_TRY(cudaMemsetAsync(out_idx, 0xFF, (size_t)nq * kq * 4, st));
    Scratch sx(st);
    int32_t* unsafe_cnt = nullptr;
    if ((rc = sx.alloc((void**)&unsafe_cnt, 8))) return rc;
    CU_TRY(cudaMemsetAsync(unsafe_cnt, 0, 8, st));
    int dpad = 4;
    while (dpad < d) dpad <<= 1;
    const int L = kq <= 26 ? 32 : 64;
    const bool fast = dpad <= 64 && kq <= 56;
    const double eps_ffma = (2.0 * d + 8.0) * 5.9604644775390625e-08 * 1.01;
    const char* tc_env = getenv("GEOCB_KNN_TC");
    const bool use_tc = d <= 64 && kq <= 26 && !(tc_env && tc_env[0] == '0');
    if (use_tc) {
        // tcgen05 filter (knn_tc.cu) -> same exact re-rank / certificate / fallback as the FFMA2 path
        const int64_t Mp = knn_tc_padded_candidates(M), nqp = knn_tc_padded_queries(nq, d);
        float *bimg, *aimg, *cand_thr;
        int32_t *cand_idx, *unsafe_rows;
        if ((rc = sx.alloc((void**)&bimg, (size_t)knn_tc_image_floats(Mp, d) * 4)) ||
            (rc = sx.alloc((void**)&aimg, (size_t)knn_tc_image_floats(nqp, d) * 4)) ||
            (rc = sx.alloc((void**)&cand_idx, (size_t)nq * 32 * 4)) || (rc = sx.alloc((void**)&cand_thr, (size_t)nq * 4)) ||
            (rc = sx.alloc((void**)&unsafe_rows, (size_t)nq * 4)))
            return rc;
        unsigned int* max_nrm = (unsigned int*)(unsafe_cnt + 1);
        if ((rc = knn_filter_tc(z, zq, M, d, nq, bimg, aimg, max_nrm, cand_idx, cand_thr, st))) return rc;
        const unsigned rb = (unsigned)((nq * 32 + 255) / 256);
        knn_rerank_kernel<32><<<rb, 256, 0, st>>>(z, zq, M, d, nq, kq, cand_idx, cand_thr, max_nrm,
                                                   knn_tc_eps_unit(), out_idx, out_dist, unsafe_rows,
                                                   unsafe_cnt);
        KERNEL_CHECK();
        knn_exact_kernel<<<di.sm_count * 4, 256, 0, st>>>(z, zq, M, d, kq, unsafe_rows, unsafe_cnt, 0, out_idx,
                                                           out_dist);
        KERNEL_CHECK();
    } else if (fast) {
        const int tile_c = kTileBytes / (dpad * 4);
        const int64_t Mp = align_up(M, tile_c);
        float *zt, *nrm, *cand_thr;
        int32_t *cand_idx, *unsafe_rows;
        if ((rc = sx.alloc((void**)&zt, (size_t)Mp * dpad * 4)) || (rc = sx.alloc((void**)&nrm, (size_t)Mp * 4)) ||
            (rc = sx.alloc((void**)&cand_idx, (size_t)nq * L * 4)) || (rc = sx.alloc((void**)&cand_thr, (size_t)nq * 4)) ||
            (rc = sx.alloc((void**)&unsafe_rows, (size_t)nq * 4)))
            return rc;
        unsigned int* max_nrm = (unsigned int*)(unsafe_cnt + 1);
        knn_prep_kernel<<<(unsigned)((Mp + 255) / 256), 256, 0, st>>>(z, M, d, dpad, Mp, zt, nrm, max_nrm);
        KERNEL_CHECK();
        rc = L == 32 ? dispatch_filter<32>(dpad, zq, d, zt, nrm, Mp, nq, cand_idx, cand_thr, st)
                     : dispatch_filter<64>(dpad, zq, d, zt, nrm, Mp, nq, cand_idx, cand_thr, st);
        if (rc) return rc;
        const unsigned rb = (unsigned)((nq * 32 + 255) / 256);
        if (L == 32)
            knn_rerank_kernel<32><<<rb, 256, 0, st>>>(z, zq, M, d, nq, kq, cand_idx, cand_thr, max_nrm, eps_ffma,
                                                       out_idx, out_dist, unsafe_rows, unsafe_cnt);
        else
            knn_rerank_kernel<64><<<rb, 256, 0, st>>>(z, zq, M, d, nq, kq, cand_idx, cand_thr, max_nrm, eps_ffma,
                                                       out_idx, out_dist, unsafe_rows, unsafe_cnt);
        KERNEL_CHECK();
        knn_exact_kernel<<<di.sm_count * 4, 256, 0, st>>>(z, zq, M, d, kq, unsafe_rows, unsafe_cnt, 0, out_idx,
                                                           out_dist);
        KERNEL_CHECK();
    } else {
        ARG_CHECK(nq < (1LL << 31));
        knn_exact_kernel<<<di.sm_count * 4, 256, 0, st>>>(z, zq, M, d, kq, nullptr, nullptr, (int)nq, out_idx,
                                                           out_dist);
        KERNEL_CHECK();
        // report every row as having taken the exact path
        int32_t nqi = (int32_t)nq;
        CU_TRY(cudaMemcpyAsync(unsafe_cnt, &nqi, 4, cudaMemcpyHostToDevice, st));
    }
    if (n_unsafe) CU_TRY(cudaMemcpyAsync(n_unsafe, unsafe_cnt, 4, cudaMemcpyDeviceToDevice, st));
    return 0;
}

GEOCB_API int geocb_knn_topk(const float* z, int64_t M, int d, int64_t q_lo, int64_t q_hi, int kq, int32_t* out_idx,
                             float* out_dist, int32_t* n_unsafe, void* stream) {
    ARG_CHECK(z && out_idx && M > 0 && d > 0 && q_lo >= 0 && q_hi <= M && q_lo < q_hi && kq > 0 && kq <= M);
    ARG_CHECK(M < (1LL << 31) - 2048);
    return knn_query_impl(z, z + q_lo * d, M, d, q_hi - q_lo, kq, out_idx, out_dist, n_unsafe, (cudaStream_t)stream);
}

GEOCB_API int geocb_knn_query(const float* queries, int64_t nq, const float* cand, int64_t M, int d, int kq,
                              int32_t* out_idx, float* out_dist, int32_t* n_unsafe, void* stream) {
    ARG_CHECK(queries && cand && out_idx && nq > 0 && M > 0 && d > 0 && kq > 0 && kq <= M);
    ARG_CHECK(M < (1LL << 31) - 2048 && nq < (1LL << 31));
    return knn_query_impl(cand, queries, M, d, nq, kq, out_idx, out_dist, n_unsafe, (cudaStream_t)stream);
}
