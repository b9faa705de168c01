// bn_jvp.cu -- BatchNorm2d (train mode) forward-mode JVP fused with the following ReLU, in place.
//
// Part of the Riemannian edge-length stage (reference: src/geo/riemannian_metric.py:12-35 differentiates
// sigmoid(decoder(z)) with torch.autograd.functional.jvp; the reference CLI leaves the decoder in train mode,
// build_codebook.py:27-29, so BatchNorm normalises with the statistics of each 512-edge batch and the JVP
// differentiates through them).  For a batch ("group") of R rows and channel c with HW positions:
//     mu = mean(x)  var = mean((x-mu)^2)  r = 1/sqrt(var+eps)  xh = (x-mu) r
//     y  = xh w + b                         ty = w r ( t - mean(t) - xh mean(xh t) )
//     ReLU: ty = y > 0 ? ty : 0 ; y = max(y, 0)
// The PyTorch formulation makes ~14 passes over the (rows, C*HW) activations; this kernel reads x and t once and
// writes them once.  One CTA = one group x one 16-float column slab (16/HW channels): HBM-bound, 16 B per float.
#include "common.cuh"

namespace {
constexpr int kBnThreads = 256;

template <int NRED>
__device__ __forceinline__ void block_sum(float (&v)[NRED], float* s_red /* [8][NRED] */) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < NRED; ++i)
#pragma unroll
        for (int o = 16; o; o >>= 1) v[i] += __shfl_xor_sync(0xffffffffu, v[i], o);
    __syncthreads();  // previous use of s_red is over
    if (lane == 0)
#pragma unroll
        for (int i = 0; i < NRED; ++i) s_red[warp * NRED + i] = v[i];
    __syncthreads();
#pragma unroll
    for (int i = 0; i < NRED; ++i) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < kBnThreads / 32; ++w) s += s_red[w * NRED + i];
        v[i] = s;
    }
}

// RPT rows per thread, CPB channels per slab (HW = 16 / CPB).
// KIN > 0: the linear map in front of this BatchNorm is fused in -- x / t are then its INPUT rows [rows][KIN] (primal
// and tangent latents) and w_in [KIN][ld] its matrix: the slab's pre-activations are KIN sequential FMAs per value (k
// ascending, the order of a SIMT SGEMM), so the [rows][ld] pre-activation matrix is never written or read.
template <int RPT, int CPB, int KIN>
__global__ void __launch_bounds__(kBnThreads) bn_relu_jvp_kernel(float* __restrict__ x, float* __restrict__ t, int R,
                                                                 int ld, const float* __restrict__ pre_bias,
                                                                 const float* __restrict__ weight,
                                                                 const float* __restrict__ bias, float eps, int relu,
                                                                 float* __restrict__ img, int64_t img_row_t,
                                                                 const float* __restrict__ w_in) {
    constexpr int HW = 16 / CPB;
    __shared__ float s_red[(kBnThreads / 32) * 2 * CPB];
    __shared__ float s_w[(KIN > 0 ? KIN : 1) * 16];
    const int col0 = blockIdx.x * 16;
    const int64_t row0 = (int64_t)blockIdx.y * R;
    float xv[RPT][16], tv[RPT][16];
    float pb[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) pb[i] = pre_bias ? pre_bias[col0 + i] : 0.f;
    if (KIN > 0) {
        for (int i = threadIdx.x; i < KIN * 16; i += kBnThreads) s_w[i] = w_in[(int64_t)(i >> 4) * ld + col0 + (i & 15)];
        __syncthreads();
    }
#pragma unroll
    for (int k = 0; k < RPT; ++k) {
        const int r = threadIdx.x + k * kBnThreads;
        if (KIN > 0) {
#pragma unroll
            for (int i = 0; i < 16; ++i) xv[k][i] = 0.f, tv[k][i] = 0.f;
            if (r < R) {
                const float4* pz = (const float4*)(x + (row0 + r) * KIN);
                const float4* pd = (const float4*)(t + (row0 + r) * KIN);
#pragma unroll
                for (int j4 = 0; j4 < KIN / 4; ++j4) {
                    const float4 a = pz[j4], b = pd[j4];
                    const float za[4] = {a.x, a.y, a.z, a.w}, zb[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
                    for (int jj = 0; jj < 4; ++jj)
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            const float w = s_w[(4 * j4 + jj) * 16 + i];
                            xv[k][i] = __fmaf_rn(za[jj], w, xv[k][i]);
                            tv[k][i] = __fmaf_rn(zb[jj], w, tv[k][i]);
                        }
                }
#pragma unroll
                for (int i = 0; i < 16; ++i) xv[k][i] += pb[i];
            }
        } else if (r < R) {
            const float4* px = (const float4*)(x + (row0 + r) * ld + col0);
            const float4* pt = (const float4*)(t + (row0 + r) * ld + col0);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float4 a = px[q], b = pt[q];
                xv[k][4 * q] = a.x + pb[4 * q], xv[k][4 * q + 1] = a.y + pb[4 * q + 1];
                xv[k][4 * q + 2] = a.z + pb[4 * q + 2], xv[k][4 * q + 3] = a.w + pb[4 * q + 3];
                tv[k][4 * q] = b.x, tv[k][4 * q + 1] = b.y, tv[k][4 * q + 2] = b.z, tv[k][4 * q + 3] = b.w;
            }
        } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) xv[k][i] = 0.f, tv[k][i] = 0.f;
        }
    }
    const float inv_cnt = 1.0f / (float)((int64_t)R * HW);
    // ---- mean ----
    float acc[2 * CPB];
#pragma unroll
    for (int c = 0; c < CPB; ++c) {
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < RPT; ++k)
#pragma unroll
            for (int h = 0; h < HW; ++h) s += xv[k][c * HW + h];
        acc[c] = s;
        acc[CPB + c] = 0.f;
    }
    block_sum<2 * CPB>(acc, s_red);
    float mu[CPB];
#pragma unroll
    for (int c = 0; c < CPB; ++c) mu[c] = acc[c] * inv_cnt;
    // ---- biased variance of the centred values ----
#pragma unroll
    for (int c = 0; c < CPB; ++c) {
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < RPT; ++k) {
            const bool live = threadIdx.x + k * kBnThreads < R;
#pragma unroll
            for (int h = 0; h < HW; ++h) {
                const float d = xv[k][c * HW + h] - mu[c];
                xv[k][c * HW + h] = d;
                s += live ? d * d : 0.f;
            }
        }
        acc[c] = s;
        acc[CPB + c] = 0.f;
    }
    block_sum<2 * CPB>(acc, s_red);
    float rinv[CPB];
#pragma unroll
    for (int c = 0; c < CPB; ++c) rinv[c] = 1.0f / sqrtf(acc[c] * inv_cnt + eps);
    // ---- tangent moments: mean(t), mean(xh t) ----
#pragma unroll
    for (int c = 0; c < CPB; ++c) {
        float s0 = 0.f, s1 = 0.f;
#pragma unroll
        for (int k = 0; k < RPT; ++k)
#pragma unroll
            for (int h = 0; h < HW; ++h) {
                const float xh = xv[k][c * HW + h] * rinv[c];
                xv[k][c * HW + h] = xh;
                s0 += tv[k][c * HW + h];
                s1 += xh * tv[k][c * HW + h];
            }
        acc[c] = s0;
        acc[CPB + c] = s1;
    }
    block_sum<2 * CPB>(acc, s_red);
    // ---- normalise, affine, ReLU, write back ----
#pragma unroll
    for (int k = 0; k < RPT; ++k) {
        const int r = threadIdx.x + k * kBnThreads;
        if (r >= R) continue;
        float ox[16], ot[16];
#pragma unroll
        for (int c = 0; c < CPB; ++c) {
            const int ch = blockIdx.x * CPB + c;
            const float w = weight ? weight[ch] : 1.f, b = bias ? bias[ch] : 0.f;
            const float mt = acc[c] * inv_cnt, m2 = acc[CPB + c] * inv_cnt;
#pragma unroll
            for (int h = 0; h < HW; ++h) {
                const float xh = xv[k][c * HW + h];
                float y = xh * w + b;
                float ty = (rinv[c] * ((tv[k][c * HW + h] - mt) - xh * m2)) * w;
                if (relu) {
                    ty = y > 0.f ? ty : 0.f;
                    y = fmaxf(y, 0.f);
                }
                ox[c * HW + h] = y;
                ot[c * HW + h] = ty;
            }
        }
        if (img) {
            // The next layer is the tcgen05 3xTF32 GEMM (jvp_gemm.cu): write its operand image directly -- tf32 hi | lo
            // halves of this 16-float slab (= one K-chunk) in the UMMA canonical layout -- instead of the row-major
            // activations.  Primal rows sit at [0, rows), tangent rows at [img_row_t, img_row_t + rows) of the GEMM's A.
            const int nchunks = ld / 16;
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                const int64_t arow = (half ? img_row_t : 0) + row0 + r;
                float* base = img + (((arow >> 7) * nchunks + blockIdx.x) * 2) * 2048 + (arow & 127) * 4;
                const float* v = half ? ot : ox;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    float hi[4], lo[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        uint32_t hb;
                        asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hb) : "f"(v[4 * q + i]));
                        hi[i] = __uint_as_float(hb);
                        lo[i] = v[4 * q + i] - hi[i];
                    }
                    *(float4*)(base + q * 512) = make_float4(hi[0], hi[1], hi[2], hi[3]);
                    *(float4*)(base + q * 512 + 2048) = make_float4(lo[0], lo[1], lo[2], lo[3]);
                }
            }
            continue;
        }
        float4* px = (float4*)(x + (row0 + r) * ld + col0);
        float4* pt = (float4*)(t + (row0 + r) * ld + col0);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            px[q] = make_float4(ox[4 * q], ox[4 * q + 1], ox[4 * q + 2], ox[4 * q + 3]);
            pt[q] = make_float4(ot[4 * q], ot[4 * q + 1], ot[4 * q + 2], ot[4 * q + 3]);
        }
    }
}

template <int RPT, int KIN>
int launch_bn(int cpb, dim3 grid, cudaStream_t st, float* x, float* t, int R, int ld, const float* pre_bias,
              const float* weight, const float* bias, float eps, int relu, float* img, int64_t img_row_t,
              const float* w_in) {
    switch (cpb) {
        case 1: bn_relu_jvp_kernel<RPT, 1, KIN><<<grid, kBnThreads, 0, st>>>(x, t, R, ld, pre_bias, weight, bias, eps, relu, img, img_row_t, w_in); break;
        case 2: bn_relu_jvp_kernel<RPT, 2, KIN><<<grid, kBnThreads, 0, st>>>(x, t, R, ld, pre_bias, weight, bias, eps, relu, img, img_row_t, w_in); break;
        case 4: bn_relu_jvp_kernel<RPT, 4, KIN><<<grid, kBnThreads, 0, st>>>(x, t, R, ld, pre_bias, weight, bias, eps, relu, img, img_row_t, w_in); break;
        case 8: bn_relu_jvp_kernel<RPT, 8, KIN><<<grid, kBnThreads, 0, st>>>(x, t, R, ld, pre_bias, weight, bias, eps, relu, img, img_row_t, w_in); break;
        case 16: bn_relu_jvp_kernel<RPT, 16, KIN><<<grid, kBnThreads, 0, st>>>(x, t, R, ld, pre_bias, weight, bias, eps, relu, img, img_row_t, w_in); break;
        default: return GEOCB_ERR_UNSUPPORTED;
    }
    KERNEL_CHECK();
    return 0;
}
}  // namespace

static int bn_impl(float* x, float* t, int64_t rows, int C, int HW, int rows_per_group, const float* pre_bias,
                   const float* weight, const float* bias, float eps, int relu, float* img, int64_t img_row_t, void* stream,
                   const float* w_in = nullptr, int kin = 0) {
    ARG_CHECK(x && t && rows > 0 && C > 0 && HW > 0 && rows_per_group > 0);
    if (!(HW == 1 || HW == 2 || HW == 4 || HW == 8 || HW == 16) || (C * HW) % 16 != 0 || rows % rows_per_group != 0 ||
        rows_per_group > 2 * kBnThreads || rows / rows_per_group > 65535) {
        geocb_set_error("bn_relu_jvp: unsupported shape (C=%d HW=%d rows=%lld group=%d)", C, HW, (long long)rows,
                        rows_per_group);
        return GEOCB_ERR_UNSUPPORTED;
    }
    const dim3 grid((unsigned)(C * HW / 16), (unsigned)(rows / rows_per_group));
    cudaStream_t st = (cudaStream_t)stream;
    const int cpb = 16 / HW, ld = C * HW;
    const bool one = rows_per_group <= kBnThreads;
#define BN_LAUNCH(KIN)                                                                                                     \
    (one ? launch_bn<1, KIN>(cpb, grid, st, x, t, rows_per_group, ld, pre_bias, weight, bias, eps, relu, img, img_row_t, w_in) \
         : launch_bn<2, KIN>(cpb, grid, st, x, t, rows_per_group, ld, pre_bias, weight, bias, eps, relu, img, img_row_t, w_in))
    if (kin == 0) return BN_LAUNCH(0);
    if (kin == 16) return BN_LAUNCH(16);
    if (kin == 32) return BN_LAUNCH(32);
#undef BN_LAUNCH
    geocb_set_error("lin_bn_relu_jvp: fused input width %d not supported (16 or 32)", kin);
    return GEOCB_ERR_UNSUPPORTED;
}

GEOCB_API int geocb_bn_relu_jvp(float* x, float* t, int64_t rows, int C, int HW, int rows_per_group,
                                const float* pre_bias, const float* weight, const float* bias, float eps, int relu,
                                void* stream) {
    return bn_impl(x, t, rows, C, HW, rows_per_group, pre_bias, weight, bias, eps, relu, nullptr, 0, stream);
}

GEOCB_API int geocb_bn_relu_jvp_to_image(const float* x, const float* t, int64_t rows, int C, int HW, int rows_per_group,
                                         const float* pre_bias, const float* weight, const float* bias, float eps,
                                         int relu, float* aimg, int64_t img_row_t, void* stream) {
    ARG_CHECK(aimg && img_row_t >= rows && rows % 128 == 0 && img_row_t % 128 == 0);
    return bn_impl((float*)x, (float*)t, rows, C, HW, rows_per_group, pre_bias, weight, bias, eps, relu, aimg, img_row_t,
                   stream);
}

// The linear map in front of the BatchNorm fused in (kin = 16 or 32 input features, e.g. the latent dimension):
// zin / tin [rows][kin] primal and tangent inputs, w_in [kin][C*HW] row-major; writes the next GEMM's operand image.
GEOCB_API int geocb_lin_bn_relu_jvp_to_image(const float* zin, const float* tin, int kin, const float* w_in, int64_t rows,
                                             int C, int HW, int rows_per_group, const float* pre_bias, const float* weight,
                                             const float* bias, float eps, int relu, float* aimg, int64_t img_row_t,
                                             void* stream) {
    ARG_CHECK(zin && tin && w_in && aimg && img_row_t >= rows && rows % 128 == 0 && img_row_t % 128 == 0);
    return bn_impl((float*)zin, (float*)tin, rows, C, HW, rows_per_group, pre_bias, weight, bias, eps, relu, aimg,
                   img_row_t, stream, w_in, kin);
}

// ---------------------------------------------------------------- last layer ----
// out[r][p] = s (1 - s) (t[r,:] . W[:,p]),  s = sigmoid(x[r,:] . W[:,p] + b[p])      (riemannian_metric.py:29-35: the
// tangent of sigmoid(decoder(z)) through the decoder's last linear map) for a narrow output (P = 16: the 4x4 patch a
// 1x1 latent cell decodes to).  A [rows x K] x [K x 16] product is memory-bound (it reads x and t once: 16 FMAs per
// loaded float); the library's SIMT SGEMM needs 195 us for 131 072 x 1024 x 16 where the reads take 82 us, and it is
// the only library GEMM left on the path (the handle alone costs 0.1 s per process).
namespace {
constexpr int kFinP = 16, kFinRows = 64, kFinKC = 128, kFinThreads = 256, kFinLd = kFinKC + 4;

// CTA = 64 rows, 256 threads: thread = 4 rows x 1 output column, operands staged through shared memory in K-chunks of
// 128, W transposed there so that both operands are read 16 bytes at a time; four partial sums per value (k mod 4),
// combined pairwise.  223 us for 65 536 x 1024 rows of x and t, the same as the library SGEMM + its three elementwise
// kernels; other tilings measured slower (4 x 2 outputs per thread, 39 KB CTAs: 291 us; a quad of lanes per row with
// all 16 outputs in registers and W whole in shared memory: 506 us) - the shared-memory reads bound all of them.
__global__ void __launch_bounds__(kFinThreads) lin_sigmoid_jvp_kernel(const float* __restrict__ x, const float* __restrict__ t,
                                                                       int64_t rows, int K, const float* __restrict__ w,
                                                                       const float* __restrict__ bias,
                                                                       float* __restrict__ out) {
    extern __shared__ __align__(16) float s_fin[];
    float* xs = s_fin;                       // [64][132]
    float* ts = xs + kFinRows * kFinLd;      // [64][132]
    float* wt = ts + kFinRows * kFinLd;      // [16][132]
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const int64_t row0 = (int64_t)blockIdx.x * kFinRows;
    float4 ax[4], at[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) ax[i] = at[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int k0 = 0; k0 < K; k0 += kFinKC) {
        __syncthreads();  // the previous chunk has been consumed
#pragma unroll
        for (int j = 0; j < (kFinRows * kFinKC / 4) / kFinThreads; ++j) {
            const int idx = threadIdx.x + j * kFinThreads;
            const int r = idx >> 5, c4 = idx & 31;
            float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
            if (row0 + r < rows) {
                a = *(const float4*)(x + (row0 + r) * K + k0 + 4 * c4);
                b = *(const float4*)(t + (row0 + r) * K + k0 + 4 * c4);
            }
            *(float4*)(xs + r * kFinLd + 4 * c4) = a;
            *(float4*)(ts + r * kFinLd + 4 * c4) = b;
        }
#pragma unroll
        for (int j = 0; j < (kFinKC * kFinP) / kFinThreads; ++j) {
            const int idx = threadIdx.x + j * kFinThreads;  // = k * 16 + p
            wt[(idx & 15) * kFinLd + (idx >> 4)] = w[(int64_t)k0 * kFinP + idx];
        }
        __syncthreads();
#pragma unroll 8
        for (int k4 = 0; k4 < kFinKC / 4; ++k4) {
            const float4 wv = *(const float4*)(wt + tx * kFinLd + 4 * k4);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float4 a = *(const float4*)(xs + (4 * ty + i) * kFinLd + 4 * k4);
                const float4 b = *(const float4*)(ts + (4 * ty + i) * kFinLd + 4 * k4);
                ax[i].x = __fmaf_rn(a.x, wv.x, ax[i].x), ax[i].y = __fmaf_rn(a.y, wv.y, ax[i].y);
                ax[i].z = __fmaf_rn(a.z, wv.z, ax[i].z), ax[i].w = __fmaf_rn(a.w, wv.w, ax[i].w);
                at[i].x = __fmaf_rn(b.x, wv.x, at[i].x), at[i].y = __fmaf_rn(b.y, wv.y, at[i].y);
                at[i].z = __fmaf_rn(b.z, wv.z, at[i].z), at[i].w = __fmaf_rn(b.w, wv.w, at[i].w);
            }
        }
    }
    const float bp = bias ? bias[tx] : 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int64_t r = row0 + 4 * ty + i;
        if (r >= rows) continue;
        const float y = ((ax[i].x + ax[i].y) + (ax[i].z + ax[i].w)) + bp;
        const float tt = (at[i].x + at[i].y) + (at[i].z + at[i].w);
        const float s = 1.0f / (1.0f + expf(-y));
        out[r * kFinP + tx] = (s * (1.0f - s)) * tt;
    }
}
}  // namespace

GEOCB_API int geocb_lin_sigmoid_jvp(const float* x, const float* t, int64_t rows, int K, int P, const float* w,
                                    const float* bias, float* out, void* stream) {
    ARG_CHECK(x && t && w && out && rows > 0 && K > 0);
    if (P != kFinP || K % kFinKC != 0) {
        geocb_set_error("lin_sigmoid_jvp: P = %d, K = %d not supported (P = 16, K a multiple of 128)", P, K);
        return GEOCB_ERR_UNSUPPORTED;
    }
    const size_t smem = (size_t)(2 * kFinRows + kFinP) * kFinLd * sizeof(float);
    CU_TRY(cudaFuncSetAttribute(lin_sigmoid_jvp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    lin_sigmoid_jvp_kernel<<<(unsigned)((rows + kFinRows - 1) / kFinRows), kFinThreads, smem, (cudaStream_t)stream>>>(
        x, t, rows, K, w, bias, out);
    KERNEL_CHECK();
    return 0;
}
