// edge.cu -- Riemannian edge-length reduce + scatter into both CSR directions.
// Replaces torch.linalg.vector_norm(dim=1) x2 and 0.5*(a+b) (src/geo/riemannian_metric.py:60-63)
// plus csr_matrix((L,(i,j))) + W.T of src/scripts/build_codebook.py:53-54.
// HBM-bound: reads 2*P*4 B per edge, writes 12 B.  One warp per edge row (P = 16 or 192 here).
#include "common.cuh"

namespace {
__global__ void edge_len_kernel(const float* __restrict__ ja, const float* __restrict__ jb, int64_t E, int P,
                                int64_t e_lo, const int64_t* pos_ij, const int64_t* pos_ji, float* len_out,
                                float* data) {
    const int lanes = P <= 16 ? 8 : 32;  // sub-warp per edge for short rows
    const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t e = gid / lanes;
    const int lane = (int)(gid % lanes);
    float sa = 0.f, sb = 0.f;
    if (e < E) {
        const float* a = ja + e * P;
        const float* b = jb + e * P;
        for (int i = lane; i < P; i += lanes) {
            sa = fmaf(a[i], a[i], sa);
            sb = fmaf(b[i], b[i], sb);
        }
    }
    for (int o = lanes / 2; o; o >>= 1) {
        sa += __shfl_xor_sync(0xffffffffu, sa, o);
        sb += __shfl_xor_sync(0xffffffffu, sb, o);
    }
    if (e < E && lane == 0) {
        const float L = 0.5f * (sqrtf(sa) + sqrtf(sb));
        if (len_out) len_out[e_lo + e] = L;
        if (data && pos_ij) {
            data[pos_ij[e_lo + e]] = L;
            const int64_t pj = pos_ji[e_lo + e];
            if (pj >= 0) data[pj] = L;
        }
    }
}
}  // namespace

GEOCB_API int geocb_edge_len_scatter(const float* ja, const float* jb, int64_t E, int P, int64_t e_lo,
                                     const int64_t* pos_ij, const int64_t* pos_ji, float* len_out, float* data,
                                     void* stream) {
    ARG_CHECK(ja && jb && E >= 0 && P > 0 && e_lo >= 0 && (len_out || data));
    ARG_CHECK(!data || (pos_ij && pos_ji));
    if (E == 0) return 0;
    const int lanes = P <= 16 ? 8 : 32;
    const int64_t threads = E * lanes;
    edge_len_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, (cudaStream_t)stream>>>(ja, jb, E, P, e_lo, pos_ij,
                                                                                         pos_ji, len_out, data);
    KERNEL_CHECK();
    return 0;
}
