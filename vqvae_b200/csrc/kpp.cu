// kpp.cu -- k-means++ draw and NumPy-exact float32 reductions.
//
// Replaces the NumPy arithmetic of src/geo/kmeans_optimized.py:47-61 (one k-means++ draw) and
// :129-134 (quantisation error).  Bit-exactness matters: a draw is  searchsorted(cumsum(p), u)
// over n boundaries, so a 1-ulp difference in the normaliser can move the chosen centre and every
// later one.  Reproduced exactly:
//   * np.sum(float32[n])  = FLOAT_pairwise_sum: recursive halving (n2 = n/2 rounded down to a
//     multiple of 8) down to blocks of <=128 elements summed with 8 strided accumulators combined
//     as ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)), remainder added sequentially.  The host builds the
//     tree once per n ("plan"); leaves are summed 8 lanes per leaf, inner nodes level by level.
//   * RandomState.choice(p): cdf = cumsum(float64(p)) sequential; cdf /= cdf[-1];
//     searchsorted(u, 'right').  A parallel fp64 scan locates the candidate; if u*cdf[-1] is within
//     3 n 2^-53 cdf[-1] of the neighbouring boundaries the draw is recomputed by a strictly
//     sequential kernel (same operation order as NumPy).
#include <vector>
#include "common.cuh"

namespace {

constexpr int kLeafMax = 128;  // NumPy PW_BLOCKSIZE

// ------------------------------------------------------------- plan (host) ----
struct PlanBuilder {
    std::vector<int64_t> leaf_off, leaf_len;          // leaves
    std::vector<int64_t> node_l, node_r, node_h;      // internal nodes (ids: leaf i -> i ; node j -> -(j+1))
    int64_t build(int64_t off, int64_t n, int64_t* height) {
        if (n <= kLeafMax) {
            leaf_off.push_back(off);
            leaf_len.push_back(n);
            *height = 0;
            return (int64_t)leaf_off.size() - 1;
        }
        int64_t n2 = n / 2;
        n2 -= n2 % 8;
        int64_t hl, hr;
        int64_t l = build(off, n2, &hl);
        int64_t r = build(off + n2, n - n2, &hr);
        node_l.push_back(l);
        node_r.push_back(r);
        int64_t h = 1 + (hl > hr ? hl : hr);
        node_h.push_back(h);
        *height = h;
        return -(int64_t)node_l.size();
    }
};
// plan words: [0]=n [1]=n_leaves [2]=n_nodes [3]=n_levels [4]=root slot
//   then leaves (off,len) x n_leaves ; level offsets (n_levels+1) ; nodes sorted by height (l slot, r slot)
//   value slots: leaf i -> i ; sorted node j -> n_leaves + j
int64_t plan_words(int64_t n) {
    if (n <= 0) return 8;
    PlanBuilder b;
    int64_t h;
    b.build(0, n, &h);
    return 5 + 2 * (int64_t)b.leaf_off.size() + (h + 2) + 2 * (int64_t)b.node_l.size();
}
void plan_fill(int64_t n, int64_t* out) {
    PlanBuilder b;
    int64_t h = 0;
    int64_t root = n > 0 ? b.build(0, n, &h) : 0;
    const int64_t nl = b.leaf_off.size(), nn = b.node_l.size();
    // order nodes by height (stable): level k (0-based) holds the nodes of height k+1
    std::vector<int64_t> order(nn), slot_of(nn), start(h + 1, 0), cur(h + 1, 0);
    for (int64_t j = 0; j < nn; ++j) start[b.node_h[j]]++;          // start[k] = #nodes of height k (k>=1)
    {
        int64_t run = 0;
        for (int64_t k = 1; k <= h; ++k) {
            int64_t c = start[k];
            start[k - 1] = run;  // first position of height k
            run += c;
        }
    }
    for (int64_t k = 1; k <= h; ++k) cur[k] = start[k - 1];
    for (int64_t j = 0; j < nn; ++j) {
        int64_t pos = cur[b.node_h[j]]++;
        order[pos] = j;
        slot_of[j] = nl + pos;
    }
    auto slot = [&](int64_t id) { return id >= 0 ? id : slot_of[-id - 1]; };
    out[0] = n;
    out[1] = nl;
    out[2] = nn;
    out[3] = h;
    out[4] = n > 0 ? slot(root) : 0;
    int64_t* p = out + 5;
    for (int64_t i = 0; i < nl; ++i) {
        *p++ = b.leaf_off[i];
        *p++ = b.leaf_len[i];
    }
    for (int64_t k = 1; k <= h; ++k) *p++ = start[k - 1];
    *p++ = nn;
    for (int64_t pos = 0; pos < nn; ++pos) {
        int64_t j = order[pos];
        *p++ = slot(b.node_l[j]);
        *p++ = slot(b.node_r[j]);
    }
}

// ----------------------------------------------------------- device kernels ----
// elementwise source of the summed values
enum { SRC_PLAIN = 0, SRC_SQUARE = 1, SRC_KPP = 2 };
struct ElemSrc {
    int mode;
    const float* x;            // PLAIN / SQUARE: the array ; KPP: dmin32
    const uint8_t* is_center;  // KPP
    const unsigned int* max_finite_bits;  // KPP: max finite dmin32 (bits), 0xFFFFFFFF if none finite
    float* p_out;              // KPP: un-normalised probabilities written here
};
__device__ __forceinline__ float elem(const ElemSrc& s, int64_t i) {
    if (s.mode == SRC_PLAIN) return s.x[i];
    if (s.mode == SRC_SQUARE) return __fmul_rn(s.x[i], s.x[i]);
    // kmeans_optimized.py:47-57
    const unsigned mb = *s.max_finite_bits;
    float d = s.x[i], safe;
    if (mb == 0xFFFFFFFFu) safe = 1.0f;                                       // no finite distance: ones_like
    else safe = isfinite(d) ? d : __fmul_rn(__uint_as_float(mb), 2.0f);      // inf -> 2*max_finite
    float p = s.is_center[i] ? 0.0f : __fmul_rn(safe, safe);
    s.p_out[i] = p;
    return p;
}

__global__ void max_finite_kernel(const float* d, int64_t n, unsigned int* out_bits, unsigned int* any_finite) {
    unsigned m = 0;
    bool any = false;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        float v = d[i];
        if (isfinite(v)) {
            any = true;
            m = max(m, __float_as_uint(v));  // distances are >= 0: bit order == value order
        }
    }
    for (int o = 16; o; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    unsigned a = __ballot_sync(0xffffffffu, any);
    if ((threadIdx.x & 31) == 0 && a) {
        atomicMax(out_bits, m);
        atomicOr(any_finite, 1u);
    }
}
__global__ void finalize_max_kernel(unsigned int* out_bits, const unsigned int* any_finite) {
    if (!*any_finite) *out_bits = 0xFFFFFFFFu;
}

// 8 lanes per leaf: lane j owns NumPy's accumulator r[j]
__global__ void leaf_sum_kernel(ElemSrc src, const int64_t* plan, float* vals) {
    const int64_t nl = plan[1];
    const int64_t leaf = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3;
    const int j = threadIdx.x & 7;
    const bool live = leaf < nl;
    const int64_t off = live ? plan[5 + 2 * leaf] : 0;
    const int64_t len = live ? plan[6 + 2 * leaf] : 0;
    // all lanes run the shuffles (groups of a warp may hold leaves of different shapes)
    const int64_t body = len >= 8 ? len - (len % 8) : 0;
    float r = 0.f;
    if (body) {
        r = elem(src, off + j);
        for (int64_t i = 8; i < body; i += 8) r = __fadd_rn(r, elem(src, off + i + j));
    }
    // ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7))
    float t = __fadd_rn(r, __shfl_down_sync(0xffffffffu, r, 1, 8));  // valid on even lanes
    float q = __fadd_rn(t, __shfl_down_sync(0xffffffffu, t, 2, 8));  // valid on lanes 0,4
    float res = __fadd_rn(q, __shfl_down_sync(0xffffffffu, q, 4, 8));  // valid on lane 0
    if (!body) res = 0.f;  // n < 8: plain sequential loop from 0
    if (j == 0)
        for (int64_t i = body; i < len; ++i) res = __fadd_rn(res, elem(src, off + i));
    if (live && j == 0) vals[leaf] = res;
}
// inner nodes, one level after another (single CTA)
__global__ void tree_sum_kernel(const int64_t* plan, float* vals, float* total_out) {
    const int64_t nl = plan[1], nlev = plan[3];
    const int64_t* lvl = plan + 5 + 2 * nl;
    const int64_t* nodes = lvl + nlev + 1;
    for (int64_t k = 0; k < nlev; ++k) {
        for (int64_t pos = lvl[k] + threadIdx.x; pos < lvl[k + 1]; pos += blockDim.x)
            vals[nl + pos] = __fadd_rn(vals[nodes[2 * pos]], vals[nodes[2 * pos + 1]]);
        __syncthreads();
    }
    if (threadIdx.x == 0) *total_out = plan[0] > 0 ? vals[plan[4]] : 0.f;
}

// ---- the draw ----
constexpr int kCdfThreads = 256;
constexpr int kCdfItems = 8;
constexpr int kCdfTile = kCdfThreads * kCdfItems;

__device__ __forceinline__ double block_scan_f64(double v, double* total) {
    __shared__ double ws[kCdfThreads / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double inc = v;
    for (int o = 1; o < 32; o <<= 1) {
        double t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) ws[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        double w = lane < kCdfThreads / 32 ? ws[lane] : 0.0;
        for (int o = 1; o < kCdfThreads / 32; o <<= 1) {
            double t = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o) w += t;
        }
        if (lane < kCdfThreads / 32) ws[lane] = w;
    }
    __syncthreads();
    double base = warp ? ws[warp - 1] : 0.0;
    *total = ws[kCdfThreads / 32 - 1];
    __syncthreads();
    return base + inc;  // inclusive
}
// q[i] = float64( p[i] / total )  (float32 divide, kmeans_optimized.py:60) ; per-tile fp64 sums
// `fine` counts the elements that can make the sequential fp64 cumsum round at all: q is a float32, i.e. a
// multiple of 2^-23*2^floor(log2 q); while every partial sum stays below 2 its ulp is at most 2^-52, so adding
// q >= 2^-28 to a multiple of 2^-52 is exact.  Only q < 2^-28 (and the <= 64 binade crossings of the running
// sum) can round, each by at most 2^-53: that gives the rigorous tolerance cdf_locate_kernel uses.
__global__ void cdf_tile_sums_kernel(const float* p, const float* total, int64_t n, double* tile_sums,
                                     unsigned long long* fine) {
    const float tot = *total;
    const int64_t base = (int64_t)blockIdx.x * kCdfTile + (int64_t)threadIdx.x * kCdfItems;
    double s = 0.0;
    unsigned nf = 0;
    for (int i = 0; i < kCdfItems; ++i)
        if (base + i < n) {
            const double q = (double)__fdiv_rn(p[base + i], tot);
            s += q;
            nf += (q > 0.0 && q < 3.7252902984619141e-09) ? 1u : 0u;  // 2^-28
        }
    nf = __reduce_add_sync(0xffffffffu, nf);
    if ((threadIdx.x & 31) == 0 && nf) atomicAdd(fine, (unsigned long long)nf);
    double t;
    block_scan_f64(s, &t);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = t;
}
// result: [0] index, [1] ambiguous flag
__global__ void cdf_locate_kernel(const float* p, const float* total, int64_t n, double u, double* tile_sums,
                                  int64_t ntiles, long long* result, const unsigned long long* fine) {
    __shared__ double s_excl, s_last;
    __shared__ long long s_tile;
    const float tot = *total;
    if (!(tot > 0.f)) {
        if (threadIdx.x == 0) {
            result[0] = -1;
            result[1] = 0;
        }
        return;
    }
    {
        // block-wide scan of the tile sums (each thread a contiguous run of tiles; one thread walking all ntiles twice
        // cost 57 us per draw at n = 2.94 M).  Any summation order of these non-negative terms is within the tolerance
        // below; a tile chosen one off at an exact tie ends in "not found" / "ambiguous" and the exact kernel decides.
        const int64_t per = (ntiles + kCdfThreads - 1) / kCdfThreads;
        const int64_t ta = (int64_t)threadIdx.x * per, tb = ta + per < ntiles ? ta + per : ntiles;
        double loc = 0.0;
        for (int64_t t = ta; t < tb; ++t) loc += tile_sums[t];
        double all;
        const double incl = block_scan_f64(loc, &all);
        if (threadIdx.x == 0) {
            s_last = all;
            s_tile = ntiles;  // "not found"
        }
        __syncthreads();
        const double target0 = u * all;
        double run = incl - loc;  // prefix before this thread's first tile
        for (int64_t t = ta; t < tb; ++t) {
            const double nxt = run + tile_sums[t];
            if (nxt > target0) {
                atomicMin((unsigned long long*)&s_tile, (unsigned long long)t);  // (prefixes are non-decreasing: first hit wins)
                break;
            }
            run = nxt;
        }
        __syncthreads();
        const long long tsel = s_tile;
        __syncthreads();
        if (tsel >= ntiles) {  // u * total at / after the last boundary: last tile, prefix before it
            if (threadIdx.x == (int)((ntiles - 1) / per)) {
                double r2 = incl - loc;
                for (int64_t t = ta; t < ntiles - 1; ++t) r2 += tile_sums[t];
                s_excl = r2;
                s_tile = ntiles - 1;
            }
        } else if (tsel >= ta && tsel < tb) {
            double r2 = incl - loc;
            for (int64_t t = ta; t < tsel; ++t) r2 += tile_sums[t];
            s_excl = r2;
        }
    }
    __syncthreads();
    const double last = s_last, target = u * last, excl = s_excl;
    // |sequential - exact| <= (fine + 64) * 2^-53 and |this scan - exact| <= (ntiles + 64) * 2^-53 while last < 2
    // (every addition of non-negative terms rounds by at most half an ulp of a sum below 2); the comparison
    // fl(c_i / c_last) > u differs from c_i > u * c_last by a few more ulps.  Doubled for slack.
    double tol = 3.0 * (double)n * 1.1102230246251565e-16 * last;  // worst-case bound, used if last >= 2
    if (last < 2.0) tol = 4.0 * ((double)*fine + (double)ntiles + 144.0) * 1.1102230246251565e-16;
    const int64_t base = s_tile * kCdfTile + (int64_t)threadIdx.x * kCdfItems;
    double q[kCdfItems], s = 0.0;
    for (int i = 0; i < kCdfItems; ++i) {
        q[i] = base + i < n ? (double)__fdiv_rn(p[base + i], tot) : 0.0;
        s += q[i];
    }
    double t;
    double run = excl + block_scan_f64(s, &t) - s;  // prefix before this thread's first item
    __shared__ long long s_idx;
    __shared__ int s_amb;
    if (threadIdx.x == 0) {
        s_idx = n;  // "not found"
        s_amb = 0;
    }
    __syncthreads();
    for (int i = 0; i < kCdfItems; ++i) {
        const double prev = run;
        run += q[i];
        if (base + i < n) {
            if (prev <= target && run > target) atomicMin(&s_idx, (long long)(base + i));
            // any boundary of this tile closer than tol to the target makes the draw ambiguous
            if (fabs(run - target) <= tol) atomicOr(&s_amb, 1);
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int amb = s_amb || fabs(excl - target) <= tol;
        if (s_idx >= n) amb = 1;  // u*last at/after the final boundary: let the exact kernel decide
        result[0] = s_idx >= n ? n - 1 : s_idx;
        result[1] = amb;
    }
}
// strictly sequential NumPy order; one warp streams values through shared memory, lane 0 adds
__global__ void cdf_exact_kernel(const float* p, const float* total, int64_t n, double u, long long* result) {
    __shared__ double buf[1024];
    const float tot = *total;
    double last = 0.0;
    for (int pass = 0; pass < 2; ++pass) {
        double run = 0.0;
        long long found = -1;
        for (int64_t base = 0; base < n && found < 0; base += 1024) {
            for (int i = threadIdx.x; i < 1024; i += blockDim.x)
                buf[i] = base + i < n ? (double)__fdiv_rn(p[base + i], tot) : 0.0;
            __syncthreads();
            if (threadIdx.x == 0) {
                const int m = (int)((n - base) < 1024 ? (n - base) : 1024);
                for (int i = 0; i < m; ++i) {
                    run += buf[i];
                    if (pass == 1 && found < 0 && run / last > u) found = base + i;
                }
                buf[0] = (double)found;
            }
            __syncthreads();
            found = (long long)buf[0];
            __syncthreads();
        }
        if (pass == 0) {
            if (threadIdx.x == 0) buf[1] = run;
            __syncthreads();
            last = buf[1];
            __syncthreads();
        } else if (threadIdx.x == 0) {
            result[0] = found < 0 ? n : found;  // searchsorted returns n when u >= cdf[-1] (cannot happen for u<1)
            result[1] = 2;
        }
    }
}

// ---- compaction of finite entries (QE path) ----
__global__ void finite_flag_kernel(const float* x, int64_t n, uint8_t* flag) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flag[i] = isfinite(x[i]) ? 1 : 0;
}
__global__ void finite_compact_kernel(const float* x, const uint8_t* flag, const int64_t* pre, int64_t n, float* out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && flag[i]) out[pre[i]] = x[i];
}

struct KppWs {
    int64_t p, vals, tile_sums, scalars, total;
};
KppWs kpp_layout(int64_t n) {
    KppWs w;
    int64_t off = 0;
    w.p = off;
    off += align_up(n * 4, 256);
    w.vals = off;
    off += align_up((2 * (n / 8 + 2) + 16) * 4, 256);  // leaves + inner nodes (leaves hold >= 8... <=128 elements)
    w.tile_sums = off;
    off += align_up(((n + kCdfTile - 1) / kCdfTile + 1) * 8, 256);
    w.scalars = off;
    off += 256;
    w.total = off;
    return w;
}
}  // namespace

int scan_exclusive_u8(const uint8_t* in, int64_t n, int64_t* out, cudaStream_t st);

GEOCB_API int64_t geocb_pairwise_plan_size(int64_t n) { return n < 0 ? -1 : plan_words(n); }
GEOCB_API int geocb_pairwise_plan_host(int64_t n, int64_t* plan_host) {
    ARG_CHECK(n >= 0 && plan_host);
    if (n == 0) {
        memset(plan_host, 0, 8 * 8);
        return 0;
    }
    plan_fill(n, plan_host);
    return 0;
}
GEOCB_API int64_t geocb_kpp_workspace_bytes(int64_t n) { return n < 0 ? -1 : kpp_layout(n).total; }

static int run_pairwise(ElemSrc src, const int64_t* plan_dev, int64_t n_leaves, float* vals, float* total_dev,
                        cudaStream_t st) {
    const int64_t threads = n_leaves * 8;
    leaf_sum_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(src, plan_dev, vals);
    tree_sum_kernel<<<1, 1024, 0, st>>>(plan_dev, vals, total_dev);
    KERNEL_CHECK();
    return 0;
}

GEOCB_API int geocb_kpp_draw(const float* dmin32, const uint8_t* is_center, int64_t n, double u, const int64_t* plan,
                             void* workspace, int64_t workspace_bytes, int64_t* result_host, void* stream) {
    ARG_CHECK(dmin32 && is_center && plan && workspace && result_host && n > 0);
    KppWs w = kpp_layout(n);
    if (workspace_bytes < w.total) {
        geocb_set_error("kpp workspace too small: %lld < %lld", (long long)workspace_bytes, (long long)w.total);
        return GEOCB_ERR_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    char* ws = (char*)workspace;
    float* p = (float*)(ws + w.p);
    float* vals = (float*)(ws + w.vals);
    double* tile_sums = (double*)(ws + w.tile_sums);
    unsigned int* maxbits = (unsigned int*)(ws + w.scalars);   // [0] max finite bits [1] any finite
    float* total = (float*)(ws + w.scalars + 16);
    long long* result = (long long*)(ws + w.scalars + 32);      // [0] idx [1] flag
    CU_TRY(cudaMemsetAsync(ws + w.scalars, 0, 64, st));
    max_finite_kernel<<<592, 256, 0, st>>>(dmin32, n, maxbits, maxbits + 1);
    finalize_max_kernel<<<1, 1, 0, st>>>(maxbits, maxbits + 1);
    ElemSrc src;
    src.mode = SRC_KPP;
    src.x = dmin32;
    src.is_center = is_center;
    src.max_finite_bits = maxbits;
    src.p_out = p;
    // number of leaves: read from the host-known formula (leaves hold >8 elements unless n<=128)
    const int64_t n_leaves_upper = n / 8 + 2;
    int rc = run_pairwise(src, plan, n_leaves_upper, vals, total, st);
    if (rc) return rc;
    const int64_t ntiles = (n + kCdfTile - 1) / kCdfTile;
    unsigned long long* fine = (unsigned long long*)(ws + w.scalars + 48);
    cdf_tile_sums_kernel<<<(unsigned)ntiles, kCdfThreads, 0, st>>>(p, total, n, tile_sums, fine);
    cdf_locate_kernel<<<1, kCdfThreads, 0, st>>>(p, total, n, u, tile_sums, ntiles, result, fine);
    KERNEL_CHECK();
    long long host[2];
    float tot_host;
    CU_TRY(cudaMemcpyAsync(host, result, 16, cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaMemcpyAsync(&tot_host, total, 4, cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaStreamSynchronize(st));
    if (host[1] == 1) {  // ambiguous: redo in NumPy's exact sequential order
        cdf_exact_kernel<<<1, 256, 0, st>>>(p, total, n, u, result);
        KERNEL_CHECK();
        CU_TRY(cudaMemcpyAsync(host, result, 16, cudaMemcpyDeviceToHost, st));
        CU_TRY(cudaStreamSynchronize(st));
        host[1] = 1;
    }
    result_host[0] = host[0];
    result_host[1] = host[1];
    uint32_t bits;
    memcpy(&bits, &tot_host, 4);
    result_host[2] = bits;
    result_host[3] = 0;
    return 0;
}

GEOCB_API int geocb_pairwise_sum_f32(const float* x, int64_t n, int square, int finite_only, void* workspace,
                                     int64_t workspace_bytes, float* out_host, int64_t* n_used_host, void* stream) {
    ARG_CHECK(x && out_host && n >= 0);
    cudaStream_t st = (cudaStream_t)stream;
    (void)workspace;
    (void)workspace_bytes;
    if (n_used_host) *n_used_host = n;
    if (n == 0) {
        *out_host = 0.f;
        return 0;
    }
    const float* src_x = x;
    int64_t m = n;
    void *flag = nullptr, *pre = nullptr, *comp = nullptr;
    if (finite_only) {
        CU_TRY(cudaMallocAsync(&flag, (size_t)n, st));
        CU_TRY(cudaMallocAsync(&pre, (size_t)(n + 1) * 8, st));
        finite_flag_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(x, n, (uint8_t*)flag);
        int rc = scan_exclusive_u8((uint8_t*)flag, n, (int64_t*)pre, st);
        if (rc) return rc;
        CU_TRY(cudaMemcpyAsync(&m, (int64_t*)pre + n, 8, cudaMemcpyDeviceToHost, st));
        CU_TRY(cudaStreamSynchronize(st));
        if (m < n) {
            CU_TRY(cudaMallocAsync(&comp, (size_t)(m ? m : 1) * 4, st));
            finite_compact_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(x, (uint8_t*)flag, (int64_t*)pre, n,
                                                                               (float*)comp);
            src_x = (const float*)comp;
        }
        if (n_used_host) *n_used_host = m;
    }
    int rc = 0;
    if (m == 0) {
        *out_host = 0.f;
    } else {
        std::vector<int64_t> plan((size_t)plan_words(m));
        plan_fill(m, plan.data());
        void *plan_dev = nullptr, *vals = nullptr;
        CU_TRY(cudaMallocAsync(&plan_dev, plan.size() * 8, st));
        CU_TRY(cudaMallocAsync(&vals, (size_t)(plan[1] + plan[2] + 2) * 4, st));
        CU_TRY(cudaMemcpyAsync(plan_dev, plan.data(), plan.size() * 8, cudaMemcpyHostToDevice, st));
        ElemSrc src;
        src.mode = square ? SRC_SQUARE : SRC_PLAIN;
        src.x = src_x;
        src.is_center = nullptr;
        src.max_finite_bits = nullptr;
        src.p_out = nullptr;
        float* total = (float*)vals + plan[1] + plan[2];
        rc = run_pairwise(src, (const int64_t*)plan_dev, plan[1], (float*)vals, total, st);
        if (!rc) {
            CU_TRY(cudaMemcpyAsync(out_host, total, 4, cudaMemcpyDeviceToHost, st));
            CU_TRY(cudaStreamSynchronize(st));
        }
        cudaFreeAsync(plan_dev, st);
        cudaFreeAsync(vals, st);
    }
    if (flag) cudaFreeAsync(flag, st);
    if (pre) cudaFreeAsync(pre, st);
    if (comp) cudaFreeAsync(comp, st);
    return rc;
}

// ------------------------------------------------ allocation-free pairwise sum ----
// geocb_pairwise_sum_f32_static: geocb_pairwise_sum_f32 (finite_only = 0) for callers that must not allocate device
// memory (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.  Its scratch -- the device copy of the summation plan and the
// partial sums of the leaves / inner nodes -- lives in a file-scope device array of fixed size, resident on every GPU
// that loads the library until the process exits.  All calls on a device share that array: calls that overlap on the
// GPU (different streams, different host threads) race.  Nothing in it has to start at zero: the plan is re-uploaded and
// every partial-sum slot the tree reads is written by the same call's leaf pass first.
#define GEOCB_STATIC_SUM_MAX_N (4LL << 20)           /* elements; keep in step with include/geocb.h */
#define GEOCB_STATIC_SUM_SCRATCH_BYTES (3 << 20)     /* plan <= 2 097 368 B + partial sums <= 524 300 B at the maximum n */
static __device__ __align__(256) unsigned char g_static_sum_scratch[GEOCB_STATIC_SUM_SCRATCH_BYTES];

namespace {
constexpr int kMaxDevices = 64;
void* g_static_sum_ptr[kMaxDevices];  // device address of g_static_sum_scratch, per device (0 = not looked up yet)

// The array's name is a host-side stand-in in host code, not a device address: look the address up once per device,
// with that device current.
int static_sum_scratch(void** out) {
    int dev = -1;
    CU_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) {
        geocb_set_error("device ordinal %d outside the %d supported by geocb_pairwise_sum_f32_static", dev, kMaxDevices);
        return GEOCB_ERR_UNSUPPORTED;
    }
    void* p = __atomic_load_n(&g_static_sum_ptr[dev], __ATOMIC_ACQUIRE);
    if (!p) {
        CU_TRY(cudaGetSymbolAddress(&p, g_static_sum_scratch));
        if (!p) {
            geocb_set_error("cudaGetSymbolAddress returned a null address for the static scratch array");
            return GEOCB_ERR_UNSUPPORTED;
        }
        __atomic_store_n(&g_static_sum_ptr[dev], p, __ATOMIC_RELEASE);  // (same value from every thread that races here)
    }
    *out = p;
    return 0;
}
}  // namespace

GEOCB_API int64_t geocb_pairwise_sum_f32_static_max_n(void) { return GEOCB_STATIC_SUM_MAX_N; }

GEOCB_API int geocb_pairwise_sum_f32_static(const float* x, int64_t n, int square, float* out_host, void* stream) {
    ARG_CHECK(x && out_host && n >= 0);
    if (n > GEOCB_STATIC_SUM_MAX_N) {
        geocb_set_error("geocb_pairwise_sum_f32_static: n = %lld exceeds the static scratch limit of %lld elements "
                        "(use geocb_pairwise_sum_f32)", (long long)n, (long long)GEOCB_STATIC_SUM_MAX_N);
        return GEOCB_ERR_WORKSPACE;
    }
    if (n == 0) {
        *out_host = 0.f;
        return 0;
    }
    cudaStream_t st = (cudaStream_t)stream;
    void* scratch = nullptr;
    int rc = static_sum_scratch(&scratch);
    if (rc) return rc;
    std::vector<int64_t> plan((size_t)plan_words(n));  // (host memory; the device side allocates nothing)
    plan_fill(n, plan.data());
    const int64_t plan_bytes = align_up((int64_t)plan.size() * 8, 256);
    const int64_t vals_bytes = (plan[1] + plan[2] + 2) * 4;
    if (plan_bytes + vals_bytes > GEOCB_STATIC_SUM_SCRATCH_BYTES) {  // cannot happen for n <= the maximum; belt and braces
        geocb_set_error("geocb_pairwise_sum_f32_static: n = %lld needs %lld B of scratch, the static array holds %d",
                        (long long)n, (long long)(plan_bytes + vals_bytes), GEOCB_STATIC_SUM_SCRATCH_BYTES);
        return GEOCB_ERR_WORKSPACE;
    }
    int64_t* plan_dev = (int64_t*)scratch;
    float* vals = (float*)((char*)scratch + plan_bytes);
    CU_TRY(cudaMemcpyAsync(plan_dev, plan.data(), plan.size() * 8, cudaMemcpyHostToDevice, st));
    ElemSrc src;
    src.mode = square ? SRC_SQUARE : SRC_PLAIN;
    src.x = x;
    src.is_center = nullptr;
    src.max_finite_bits = nullptr;
    src.p_out = nullptr;
    float* total = vals + plan[1] + plan[2];
    rc = run_pairwise(src, plan_dev, plan[1], vals, total, st);
    if (rc) return rc;
    CU_TRY(cudaMemcpyAsync(out_host, total, 4, cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaStreamSynchronize(st));
    return 0;
}
