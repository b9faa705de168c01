// sssp_vm32.cu -- source-vectorised SSSP, second generation: 32 sources per launch, bf16 upper-bound table.
//
// Same contract and arithmetic as sssp.cu / sssp_vm.cu (replaces scipy.sparse.csgraph.dijkstra as called at
// src/geo/geo_shortest_paths.py:36-49; fp64 path sums as uint64 bit patterns, result bit-identical to Dijkstra).
// What the ncu capture of sssp_vm.cu showed (profiles/r2_ncu_sssp_vm.txt): an edge visit served only 3.8 of 16 sources,
// cost 20 warp instructions and a 64-byte gather from a 188 MB DRAM-resident table.  This variant attacks all three:
//   * 32 sources side by side per vertex: the union frontier of 32 sources is the same set of vertices as that of 16
//     (in the wide rounds: all of them), so twice the relaxations ride on the same edge visits;
//   * the gathered table is bf16 (ub[v][32] = 64 B: the same gather, twice the sources; an exact filter because
//     ub = round_up_bf16(dist) and a candidate is dropped iff round_down_bf16(cand) >= ub);
//     one `red.global.v2.bf16x2.min` updates four sources' bounds at once;
//   * a lane handles 8 sources of one edge (quad per edge, one LDG.128 per lane), operands come from shared memory
//     through precomputed byte offsets, 2 CTAs x 128 registers instead of 3 x 80: no address re-materialisation.
// Layout: dist[v][32] u64, proc[v][32] u64, ub[v][32] bf16, one 32-bit frontier / far / active mask word per vertex.
// Rounds: A0 pop pass (lane = vertex), A1 edge pass (static warp striding over (block, 32-vertex chunk) items), near-far
// control + threshold scan, exactly as in sssp_vm.cu.
#include <stdlib.h>
#include "common.cuh"

namespace {

constexpr int G = 32;  // slots per launch
constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr unsigned FULL = 0xffffffffu;
constexpr int kStatsBytes = 2048;  // u64 words: 8 counters, then (start time, popped) per round for 120 rounds
constexpr int kFailedOff = 2000;
constexpr uint32_t UB16_INF = 0x7F80u;
typedef unsigned long long u64;

struct VmCtl {
    double thr2[G][2];
    uint32_t scan_min[G][3];  // min upper bound (as f32 bits) over the vertices a scan left dirty, by round % 3
    int scan_pushed[G][3];
    int active[G][3];
    u64 nact[3];
};

struct VmParams {
    const uint32_t* blk;  // [(B+1)][n], see geocb_sssp_vm_blocks
    const uint2* adj;
    int64_t n;
    int B;
    int nslots;
    int max_rounds;
    int* failed;
    double delta;
    int64_t blk_min_active;
    u64* dist;        // [n][G] fp64 bits
    u64* proc;        // [n][G]
    uint16_t* ub;     // [n][G] bf16 bits, ub = round_up_bf16(dist)
    uint32_t* mask[2];
    uint32_t* dirty;
    uint32_t* act;
    VmCtl* ctl;
    u64* stats;
    const int32_t* sources;
    float* out_d32;   // [nslots][n]
};

__device__ __forceinline__ uint64_t policy_evict_last() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t policy_evict_first() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
// bf16 bounds of a non-negative double: floor / ceil on the bf16 grid, as the upper 16 bits of the f32 pattern
__device__ __forceinline__ uint32_t f32_floor_bits(double c) { return __float_as_uint(__double2float_rd(c)); }
__device__ __forceinline__ uint32_t ub16_ceil(double c) { return (__float_as_uint(__double2float_ru(c)) + 0xFFFFu) >> 16; }
__device__ __forceinline__ uint4 ld_ub8(const uint16_t* p, uint64_t pol) {  // 8 consecutive bf16 bounds
    uint4 v;
    asm volatile("ld.global.cg.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p), "l"(pol));
    return v;
}
// min into four consecutive bf16 bounds (8-byte aligned): lanes that must not change carry +inf
__device__ __forceinline__ void red_min_ub4(uint16_t* p, uint32_t b0, uint32_t b1) {
    asm volatile("red.global.v2.bf16x2.min.noftz [%0], {%1, %2};" ::"l"(p), "r"(b0), "r"(b1) : "memory");
}
__device__ __forceinline__ uint2 ld_edge(const uint2* p, uint64_t pol) {
    uint2 v;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.u32 {%0,%1}, [%2], %3;"
                 : "=r"(v.x), "=r"(v.y)
                 : "l"(p), "l"(pol));
    return v;
}
__device__ __forceinline__ void red_min_u64(u64* p, u64 v) {
    asm volatile("red.global.min.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void red_or_u32(uint32_t* p, uint32_t v) {
    asm volatile("red.global.or.b32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void red_and_u32(uint32_t* p, uint32_t v) {
    asm volatile("red.global.and.b32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_cg_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.global.cg.u32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ double ld_cg_f64(const double* p) {
    return __longlong_as_double((long long)ld_cg_u64((const u64*)p));
}
__device__ __forceinline__ u64 now_ns() {
    u64 t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

constexpr int kDuStride = G + 2;  // u64 words per staged vertex row (272 B: 16-byte aligned, spreads banks)

// One lane's share of one edge visit: 8 sources (slots 8q..8q+7).  u4 = their bf16 bounds at the target, du = the
// popped vertex's own distances for these slots (shared memory), w = edge weight.  Returns the pass mask; improving
// candidates are left in cand[] for the caller's slow path.
__device__ __forceinline__ unsigned test8(const uint4 u4, const u64* du, double w, double (&cand)[8]) {
    // staged row order: slot 8q+k sits at word (k>>1)*8 + 2q + (k&1), so the four lanes of a quad read 64 contiguous
    // bytes with each of their four 16-byte loads (du already points at word 2q)
    const ulonglong2 d01 = *reinterpret_cast<const ulonglong2*>(du);
    const ulonglong2 d23 = *reinterpret_cast<const ulonglong2*>(du + 8);
    const ulonglong2 d45 = *reinterpret_cast<const ulonglong2*>(du + 16);
    const ulonglong2 d67 = *reinterpret_cast<const ulonglong2*>(du + 24);
    cand[0] = __longlong_as_double((long long)d01.x) + w;
    cand[1] = __longlong_as_double((long long)d01.y) + w;
    cand[2] = __longlong_as_double((long long)d23.x) + w;
    cand[3] = __longlong_as_double((long long)d23.y) + w;
    cand[4] = __longlong_as_double((long long)d45.x) + w;
    cand[5] = __longlong_as_double((long long)d45.y) + w;
    cand[6] = __longlong_as_double((long long)d67.x) + w;
    cand[7] = __longlong_as_double((long long)d67.y) + w;
    // cand cannot improve  <=>  round_down_bf16(cand) >= ub  <=>  f32_floor_bits(cand) >= ub << 16
    unsigned pm = 0;
    pm |= f32_floor_bits(cand[0]) < (u4.x << 16) ? 1u : 0u;
    pm |= f32_floor_bits(cand[1]) < (u4.x & 0xFFFF0000u) ? 2u : 0u;
    pm |= f32_floor_bits(cand[2]) < (u4.y << 16) ? 4u : 0u;
    pm |= f32_floor_bits(cand[3]) < (u4.y & 0xFFFF0000u) ? 8u : 0u;
    pm |= f32_floor_bits(cand[4]) < (u4.z << 16) ? 16u : 0u;
    pm |= f32_floor_bits(cand[5]) < (u4.z & 0xFFFF0000u) ? 32u : 0u;
    pm |= f32_floor_bits(cand[6]) < (u4.w << 16) ? 64u : 0u;
    pm |= f32_floor_bits(cand[7]) < (u4.w & 0xFFFF0000u) ? 128u : 0u;
    return pm;
}

__global__ void __launch_bounds__(kThreads, 2) sssp_vm32_kernel(VmParams p) {
    cg::grid_group grid = cg::this_grid();
    const int64_t n = p.n;
    const int64_t gtid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t gthreads = (int64_t)gridDim.x * blockDim.x;
    const int64_t gwarp = gtid >> 5, nwarps = gthreads >> 5;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int Q = lane >> 2, q = lane & 3;  // edge pass: quad Q of the warp = one edge, lane q = slots 8q..8q+7
    const int64_t nchunks = (n + 31) >> 5;
    const int B = p.B;
    VmCtl* ctl = p.ctl;

    // popped vertices' own distances, [warp][vertex of the chunk][slot]: 68 KB, dynamic (above the 48 KB static limit)
    extern __shared__ __align__(16) unsigned char s_dyn[];
    u64(*s_du)[32][kDuStride] = reinterpret_cast<u64(*)[32][kDuStride]>(s_dyn);
    __shared__ uint4 s_rec[kWarps][2][32];  // two tiles of edges: {target*G, byte offset of the owner's s_du row, weight as fp64}
    __shared__ double s_thr[G], s_hint[G];
    __shared__ float s_thrn[G];
    __shared__ int s_epoch[G], s_done[G], s_need[G];
    __shared__ int s_flag, s_alldone;
    const uint64_t pol_keep = policy_evict_last();
    const uint64_t pol_stream = policy_evict_first();

    // ---- init ----
    {
        const uint4 inf2 = make_uint4(0u, 0x7FF00000u, 0u, 0x7FF00000u);  // two fp64 +inf
        uint4* d4 = reinterpret_cast<uint4*>(p.dist);
        uint4* p4 = reinterpret_cast<uint4*>(p.proc);
        for (int64_t i = gtid; i < n * (G / 2); i += gthreads) {
            d4[i] = inf2;
            p4[i] = inf2;
        }
        const uint4 infu = make_uint4(0x7F807F80u, 0x7F807F80u, 0x7F807F80u, 0x7F807F80u);  // eight bf16 +inf
        uint4* u4 = reinterpret_cast<uint4*>(p.ub);
        for (int64_t i = gtid; i < n * (G / 8); i += gthreads) u4[i] = infu;
        for (int64_t i = gtid; i < n; i += gthreads) {
            p.mask[0][i] = 0;
            p.mask[1][i] = 0;
            p.dirty[i] = 0;
            p.act[i] = 0;
        }
    }
    if (threadIdx.x < G) {
        s_epoch[threadIdx.x] = 0;
        s_done[threadIdx.x] = threadIdx.x >= p.nslots;  // padding slots never run
        s_hint[threadIdx.x] = 0.0;
    }
    grid.sync();
    if (gtid < G) {
        const int g = (int)gtid;
        ctl->thr2[g][0] = ctl->thr2[g][1] = p.delta;
        for (int i = 0; i < 3; ++i) ctl->scan_min[g][i] = F32_INF_BITS, ctl->scan_pushed[g][i] = 0, ctl->active[g][i] = 0;
        if (g < 3) ctl->nact[g] = 0;
        if (g < p.nslots) {
            const int64_t src = p.sources[g];
            p.dist[src * G + g] = 0ULL;
            p.ub[src * G + g] = 0;
            atomicOr(p.mask[0] + src, 1u << g);
            ctl->active[g][0] = 1;
        }
    }
    grid.sync();

    unsigned n_improve = 0, n_rounds = 0, n_scans = 0, n_blocked = 0;
    u64 n_relax = 0, n_visit = 0;
    // per-lane shared-memory bases of the edge pass (byte addresses are formed once, not per edge)
    const char* du_lane = reinterpret_cast<const char*>(&s_du[warp][0][2 * q]);

    for (int round = 0;; ++round) {
        const int c_cur = round % 3, c_next = (round + 1) % 3, c_clear = (round + 2) % 3;
        uint32_t* cur = p.mask[round & 1];
        uint32_t* next = p.mask[(round + 1) & 1];
        if (gtid < G) ctl->active[gtid][c_clear] = 0;  // last read before the previous round's last grid.sync
        if (gtid == 0) ctl->nact[c_clear] = 0;         // last read before the previous round's A1
        if (threadIdx.x < G) s_thr[threadIdx.x] = ld_cg_f64(&ctl->thr2[threadIdx.x][s_epoch[threadIdx.x] & 1]);
        __syncthreads();
        if (p.stats && gtid == 0 && round < 120) p.stats[8 + 2 * round] = now_ns();

        // ---- A0: pop (lane = vertex): active (vertex, slot) pairs = frontier bit set and distance dropped since the
        //      vertex was last relaxed for that slot (proc); act[v] is written for every vertex ----
        {
            unsigned popped = 0, nactv = 0;
            uint32_t m_nx = 0;
            if (gwarp * 32 + lane < n) m_nx = ld_cg_u32(cur + gwarp * 32 + lane);
            for (int64_t c = gwarp; c < nchunks; c += nwarps) {
                const int64_t v = c * 32 + lane;
                const uint32_t m = m_nx;
                m_nx = 0;
                if ((c + nwarps) * 32 + lane < n) m_nx = ld_cg_u32(cur + (c + nwarps) * 32 + lane);
                uint32_t my_act = 0;
                if (m) {
                    cur[v] = 0;                    // consumed
                    red_and_u32(p.dirty + v, ~m);  // popped pairs are no longer "far"
                    uint32_t mm = m;
                    while (mm) {  // four (slot) pairs per trip, their loads independent
                        int sk[4];
                        u64 d[4], pr[4];
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            sk[k] = mm ? __ffs(mm) - 1 : -1;
                            mm &= mm - 1;  // (stays 0)
                        }
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            if (sk[k] >= 0) {
                                d[k] = ld_cg_u64(p.dist + v * G + sk[k]);
                                pr[k] = ld_cg_u64(p.proc + v * G + sk[k]);
                            }
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            if (sk[k] >= 0 && d[k] < pr[k]) {
                                p.proc[v * G + sk[k]] = d[k];
                                my_act |= 1u << sk[k];
                            }
                    }
                }
                if (v < n) p.act[v] = my_act;
                if (my_act) {
                    const uint32_t deg = __ldg(p.blk + (int64_t)B * n + v) - __ldg(p.blk + v);
                    n_relax += (u64)__popc(my_act) * deg;
                    n_visit += deg;
                    popped += __popc(my_act);
                }
                nactv += __popc(__ballot_sync(FULL, my_act != 0));
            }
            if (lane == 0 && nactv) atomicAdd(&ctl->nact[c_cur], (u64)nactv);
            if (p.stats && round < 120) {
                popped = __reduce_add_sync(FULL, popped);
                if (lane == 0 && popped) atomicAdd(p.stats + 9 + 2 * round, (u64)popped);
            }
        }
        grid.sync();

        // ---- A1: relax the rows of the active vertices (optionally one destination block at a time) ----
        const int64_t nact = (int64_t)ld_cg_u64(&ctl->nact[c_cur]);
        if (nact > 0) {
            const int Beff = (B > 1 && nact >= p.blk_min_active) ? B : 1;
            n_blocked += Beff > 1;
            const int64_t nitems = (int64_t)Beff * nchunks;
            unsigned near_acc = 0;  // bit k: this lane set a near-frontier mark for slot 8q+k
            uint32_t a_nx = 0, lo_nx = 0, hi_nx = 0;
            auto fetch = [&](int64_t g, uint32_t& a, uint32_t& lo, uint32_t& hi) {
                a = lo = hi = 0;
                if (g < nitems) {
                    const int64_t b = g / nchunks, c = g - b * nchunks;
                    const int64_t v = c * 32 + lane;
                    if (v < n) {
                        a = ld_cg_u32(p.act + v);
                        const int64_t bl = Beff == 1 ? 0 : b, bh = Beff == 1 ? B : b + 1;
                        lo = __ldg(p.blk + bl * n + v);
                        hi = __ldg(p.blk + bh * n + v);
                    }
                }
            };
            fetch(gwarp, a_nx, lo_nx, hi_nx);
            for (int64_t g = gwarp; g < nitems; g += nwarps) {
                const uint32_t a = a_nx, lo = lo_nx, hi = hi_nx;
                fetch(g + nwarps, a_nx, lo_nx, hi_nx);  // next item's mask / offsets are in flight during this one
                const uint32_t len = a ? hi - lo : 0;
                if (!__ballot_sync(FULL, len != 0)) continue;
                const int64_t v = (g % nchunks) * 32 + lane;
                uint32_t incl = len;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t t = __shfl_up_sync(FULL, incl, o);
                    if (lane >= o) incl += t;
                }
                const uint32_t start = incl - len;
                const uint32_t T = __shfl_sync(FULL, incl, 31);
                const uint32_t rowoff = lo - start;  // adj index of flat position f inside this vertex's segment = rowoff + f
                auto load_edge_tile = [&](uint32_t tile, uint2& e, int& j) {
                    const uint32_t f = tile + lane;
                    j = 0;  // owner = last vertex of the chunk whose start <= f (empty segments share their successor's start)
#pragma unroll
                    for (int st = 16; st; st >>= 1) {
                        const int probe = min(j + st, 31);
                        const uint32_t t = __shfl_sync(FULL, start, probe);
                        if (j + st <= 31 && t <= f) j += st;
                    }
                    const uint32_t off = __shfl_sync(FULL, rowoff, j);
                    e = make_uint2(0u, F32_INF_BITS);  // pad: vertex 0, +inf weight -> never improves anything
                    if (f < T) e = ld_edge(p.adj + (uint32_t)(off + f), pol_stream);
                };
                uint2 e_nx;
                int j_nx;
                load_edge_tile(0, e_nx, j_nx);  // in flight while the distances are staged
                // own distances of the active (vertex, slot) pairs -> shared memory (lane = vertex); inactive: +inf
#pragma unroll
                for (int h = 0; h < G; h += 8) {
                    u64 d[8];
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        d[k] = D64_INF_BITS;
                        if (len != 0 && ((a >> (h + k)) & 1u)) d[k] = ld_cg_u64(p.dist + v * G + h + k);
                    }
#pragma unroll
                    for (int k = 0; k < 8; ++k) s_du[warp][lane][(k >> 1) * 8 + (h >> 2) + (k & 1)] = d[k];  // see test8
                }
                auto publish = [&](int buf, uint2 e, int j) {
                    const double w = (double)__uint_as_float(e.y);
                    s_rec[warp][buf][lane] = make_uint4(e.x * G, (uint32_t)(j * kDuStride * 8), (uint32_t)__double2loint(w),
                                                        (uint32_t)__double2hiint(w));
                };
                auto consume = [&](const uint4 u4, const uint4 r) {
                    double cand[8];
                    const unsigned pm = test8(u4, reinterpret_cast<const u64*>(du_lane + r.y), __hiloint2double((int)r.w, (int)r.z), cand);
                    if (__any_sync(FULL, pm != 0)) {
                        unsigned nm = 0;
                        if (pm) {
                            u64* dt = p.dist + r.x + 8 * q;
                            uint32_t b[4] = {0x7F807F80u, 0x7F807F80u, 0x7F807F80u, 0x7F807F80u};  // bf16 +inf pairs
#pragma unroll
                            for (int k = 0; k < 8; ++k)
                                if ((pm >> k) & 1u) {
                                    red_min_u64(dt + k, (u64)__double_as_longlong(cand[k]));
                                    const uint32_t c16 = ub16_ceil(cand[k]);
                                    b[k >> 1] = (k & 1) ? ((b[k >> 1] & 0x0000FFFFu) | (c16 << 16)) : ((b[k >> 1] & 0xFFFF0000u) | c16);
                                    nm |= cand[k] < s_thr[8 * q + k] ? (1u << k) : 0u;
                                }
                            uint16_t* ut = p.ub + r.x + 8 * q;
                            if (pm & 0x0Fu) red_min_ub4(ut, b[0], b[1]);
                            if (pm & 0xF0u) red_min_ub4(ut + 4, b[2], b[3]);
                            near_acc |= nm;
                            n_improve += __popc(pm);
                        }
                        // near / far marks of the 32 slots, combined over the quad
                        uint32_t nbits = nm << (8 * q), fbits = (pm & ~nm) << (8 * q);
                        nbits |= __shfl_xor_sync(FULL, nbits, 1);
                        fbits |= __shfl_xor_sync(FULL, fbits, 1);
                        nbits |= __shfl_xor_sync(FULL, nbits, 2);
                        fbits |= __shfl_xor_sync(FULL, fbits, 2);
                        if (q == 0) {
                            const uint32_t t = r.x / G;
                            if (nbits) red_or_u32(next + t, nbits);
                            if (fbits) red_or_u32(p.dirty + t, fbits);
                        }
                    }
                };
                publish(0, e_nx, j_nx);
                __syncwarp();
                uint4 u[4];
                {
                    const int ng0 = (int)min(4u, (T + 7) >> 3);
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        if (i < ng0) u[i] = ld_ub8(p.ub + s_rec[warp][0][8 * i + Q].x + 8 * q, pol_keep);
                }
                if (32 < T) load_edge_tile(32, e_nx, j_nx);
                int buf = 0;
                for (uint32_t tile = 0; tile < T; tile += 32, buf ^= 1) {
                    const bool more = tile + 32 < T;
                    if (more) publish(buf ^ 1, e_nx, j_nx);  // records of the next tile
                    __syncwarp();
                    if (tile + 64 < T) load_edge_tile(tile + 64, e_nx, j_nx);  // edges two tiles ahead are in flight
                    const int ng = (int)min(4u, (T - tile + 7) >> 3);
                    const int ngn = more ? (int)min(4u, (T - tile - 32 + 7) >> 3) : 0;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        // consume group i of this tile, then reuse its registers for group i of the next tile
                        if (i < ng) consume(u[i], s_rec[warp][buf][8 * i + Q]);
                        if (i < ngn) u[i] = ld_ub8(p.ub + s_rec[warp][buf ^ 1][8 * i + Q].x + 8 * q, pol_keep);
                    }
                    __syncwarp();
                }
            }
#pragma unroll
            for (int k = 0; k < 8; ++k)
                if ((near_acc >> k) & 1u) *(volatile int*)&ctl->active[8 * q + k][c_next] = 1;
        }
        grid.sync();
        ++n_rounds;

        // ---- control: which slots ran dry? (uniform across the grid) ----
        if (gtid < G) {  // scan accumulators of the previous round: every block read them before this sync
            ctl->scan_min[gtid][c_clear] = F32_INF_BITS;
            ctl->scan_pushed[gtid][c_clear] = 0;
        }
        if (threadIdx.x == 0) s_flag = 0;
        __syncthreads();
        if (threadIdx.x < G) {
            const int g = threadIdx.x;
            const int need = !s_done[g] && ld_cg_s32(&ctl->active[g][c_next]) == 0;
            s_need[g] = need;
            if (need) s_flag = 1;
            double thr_new = s_thr[g] + p.delta;
            if (s_hint[g] > thr_new) thr_new = s_hint[g];
            // the scan compares bf16 bounds: which far vertices it releases is a scheduling choice, not a correctness
            // one (every dirty vertex is released eventually: ceil_bf16(thr_new) > any bf16 value below thr_new)
            s_thrn[g] = __uint_as_float(ub16_ceil(thr_new) << 16);
            if (need && blockIdx.x == 0) ctl->thr2[g][(s_epoch[g] + 1) & 1] = thr_new;
        }
        __syncthreads();
        if (s_flag) {
            ++n_scans;
            // ---- phase B (quad = vertex, lane q = slots 8q..8q+7): raise the threshold of every dry slot, move its
            //      dirty vertices below it to the frontier; remember the smallest bound left behind ----
            uint32_t need8 = 0, thr16[8], rem[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                need8 |= s_need[8 * q + k] ? (1u << k) : 0u;
                thr16[k] = __float_as_uint(s_thrn[8 * q + k]) >> 16;
                rem[k] = UB16_INF;
            }
            uint32_t pushed8 = 0;
            const int64_t nv8 = (n + 7) >> 3;
            for (int64_t c = gwarp; c < nv8; c += nwarps) {
                const int64_t v = c * 8 + Q;
                uint32_t dfull = 0;
                if (v < n) dfull = ld_cg_u32(p.dirty + v);
                const uint32_t dm = (dfull >> (8 * q)) & need8;
                uint32_t taken = 0;
                if (dm) {
                    const uint4 x = ld_ub8(p.ub + v * G + 8 * q, pol_keep);
                    const uint32_t ur[8] = {x.x & 0xFFFFu, x.x >> 16, x.y & 0xFFFFu, x.y >> 16,
                                            x.z & 0xFFFFu, x.z >> 16, x.w & 0xFFFFu, x.w >> 16};
#pragma unroll
                    for (int k = 0; k < 8; ++k)
                        if ((dm >> k) & 1u) {
                            if (ur[k] < thr16[k]) taken |= 1u << k;
                            else rem[k] = min(rem[k], ur[k]);
                        }
                }
                if (__any_sync(FULL, taken != 0)) {
                    pushed8 |= taken;
                    uint32_t tk = taken << (8 * q);
                    tk |= __shfl_xor_sync(FULL, tk, 1);
                    tk |= __shfl_xor_sync(FULL, tk, 2);
                    if (q == 0 && tk) {
                        p.dirty[v] = dfull & ~tk;  // nothing else touches dirty[v] / next[v] during a scan
                        red_or_u32(next + v, tk);
                    }
                }
            }
#pragma unroll
            for (int o = 4; o < 32; o <<= 1) pushed8 |= __shfl_xor_sync(FULL, pushed8, o);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                uint32_t r = rem[k];
#pragma unroll
                for (int o = 4; o < 32; o <<= 1) r = min(r, __shfl_xor_sync(FULL, r, o));
                // (active[c_next] must not be touched here: slower blocks may still be reading it to decide whether
                //  to enter this phase, and its grid.sync, at all)
                if (Q == 0) {
                    if (r != UB16_INF) atomicMin(&ctl->scan_min[8 * q + k][c_cur], r << 16);
                    if ((pushed8 >> k) & 1u) atomicOr(&ctl->scan_pushed[8 * q + k][c_cur], 1);
                }
            }
            grid.sync();
            if (threadIdx.x < G && s_need[threadIdx.x]) {  // identical in every block
                const int g = threadIdx.x;
                const int pu = ld_cg_s32(&ctl->scan_pushed[g][c_cur]);
                const uint32_t mn = ld_cg_u32(&ctl->scan_min[g][c_cur]);
                s_epoch[g] += 1;
                s_hint[g] = 0.0;
                if (pu == 0) {
                    if (mn == F32_INF_BITS) s_done[g] = 1;  // frontier empty and nothing dirty: converged
                    else s_hint[g] = (double)__uint_as_float(mn) + p.delta;  // jump over the gap
                }
            }
            __syncthreads();
        }
        // ---- termination test (uniform) ----
        if (threadIdx.x == 0) s_alldone = 1;
        __syncthreads();
        if (threadIdx.x < G && !s_done[threadIdx.x]) s_alldone = 0;
        __syncthreads();
        if (s_alldone) break;
        if (round >= p.max_rounds) {  // uniform
            if (gtid == 0) *p.failed = 1;
            break;
        }
    }

    // ---- finalize: out_d32[s][v] = f32(dist[v][s]), transposed through shared memory (coalesced both ways) ----
    {
        float* tile = reinterpret_cast<float*>(&s_du[warp][0][0]);  // [G][33] floats = 4224 B of this warp's 8704
        for (int64_t c = gwarp; c < nchunks; c += nwarps) {
#pragma unroll 8
            for (int j = 0; j < 32; ++j) {
                const int64_t v = c * 32 + j;
                float d = 0.f;
                if (v < n) d = __double2float_rn(__longlong_as_double((long long)ld_cg_u64(p.dist + v * G + lane)));
                tile[lane * 33 + j] = d;
            }
            __syncwarp();
            const int64_t v = c * 32 + lane;
            if (v < n)
                for (int g = 0; g < p.nslots; ++g) p.out_d32[(int64_t)g * n + v] = tile[g * 33 + lane];
            __syncwarp();
        }
    }
    if (p.stats) {
        if (gtid == 0) {
            p.stats[6] = now_ns();
            p.stats[7] = (u64)n_rounds;
        }
        n_improve = __reduce_add_sync(FULL, n_improve);
        for (int o = 16; o; o >>= 1) {
            n_relax += __shfl_xor_sync(FULL, n_relax, o);
            n_visit += __shfl_xor_sync(FULL, n_visit, o);
        }
        if (lane == 0) {
            atomicAdd(p.stats + 0, n_relax);
            atomicAdd(p.stats + 3, (u64)n_improve);
            atomicAdd(p.stats + 4, n_visit);
        }
        if (gtid == 0) {
            atomicAdd(p.stats + 1, (u64)n_rounds);
            atomicAdd(p.stats + 2, (u64)n_scans);
            atomicAdd(p.stats + 5, (u64)n_blocked);
        }
    }
}

constexpr int kDynSmem = kWarps * 32 * kDuStride * 8;

struct VmLayout {
    int64_t dist, proc, ub, mask0, mask1, dirty, act, ctl, stats, total;
};
VmLayout vm_layout(int64_t n) {
    VmLayout w;
    int64_t off = 0;
    w.dist = off;
    off += align_up(n * G * 8, 256);
    w.proc = off;
    off += align_up(n * G * 8, 256);
    w.ub = off;
    off += align_up(n * G * 2, 256);
    w.mask0 = off;
    off += align_up(n * 4, 256);
    w.mask1 = off;
    off += align_up(n * 4, 256);
    w.dirty = off;
    off += align_up(n * 4, 256);
    w.act = off;
    off += align_up(n * 4, 256);
    w.ctl = off;
    off += align_up((int64_t)sizeof(VmCtl), 256);
    w.stats = off;
    off += kStatsBytes;
    w.total = off;
    return w;
}

int vm_grid(int* grid_out) {
    DevInfo di;
    int rc = geocb_devinfo(&di);
    if (rc) return rc;
    CU_TRY(cudaFuncSetAttribute(sssp_vm32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kDynSmem));
    int occ = 0;
    CU_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, sssp_vm32_kernel, kThreads, kDynSmem));
    if (occ < 1) {
        geocb_set_error("sssp_vm32_kernel cannot be resident (occupancy 0)");
        return GEOCB_ERR_UNSUPPORTED;
    }
    *grid_out = di.sm_count * occ;
    return 0;
}

}  // namespace

GEOCB_API int64_t geocb_sssp_vm32_workspace_bytes(int64_t n) {
    if (n < 0) return -1;
    return vm_layout(n).total;
}

// 32 sources per launch; blk from geocb_sssp_vm_blocks.  stats_host: 8 int64 (relaxations, rounds, scans,
// improvements, edge visits, blocked rounds, end time, rounds of the last launch) or NULL.
GEOCB_API int geocb_sssp_multi_vm32(const void* adj, const uint32_t* blk, int B, int64_t n, const int32_t* sources,
                                    int S, double delta, int64_t blk_min_active, void* workspace,
                                    int64_t workspace_bytes, float* out_d32, int64_t* stats_host, void* stream) {
    ARG_CHECK(adj && blk && sources && out_d32 && workspace);
    ARG_CHECK(n > 0 && n <= (1LL << 26) && S > 0 && B >= 1 && B <= 256 && delta > 0);
    VmLayout w = vm_layout(n);
    if (workspace_bytes < w.total) {
        geocb_set_error("sssp_vm32 workspace too small: %lld < %lld", (long long)workspace_bytes, (long long)w.total);
        return GEOCB_ERR_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    char* ws = (char*)workspace;
    int grid = 0;
    int rc = vm_grid(&grid);
    if (rc) return rc;
    CU_TRY(cudaMemsetAsync(ws + w.stats, 0, kStatsBytes, st));
    const char* env = getenv("GEOCB_SSSP_MAX_ROUNDS");
    const int64_t lim = env ? atoll(env) : 8 * n + 4096;
    for (int s0 = 0; s0 < S; s0 += G) {
        VmParams p;
        memset(&p, 0, sizeof(p));
        p.blk = blk;
        p.adj = (const uint2*)adj;
        p.n = n;
        p.B = B;
        p.nslots = (S - s0) < G ? (S - s0) : G;
        p.max_rounds = (int)(lim < 0x7fffff00 ? lim : 0x7fffff00);
        p.failed = (int*)(ws + w.stats + kFailedOff);
        p.delta = delta;
        p.blk_min_active = blk_min_active;
        p.dist = (u64*)(ws + w.dist);
        p.proc = (u64*)(ws + w.proc);
        p.ub = (uint16_t*)(ws + w.ub);
        p.mask[0] = (uint32_t*)(ws + w.mask0);
        p.mask[1] = (uint32_t*)(ws + w.mask1);
        p.dirty = (uint32_t*)(ws + w.dirty);
        p.act = (uint32_t*)(ws + w.act);
        p.ctl = (VmCtl*)(ws + w.ctl);
        p.stats = (u64*)(ws + w.stats);
        p.sources = sources + s0;
        p.out_d32 = out_d32 + (int64_t)s0 * n;
        if (s0 > 0) CU_TRY(cudaMemsetAsync(ws + w.stats + 64, 0, kFailedOff - 64, st));  // round trace = last launch
        void* args[] = {&p};
        CU_TRY(cudaLaunchCooperativeKernel((void*)sssp_vm32_kernel, dim3(grid), dim3(kThreads), args, kDynSmem, st));
    }
    if (stats_host) CU_TRY(cudaMemcpyAsync(stats_host, ws + w.stats, 64, cudaMemcpyDeviceToHost, st));
    int failed = 0;
    CU_TRY(cudaMemcpyAsync(&failed, ws + w.stats + kFailedOff, 4, cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaStreamSynchronize(st));
    if (failed) {
        geocb_set_error("sssp_vm32_kernel did not converge within the round limit");
        return GEOCB_ERR_OVERFLOW;
    }
    return 0;
}
