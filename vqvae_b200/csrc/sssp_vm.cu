// sssp_vm.cu -- source-vectorised ("vertex-major") multi-source SSSP with destination blocking.
//
// Same contract and the same arithmetic as sssp.cu (replaces scipy.sparse.csgraph.dijkstra as called at
// src/geo/geo_shortest_paths.py:36-49): fp64 path sums kept as uint64 bit patterns, label-correcting near-far
// relaxation, result bit-identical to Dijkstra.  What changes is the data layout and who shares what:
//
//   sssp.cu      one table per source ("slot"); a relaxation of slot s gathers ub_s[v] -- one L1 tag lookup and one
//                32-byte sector for ONE (slot, edge) pair -- and every slot re-streams the whole adjacency.
//   this file    the 16 sources of a launch live side by side per vertex:
//                  dist[v][16] u64 (one 128-byte line), proc[v][16] u64, ub[v][16] f32 bits (two sectors),
//                  frontier / far marks as one 32-bit mask word per vertex (bit s = slot s).
//                lane & 15 = slot, a half-warp = one edge.  A vertex is popped once per round for all the slots
//                whose distance dropped; each of its edges is loaded once, ONE coalesced 64-byte gather fetches the
//                upper bounds of all 16 slots of the target, and every lane tests its own slot from registers.
//                In the wide rounds (where 30-60 % of the slots want the same vertex) that is one tag lookup and one
//                edge load per 5-10 relaxations instead of per relaxation.
//
//   destination blocking: with 16 slots the randomly accessed state (ub + dist + masks = 200 B per vertex) is 0.6 GB at
//                2.94M vertices, far beyond L2.  Rows are sorted by column, so the vertex range is cut into B blocks and
//                blk[b][v] holds where row v enters block b; a round sweeps the frontier once per block, touching only
//                n/B vertices' state (~40 MB, L2 resident) with its gathers and REDs.  The price is re-reading the
//                popped vertices' own distances B times (sequential, coalesced).  Narrow rounds (few popped vertices)
//                run unblocked.
//
// Per round: A0 pop pass (which (v, slot) pairs really dropped since they were last relaxed -> act[v], proc),
// A1 edge pass (static warp striding over (block, 32-vertex chunk) items, next item's masks/offsets prefetched),
// control + threshold scan for the slots whose near frontier ran dry (same near-far logic as sssp.cu).
#include <stdlib.h>
#include "common.cuh"

namespace {

constexpr int G = 16;  // slots per launch; lane & 15 = slot
constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr unsigned FULL = 0xffffffffu;
constexpr int kStatsBytes = 2048;  // u64 words: 8 counters, then (start time, popped) per round for 120 rounds
constexpr int kFailedOff = 2000;   // byte offset of the 'hit the round limit' flag inside the stats block
typedef unsigned long long u64;

struct VmCtl {
    double thr2[G][2];        // near/far threshold per slot, double-buffered by the slot's scan count
    uint32_t scan_min[G][3];  // min ub (f32 bits) over the vertices a scan left dirty, by round % 3
    int scan_pushed[G][3];
    int active[G][3];         // "slot has near-frontier bits for the next round", index (r+1)%3 written in round r
    u64 nact[3];              // vertices with a non-empty active mask in round r (index r%3)
};

struct VmParams {
    const uint32_t* blk;  // [(B+1)][n] pair offset where row v enters destination block b; blk[B] = end of the real entries
    const uint2* adj;
    int64_t n;
    int B;
    int nslots;
    int max_rounds;
    int* failed;
    double delta;
    int64_t blk_min_active;  // rounds popping fewer vertices than this run unblocked (one pass over whole rows)
    u64* dist;               // [n][G] fp64 bits
    u64* proc;               // [n][G]
    uint32_t* ub;            // [n][G] f32 bits, ub = round_up_f32(dist)
    uint32_t* mask[2];       // [n] frontier masks, double-buffered by round parity
    uint32_t* dirty;         // [n] "improved to a value >= threshold" marks
    uint32_t* act;           // [n] slots whose distance dropped since the vertex was last relaxed (this round)
    VmCtl* ctl;
    u64* stats;
    const int32_t* sources;  // [nslots]
    float* out_d32;          // [nslots][n]
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
__device__ __forceinline__ uint32_t ub_floor(double c) { return __float_as_uint(__double2float_rd(c)); }
__device__ __forceinline__ uint32_t ub_ceil(double c) { return __float_as_uint(__double2float_ru(c)); }
__device__ __forceinline__ void red_min_ub(uint32_t* p, uint32_t val, uint64_t pol) {
    asm volatile("red.global.min.L2::cache_hint.u32 [%0], %1, %2;" ::"l"(p), "r"(val), "l"(pol) : "memory");
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

#ifndef GEOCB_VM_MINBLOCKS
#define GEOCB_VM_MINBLOCKS 3
#endif
constexpr int kDuStride = G + 2;  // u64 words per staged vertex row: 144 B keeps 16-byte alignment, spreads banks

__device__ __forceinline__ uint4 ld_ub4(const uint32_t* p, uint64_t pol) {
    uint4 v;
    asm volatile("ld.global.cg.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p), "l"(pol));
    return v;
}

__global__ void __launch_bounds__(kThreads, GEOCB_VM_MINBLOCKS) sssp_vm_kernel(VmParams p) {
    cg::grid_group grid = cg::this_grid();
    const int64_t n = p.n;
    const int64_t gtid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t gthreads = (int64_t)gridDim.x * blockDim.x;
    const int64_t gwarp = gtid >> 5, nwarps = gthreads >> 5;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int Q = lane >> 2, q = lane & 3;  // edge pass: quad Q of the warp = one edge, lane q of the quad = slots 4q..4q+3
    const int64_t nchunks = (n + 31) >> 5;
    const int B = p.B;
    VmCtl* ctl = p.ctl;

    __shared__ __align__(16) u64 s_du[kWarps][32][kDuStride];  // popped vertices' own distances (fp64 bits), [vertex of the chunk][slot]
    __shared__ uint4 s_rec[kWarps][2][32];  // two tiles of edges: {target*G, owner vertex of the chunk, weight as fp64}
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
        const uint4 infu = make_uint4(F32_INF_BITS, F32_INF_BITS, F32_INF_BITS, F32_INF_BITS);
        uint4* u4 = reinterpret_cast<uint4*>(p.ub);
        for (int64_t i = gtid; i < n * (G / 4); i += gthreads) u4[i] = infu;
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
            p.ub[src * G + g] = 0u;
            atomicOr(p.mask[0] + src, 1u << g);
            ctl->active[g][0] = 1;
        }
    }
    grid.sync();

    unsigned n_improve = 0, n_rounds = 0, n_scans = 0, n_blocked = 0;
    u64 n_relax = 0, n_visit = 0;

    for (int round = 0;; ++round) {
        const int c_cur = round % 3, c_next = (round + 1) % 3, c_clear = (round + 2) % 3;
        uint32_t* cur = p.mask[round & 1];
        uint32_t* next = p.mask[(round + 1) & 1];
        if (gtid < G) ctl->active[gtid][c_clear] = 0;  // last read before the previous round's last grid.sync
        if (gtid == 0) ctl->nact[c_clear] = 0;         // last read before the previous round's A1
        if (threadIdx.x < G) s_thr[threadIdx.x] = ld_cg_f64(&ctl->thr2[threadIdx.x][s_epoch[threadIdx.x] & 1]);
        __syncthreads();
        if (p.stats && gtid == 0 && round < 120) p.stats[8 + 2 * round] = now_ns();

        // ---- A0: pop (lane = vertex).  A (vertex, slot) pair is active iff its frontier bit is set and its distance
        //      dropped since the vertex was last relaxed for that slot (proc); act[v] is written for every vertex. ----
        {
            unsigned popped = 0, nactv = 0;
            uint32_t m_nx = 0;
            if (gwarp * 32 + lane < n) m_nx = ld_cg_u32(cur + gwarp * 32 + lane);
            for (int64_t c = gwarp; c < nchunks; c += nwarps) {
                const int64_t v = c * 32 + lane;
                const uint32_t m = m_nx;
                m_nx = 0;
                if ((c + nwarps) * 32 + lane < n) m_nx = ld_cg_u32(cur + (c + nwarps) * 32 + lane);  // next chunk's masks in flight
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

        // ---- A1: relax the rows of the active vertices, one destination block at a time ----
        const int64_t nact = (int64_t)ld_cg_u64(&ctl->nact[c_cur]);
        if (nact > 0) {
            const int Beff = (B > 1 && nact >= p.blk_min_active) ? B : 1;
            n_blocked += Beff > 1;
            const int64_t nitems = (int64_t)Beff * nchunks;
            const double thr0 = s_thr[4 * q], thr1 = s_thr[4 * q + 1], thr2 = s_thr[4 * q + 2], thr3 = s_thr[4 * q + 3];
            unsigned near_acc = 0;
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
                // lane -> flat edge tile+lane; its owner is the last vertex of the chunk whose start <= position
                // (an empty segment shares its start with its successor, so it is never selected while f < T)
                auto load_edge_tile = [&](uint32_t tile, uint2& e, int& j) {
                    const uint32_t f = tile + lane;
                    j = 0;
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
                    for (int k = 0; k < 8; ++k) s_du[warp][lane][h + k] = d[k];
                }
                auto publish = [&](int buf, uint2 e, int j) {
                    const double w = (double)__uint_as_float(e.y);
                    s_rec[warp][buf][lane] = make_uint4(e.x * G, (uint32_t)j, (uint32_t)__double2loint(w), (uint32_t)__double2hiint(w));
                };
                auto consume = [&](const uint4 u4, const uint4 r) {
                    const ulonglong2* dp = reinterpret_cast<const ulonglong2*>(&s_du[warp][r.y][4 * q]);
                    const ulonglong2 d01 = dp[0], d23 = dp[1];
                    const double w = __hiloint2double((int)r.w, (int)r.z);
                    const double c0 = __longlong_as_double((long long)d01.x) + w, c1 = __longlong_as_double((long long)d01.y) + w;
                    const double c2 = __longlong_as_double((long long)d23.x) + w, c3 = __longlong_as_double((long long)d23.y) + w;
                    // cand >= ub[t][s] (>= dist[t][s])  <=>  round_down_f32(cand) >= ub: nothing to do
                    const bool p0 = ub_floor(c0) < u4.x, p1 = ub_floor(c1) < u4.y, p2 = ub_floor(c2) < u4.z, p3 = ub_floor(c3) < u4.w;
                    const unsigned pm = (unsigned)p0 | ((unsigned)p1 << 1) | ((unsigned)p2 << 2) | ((unsigned)p3 << 3);
                    if (__any_sync(FULL, pm != 0)) {
                        unsigned nm = 0;
                        u64* dt = p.dist + r.x + 4 * q;
                        uint32_t* ut = p.ub + r.x + 4 * q;
                        if (p0) { red_min_u64(dt + 0, (u64)__double_as_longlong(c0)); red_min_ub(ut + 0, ub_ceil(c0), pol_keep); nm |= c0 < thr0 ? 1u : 0u; }
                        if (p1) { red_min_u64(dt + 1, (u64)__double_as_longlong(c1)); red_min_ub(ut + 1, ub_ceil(c1), pol_keep); nm |= c1 < thr1 ? 2u : 0u; }
                        if (p2) { red_min_u64(dt + 2, (u64)__double_as_longlong(c2)); red_min_ub(ut + 2, ub_ceil(c2), pol_keep); nm |= c2 < thr2 ? 4u : 0u; }
                        if (p3) { red_min_u64(dt + 3, (u64)__double_as_longlong(c3)); red_min_ub(ut + 3, ub_ceil(c3), pol_keep); nm |= c3 < thr3 ? 8u : 0u; }
                        near_acc |= nm;
                        n_improve += __popc(pm);
                        // near marks in bits 0..15, far marks in bits 16..31, combined over the quad
                        unsigned packed = (nm | ((pm & ~nm) << 16)) << (4 * q);
                        packed |= __shfl_xor_sync(FULL, packed, 1);
                        packed |= __shfl_xor_sync(FULL, packed, 2);
                        if (q == 0 && packed) {
                            const uint32_t t = r.x / G;
                            if (packed & 0xffffu) red_or_u32(next + t, packed & 0xffffu);
                            if (packed >> 16) red_or_u32(p.dirty + t, packed >> 16);
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
                        if (i < ng0) u[i] = ld_ub4(p.ub + s_rec[warp][0][8 * i + Q].x + 4 * q, pol_keep);
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
                        // consume group i of this tile, then reuse its registers for group i of the next tile:
                        // three to four gather groups (24-32 edges x 16 slots) stay in flight per warp
                        if (i < ng) consume(u[i], s_rec[warp][buf][8 * i + Q]);
                        if (i < ngn) u[i] = ld_ub4(p.ub + s_rec[warp][buf ^ 1][8 * i + Q].x + 4 * q, pol_keep);
                    }
                    __syncwarp();
                }
            }
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if ((near_acc >> k) & 1u) *(volatile int*)&ctl->active[4 * q + k][c_next] = 1;
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
            // ub (f32) < thr_new (fp64)  <=>  ub < round_up_f32(thr_new)
            s_thrn[g] = __double2float_ru(thr_new);
            if (need && blockIdx.x == 0) ctl->thr2[g][(s_epoch[g] + 1) & 1] = thr_new;
        }
        __syncthreads();
        if (s_flag) {
            ++n_scans;
            // ---- phase B (lane = vertex): raise the threshold of every dry slot, move its dirty vertices below it to
            //      the frontier; remember the smallest bound left behind (to jump over empty buckets) ----
            uint32_t needmask = 0;
#pragma unroll
            for (int g = 0; g < G; ++g) needmask |= s_need[g] ? (1u << g) : 0u;
            uint32_t rem[G];
#pragma unroll
            for (int g = 0; g < G; ++g) rem[g] = F32_INF_BITS;
            uint32_t pushed_mask = 0;
            for (int64_t c = gwarp; c < nchunks; c += nwarps) {
                const int64_t v = c * 32 + lane;
                uint32_t dfull = 0;
                if (v < n) dfull = ld_cg_u32(p.dirty + v);
                const uint32_t dm = dfull & needmask;
                if (dm) {
                    const uint32_t* row = p.ub + v * G;
                    uint32_t ur[G];
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const uint4 x = ld_ub4(row + 4 * i, pol_keep);
                        ur[4 * i] = x.x, ur[4 * i + 1] = x.y, ur[4 * i + 2] = x.z, ur[4 * i + 3] = x.w;
                    }
                    uint32_t taken = 0;
#pragma unroll
                    for (int g = 0; g < G; ++g) {
                        if ((dm >> g) & 1u) {
                            if (__uint_as_float(ur[g]) < s_thrn[g]) taken |= 1u << g;
                            else rem[g] = min(rem[g], ur[g]);
                        }
                    }
                    if (taken) {
                        p.dirty[v] = dfull & ~taken;  // nothing else touches dirty[v] / next[v] during a scan
                        red_or_u32(next + v, taken);
                        pushed_mask |= taken;
                    }
                }
            }
            pushed_mask = __reduce_or_sync(FULL, pushed_mask);
#pragma unroll
            for (int g = 0; g < G; ++g) {
                const uint32_t r = __reduce_min_sync(FULL, rem[g]);
                // (active[c_next] must not be touched here: slower blocks may still be reading it to decide whether
                //  to enter this phase, and its grid.sync, at all)
                if (lane == 0 && r != F32_INF_BITS) atomicMin(&ctl->scan_min[g][c_cur], r);
            }
            if (lane < G && ((pushed_mask >> lane) & 1u)) atomicOr(&ctl->scan_pushed[lane][c_cur], 1);
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
        float* tile = reinterpret_cast<float*>(&s_du[warp][0][0]);  // [G][33] floats = 2112 B of this warp's 4608
        const int s = lane & 15, half = lane >> 4;
        for (int64_t c = gwarp; c < nchunks; c += nwarps) {
#pragma unroll 4
            for (int k = 0; k < 16; ++k) {
                const int j = 2 * k + half;
                const int64_t v = c * 32 + j;
                float d = 0.f;
                if (v < n) d = __double2float_rn(__longlong_as_double((long long)ld_cg_u64(p.dist + v * G + s)));
                tile[s * 33 + j] = d;
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

// blk[b][v] = first pair of row v whose column is >= b*nb (b = 0..B-1), blk[B][v] = end of the real entries (the
// +inf pad that keeps rows 16-byte aligned is never visited)
__global__ void vm_blocks_kernel(const int64_t* row_ptr, const uint2* adj, int64_t n, int B, int64_t nb, uint32_t* blk) {
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    const int64_t b0 = row_ptr[v];
    int64_t e = row_ptr[v + 1];
    if (e > b0) {
        const uint2 last = adj[e - 1];
        if (last.y == F32_INF_BITS && (int64_t)last.x == v) --e;
    }
    int b = 0;
    blk[v] = (uint32_t)b0;
    for (int64_t q = b0; q < e; ++q) {
        const int cb = (int)((int64_t)adj[q].x / nb);
        while (b < cb && b < B) {
            ++b;
            blk[(int64_t)b * n + v] = (uint32_t)q;
        }
    }
    while (b < B) {
        ++b;
        blk[(int64_t)b * n + v] = (uint32_t)e;
    }
}

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
    off += align_up(n * G * 4, 256);
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
    int occ = 0;
    CU_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, sssp_vm_kernel, kThreads, 0));
    if (occ < 1) {
        geocb_set_error("sssp_vm_kernel cannot be resident (occupancy 0)");
        return GEOCB_ERR_UNSUPPORTED;
    }
    *grid_out = di.sm_count * occ;
    return 0;
}

}  // namespace

GEOCB_API int geocb_sssp_vm_slots(void) { return G; }

GEOCB_API int64_t geocb_sssp_vm_workspace_bytes(int64_t n) {
    if (n < 0) return -1;
    return vm_layout(n).total;
}

GEOCB_API int geocb_sssp_vm_blocks(const int64_t* row_ptr, const void* adj, int64_t n, int B, uint32_t* blk,
                                   void* stream) {
    ARG_CHECK(row_ptr && adj && blk && n > 0 && B >= 1 && B <= 256);
    cudaStream_t st = (cudaStream_t)stream;
    int64_t npairs = 0;
    CU_TRY(cudaMemcpyAsync(&npairs, row_ptr + n, 8, cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaStreamSynchronize(st));
    if (npairs >= (1LL << 32)) {
        geocb_set_error("graph has %lld adjacency pairs; the SSSP kernels address pairs with 32 bits", (long long)npairs);
        return GEOCB_ERR_UNSUPPORTED;
    }
    const int64_t nb = (n + B - 1) / B;
    vm_blocks_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(row_ptr, (const uint2*)adj, n, B, nb, blk);
    KERNEL_CHECK();
    return 0;
}

GEOCB_API int geocb_sssp_multi_vm(const void* adj, const uint32_t* blk, int B, int64_t n, const int32_t* sources,
                                  int S, double delta, int64_t blk_min_active, void* workspace, int64_t workspace_bytes,
                                  float* out_d32, int64_t* stats_host, void* stream) {
    ARG_CHECK(adj && blk && sources && out_d32 && workspace);
    ARG_CHECK(n > 0 && n <= (1LL << 27) && S > 0 && B >= 1 && B <= 256 && delta > 0);
    VmLayout w = vm_layout(n);
    if (workspace_bytes < w.total) {
        geocb_set_error("sssp_vm workspace too small: %lld < %lld", (long long)workspace_bytes, (long long)w.total);
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
        p.ub = (uint32_t*)(ws + w.ub);
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
        CU_TRY(cudaLaunchCooperativeKernel((void*)sssp_vm_kernel, dim3(grid), dim3(kThreads), args, 0, st));
    }
    if (stats_host) CU_TRY(cudaMemcpyAsync(stats_host, ws + w.stats, 64, cudaMemcpyDeviceToHost, st));
    int failed = 0;
    CU_TRY(cudaMemcpyAsync(&failed, ws + w.stats + kFailedOff, 4, cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaStreamSynchronize(st));
    if (failed) {
        geocb_set_error("sssp_vm_kernel did not converge within the round limit");
        return GEOCB_ERR_OVERFLOW;
    }
    return 0;
}
