// sssp.cu -- multi-source / incremental-Voronoi shortest paths on the adjacency-pair graph.
//
// Replaces scipy.sparse.csgraph.dijkstra as called by the reference at
// src/geo/geo_shortest_paths.py:36-49 (and through it kmeans_optimized.py:42,97,125).
//
// Algorithm: near-far (delta-stepping) frontier relaxation, push style, fp64-exact.
//   dist[v] is an fp64 kept as its uint64 bit pattern (non-negative doubles order as unsigned integers) so a
//   u64 min IS an fp64 min.  A relaxation computes fl64(dist[u] + (double)w) -- the same rounded addition, in
//   the same source->target order, scipy's Dijkstra performs -- and because rounded addition is monotone the
//   label-correcting fixed point is bit-identical to Dijkstra's result.
//
// What the hardware allows (tools/ubench/gather.cu, B200): 275 G random 4/8-byte gathers/s when the gathered
// array stays in L2 (<= ~60 MB), ~50 G/s when it comes from DRAM.  One relaxation = one gather of dist[v], so
// the design keeps the gathered array small and everything else sequential or fire-and-forget:
//   ub[v]    float32 upper bound, ub = round_up_f32(dist) -- the ONLY array gathered per relaxation (4 B/vertex/
//            source; L2 evict_last hint -- a persisting-L2 carve-out made things worse and is not used).
//            cand cannot improve v  <=>  round_down_f32(cand) >= ub[v]  (exact);
//            otherwise red.min.u64 dist[v], red.min.u32 ub[v] and a frontier bit are fired without waiting.
//   inq[2]   frontier bitmaps (1 bit per vertex and source, double-buffered by round parity): improvements
//            below the slot threshold set inq[next]; a round sweeps inq[cur] word by word, so popped vertices
//            are consecutive -- dist/proc/row_ptr reads coalesce and adjacency rows stream sequentially.
//   dirty    bitmap of vertices improved to a value >= threshold ("far"); when a slot's frontier runs dry the
//            threshold rises by delta and dirty vertices with ub below it move to the frontier.
//   proc[v]  distance v's row was last relaxed with; a popped vertex is relaxed only if dist[v] < proc[v], which
//            drops the (rare) spurious frontier bits of the float filter and stale far marks exactly.
//   One persistent cooperative kernel runs all rounds of `slots` concurrent sources; rounds are separated by
//   grid.sync.  A warp flattens the rows of its <=32 popped vertices into one edge range and walks it 128 edges
//   per trip (positions interleaved across lanes, so each edge load instruction reads 32 consecutive pairs = 8
//   sectors), the next trip's edge loads issued before this trip's gathers are consumed.
//   What bounds it (ncu): the L1 tag stage -- a gather of 32 random addresses is 32 tag lookups, one
//   per clock per SM, 291 G lookups/s chip-wide -- not DRAM bandwidth (42 % of peak) and not latency (16 to 24
//   resident warps per SM give the same throughput).
// HBM model: 8 B edge pair + 8 B distance per relaxation = 16*(nnz+n) algorithmic bytes per source.
#include <stdlib.h>
#include "common.cuh"

namespace {

constexpr int kMaxSlots = 32;
#ifndef GEOCB_SSSP_THREADS
#define GEOCB_SSSP_THREADS 256
#endif
#ifndef GEOCB_SSSP_MINBLOCKS
#define GEOCB_SSSP_MINBLOCKS 3
#endif
constexpr int kThreads = GEOCB_SSSP_THREADS;
#ifndef GEOCB_SSSP_PER
#define GEOCB_SSSP_PER 4
#endif
constexpr int kPer = GEOCB_SSSP_PER;  // edges (= independent ub gathers in flight) per lane per trip: 4 or 8
constexpr int kTile = 32 * kPer;      // flat edges per warp trip
typedef unsigned long long u64;
// stats block at the end of the workspace: 4 counters + per-round (start time, popped) trace of the last launch;
// GEOCB_SSSP_TRACE_PHASES (profiling builds only) adds per-round end-of-sweep / end-of-scan times at u64 [256+r], [384+r]
#ifdef GEOCB_SSSP_TRACE_PHASES
constexpr int kStatsBytes = 4096;
#else
constexpr int kStatsBytes = 2048;
#endif

#ifdef GEOCB_SSSP_UB16
typedef uint16_t ub_t;  // bf16 bits
#else
typedef uint32_t ub_t;  // f32 bits
#endif

struct SlotCtl {
    double thr2[2];         // near/far threshold, double-buffered by the slot's scan count (block-local epoch)
    uint32_t scan_min[3];   // min ub (f32 bits) over vertices left dirty by the scan of round r (index r%3)
    int scan_pushed[3];
    int active[3];          // "the slot has frontier bits for the next round", index (r+1)%3 written in round r
    int pad;
    u64 work[3];            // (slot 0 only) phase-A grab cursor: round r uses [r%3], clears [(r+2)%3]
};

struct SsspParams {
    const int64_t* row_ptr;
    const uint2* adj;
    int64_t n;
    int64_t nwords;  // bitmap words per slot (multiple of 32)
    int nslots;
    int max_rounds;  // safety net: a kernel that cannot converge must not hang the device
    int* failed;     // device flag, set when max_rounds was hit
    double delta;
    u64* dist;         // [nslots][n] fp64 bits
    u64* proc;         // [nslots][n] fp64 bits
    ub_t* ub;          // [nslots][nub] upper bounds (f32 or bf16 bits), nub = n rounded up to a multiple of 4
    int64_t nub;
    uint32_t* inq[2];  // [nslots][nwords]
    uint32_t* dirty;   // [nslots][nwords]
    SlotCtl* ctl;      // [nslots]
    u64* stats;        // [4] counters + round trace; nullable
    // full-SSSP mode: nsources sources stream through the nslots tables (a table that converges writes its row out
    // and takes the next source at once); table i is first seeded in round i*stagger/nslots
    const int32_t* sources;  // [nsources]
    float* out_d32;          // [nsources][n]
    int nsources;
    int stagger;
    // voronoi mode
    int voronoi;
    int32_t center, code;
    float* dmin32;
    int32_t* label;
};

// ---- L2 residency control: the gathered ub array stays (evict_last), edge pairs stream through (evict_first) ----
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
// ---- the gathered upper-bound table: float32 bits (default) or bf16 bits (GEOCB_SSSP_UB16: half the footprint, so
//      twice as many slots stay L2 resident; the filter gets ~0.8 % looser, results are unchanged) ----
#ifdef GEOCB_SSSP_UB16
constexpr ub_t UB_INF = 0x7F80u;
__device__ __forceinline__ ub_t ub_floor(double c) { return (ub_t)(__float_as_uint(__double2float_rd(c)) >> 16); }
__device__ __forceinline__ ub_t ub_ceil(double c) { return (ub_t)((__float_as_uint(__double2float_ru(c)) + 0xFFFFu) >> 16); }
__device__ __forceinline__ uint32_t ub_f32bits(ub_t u) { return (uint32_t)u << 16; }
__device__ __forceinline__ ub_t ld_ub(const ub_t* p, uint64_t pol) {
    uint16_t v;
    asm volatile("ld.global.cg.L2::cache_hint.u16 %0, [%1], %2;" : "=h"(v) : "l"(p), "l"(pol));
    return v;
}
// min into element v of a slot's table: one 8-byte BF16x4 min whose other three lanes carry +inf (non-negative bf16
// values order like their bit patterns, and min with +inf leaves a lane unchanged)
__device__ __forceinline__ void red_min_ub(ub_t* slot_base, uint32_t v, ub_t val, uint64_t pol) {
    uint32_t b0 = 0x7F807F80u, b1 = 0x7F807F80u;
    const uint32_t sh = (v & 1u) * 16u;
    const uint32_t w = (0x7F807F80u & ~(0xFFFFu << sh)) | ((uint32_t)val << sh);
    if (v & 2u) b1 = w; else b0 = w;
    asm volatile("red.global.v2.bf16x2.min.noftz [%0], {%1, %2};" ::"l"(slot_base + (v & ~3u)), "r"(b0), "r"(b1) : "memory");
}
#else
constexpr ub_t UB_INF = F32_INF_BITS;
__device__ __forceinline__ ub_t ub_floor(double c) { return __float_as_uint(__double2float_rd(c)); }
__device__ __forceinline__ ub_t ub_ceil(double c) { return __float_as_uint(__double2float_ru(c)); }
__device__ __forceinline__ uint32_t ub_f32bits(ub_t u) { return u; }
__device__ __forceinline__ ub_t ld_ub(const ub_t* p, uint64_t pol) {
    uint32_t v;
#ifdef GEOCB_SSSP_NOHINT_UB
    asm volatile("ld.global.cg.u32 %0, [%1];" : "=r"(v) : "l"(p));
#else
    asm volatile("ld.global.cg.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(v) : "l"(p), "l"(pol));
#endif
    return v;
}
__device__ __forceinline__ void red_min_ub(ub_t* slot_base, uint32_t v, ub_t val, uint64_t pol) {
#ifdef GEOCB_SSSP_NOHINT_UB
    asm volatile("red.global.min.u32 [%0], %1;" ::"l"(slot_base + v), "r"(val) : "memory");
#else
    asm volatile("red.global.min.L2::cache_hint.u32 [%0], %1, %2;" ::"l"(slot_base + v), "r"(val), "l"(pol) : "memory");
#endif
}
#endif
__device__ __forceinline__ uint2 ld_edge(const uint2* p, uint64_t pol) {
    uint2 v;
#ifdef GEOCB_SSSP_NOHINT_EDGE
    asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p));
#else
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.u32 {%0,%1}, [%2], %3;"
                 : "=r"(v.x), "=r"(v.y)
                 : "l"(p), "l"(pol));
#endif
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

// Item (= lane index) owning each of this lane's kPer flat positions tile + 32*k + lane.  Positions are interleaved
// across the warp so that one load instruction covers 32 consecutive edge pairs (8 sectors, not 32).  Each lane holds
// one item [start, start+len); starts are non-decreasing, so the owner of p is the last lane whose start <= p (an
// empty item shares its start with its successor and can never be that lane while p is inside the flattened range).
__device__ __forceinline__ void locate(uint32_t tile, uint32_t start, int wl, int j[kPer]) {
#pragma unroll
    for (int k = 0; k < kPer; ++k) j[k] = 0;
#pragma unroll
    for (int st = 16; st; st >>= 1) {
#pragma unroll
        for (int k = 0; k < kPer; ++k) {
            const int probe = min(j[k] + st, 31);
            const uint32_t t = __shfl_sync(0xffffffffu, start, probe);
            if (j[k] + st <= 31 && t <= tile + 32u * k + wl) j[k] += st;
        }
    }
}

__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

__global__ void __launch_bounds__(kThreads, GEOCB_SSSP_MINBLOCKS) sssp_kernel(SsspParams p) {
    cg::grid_group grid = cg::this_grid();
    const int64_t n = p.n;
    const int64_t nwords = p.nwords;
    const int B = p.nslots;
    const int64_t gtid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t gthreads = (int64_t)gridDim.x * blockDim.x;
    const int64_t gwarp = gtid >> 5, nwarps = gthreads >> 5;
    const int wlane = threadIdx.x & 31;
    __shared__ double s_thr[kMaxSlots], s_hint[kMaxSlots];
    __shared__ int s_epoch[kMaxSlots], s_done[kMaxSlots], s_need[kMaxSlots];
    __shared__ int s_src[kMaxSlots], s_fin[kMaxSlots], s_seed[kMaxSlots];  // source index a table works on / hands over
    __shared__ int s_flag, s_alldone, s_next, s_any;
    const uint64_t pol_keep = policy_evict_last();
    const uint64_t pol_stream = policy_evict_first();

    // ---- init ----
    if (!p.voronoi) {
        for (int64_t i = gtid; i < (int64_t)B * n; i += gthreads) {
            p.dist[i] = D64_INF_BITS;
            p.proc[i] = D64_INF_BITS;
        }
        for (int64_t i = gtid; i < (int64_t)B * p.nub; i += gthreads) p.ub[i] = UB_INF;
        for (int64_t i = gtid; i < (int64_t)B * nwords; i += gthreads) {
            p.inq[0][i] = 0;
            p.inq[1][i] = 0;
            p.dirty[i] = 0;
        }
    }
    if (threadIdx.x < kMaxSlots) {
        s_epoch[threadIdx.x] = 0;
        s_done[threadIdx.x] = !p.voronoi;  // full-SSSP tables start idle and are seeded by the hand-over step
        s_hint[threadIdx.x] = 0.0;
        s_src[threadIdx.x] = -1;
    }
    if (threadIdx.x == 0) s_next = 0;
    grid.sync();
    if (gtid < B) {
        const int s = (int)gtid;
        SlotCtl c;
        c.thr2[0] = c.thr2[1] = p.delta;
        for (int i = 0; i < 3; ++i) c.scan_min[i] = F32_INF_BITS, c.scan_pushed[i] = 0, c.active[i] = 0;
        c.pad = 0;
        c.work[0] = c.work[1] = c.work[2] = 0;
        if (p.voronoi) {
            const int src = p.center;
            const u64 old = atomicMin(p.dist + src, 0ULL);
            if (old != 0ULL) {  // (re-inserting an existing centre changes nothing)
                p.ub[src] = 0;
                atomicOr(p.inq[0] + (src >> 5), 1u << (src & 31));
            }
        }
        p.ctl[s] = c;
    }
    grid.sync();

    unsigned n_relax = 0, n_improve = 0, n_rounds = 0, n_scans = 0, n_slot_rounds = 0;
    const uint2 pad_edge = make_uint2(0u, F32_INF_BITS);
    const int64_t ngroups = nwords / 32;  // grabs per slot

    for (int round = 0;; ++round) {
        const int c_cur = round % 3, c_next = (round + 1) % 3, c_clear = (round + 2) % 3;
        uint32_t* inq_cur = (round & 1) ? p.inq[1] : p.inq[0];
        uint32_t* inq_next = (round & 1) ? p.inq[0] : p.inq[1];
        if (!p.voronoi) {
            // ---- hand-over (decided identically in every block): an idle table writes the row of the source it has
            //      finished and is seeded with the next one.  A finished table's frontier and far bitmaps are all-zero
            //      (that is what "converged" means), so only dist / proc / ub are reset, in the same pass. ----
            if (threadIdx.x == 0) {
                int any = 0, next = s_next, live = 0;
                for (int s = 0; s < B; ++s) live += !s_done[s];
                for (int s = 0; s < B; ++s) {
                    s_fin[s] = s_seed[s] = -1;
                    if (!s_done[s]) continue;
                    s_fin[s] = s_src[s];
                    const bool due = s_src[s] >= 0 || round >= (int)(((int64_t)s * p.stagger) / B) || live == 0;
                    if (next < p.nsources && due) {
                        s_seed[s] = next++;
                        ++live;
                    }
                    any |= (s_fin[s] >= 0) | (s_seed[s] >= 0);
                }
                s_next = next;
                s_any = any;
            }
            __syncthreads();
            if (s_any) {
                for (int s = 0; s < B; ++s) {
                    const int fin = s_fin[s], seed = s_seed[s];
                    if (fin < 0 && seed < 0) continue;
                    u64* dist_s = p.dist + (int64_t)s * n;
                    u64* proc_s = p.proc + (int64_t)s * n;
                    ub_t* ub_s = p.ub + (int64_t)s * p.nub;
                    const int64_t srcv = seed >= 0 ? (int64_t)p.sources[seed] : -1;
                    if (fin >= 0) {
                        float* out = p.out_d32 + (int64_t)fin * n;
                        for (int64_t v = gtid; v < n; v += gthreads) {
                            // (.cg: the previous hand-over of this table read the same addresses)
                            out[v] = __double2float_rn(__longlong_as_double((long long)ld_cg_u64(dist_s + v)));
                            if (seed >= 0) {
                                dist_s[v] = v == srcv ? 0ULL : D64_INF_BITS;
                                proc_s[v] = D64_INF_BITS;
                            }
                        }
                        if (seed >= 0)
                            for (int64_t v = gtid; v < p.nub; v += gthreads) ub_s[v] = v == srcv ? (ub_t)0 : UB_INF;
                    } else if (gtid == 0) {  // first use of the table: everything is +inf already
                        dist_s[srcv] = 0ULL;
                        ub_s[srcv] = 0;
                    }
                    if (seed >= 0 && gtid == 0) {
                        atomicOr(inq_cur + (int64_t)s * nwords + (srcv >> 5), 1u << (srcv & 31));
                        p.ctl[s].thr2[0] = p.ctl[s].thr2[1] = p.delta;
                    }
                }
                __syncthreads();
                if (threadIdx.x < B && (s_fin[threadIdx.x] >= 0 || s_seed[threadIdx.x] >= 0)) {
                    const int s = threadIdx.x;
                    s_src[s] = s_seed[s];
                    if (s_seed[s] >= 0) s_done[s] = 0, s_epoch[s] = 0, s_hint[s] = 0.0;
                }
                grid.sync();
            }
        }
        // ---- termination test (uniform) ----
        if (threadIdx.x == 0) s_alldone = 1;
        __syncthreads();
        if (threadIdx.x < B && !s_done[threadIdx.x]) s_alldone = 0;
        __syncthreads();
        if (s_alldone) break;
        if (round > p.max_rounds) {  // uniform
            if (gtid == 0) *p.failed = 1;
            break;
        }
        if (threadIdx.x == 0)
            for (int s = 0; s < B; ++s) n_slot_rounds += !s_done[s];
        if (gtid < B) p.ctl[gtid].active[c_clear] = 0;  // last read before the previous round's grid.sync
        if (gtid == 0) p.ctl[0].work[c_clear] = 0;     // last used before the previous round's grid.sync
        if (threadIdx.x < B) s_thr[threadIdx.x] = ld_cg_f64(&p.ctl[threadIdx.x].thr2[s_epoch[threadIdx.x] & 1]);
        __syncthreads();
        if (p.stats && gtid == 0 && round < 120) {  // round trace for tools/prof_sssp.py
            u64 t;
            asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
            p.stats[8 + 2 * round] = t;
        }
        // ---- phase A: sweep the frontier bitmaps, relax the rows of the popped vertices ----
        // A warp grabs 32 frontier words (1024 vertices of one slot) at a time, compacts their set bits and pops
        // them 32 per chunk, so chunks stay full on sparse frontiers; the next grab is in flight meanwhile.
        unsigned popped = 0;
        const int64_t ngrab = (int64_t)B * ngroups;
        u64* wcursor = &p.ctl[0].work[c_cur];
        int64_t g = 0;
        if (wlane == 0) g = (int64_t)atomicAdd(wcursor, 1ULL);
        g = __shfl_sync(0xffffffffu, g, 0);
        while (g < ngrab) {
            int64_t g_next = 0;
            if (wlane == 0) g_next = (int64_t)atomicAdd(wcursor, 1ULL);
            const int s = (int)(g / ngroups);
            const int64_t w0 = (g - (int64_t)s * ngroups) * 32;
            uint32_t word = 0;
            if (!s_done[s]) word = ld_cg_u32(inq_cur + (int64_t)s * nwords + w0 + wlane);
            if (!__ballot_sync(0xffffffffu, word != 0)) {
                g = __shfl_sync(0xffffffffu, g_next, 0);
                continue;
            }
            u64* dist_s = p.dist + (int64_t)s * n;
            u64* proc_s = p.proc + (int64_t)s * n;
            ub_t* ub_s = p.ub + (int64_t)s * p.nub;
            uint32_t* inqn_s = inq_next + (int64_t)s * nwords;
            uint32_t* dirty_s = p.dirty + (int64_t)s * nwords;
            if (word) {
                inq_cur[(int64_t)s * nwords + w0 + wlane] = 0;  // consumed
                red_and_u32(dirty_s + w0 + wlane, ~word);      // popped vertices are no longer "far"
            }
            const double thr = s_thr[s];
            bool near_any = false;
            uint32_t wincl = __popc(word);
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, wincl, o);
                if (wlane >= o) wincl += t;
            }
            const uint32_t wexcl = wincl - __popc(word);
            const uint32_t nitems = __shfl_sync(0xffffffffu, wincl, 31);
            for (uint32_t c0 = 0; c0 < nitems; c0 += 32) {
                // ---- pop: lane = k-th set bit of the 32 words; relax its row only if its distance dropped since
                //      it was last relaxed ----
                const uint32_t k = c0 + wlane;
                int wi = 0;  // word holding item k: first lane whose inclusive count exceeds k
#pragma unroll
                for (int st = 16; st; st >>= 1) {
                    const uint32_t t = __shfl_sync(0xffffffffu, wincl, wi + st - 1);
                    if (t <= k) wi += st;
                }
                wi = min(wi, 31);
                const uint32_t wv = __shfl_sync(0xffffffffu, word, wi);
                const uint32_t we = __shfl_sync(0xffffffffu, wexcl, wi);
                uint32_t len = 0, beg = 0;
                double du = 0.0;
                if (k < nitems) {
                    const int64_t u = (w0 + wi) * 32 + (int)__fns(wv, 0, (int)(k - we) + 1);
                    const u64 cur = ld_cg_u64(dist_s + u);
                    const u64 pr = ld_cg_u64(proc_s + u);
                    const int64_t b64 = __ldg(p.row_ptr + u), e64 = __ldg(p.row_ptr + u + 1);
                    if (cur < pr) {
                        proc_s[u] = cur;
                        len = (uint32_t)(e64 - b64);
                        beg = (uint32_t)b64;
                        du = __longlong_as_double((long long)cur);
                        ++popped;
                        // the row is streamed from HBM exactly once: start it towards L2 now, the trips below
                        // then see L2 latency (rows longer than 4 lines are covered by the trip pipeline)
                        const char* row = (const char*)(p.adj + b64);
                        const int rb = (int)(len * 8u);
#pragma unroll
                        for (int o = 0; o < 512; o += 128)
                            if (o < rb) prefetch_l2(row + o);
                    }
                }
                uint32_t incl = len;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
                    if (wlane >= o) incl += t;
                }
                const uint32_t start = incl - len;
                const uint32_t T = __shfl_sync(0xffffffffu, incl, 31);
                const uint32_t rowoff = beg - start;  // adj index of flat position f inside this item = rowoff + f (mod 2^32)
                if (T == 0) continue;
                int jj[kPer];
                uint2 e[kPer];
                locate(0, start, wlane, jj);
#pragma unroll
                for (int k = 0; k < kPer; ++k) {
                    const uint32_t f = 32u * k + wlane;
                    const uint32_t off = __shfl_sync(0xffffffffu, rowoff, jj[k]);
                    e[k] = f < T ? ld_edge(p.adj + (uint32_t)(off + f), pol_stream) : pad_edge;
                }
                for (uint32_t tile = 0; tile < T; tile += kTile) {
                    double duk[kPer];
                    ub_t ubv[kPer];
#pragma unroll
                    for (int k = 0; k < kPer; ++k) duk[k] = __shfl_sync(0xffffffffu, du, jj[k]);
#pragma unroll
                    for (int k = 0; k < kPer; ++k) ubv[k] = ld_ub(ub_s + e[k].x, pol_keep);
#pragma unroll
                    for (int k = 0; k < kPer; ++k) n_relax += tile + 32u * k + wlane < T;
                    // next trip's edge pairs are requested before this trip's gathers are consumed
                    uint2 en[kPer];
                    int jn[kPer];
#pragma unroll
                    for (int k = 0; k < kPer; ++k) en[k] = pad_edge, jn[k] = 0;
                    if (tile + kTile < T) {
                        locate(tile + kTile, start, wlane, jn);
#pragma unroll
                        for (int k = 0; k < kPer; ++k) {
                            const uint32_t f = tile + kTile + 32u * k + wlane;
                            const uint32_t off = __shfl_sync(0xffffffffu, rowoff, jn[k]);
                            if (f < T) en[k] = ld_edge(p.adj + (uint32_t)(off + f), pol_stream);
                        }
                    }
#pragma unroll
                    for (int k = 0; k < kPer; ++k) {
                        const uint32_t v = e[k].x;
                        const double cand = duk[k] + (double)__uint_as_float(e[k].y);
                        // cand >= ub[v] (>= dist[v])  <=>  round_down_f32(cand) >= ub[v]: nothing to do
                        if (ub_floor(cand) < ubv[k]) {
                            red_min_u64(dist_s + v, (u64)__double_as_longlong(cand));
                            red_min_ub(ub_s, v, ub_ceil(cand), pol_keep);
                            ++n_improve;
                            if (cand < thr) {
                                red_or_u32(inqn_s + (v >> 5), 1u << (v & 31));
                                near_any = true;
                            } else {
                                red_or_u32(dirty_s + (v >> 5), 1u << (v & 31));
                            }
                        }
                    }
#pragma unroll
                    for (int k = 0; k < kPer; ++k) e[k] = en[k], jj[k] = jn[k];
                }
            }
            if (__any_sync(0xffffffffu, near_any) && wlane == 0) *(volatile int*)&p.ctl[s].active[c_next] = 1;
            g = __shfl_sync(0xffffffffu, g_next, 0);
        }
        grid.sync();
        ++n_rounds;
#ifdef GEOCB_SSSP_TRACE_PHASES
        if (p.stats && gtid == 0 && round < 120) {  // every block has finished its sweep
            u64 t;
            asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
            p.stats[256 + round] = t;
        }
#endif
        if (p.stats && round < 120) {
            popped = __reduce_add_sync(0xffffffffu, popped);
            if (wlane == 0 && popped) atomicAdd(p.stats + 9 + 2 * round, (u64)popped);
        }
        // ---- control: which slots ran dry? (uniform across the grid) ----
        if (gtid < B) {  // scan accumulators of round-1 (index c_clear): every block read them before this sync
            p.ctl[gtid].scan_min[c_clear] = F32_INF_BITS;
            p.ctl[gtid].scan_pushed[c_clear] = 0;
        }
        if (threadIdx.x == 0) s_flag = 0;
        __syncthreads();
        if (threadIdx.x < B) {
            const int s = threadIdx.x;
            const int need = !s_done[s] && ld_cg_s32(&p.ctl[s].active[c_next]) == 0;
            s_need[s] = need;
            if (need) s_flag = 1;
        }
        __syncthreads();
        if (s_flag) {
            ++n_scans;
            // ---- phase B: raise the threshold of every dry slot and move its dirty vertices below it to the frontier ----
            for (int s = 0; s < B; ++s) {
                if (!s_need[s]) continue;
                double thr_new = s_thr[s] + p.delta;
                if (s_hint[s] > thr_new) thr_new = s_hint[s];
                uint32_t* dirty_s = p.dirty + (int64_t)s * nwords;
                uint32_t* inqn_s = inq_next + (int64_t)s * nwords;
                const ub_t* ub_s = p.ub + (int64_t)s * p.nub;
                uint32_t rem_min = F32_INF_BITS;
                int pushed = 0;
                for (int64_t w0 = gwarp * 32; w0 < nwords; w0 += nwarps * 32) {
                    const uint32_t word = ld_cg_u32(dirty_s + w0 + wlane);
                    uint32_t taken = 0;
                    unsigned any = __ballot_sync(0xffffffffu, word != 0);
                    while (any) {
                        int src[4];
                        uint32_t d[4];
                        bool bit[4];
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            src[k] = any ? __ffs(any) - 1 : -1;
                            any &= any - 1;
                            const uint32_t wv = __shfl_sync(0xffffffffu, word, src[k] < 0 ? 0 : src[k]);
                            bit[k] = src[k] >= 0 && ((wv >> wlane) & 1u);
                            d[k] = F32_INF_BITS;
                            if (bit[k]) d[k] = ub_f32bits(ld_ub(ub_s + (w0 + src[k]) * 32 + wlane, pol_keep));
                        }
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const bool take = bit[k] && (double)__uint_as_float(d[k]) < thr_new;
                            const unsigned bal = __ballot_sync(0xffffffffu, take);
                            if (wlane == src[k]) taken |= bal;
                            pushed += __popc(bal);
                            if (bit[k] && !take) rem_min = min(rem_min, d[k]);
                        }
                    }
                    if (taken) {
                        dirty_s[w0 + wlane] = word & ~taken;
                        red_or_u32(inqn_s + w0 + wlane, taken);
                    }
                }
                rem_min = __reduce_min_sync(0xffffffffu, rem_min);
                if (wlane == 0) {
                    if (rem_min != F32_INF_BITS) atomicMin(&p.ctl[s].scan_min[c_cur], rem_min);
                    // (active[c_next] must not be touched here: slower blocks may still be reading it to decide
                    //  whether to enter this phase, and its grid.sync, at all)
                    if (pushed) atomicAdd(&p.ctl[s].scan_pushed[c_cur], pushed);
                }
                if (gtid == 0) p.ctl[s].thr2[(s_epoch[s] + 1) & 1] = thr_new;
            }
            grid.sync();
            if (threadIdx.x < B && s_need[threadIdx.x]) {  // identical in every block
                const int s = threadIdx.x;
                const int pushed = ld_cg_s32(&p.ctl[s].scan_pushed[c_cur]);
                const uint32_t mn = ld_cg_u32(&p.ctl[s].scan_min[c_cur]);
                s_epoch[s] += 1;
                s_hint[s] = 0.0;
                if (pushed == 0) {
                    if (mn == F32_INF_BITS) s_done[s] = 1;  // frontier empty and nothing dirty: converged
                    else s_hint[s] = (double)__uint_as_float(mn) + p.delta;  // jump over the gap
                }
            }
            __syncthreads();
        }
#ifdef GEOCB_SSSP_TRACE_PHASES
        if (p.stats && gtid == 0 && round < 120) {  // control + (if any slot ran dry) threshold scan done
            u64 t;
            asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
            p.stats[384 + round] = t;
        }
#endif
    }

    // ---- finalize (full-SSSP rows were written by the hand-over step) ----
    if (p.voronoi) {
        // dmin32 = f32(dmin64); label = code iff the f32 value strictly decreased
        // (reference: argmin over k of float32(D[k,v]), lowest k on ties, kmeans_optimized.py:100)
        for (int64_t v = gtid; v < n; v += gthreads) {
            float nd = __double2float_rn(__longlong_as_double((long long)p.dist[v]));
            if (nd < p.dmin32[v]) {
                p.dmin32[v] = nd;
                p.label[v] = p.code;
            }
        }
    }
    if (p.stats) {
        if (gtid == 0) {
            u64 t;
            asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
            p.stats[6] = t;
            p.stats[7] = (u64)n_rounds;
        }
        n_relax = __reduce_add_sync(0xffffffffu, n_relax);
        n_improve = __reduce_add_sync(0xffffffffu, n_improve);
        if (wlane == 0) {
            atomicAdd(p.stats + 0, (u64)n_relax);
            atomicAdd(p.stats + 3, (u64)n_improve);
        }
        if (gtid == 0) {
            atomicAdd(p.stats + 1, (u64)n_rounds);
            atomicAdd(p.stats + 2, (u64)n_scans);
            atomicAdd(p.stats + 4, (u64)n_slot_rounds);  // sum over sources of the rounds they were live
        }
    }
}

// Predecessors from the converged fp64 distances: pred[s][v] = lowest-index u with an edge u->v and
// fl64(dist[u] + w(u,v)) == dist[v]  (-9999 for the source and unreachable vertices, scipy's convention).
// The scan runs over v's IN-edges, so a directed graph passes the adjacency of its transpose (in_row_ptr / in_adj);
// a symmetric graph passes its own.  8 lanes per vertex.
__global__ void sssp_pred_kernel(const int64_t* in_row_ptr, const uint2* in_adj, int64_t n, int B, const u64* dist,
                                 const int32_t* sources, int32_t* out_pred) {
    constexpr int kGroup = 8;  // shuffles name only the group's own lanes
    const int64_t gtid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t gthreads = (int64_t)gridDim.x * blockDim.x;
    const int lane = threadIdx.x & (kGroup - 1);
    const unsigned gmask = ((1u << kGroup) - 1u) << (((threadIdx.x & 31) / kGroup) * kGroup);
    const int64_t ggroup = gtid / kGroup, ngroupsp = gthreads / kGroup;
    for (int64_t item = ggroup; item < (int64_t)B * n; item += ngroupsp) {
        const int s = (int)(item / n);
        const int64_t v = item - (int64_t)s * n;
        const u64* dist_s = dist + (int64_t)s * n;
        const u64 dv = dist_s[v];
        int best = 0x7fffffff;
        if (dv != D64_INF_BITS && v != sources[s]) {
            const int64_t beg = in_row_ptr[v], end = in_row_ptr[v + 1];
            for (int64_t i = beg + lane; i < end; i += kGroup) {
                const uint2 e = in_adj[i];
                const double c0 = __longlong_as_double((long long)dist_s[e.x]) + (double)__uint_as_float(e.y);
                if ((u64)__double_as_longlong(c0) == dv && (int)e.x < best) best = (int)e.x;
            }
        }
        for (int o = kGroup / 2; o; o >>= 1) best = min(best, __shfl_xor_sync(gmask, best, o, kGroup));
        if (lane == 0) out_pred[item] = best == 0x7fffffff ? -9999 : best;
    }
}

int coop_grid(int* grid_out) {
    DevInfo di;
    int rc = geocb_devinfo(&di);
    if (rc) return rc;
    int occ = 0;
    CU_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, sssp_kernel, kThreads, 0));
    if (occ < 1) {
        geocb_set_error("sssp_kernel cannot be resident (occupancy 0)");
        return GEOCB_ERR_UNSUPPORTED;
    }
    *grid_out = di.sm_count * occ;
    return 0;
}

struct WsLayout {
    int64_t dist, proc, ub, inq0, inq1, dirty, ctl, stats, total, nwords;
};
WsLayout ws_layout(int64_t n, int slots, bool own_dist) {
    WsLayout w;
    int64_t off = 0;
    w.nwords = align_up((n + 31) / 32, 32);
    w.dist = off;
    if (own_dist) off += align_up((int64_t)slots * n * 8, 256);
    w.proc = off;
    off += align_up((int64_t)slots * n * 8, 256);
    w.ub = off;
    off += align_up((int64_t)slots * align_up(n, 4) * (int64_t)sizeof(ub_t), 256);
    w.inq0 = off;
    off += align_up((int64_t)slots * w.nwords * 4, 256);
    w.inq1 = off;
    off += align_up((int64_t)slots * w.nwords * 4, 256);
    w.dirty = off;
    off += align_up((int64_t)slots * w.nwords * 4, 256);
    w.ctl = off;
    off += align_up((int64_t)kMaxSlots * sizeof(SlotCtl), 256);
    w.stats = off;
    off += kStatsBytes;  // 4 counters + a (time, popped vertices) trace of the last launch's first 120 rounds
    w.total = off;
    return w;
}

void fill_params(SsspParams& p, const WsLayout& w, char* ws, const int64_t* row_ptr, const void* adj, int64_t n,
                 double delta) {
    memset(&p, 0, sizeof(p));
    p.row_ptr = row_ptr;
    p.adj = (const uint2*)adj;
    p.n = n;
    p.nwords = w.nwords;
    p.delta = delta;
    p.dist = (u64*)(ws + w.dist);
    p.proc = (u64*)(ws + w.proc);
    p.ub = (ub_t*)(ws + w.ub);
    p.nub = align_up(n, 4);
    p.inq[0] = (uint32_t*)(ws + w.inq0);
    p.inq[1] = (uint32_t*)(ws + w.inq1);
    p.dirty = (uint32_t*)(ws + w.dirty);
    p.ctl = (SlotCtl*)(ws + w.ctl);
    p.failed = (int*)(ws + w.stats + 40);
    const char* env = getenv("GEOCB_SSSP_MAX_ROUNDS");
    const int64_t lim = env ? atoll(env) : 8 * n + 4096;
    p.max_rounds = (int)(lim < 0x7fffff00 ? lim : 0x7fffff00);
}

__global__ void voronoi_init_kernel(int64_t n, int64_t nwords, u64* d64, float* d32, int32_t* label, u64* proc,
                                    ub_t* ub, int64_t nub, uint32_t* inq0, uint32_t* inq1, uint32_t* dirty) {
    const int64_t t0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nt = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = t0; i < n; i += nt) {
        d64[i] = D64_INF_BITS;
        d32[i] = __uint_as_float(F32_INF_BITS);
        label[i] = -1;
        proc[i] = D64_INF_BITS;
    }
    for (int64_t i = t0; i < nub; i += nt) ub[i] = UB_INF;
    for (int64_t i = t0; i < nwords; i += nt) inq0[i] = inq1[i] = dirty[i] = 0;
}

}  // namespace

GEOCB_API int64_t geocb_sssp_workspace_bytes(int64_t n, int slots) {
    if (n < 0 || slots < 1 || slots > kMaxSlots) return -1;
    return ws_layout(n, slots, true).total;
}

GEOCB_API int64_t geocb_voronoi_workspace_bytes(int64_t n) {
    if (n < 0) return -1;
    return ws_layout(n, 1, false).total;
}

GEOCB_API int geocb_sssp_multi(const int64_t* row_ptr, const void* adj, int64_t n, const int32_t* sources, int S,
                               double delta, int slots, int stagger_rounds, void* workspace, int64_t workspace_bytes,
                               float* out_d32, int32_t* out_pred, const int64_t* in_row_ptr, const void* in_adj,
                               int64_t* stats_host, void* stream) {
    ARG_CHECK(row_ptr && adj && sources && out_d32 && workspace);
    ARG_CHECK((in_row_ptr == nullptr) == (in_adj == nullptr));
    ARG_CHECK(n > 0 && n < (1LL << 31) - 64 && S > 0 && slots >= 1 && slots <= kMaxSlots && delta > 0);
    ARG_CHECK(stagger_rounds >= 0 && stagger_rounds <= 4096);
    // (adjacency offsets are 32-bit inside the kernel: geocb_adj_count refuses graphs with >= 2^32 pairs)
    WsLayout w = ws_layout(n, slots, true);
    if (workspace_bytes < w.total) {
        geocb_set_error("sssp workspace too small: %lld < %lld", (long long)workspace_bytes, (long long)w.total);
        return GEOCB_ERR_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    char* ws = (char*)workspace;
    int grid = 0;
    int rc = coop_grid(&grid);
    if (rc) return rc;
    CU_TRY(cudaMemsetAsync(ws + w.stats, 0, kStatsBytes, st));
    // Distances only: ONE launch, the S sources stream through `slots` live tables.  With predecessors: batches of
    // `slots` sources, because the predecessor pass reads a table's converged fp64 distances after the launch.
    const int batch = out_pred ? slots : S;
    for (int s0 = 0; s0 < S; s0 += batch) {
        SsspParams p;
        fill_params(p, w, ws, row_ptr, adj, n, delta);
        p.nsources = (S - s0) < batch ? (S - s0) : batch;
        p.nslots = p.nsources < slots ? p.nsources : slots;
        p.stagger = out_pred ? 0 : stagger_rounds;
        {  // the round limit is per source; sources queue up behind each other
            const int64_t waves = (p.nsources + p.nslots - 1) / p.nslots + 1;
            const int64_t lim = (int64_t)p.max_rounds * waves + p.stagger;
            p.max_rounds = (int)(lim < 0x7fffff00 ? lim : 0x7fffff00);
        }
        p.stats = (u64*)(ws + w.stats);
        p.sources = sources + s0;
        p.out_d32 = out_d32 + (int64_t)s0 * n;
        p.voronoi = 0;
        if (s0 + batch < S || s0 > 0) CU_TRY(cudaMemsetAsync(ws + w.stats + 64, 0, kStatsBytes - 64, st));  // trace = last launch
        void* args[] = {&p};
        CU_TRY(cudaLaunchCooperativeKernel((void*)sssp_kernel, dim3(grid), dim3(kThreads), args, 0, st));
        if (out_pred) {
            sssp_pred_kernel<<<grid * 2, 256, 0, st>>>(in_row_ptr ? in_row_ptr : row_ptr,
                                                       (const uint2*)(in_adj ? in_adj : adj), n, p.nslots, p.dist,
                                                       p.sources, out_pred + (int64_t)s0 * n);
            KERNEL_CHECK();
        }
    }
    if (stats_host) CU_TRY(cudaMemcpyAsync(stats_host, ws + w.stats, 40, cudaMemcpyDeviceToHost, st));
    int failed = 0;
    CU_TRY(cudaMemcpyAsync(&failed, ws + w.stats + 40, 4, cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaStreamSynchronize(st));
    if (failed) {
        geocb_set_error("sssp_kernel did not converge within the round limit");
        return GEOCB_ERR_OVERFLOW;
    }
    return 0;
}

GEOCB_API int geocb_voronoi_init(int64_t n, double* dmin64, float* dmin32, int32_t* label, void* workspace,
                                 int64_t workspace_bytes, void* stream) {
    ARG_CHECK(n > 0 && dmin64 && dmin32 && label && workspace);
    WsLayout w = ws_layout(n, 1, false);
    if (workspace_bytes < w.total) {
        geocb_set_error("voronoi workspace too small: %lld < %lld", (long long)workspace_bytes, (long long)w.total);
        return GEOCB_ERR_WORKSPACE;
    }
    char* ws = (char*)workspace;
    CU_TRY(cudaMemsetAsync(ws + w.stats, 0, kStatsBytes, (cudaStream_t)stream));
    voronoi_init_kernel<<<592, 256, 0, (cudaStream_t)stream>>>(
        n, w.nwords, (u64*)dmin64, dmin32, label, (u64*)(ws + w.proc), (ub_t*)(ws + w.ub), align_up(n, 4), (uint32_t*)(ws + w.inq0),
        (uint32_t*)(ws + w.inq1), (uint32_t*)(ws + w.dirty));
    KERNEL_CHECK();
    return 0;
}

GEOCB_API int geocb_voronoi_insert(const int64_t* row_ptr, const void* adj, int64_t n, int32_t center, int32_t code,
                                   double delta, double* dmin64, float* dmin32, int32_t* label, void* workspace,
                                   int64_t workspace_bytes, void* stream) {
    ARG_CHECK(row_ptr && adj && dmin64 && dmin32 && label && workspace);
    ARG_CHECK(n > 0 && n < (1LL << 31) - 64 && center >= 0 && center < n && delta > 0);
    WsLayout w = ws_layout(n, 1, false);
    if (workspace_bytes < w.total) {
        geocb_set_error("voronoi workspace too small: %lld < %lld", (long long)workspace_bytes, (long long)w.total);
        return GEOCB_ERR_WORKSPACE;
    }
    char* ws = (char*)workspace;
    int grid = 0;
    int rc = coop_grid(&grid);
    if (rc) return rc;
    // workspace contract: initialised by geocb_voronoi_init and passed unchanged to every insert (proc, ub persist;
    // the frontier and dirty bitmaps are all-zero again when the kernel exits)
    SsspParams p;
    fill_params(p, w, ws, row_ptr, adj, n, delta);
    p.nslots = 1;
    p.dist = (u64*)dmin64;
    p.stats = nullptr;  // (p.failed still points into the stats block: read back by geocb_voronoi_check)
    p.voronoi = 1;
    p.center = center;
    p.code = code;
    p.dmin32 = dmin32;
    p.label = label;
    void* args[] = {&p};
    CU_TRY(cudaLaunchCooperativeKernel((void*)sssp_kernel, dim3(grid), dim3(kThreads), args, 0, (cudaStream_t)stream));
    return 0;
}

GEOCB_API int geocb_voronoi_check(int64_t n, void* workspace, int64_t workspace_bytes, void* stream) {
    ARG_CHECK(n > 0 && workspace);
    WsLayout w = ws_layout(n, 1, false);
    ARG_CHECK(workspace_bytes >= w.total);
    int failed = 0;
    CU_TRY(cudaMemcpyAsync(&failed, (char*)workspace + w.stats + 40, 4, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    CU_TRY(cudaStreamSynchronize((cudaStream_t)stream));
    if (failed) {
        geocb_set_error("a voronoi insert hit the round limit without converging");
        return GEOCB_ERR_OVERFLOW;
    }
    return 0;
}
