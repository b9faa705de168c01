// sssp.cu -- multi-source / incremental-Voronoi shortest paths on the adjacency-pair graph.
//
// Replaces scipy.sparse.csgraph.dijkstra as called by the reference at
// src/geo/geo_shortest_paths.py:36-49 (and through it kmeans_optimized.py:42,97,125).
//
// Algorithm: near-far (delta-stepping) frontier relaxation, push style.
//   dist[v] is an fp64 kept as its uint64 bit pattern (non-negative doubles order as
//   unsigned integers) so `atomicMin(u64)` is an exact fp64 min.  A relaxation computes
//   fl64(dist[u] + (double)w) -- the same rounded addition, in the same source->target
//   order, scipy's Dijkstra performs -- and because rounded addition is monotone the
//   label-correcting fixed point is bit-identical to Dijkstra's result.
//   One persistent cooperative kernel runs all rounds of `slots` concurrent sources:
//     phase A  every queued (slot,u): 16 lanes stream u's (col,w) pairs with 128-bit loads,
//              gather dist[v] from L2, atomicMin the improvements and queue v
//              (near: cand < thr[slot]) or mark it dirty (far);
//     phase B  for slots whose near queue ran dry: thr = min(dirty dist)+delta, queue dirty<thr.
// HBM model: 8 B per relaxation streamed (the pair) + 8 B dist gather (L2 resident
// when slots*n*8 B fits) ; algorithmic bytes = 16*(nnz+n) per source.
#include "common.cuh"

namespace {

constexpr int kMaxSlots = 32;
constexpr int kThreads = 256;
#ifndef GEOCB_GROUP
#define GEOCB_GROUP 8
#endif
constexpr int kGroup = GEOCB_GROUP;  // lanes cooperating on one vertex (each lane: 2 x uint4 = 4 pairs per step)
constexpr int kGroupsPerWarp = 32 / kGroup;

struct SlotCtl {
    double thr;                  // current near/far threshold
    unsigned long long far_min;  // min dist bits over dirty vertices (phase B1)
    int active[3];               // "something was queued for this slot", rotating like GlobalCtl::qcount
    int done;
    int pad[2];
};
struct GlobalCtl {
    int qcount[3];  // combined queue sizes: round r reads [r%3], fills [(r+1)%3], clears [(r+2)%3]
    int work[3];    // dynamic work cursor of phase A, same rotation
};

struct SsspParams {
    const int64_t* row_ptr;
    const uint4* adj4;
    int64_t n;
    int nslots;
    double delta;
    unsigned long long* dist;  // [nslots][n]
    int* state;                // [nslots][n]  0 clean, 1 dirty (far), 2 queued (near)
    uint32_t* queue[2];        // combined work lists, entries (slot << 27) | vertex, capacity nslots*n each
    SlotCtl* ctl;              // [nslots]
    GlobalCtl* gctl;
    unsigned long long* stats; // [4]
    // full-SSSP mode
    const int32_t* sources;  // [nslots]
    float* out_d32;          // [nslots][n]
    int32_t* out_pred;       // nullable
    // voronoi mode
    int voronoi;
    int32_t center, code;
    float* dmin32;
    int32_t* label;
};

constexpr int kVertexBits = 27;
constexpr uint32_t kVertexMask = (1u << kVertexBits) - 1;
constexpr int kStage = 128;  // per-warp staging entries before one aggregated global atomicAdd

// queue push through the warp's shared-memory staging buffer (same-address global atomics on the queue
// tail serialise in L2: ~1 cycle each, which capped the first version at 5 % of roofline)
__device__ __forceinline__ void stage_push(uint32_t entry, uint32_t* wbuf, int* wcnt, uint32_t* qnext, int* qtail) {
    const int pos = atomicAdd(wcnt, 1);
    if (pos < kStage) wbuf[pos] = entry;
    else qnext[atomicAdd(qtail, 1)] = entry;  // staging full (rare): straight to the global list
}
// warp-synchronous flush; every lane of the warp must call it
__device__ __forceinline__ void stage_flush(uint32_t* wbuf, int* wcnt, uint32_t* qnext, int* qtail, int min_fill) {
    __syncwarp();
    int cnt = *(volatile int*)wcnt;
    if (cnt >= min_fill && cnt > 0) {
        cnt = cnt < kStage ? cnt : kStage;
        const int wl = threadIdx.x & 31;
        int base = 0;
        if (wl == 0) base = atomicAdd(qtail, cnt);
        base = __shfl_sync(0xffffffffu, base, 0);
        for (int i = wl; i < cnt; i += 32) qnext[base + i] = wbuf[i];
        __syncwarp();
        if (wl == 0) *(volatile int*)wcnt = 0;
    }
    __syncwarp();
}

__device__ __forceinline__ void red_min_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("red.global.min.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void red_max_s32(int* p, int v) {
    asm volatile("red.global.max.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// The update half of a relaxation; the caller has already gathered `cur` = dist[v], so a lane keeps several
// independent gathers in flight.  No returning atomic sits on the critical path: the distance is lowered with
// a fire-and-forget red.min (a stale `cur` only costs a redundant red), and the queue flag is read first and
// exchanged only when the vertex is not queued yet (about once per vertex per source).
__device__ __forceinline__ void relax_update(uint32_t v, double cand, unsigned long long cur,
                                             unsigned long long* dist_s, int* state_s, double thr, uint32_t slot_tag,
                                             unsigned& active_mask, uint32_t* wbuf, int* wcnt, uint32_t* qnext,
                                             int* qtail, unsigned& n_improve) {
    const unsigned long long cb = (unsigned long long)__double_as_longlong(cand);
    if (cb < cur) {
        red_min_u64(dist_s + v, cb);
        ++n_improve;
        const int st = ld_cg_s32(state_s + v);
        if (cand < thr) {
            if (st != 2 && atomicExch(state_s + v, 2) != 2) {
                active_mask |= 1u << (slot_tag >> kVertexBits);
                stage_push(slot_tag | v, wbuf, wcnt, qnext, qtail);
            }
        } else if (st == 0) {
            red_max_s32(state_s + v, 1);
        }
    }
}

#ifndef GEOCB_MINBLOCKS
#define GEOCB_MINBLOCKS 4
#endif
__global__ void __launch_bounds__(kThreads, GEOCB_MINBLOCKS) sssp_kernel(SsspParams p) {
    cg::grid_group grid = cg::this_grid();
    const int64_t n = p.n;
    const int B = p.nslots;
    const int64_t gtid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t gthreads = (int64_t)gridDim.x * blockDim.x;
    __shared__ int s_need[kMaxSlots];
    __shared__ int s_flag, s_alldone;
    __shared__ unsigned long long s_red[kThreads / 32];
    __shared__ uint32_t s_stage[kThreads / 32][kStage];
    __shared__ int s_stage_n[kThreads / 32];

    // ---- init ----
    if (!p.voronoi) {
        for (int64_t i = gtid; i < (int64_t)B * n; i += gthreads) {
            p.dist[i] = D64_INF_BITS;
            p.state[i] = 0;
        }
    }
    if (threadIdx.x < kThreads / 32) s_stage_n[threadIdx.x] = 0;
    if (gtid == 0) {
        p.gctl->qcount[0] = 0;
        p.gctl->qcount[1] = 0;
        p.gctl->qcount[2] = 0;
        p.gctl->work[0] = p.gctl->work[1] = p.gctl->work[2] = 0;
    }
    grid.sync();
    if (gtid < B) {
        int s = (int)gtid;
        int src = p.voronoi ? p.center : p.sources[s];
        SlotCtl c;
        c.thr = p.delta;
        c.far_min = D64_INF_BITS;
        c.active[0] = c.active[1] = c.active[2] = 0;
        c.done = 0;
        unsigned long long old = atomicMin(p.dist + (int64_t)s * n + src, 0ULL);
        if (old != 0ULL) {  // (voronoi: re-inserting an existing centre changes nothing)
            p.state[(int64_t)s * n + src] = 2;
            p.queue[0][atomicAdd(&p.gctl->qcount[0], 1)] = ((uint32_t)s << kVertexBits) | (uint32_t)src;
            c.active[0] = 1;
        }
        p.ctl[s] = c;
    }
    grid.sync();

    unsigned n_relax = 0, n_improve = 0, n_rounds = 0, n_scans = 0;
    const int lane = threadIdx.x & (kGroup - 1);
    const int wlane = threadIdx.x & 31;
    const int grp = (threadIdx.x & 31) / kGroup;
    // the two 16-lane groups of a warp may hold different items: shuffles name only their own lanes
    const unsigned gmask = (kGroup == 32 ? 0xFFFFFFFFu : ((1u << kGroup) - 1u)) << (grp * kGroup);
    const int64_t nwarps = gthreads >> 5;
    uint32_t* wbuf = s_stage[threadIdx.x >> 5];
    int* wcnt = &s_stage_n[threadIdx.x >> 5];

    for (int round = 0;; ++round) {
        const int par = round & 1;  // queue buffer parity
        const int c_cur = round % 3, c_next = (round + 1) % 3, c_clear = (round + 2) % 3;
        if (gtid < B) {  // safe: all of these were last touched before the previous grid.sync
            p.ctl[gtid].active[c_clear] = 0;
            p.ctl[gtid].far_min = D64_INF_BITS;
        }
        if (gtid == 0) {
            p.gctl->qcount[c_clear] = 0;
            p.gctl->work[c_clear] = 0;
        }
        // ---- phase A0: un-queue every listed vertex (state 2 -> 0) before any distance is read, so that a
        //      later improvement of the same vertex re-queues it; the grid.sync replaces a per-item fence ----
        const int total = ld_cg_s32(&p.gctl->qcount[c_cur]);
        for (int64_t i = gtid; i < total; i += gthreads) {
            const uint32_t entry = p.queue[par][i];
            p.state[(int64_t)(entry >> kVertexBits) * n + (entry & kVertexMask)] = 0;
        }
        grid.sync();
        // ---- phase A: relax every queued (slot, vertex).  Each 16-lane group prefetches a chunk of `chunk`
        //      items (entry, dist[u], row bounds, threshold: independent loads, one per lane) and then walks
        //      them; the trip count is uniform per warp so the staging flush can be warp-synchronous ----
        uint32_t* qnext = p.queue[par ^ 1];
        int* qtail = &p.gctl->qcount[c_next];
        unsigned active_mask = 0;
        int chunk = 1;
        while (chunk < kGroup && (int64_t)chunk * 4 * nwarps * kGroupsPerWarp <= total) chunk <<= 1;
        const int per_warp = kGroupsPerWarp * chunk;
        const uint4 pad4 = make_uint4(0, F32_INF_BITS, 0, F32_INF_BITS);
        for (;;) {
            // warps grab work dynamically: rows differ in length and rounds end at a grid barrier
            int wbase = 0;
            if (wlane == 0) wbase = atomicAdd(&p.gctl->work[c_cur], per_warp);
            wbase = __shfl_sync(0xffffffffu, wbase, 0);
            if (wbase >= total) break;
            const int my_item = wbase + grp * chunk + lane;
            uint32_t entry = 0, beg = 0, end = 0;
            double du = 0.0, thr = 0.0;
            if (lane < chunk && my_item < total) {
                entry = p.queue[par][my_item];
                const int s = (int)(entry >> kVertexBits);
                const uint32_t u = entry & kVertexMask;
                du = __longlong_as_double((long long)ld_cg_u64(p.dist + (int64_t)s * n + u));
                thr = __longlong_as_double((long long)ld_cg_u64((const unsigned long long*)&p.ctl[s].thr));
                beg = (uint32_t)(__ldg(p.row_ptr + u) >> 1);
                end = (uint32_t)(__ldg(p.row_ptr + u + 1) >> 1);
            }
            // software pipeline: the pair loads of item j+1 are issued before item j's gathers are consumed
            uint32_t nbeg = __shfl_sync(gmask, beg, 0, kGroup), nend = __shfl_sync(gmask, end, 0, kGroup);
            uint4 n0 = pad4, n1 = pad4;
            if (nbeg + lane < nend) n0 = ld_stream_v4(p.adj4 + nbeg + lane);
            if (nbeg + lane + kGroup < nend) n1 = ld_stream_v4(p.adj4 + nbeg + lane + kGroup);
            for (int j = 0; j < chunk; ++j) {
                const uint32_t e_j = __shfl_sync(gmask, entry, j, kGroup);
                const double du_j = __shfl_sync(gmask, du, j, kGroup);
                const double thr_j = __shfl_sync(gmask, thr, j, kGroup);
                const uint32_t beg_j = nbeg, end_j = nend;
                uint4 e0 = n0, e1 = n1;
                if (j + 1 < chunk) {
                    nbeg = __shfl_sync(gmask, beg, j + 1, kGroup);
                    nend = __shfl_sync(gmask, end, j + 1, kGroup);
                    n0 = pad4;
                    n1 = pad4;
                    if (nbeg + lane < nend) n0 = ld_stream_v4(p.adj4 + nbeg + lane);
                    if (nbeg + lane + kGroup < nend) n1 = ld_stream_v4(p.adj4 + nbeg + lane + kGroup);
                }
                const int64_t sbase = (int64_t)(e_j >> kVertexBits) * n;
                unsigned long long* dist_s = p.dist + sbase;
                int* state_s = p.state + sbase;
                const uint32_t tag = e_j & ~kVertexMask;
                for (uint32_t i = beg_j + lane;;) {
                    // four independent distance gathers in flight, then the (rare) updates
                    const unsigned long long d0 = ld_cg_u64(dist_s + e0.x), d1 = ld_cg_u64(dist_s + e0.z);
                    const unsigned long long d2 = ld_cg_u64(dist_s + e1.x), d3 = ld_cg_u64(dist_s + e1.z);
                    n_relax += (i < end_j ? 2 : 0) + (i + kGroup < end_j ? 2 : 0);
                    relax_update(e0.x, du_j + (double)__uint_as_float(e0.y), d0, dist_s, state_s, thr_j, tag, active_mask,
                                 wbuf, wcnt, qnext, qtail, n_improve);
                    relax_update(e0.z, du_j + (double)__uint_as_float(e0.w), d1, dist_s, state_s, thr_j, tag, active_mask,
                                 wbuf, wcnt, qnext, qtail, n_improve);
                    relax_update(e1.x, du_j + (double)__uint_as_float(e1.y), d2, dist_s, state_s, thr_j, tag, active_mask,
                                 wbuf, wcnt, qnext, qtail, n_improve);
                    relax_update(e1.z, du_j + (double)__uint_as_float(e1.w), d3, dist_s, state_s, thr_j, tag, active_mask,
                                 wbuf, wcnt, qnext, qtail, n_improve);
                    i += 2 * kGroup;
                    if (i >= end_j) break;  // rows longer than 4*kGroup pairs take further trips
                    e0 = ld_stream_v4(p.adj4 + i);
                    e1 = i + kGroup < end_j ? ld_stream_v4(p.adj4 + i + kGroup) : pad4;
                }
                stage_flush(wbuf, wcnt, qnext, qtail, kStage / 2);
            }
        }
        stage_flush(wbuf, wcnt, qnext, qtail, 1);
        for (int o = 16; o; o >>= 1) active_mask |= __shfl_xor_sync(0xffffffffu, active_mask, o);
        if (wlane == 0)
            for (unsigned m = active_mask; m; m &= m - 1) *(volatile int*)&p.ctl[__ffs(m) - 1].active[c_next] = 1;
        grid.sync();
        ++n_rounds;
        // ---- control: which slots ran dry? (uniform across the grid) ----
        if (threadIdx.x == 0) s_flag = 0;
        __syncthreads();
        if (threadIdx.x < B) {
            int s = threadIdx.x;
            int need = (!ld_cg_s32(&p.ctl[s].done)) && ld_cg_s32(&p.ctl[s].active[c_next]) == 0;
            s_need[s] = need;
            if (need) s_flag = 1;
        }
        __syncthreads();
        const int any_scan = s_flag;
        if (any_scan) {
            ++n_scans;
            // ---- phase B1: min dist over dirty vertices of each dry slot ----
            for (int s = 0; s < B; ++s) {
                if (!s_need[s]) continue;
                const int64_t sbase = (int64_t)s * n;
                unsigned long long m = D64_INF_BITS;
                for (int64_t v = gtid; v < n; v += 4 * gthreads) {
                    int st[4];
#pragma unroll
                    for (int k = 0; k < 4; ++k) st[k] = v + k * gthreads < n ? ld_cg_s32(p.state + sbase + v + k * gthreads) : 0;
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        if (st[k] == 1) {
                            unsigned long long d = ld_cg_u64(p.dist + sbase + v + k * gthreads);
                            m = d < m ? d : m;
                        }
                }
                for (int o = 16; o; o >>= 1) {
                    unsigned long long t = __shfl_xor_sync(0xffffffffu, m, o);
                    m = t < m ? t : m;
                }
                if (wlane == 0) s_red[threadIdx.x >> 5] = m;
                __syncthreads();
                if (threadIdx.x == 0) {
                    for (int w = 1; w < kThreads / 32; ++w) m = s_red[w] < m ? s_red[w] : m;
                    if (m != D64_INF_BITS) atomicMin(&p.ctl[s].far_min, m);
                }
                __syncthreads();
            }
            grid.sync();
            // ---- phase B2: raise thr, queue dirty vertices below it (ballot-aggregated tail bump) ----
            for (int s = 0; s < B; ++s) {
                if (!s_need[s]) continue;
                const unsigned long long fm = ld_cg_u64(&p.ctl[s].far_min);
                if (fm == D64_INF_BITS) {  // nothing dirty: this source has converged
                    if (gtid == 0) p.ctl[s].done = 1;
                    continue;
                }
                const double thr_new = __longlong_as_double((long long)fm) + p.delta;
                const int64_t sbase = (int64_t)s * n;
                const int64_t n_round = (n + 31) & ~(int64_t)31;
                for (int64_t v0 = gtid; v0 < n_round; v0 += 4 * gthreads) {
                    int st[4];
#pragma unroll
                    for (int k = 0; k < 4; ++k) st[k] = v0 + k * gthreads < n ? ld_cg_s32(p.state + sbase + v0 + k * gthreads) : 0;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const int64_t v = v0 + k * gthreads;
                        if (v - wlane >= n_round) break;  // warp-uniform
                        bool take = false;
                        if (st[k] == 1) {
                            double d = __longlong_as_double((long long)ld_cg_u64(p.dist + sbase + v));
                            take = d < thr_new;
                        }
                        const unsigned bal = __ballot_sync(0xffffffffu, take);
                        if (bal) {
                            int wbase = 0;
                            if (wlane == 0) wbase = atomicAdd(qtail, __popc(bal));
                            wbase = __shfl_sync(0xffffffffu, wbase, 0);
                            if (take) {
                                p.state[sbase + v] = 2;
                                qnext[wbase + __popc(bal & ((1u << wlane) - 1))] = ((uint32_t)s << kVertexBits) | (uint32_t)v;
                            }
                        }
                    }
                }
                if (gtid == 0) {
                    p.ctl[s].thr = thr_new;
                    p.ctl[s].active[c_next] = 1;
                }
            }
            grid.sync();
        }
        // ---- termination test (uniform: `done` only changes in B2, which ends with a grid.sync) ----
        if (threadIdx.x == 0) s_alldone = 1;
        __syncthreads();
        if (threadIdx.x < B && !ld_cg_s32(&p.ctl[threadIdx.x].done)) s_alldone = 0;
        __syncthreads();
        if (s_alldone) break;
    }

    // ---- finalize ----
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
    } else {
        for (int64_t i = gtid; i < (int64_t)B * n; i += gthreads)
            p.out_d32[i] = __double2float_rn(__longlong_as_double((long long)p.dist[i]));
        if (p.out_pred) {
            const int64_t ggroup = gtid / kGroup, ngroups = gthreads / kGroup;
            for (int64_t item = ggroup; item < (int64_t)B * n; item += ngroups) {
                const int s = (int)(item / n);
                const int64_t v = item - (int64_t)s * n;
                const unsigned long long* dist_s = p.dist + (int64_t)s * n;
                const unsigned long long dv = dist_s[v];
                int best = 0x7fffffff;
                if (dv != D64_INF_BITS && v != p.sources[s]) {
                    const int64_t beg = p.row_ptr[v] >> 1, end = p.row_ptr[v + 1] >> 1;
                    for (int64_t i = beg + lane; i < end; i += kGroup) {
                        uint4 e = ld_stream_v4(p.adj4 + i);
                        double c0 = __longlong_as_double((long long)dist_s[e.x]) + (double)__uint_as_float(e.y);
                        double c1 = __longlong_as_double((long long)dist_s[e.z]) + (double)__uint_as_float(e.w);
                        if ((unsigned long long)__double_as_longlong(c0) == dv && (int)e.x < best) best = (int)e.x;
                        if ((unsigned long long)__double_as_longlong(c1) == dv && (int)e.z < best) best = (int)e.z;
                    }
                }
                for (int o = kGroup / 2; o; o >>= 1) best = min(best, __shfl_xor_sync(gmask, best, o, kGroup));
                if (lane == 0) p.out_pred[item] = best == 0x7fffffff ? -9999 : best;
            }
        }
    }
    if (p.stats) {
        for (int o = 16; o; o >>= 1) {
            n_relax += __shfl_xor_sync(0xffffffffu, n_relax, o);
            n_improve += __shfl_xor_sync(0xffffffffu, n_improve, o);
        }
        if ((threadIdx.x & 31) == 0) {
            atomicAdd(p.stats + 0, (unsigned long long)n_relax);
            atomicAdd(p.stats + 3, (unsigned long long)n_improve);
        }
        if (gtid == 0) {
            atomicAdd(p.stats + 1, (unsigned long long)n_rounds);
            atomicAdd(p.stats + 2, (unsigned long long)n_scans);
        }
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
    int64_t dist, state, q0, q1, ctl, gctl, stats, total;
};
WsLayout ws_layout(int64_t n, int slots, bool own_dist) {
    WsLayout w;
    int64_t off = 0;
    w.dist = off;
    if (own_dist) off += align_up((int64_t)slots * n * 8, 256);
    w.state = off;
    off += align_up((int64_t)slots * n * 4, 256);
    w.q0 = off;
    off += align_up((int64_t)slots * n * 4, 256);
    w.q1 = off;
    off += align_up((int64_t)slots * n * 4, 256);
    w.ctl = off;
    off += align_up((int64_t)kMaxSlots * sizeof(SlotCtl), 256);
    w.gctl = off;
    off += 256;
    w.stats = off;
    off += 256;
    w.total = off;
    return w;
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
                               double delta, int slots, void* workspace, int64_t workspace_bytes, float* out_d32,
                               int32_t* out_pred, int64_t* stats_host, void* stream) {
    ARG_CHECK(row_ptr && adj && sources && out_d32 && workspace);
    ARG_CHECK(n > 0 && n < (1LL << kVertexBits) && S > 0 && slots >= 1 && slots <= kMaxSlots && delta > 0);
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
    CU_TRY(cudaMemsetAsync(ws + w.stats, 0, 32, st));
    for (int s0 = 0; s0 < S; s0 += slots) {
        SsspParams p;
        memset(&p, 0, sizeof(p));
        p.row_ptr = row_ptr;
        p.adj4 = (const uint4*)adj;
        p.n = n;
        p.nslots = (S - s0) < slots ? (S - s0) : slots;
        p.delta = delta;
        p.dist = (unsigned long long*)(ws + w.dist);
        p.state = (int*)(ws + w.state);
        p.queue[0] = (uint32_t*)(ws + w.q0);
        p.queue[1] = (uint32_t*)(ws + w.q1);
        p.ctl = (SlotCtl*)(ws + w.ctl);
        p.gctl = (GlobalCtl*)(ws + w.gctl);
        p.stats = (unsigned long long*)(ws + w.stats);
        p.sources = sources + s0;
        p.out_d32 = out_d32 + (int64_t)s0 * n;
        p.out_pred = out_pred ? out_pred + (int64_t)s0 * n : nullptr;
        p.voronoi = 0;
        void* args[] = {&p};
        CU_TRY(cudaLaunchCooperativeKernel((void*)sssp_kernel, dim3(grid), dim3(kThreads), args, 0, st));
    }
    if (stats_host) CU_TRY(cudaMemcpyAsync(stats_host, ws + w.stats, 32, cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaStreamSynchronize(st));
    return 0;
}

namespace {
__global__ void voronoi_init_kernel(int64_t n, unsigned long long* d64, float* d32, int32_t* label) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        d64[i] = D64_INF_BITS;
        d32[i] = __uint_as_float(F32_INF_BITS);
        label[i] = -1;
    }
}
}  // namespace

GEOCB_API int geocb_voronoi_init(int64_t n, double* dmin64, float* dmin32, int32_t* label, void* stream) {
    ARG_CHECK(n > 0 && dmin64 && dmin32 && label);
    voronoi_init_kernel<<<592, 256, 0, (cudaStream_t)stream>>>(n, (unsigned long long*)dmin64, dmin32, label);
    KERNEL_CHECK();
    return 0;
}

GEOCB_API int geocb_voronoi_insert(const int64_t* row_ptr, const void* adj, int64_t n, int32_t center, int32_t code,
                                   double delta, double* dmin64, float* dmin32, int32_t* label, void* workspace,
                                   int64_t workspace_bytes, void* stream) {
    ARG_CHECK(row_ptr && adj && dmin64 && dmin32 && label && workspace);
    ARG_CHECK(n > 0 && n < (1LL << kVertexBits) && center >= 0 && center < n && delta > 0);
    WsLayout w = ws_layout(n, 1, false);
    if (workspace_bytes < w.total) {
        geocb_set_error("voronoi workspace too small: %lld < %lld", (long long)workspace_bytes, (long long)w.total);
        return GEOCB_ERR_WORKSPACE;
    }
    char* ws = (char*)workspace;
    int grid = 0;
    int rc = coop_grid(&grid);
    if (rc) return rc;
    SsspParams p;
    memset(&p, 0, sizeof(p));
    p.row_ptr = row_ptr;
    p.adj4 = (const uint4*)adj;
    p.n = n;
    p.nslots = 1;
    p.delta = delta;
    p.dist = (unsigned long long*)dmin64;
    p.state = (int*)(ws + w.state);  // must be all-zero on entry; the kernel leaves it all-zero
    p.queue[0] = (uint32_t*)(ws + w.q0);
    p.queue[1] = (uint32_t*)(ws + w.q1);
    p.ctl = (SlotCtl*)(ws + w.ctl);
    p.gctl = (GlobalCtl*)(ws + w.gctl);
    p.stats = nullptr;
    p.voronoi = 1;
    p.center = center;
    p.code = code;
    p.dmin32 = dmin32;
    p.label = label;
    void* args[] = {&p};
    CU_TRY(cudaLaunchCooperativeKernel((void*)sssp_kernel, dim3(grid), dim3(kThreads), args, 0, (cudaStream_t)stream));
    return 0;
}
