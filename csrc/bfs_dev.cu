// bfs_dev.cu — the persistent, direction-optimising bidirectional-BFS kernel (sm_100a) and the
// parent-walk path kernel.
//
// ONE cooperative launch per rank answers a whole s-t query: level loop, frontier exchange
// between ranks, meet-in-the-middle test, termination and min-distance reduction all happen on
// the device.  The reference needs, per level, ~10 blocking runtime calls and 3 host syncs on one
// GPU (v3/bibfs_cuda_only.cu:173-203) or a 4N-byte H2D + 8N-byte D2H + MPI_Allreduce(BOR) over a
// host int[N] + two scalar allreduces with several ranks (v4/comp.cu:84-107,
// v4/mpi_bas.cpp:79-132).  What replaces what:
//
//   expand_frontier(_kernel)  (v3:13-43, comp.cu:20-38; one thread per vertex over all N)
//        -> expand_td: frontier-proportional work, warp-cooperative neighbour gathering with
//           batched independent loads/atomics; rows >= HUB_ROW are deferred to a hub phase in
//           which the whole grid splits them into 1 KB chunks (HUB_CHUNK entries);
//           expand_bu: bottom-up sweep over unvisited owned vertices against a frontier bitmap
//   check_intersect(_kernel) + host O(N) scan  (v3:45-62, comp.cu:40-47, mpi_bas.cpp:81-87)
//        -> fused: the visited bitmap word carries both sides' bits, so the atomic OR that claims
//           a vertex also returns "was it reached from the other side" (discovery-time test)
//   MPI_Allreduce(BOR) of the next frontier  (mpi_bas.cpp:107; second_try.cpp:82-104)
//        -> a remote atom.or on the OWNER's bitmap word + remote stores of the parent and of the
//           vertex into the owner's per-sender queue segment, straight over NVLink (peer pointers)
//   MPI_Allreduce(LOR/SUM) for found / changed / frontier sizes  (mpi_bas.cpp:85,124; second_try:115-124)
//        -> four tagged 8-byte words per peer (value | epoch tag << 32, st.relaxed.sys); every rank
//           polls its own slots until all tags match and sums them itself (cross_exchange)
//   the blocking collectives' implicit barrier
//        -> device-side barrier: 16 local arrive counters + a generation word; across ranks the
//           tagged words ARE the barrier (no fence.sys, no release/acquire pair on the level
//           path).  A watchdog turns a lost rank into an error instead of a hang.
//
// Scheduling: the side with the smaller (global) frontier is expanded, ties to the source side
// (v1/main-v1.cpp:51, mpi_bas.cpp:90-92).  Small local frontiers are expanded by CTA 0 alone
// ("solo"): level state and accumulators live in shared memory and levels are separated by
// __syncthreads only, while the other CTAs sleep on the generation word.  Large ones wake the
// whole grid; very large ones switch to bottom-up (Beamer et al., SC'12).
//
// Exactness: levels are always completed; the first level in which any discovered vertex is
// already marked by the other side ends the search and length = depth_s + depth_t (all hits of
// that level have the same length — SURVEY 7.4 item 6, the argument that makes v1 correct).
//
// With P "virtual ranks" hosted by one launch (blockIdx.x / cpr = rank) the very same protocol
// runs on a single GPU, which is how the multi-rank path is tested without 8 GPUs.
#include "bfs.h"
#include <cstdlib>
#include <type_traits>
#include "ptx.cuh"

namespace bibfs {

#ifdef BIBFS_FINE_TRACE
__device__ u64 g_dbg[256][8];
__device__ u64 g_dbg2[256][8];
__device__ u32 g_dbg_level;
__device__ u64 g_kt[32];      // kernel-level stamps of the leader thread: entry, init done, tier entry / exit, path walk, end
#define KT_STAMP(i)                                                                              \
    do {                                                                                         \
        if (blockIdx.x == 0 && threadIdx.x == 0) {                                               \
            u64 _t;                                                                              \
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(_t)::"memory");                     \
            g_kt[i] = _t;                                                                        \
        }                                                                                        \
    } while (0)
#define DBG_STAMP(slot, dep)                                                                     \
    do {                                                                                         \
        if (blockIdx.x == 0 && threadIdx.x == 0 && g_dbg_level < 256) {                          \
            u64 _t;                                                                              \
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(_t) : "l"((u64)(dep)) : "memory"); \
            g_dbg[g_dbg_level][slot] = _t;                                                       \
        }                                                                                        \
    } while (0)
#define DBG_STAMP2(slot, dep)                                                                    \
    do {                                                                                         \
        if (blockIdx.x == 0 && threadIdx.x == 0 && g_dbg_level > 0 && g_dbg_level <= 256) {       \
            u64 _t;                                                                              \
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(_t) : "l"((u64)(dep)) : "memory"); \
            g_dbg2[g_dbg_level - 1][slot] = _t;                                                  \
        }                                                                                        \
    } while (0)
#define DBG_STAMP3(slot, dep)                                                                    \
    do {                                                                                         \
        if (blockIdx.x == 1 && threadIdx.x == 0 && g_dbg_level > 0 && g_dbg_level <= 256) {       \
            u64 _t;                                                                              \
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(_t) : "l"((u64)(dep)) : "memory"); \
            g_dbg2[g_dbg_level - 1][slot] = _t;                                                  \
        }                                                                                        \
    } while (0)
#else
#define KT_STAMP(i) do { } while (0)
#define DBG_STAMP3(slot, dep) do { } while (0)
#define DBG_STAMP(slot, dep) do { } while (0)
#define DBG_STAMP2(slot, dep) do { } while (0)
#endif

__device__ __noinline__ void path_body(const KParams& K);   // defined with the path kernel below
__device__ __noinline__ void path_body_local(const KParams& K, const u32* par0, const u32* par1);
__device__ __noinline__ void result_to_host(const KParams& K);

namespace {

constexpr u32 FULL = 0xffffffffu;
constexpr u32 BIG_ROW = 64;          // rows at least this long are strip-mined by the whole warp
constexpr int UNR = 4;               // independent edges in flight per lane
constexpr u32 WPC = BFS_THREADS / 32;
constexpr u32 SOLO_CAP = 4096;       // entries of the shared-memory queue mirror per side (larger cluster levels read the queue itself)
constexpr u32 FILTER_MIN = 4096;
constexpr u32 STAGE_CAP = 4096;      // grid mode: discoveries a CTA stages in shared memory before one queue reservation     // frontiers at least this large pre-check the visited word

struct Env {
    u32 me, lo, P, B, n, nw;
    u32 cta, cpr;
    Ctrl* C;
    Acc* A;            // accumulators in use: &C->acc (grid) or the shared-memory copy (solo)
    Acc* sA;           // the shared-memory copy
    bool solo;
    u32* sq;           // solo mode: shared-memory mirror of the entries appended in this level (else null)
    u32 sq_base;       // queue index of sq[0]
    u32* s_hub;        // solo mode: shared flag "a hub row was deferred" (else null)
    bool filter;       // read the visited word before the atomic (pays off on large frontiers only)
    u32 tag;           // epoch tag stamped on entries sent to other ranks in the current level
    bool smem_src;     // visit_batch reads the adjacency entries from shared memory (hub chunks staged by TMA)
    // grid mode: per-CTA staging (same-address L2 atomics serialise: ~2400 warps bumping one queue cursor
    // cost 100+ us per level, 148 CTAs do not)
    u32* stage; u32* stage_cnt; u32* cta_new; unsigned long long* cta_edges;
    u32* s_sent;       // solo mode: shared copy of Ctrl::sent ([2][MAXR])
    u32* s_rnew;       // solo mode: shared counter of vertices discovered at other ranks
    u64* vis;
    u32* parent[2];
    u32* q[2];
    const u64* row_ptr;
    const u32* col;
};

__device__ __forceinline__ Ctrl* ctrl_of(const KParams& K, u32 r) { return reinterpret_cast<Ctrl*>(K.arena[r]); }
__device__ __forceinline__ u64* vis_of(const KParams& K, u32 r) { return reinterpret_cast<u64*>(K.arena[r] + K.L.off_vis); }
__device__ __forceinline__ u32* fbm_of(const KParams& K, u32 r, u32 side, u32 buf) {
    return reinterpret_cast<u32*>(K.arena[r] + K.L.off_fbm) + (size_t)(side * 2 + buf) * K.L.nw;
}
__device__ __forceinline__ u32* parent_of(const KParams& K, u32 r, u32 side) {
    return reinterpret_cast<u32*>(K.arena[r] + K.L.off_parent) + (size_t)side * K.B;
}
__device__ __forceinline__ u32* q_of(const KParams& K, u32 r, u32 side) {
    return reinterpret_cast<u32*>(K.arena[r] + K.L.off_q) + (size_t)side * K.B;
}
__device__ __forceinline__ u64* rq_of(const KParams& K, u32 r, u32 side, u32 sender) {
    return reinterpret_cast<u64*>(K.arena[r] + K.L.off_rq) + ((size_t)side * K.P + sender) * K.B;
}
__device__ __forceinline__ bibfs_level_trace* trace_of(const KParams& K, u32 r) {
    return reinterpret_cast<bibfs_level_trace*>(K.arena[r] + K.L.off_trace);
}
__device__ __forceinline__ HubEntry* hub_of(const KParams& K, u32 r) {
    return reinterpret_cast<HubEntry*>(K.arena[r] + K.L.off_hub);
}

// private scratch word of the calling thread: where the branch-free claim code sends its no-op atomics
__device__ __forceinline__ u64* dummy_of(const KParams& K, u32 r) {
    return reinterpret_cast<u64*>(K.arena[r] + K.L.off_dummy) + ((blockIdx.x * BFS_THREADS + threadIdx.x) % DUMMY_WORDS);
}

// Warp-aggregated reservation: every lane wants `c` consecutive slots; one atomic per warp on the cursor (returns the
// lane's first slot).  ADD is called by the last lane with the warp total.  Used when a level is wide enough for the
// 512 per-lane atomics on ONE shared-memory word to queue up.
template <typename AddFn>
__device__ __forceinline__ u32 warp_reserve(u32 c, AddFn add) {
    const u32 lane = threadIdx.x & 31;
    u32 incl = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const u32 t = __shfl_up_sync(FULL, incl, d);
        if (lane >= (u32)d) incl += t;
    }
    u32 base = 0;
    if (lane == 31 && incl) base = add(incl);
    base = __shfl_sync(FULL, base, 31);
    return base + incl - c;
}

// rank that owns vertex v (v / B without the 20-instruction division)
__device__ __forceinline__ u32 owner_of(const KParams& K, u32 v) { return (u32)__umul64hi(K.div_B, (u64)v); }

__device__ __forceinline__ u32 lanemask_lt() {
    u32 m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

// Append the vertices held by the lanes of `mask` (exactly those lanes call this, together) to the
// local queue of `side`.  Solo mode: shared-memory cursor, entries go straight to the queue (and its
// shared mirror).  Grid mode: entries are staged in the CTA's shared memory and written to the queue
// with ONE cursor reservation per CTA at the next synchronisation point (cta_flush); what does not
// fit in the stage falls back to a warp-sized reservation.
__device__ __forceinline__ void q_append(const Env& E, u32 side, u32 mask, u32 v) {
    const u32 lane = threadIdx.x & 31;
    const u32 leader = __ffs(mask) - 1;
    const u32 n = __popc(mask), r = __popc(mask & lanemask_lt());
    if (E.solo) {
        u32 base = 0;
        if (lane == leader) base = ptx::atoms_add(&E.sA->q_tail[side], n);
        base = __shfl_sync(mask, base, leader);
        const u32 slot = base + r;
        E.q[side][slot] = v;
        if (E.sq && slot - E.sq_base < SOLO_CAP) E.sq[slot - E.sq_base] = v;
    } else {
        u32 s0 = 0, g = 0;
        if (lane == leader) {
            s0 = ptx::atoms_add(E.stage_cnt, n);
            if (s0 + n > STAGE_CAP) g = atomicAdd(&E.A->q_tail[side], s0 + n - max(s0, STAGE_CAP));
        }
        s0 = __shfl_sync(mask, s0, leader);
        g = __shfl_sync(mask, g, leader);
        const u32 idx = s0 + r;
        if (idx < STAGE_CAP) E.stage[idx] = v;
        else E.q[side][g + (idx - max(s0, STAGE_CAP))] = v;
    }
}
__device__ __forceinline__ void acc_note_meet(const Env& E, u32 v) {
    if (E.solo) ptx::reds_min(&E.sA->meet_local, v); else atomicMin(&E.A->meet_local, v);
}
// Has this rank already seen a meeting vertex in the current level?  Every hit of a level gives the
// same path length, so the rest of the level can be skipped (what v1's `break` does, v1/main-v1.cpp:61).
__device__ __forceinline__ bool acc_meet_seen(const Env& E) {
    return (E.solo ? *reinterpret_cast<volatile u32*>(&E.sA->meet_local) : ptx::ld_relaxed_gpu(&E.A->meet_local)) != NO_V;
}
__device__ __forceinline__ void acc_flush(const Env& E, u32 cnt, u64 edges) {
    if (E.solo) {
        if (cnt) ptx::reds_add(&E.sA->new_count, cnt);
        if (edges) ptx::reds_add64(&E.sA->edges, edges);
    } else {
        if (cnt) ptx::reds_add(E.cta_new, cnt);
        if (edges) ptx::reds_add64(E.cta_edges, edges);
    }
}

// ------------------------------------------------------------------------------------------
// waiting with a watchdog
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ bool expired(Ctrl* C, u64 deadline) {
    return ptx::globaltimer_ns() > deadline || ptx::ld_relaxed_gpu(&C->abort_flag) != 0;
}

// ---- polling loops are WARP-COLLECTIVE -----------------------------------------------------
// Every lane of the calling warp runs the loop and all lanes leave it in the same iteration.
// Measured reason (profiles/latency_microbench.md): when ONE lane spins while the other lanes of
// its warp wait at the implicit reconvergence point, the hardware releases the waiting lanes after
// ~4 us (forward-progress guarantee of independent thread scheduling).  From then on the warp runs
// diverged and every later *_sync collective takes the WARPSYNC.COLLECTIVE slow path (~150 ns
// apiece): a level that had waited 4 us for a peer paid another 2.6 us for its shuffles.

__device__ __forceinline__ bool expired_warp(Ctrl* C, u64 deadline) {
    bool ex = false;
    if ((threadIdx.x & 31) == 0) ex = expired(C, deadline);
    return __shfl_sync(FULL, (int)ex, 0) != 0;
}

// whole warp.  `parked`: the CTA may wait for many levels (solo mode) -> back off; otherwise the release is imminent -> poll hard
__device__ bool wait_gen(Ctrl* C, u64 want, u64 deadline, bool parked = true) {
    const u32 lane = threadIdx.x & 31;
    u32 spins = 0;
    while (true) {
        u64 g = 0;
        if (lane == 0) g = ptx::ld_acquire_gpu(&C->bar_gen);
        g = __shfl_sync(FULL, g, 0);
        if (g >= want) return true;
        if (++spins > (parked ? 64u : 4096u)) {
            ptx::nanosleep(32);
            if ((spins & 255u) == 0 && expired_warp(C, deadline)) return false;
        }
    }
}

// leader warp: lane w < BAR_WAYS waits until every non-leader CTA mapped to way w has arrived, then resets the way
__device__ bool wait_count(Ctrl* C, u32 cpr, u64 deadline) {
    const u32 lane = threadIdx.x & 31;
    // CTAs 1..cpr-1 arrive; CTA c uses way c % BAR_WAYS
    const u32 want = (lane < BAR_WAYS && lane < cpr) ? (cpr - 1u - lane) / BAR_WAYS + (lane != 0 ? 1u : 0u) : 0u;
    bool ok = true, done = want == 0;
    u32 spins = 0;
    while (true) {
        if (!done) done = ptx::ld_acquire_gpu(&C->bar_way[lane].count) >= want;
        if (__all_sync(FULL, done)) break;
        if (++spins > 64 && (spins & 255u) == 0 && expired_warp(C, deadline)) { ok = false; break; }
    }
    if (want) ptx::st_relaxed_gpu(&C->bar_way[lane].count, 0u);
    return ok;
}

// whole warp; lanes with sl != nullptr poll the four tagged words a peer writes into this rank's slot until they carry `tag`
__device__ bool wait_slot(Ctrl* C, const LSlot* sl, u32 tag, Slot& out, u64 deadline) {
    bool done = sl == nullptr;
    u32 spins = 0;
    while (true) {
        if (!done) {
            const u64 a = ptx::ld_relaxed_sys(&sl->w[0]), b = ptx::ld_relaxed_sys(&sl->w[1]);
            const u64 c = ptx::ld_relaxed_sys(&sl->w[2]), d = ptx::ld_relaxed_sys(&sl->w[3]);
            if ((u32)(a >> 32) == tag && (u32)(b >> 32) == tag && (u32)(c >> 32) == tag && (u32)(d >> 32) == tag) {
                out.count = (u32)a; out.meet = (u32)b; out.rq_tail = (u32)c; out.edges = (u32)d;
                done = true;
            }
        }
        if (__all_sync(FULL, done)) return true;
        if (++spins > 64 && (spins & 255u) == 0 && expired_warp(C, deadline)) return false;
    }
}

// accumulators are updated with atomics that live in L2: read them around the L1
__device__ __forceinline__ void pull_acc(Acc& dst, const Acc* src) {
    dst.q_tail[0] = ptx::ld_relaxed_gpu(&src->q_tail[0]);
    dst.q_tail[1] = ptx::ld_relaxed_gpu(&src->q_tail[1]);
    dst.new_count = ptx::ld_relaxed_gpu(&src->new_count);
    dst.meet_local = ptx::ld_relaxed_gpu(&src->meet_local);
    dst.edges = ptx::ld_relaxed_gpu(reinterpret_cast<const u64*>(&src->edges));
}

// ------------------------------------------------------------------------------------------
// leader side: decisions
// ------------------------------------------------------------------------------------------
__device__ __noinline__ void decide_next(const KParams& K, LevelState& G) {
    const u32 side = G.nf[0] <= G.nf[1] ? 0u : 1u;   // ties -> source side, like v1
    G.side = side;
    if (G.nf[side] == 0) {                            // the side to expand is exhausted: no path
        G.length = -1;
        G.phase = PH_DONE;
        return;
    }
    u32 lf = G.q_hi[side] - G.q_lo[side];
    for (u32 p = 0; p < K.P; ++p) lf += G.rq_hi[side][p] - G.rq_lo[side][p];
    G.local_front = lf;
    bool bu = false;
    if (K.flags & BIBFS_F_FORCE_BU) bu = true;
    else if (!(K.flags & BIBFS_F_FORCE_TD) && K.bu_div != 0xFFFFFFFFu) bu = G.nf[side] > K.n / K.bu_div;
    if (bu) G.phase = PH_BU;
    else if (!(K.flags & BIBFS_F_FORCE_GRID) && lf <= K.solo_threshold) G.phase = PH_TD_SOLO;
    else G.phase = PH_TD_GRID;
}

__device__ __noinline__ void write_result(const KParams& K, Ctrl* C, const LevelState& G) {
    Result& R = C->res;
    R.length = G.length;
    R.meet = G.meet;
    R.depth_s = G.depth[0];
    R.depth_t = G.depth[1];
    R.levels = G.levels;
    R.levels_bu = G.levels_bu;
    R.status = (i32)G.status;
    R.path_len = 0;
    R.edges_scanned = G.edges_total;
    R.vertices_visited = G.visited_total;
    R.levels_grid = G.levels_grid;
    R.rounds_paired = G.rounds_paired;
    { const u64 dt = ptx::globaltimer_ns() - C->t_begin_ns; R.search_ns = dt > 0xFFFFFFFFull ? 0xFFFFFFFFu : (u32)dt; R.path_ns = 0; }
    C->xepoch = G.xepoch;
}

// Cross-rank handshake executed by the leader warp (lanes 0..P-1 each talk to one peer).
// Writes `mine` as four tagged words into slot[parity][me] of every peer with relaxed system-scope
// stores and polls this rank's own slots until every peer's words carry the epoch tag.  `fenced`
// adds a fence.acq_rel.sys on both sides: needed only when bulk data that is read with plain loads
// (bottom-up bitmap slices, parents at the end of the query) must be ordered with the exchange;
// queue entries are tagged themselves and need no fence.  Returns false on timeout.
template <bool MULTI>
__device__ __noinline__ bool cross_exchange(const KParams& K, const Env& E, LevelState& G, const Slot mine_for_lane, Slot& got, u64 deadline,
                                            bool fenced) {
    const u32 lane = threadIdx.x & 31;
    got = mine_for_lane;
    if (!MULTI) return true;
    const u64 epoch = G.xepoch + 1;
    const u32 par = (u32)(epoch & 1);
    const u64 tag = (u64)(u32)epoch << 32;
    bool ok = true;
    if (lane < K.P) {
        if (fenced) ptx::fence_sys();
        LSlot* dst = &ctrl_of(K, lane)->slot[par][E.me];
        ptx::st_relaxed_sys(&dst->w[0], tag | mine_for_lane.count);
        ptx::st_relaxed_sys(&dst->w[1], tag | mine_for_lane.meet);
        ptx::st_relaxed_sys(&dst->w[2], tag | mine_for_lane.rq_tail);
        ptx::st_relaxed_sys(&dst->w[3], tag | mine_for_lane.edges);
    } else {
        got.count = 0; got.meet = NO_V; got.rq_tail = 0; got.edges = 0;
    }
    ok = wait_slot(E.C, lane < K.P ? &E.C->slot[par][lane] : nullptr, (u32)epoch, got, deadline);
    if (fenced && lane < K.P) ptx::fence_sys();
    if (lane == 0) G.xepoch = epoch;
    __syncwarp();
    return ok;
}

enum SyncKind : int { SYNC_LEVEL = 0, SYNC_LOCAL = 1, SYNC_CROSS = 2 };

// Everything the leader warp (warp 0 of CTA 0 of the rank) does at a synchronisation point, after
// all participating local CTAs have arrived.  G / A are the leader's shared-memory level state and
// accumulators (pulled from global memory first when the level ran on the whole grid).
// Returns false if the query was aborted.  On return G.phase says what happens next.
template <bool MULTI>
__device__ __noinline__ bool leader_sync(const KParams& K, const Env& E, LevelState& G, Acc& A, int kind, u32 level_mode, u64 deadline,
                            bool maybe_hubs) {
    Ctrl* C = E.C;
    const u32 lane = threadIdx.x & 31;
    if (kind == SYNC_LOCAL) return true;
    const u32 side = G.side;
    const u64 t_sync0 = (K.flags & BIBFS_F_TRACE) ? ptx::globaltimer_ns() : 0ull;
    DBG_STAMP2(0, side);
    // hub rows found during the level?  Then the level is not over: the grid expands them first.
    if (kind == SYNC_LEVEL && maybe_hubs && (level_mode == PH_TD_SOLO || level_mode == PH_TD_GRID)) {
        const u64 hp = ptx::ld_relaxed_gpu(reinterpret_cast<const u64*>(&C->hub_packed));
        if (hp) {
            if (lane == 0) {
                G.hub_entries = (u32)(hp >> 40);
                G.hub_chunks = (u32)(hp & 0xFFFFFFFFFFull);
                G.prev_phase = level_mode;
                G.phase = PH_TD_HUB;
                C->hub_packed = 0;
            }
            __syncwarp();
            return true;
        }
    }
    if (kind == SYNC_LEVEL && level_mode == PH_BU && lane == 0) C->hub_packed = 0;      // (the bottom-up level's own hub list)
    Slot mine;
    mine.count = A.new_count;
    mine.meet = A.meet_local;
    mine.rq_tail = (MULTI && lane < K.P) ? ptx::ld_relaxed_gpu(&C->sent[side][lane]) : 0u;
    const u64 my_edges = A.edges;
    mine.edges = (u32)my_edges;
    Slot got;
    const bool fenced = (level_mode == PH_BU);     // bitmap slices were pushed with plain stores
    const bool ok = cross_exchange<MULTI>(K, E, G, mine, got, deadline, fenced);
    if (!MULTI && lane != 0) { got.count = 0; got.meet = NO_V; got.edges = 0; }   // single rank: lane 0 holds the totals
    if (!ok) {
        if (lane == 0) {
            atomicExch(&C->abort_flag, 1u);
            G.status = 1; G.length = -1; G.phase = PH_DONE;
            write_result(K, C, G);
        }
        __syncwarp();
        return false;
    }
    if (kind == SYNC_CROSS) return true;
    // ---- SYNC_LEVEL: reduce the slots, advance the level state, choose the next level
    u32 total = got.count, meetmin = got.meet;
    u64 etot = got.edges;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        total += __shfl_xor_sync(FULL, total, d);
        meetmin = min(meetmin, __shfl_xor_sync(FULL, meetmin, d));
        etot += __shfl_xor_sync(FULL, etot, d);
    }
    if (MULTI && lane < K.P) {
        G.rq_lo[side][lane] = G.rq_hi[side][lane];
        G.rq_hi[side][lane] = got.rq_tail;
    }
    if (MULTI && lane == 0) G.rq_tag[side] = (u32)G.xepoch;    // the tag this level's senders stamped
    __syncwarp();
    DBG_STAMP2(1, total + meetmin);
    if (lane == 0) {
        const u32 expanded = G.nf[side];
        const u32 mode = (level_mode == PH_TD_HUB) ? G.prev_phase : level_mode;
        G.nf[side] = total;
        G.depth[side] += 1;
        G.levels += 1;
        if (mode != PH_TD_SOLO || level_mode == PH_TD_HUB) G.levels_grid += 1;
        if (mode == PH_BU) { G.levels_bu += 1; G.fbm_cur[side] ^= 1u; G.fbm_valid[side] = 1; }
        else G.fbm_valid[side] = 0;
        G.q_lo[side] = G.q_hi[side];
        G.q_hi[side] = A.q_tail[side];
        G.edges_total += etot;    // all ranks' adjacency entries of this level
        G.visited_total += total;
        if ((K.flags & BIBFS_F_TRACE) && C->trace_count < TRACE_CAP) {
            bibfs_level_trace& T = trace_of(K, E.me)[C->trace_count++];
            T.side = side; T.mode = mode | (level_mode == PH_TD_HUB ? 8u : 0u); T.frontier = expanded; T.discovered = total;
            T.t_ns = ptx::globaltimer_ns(); T.edges = my_edges; T.t_expand_ns = t_sync0;
        }
        // reset the per-level accumulators
        A.new_count = 0;
        A.meet_local = NO_V;
        A.edges = 0;
        if (meetmin != NO_V) {
            G.length = (i32)(G.depth[0] + G.depth[1]);
            G.meet = meetmin;
            G.phase = PH_DONE;
        } else {
            decide_next(K, G);
        }
    }
    __syncwarp();
    if (MULTI && G.phase == PH_DONE) {
        // every rank reaches the same verdict: one fenced exchange makes all remote parent stores
        // of the query visible before any rank's path kernel walks them
        Slot z; z.count = 0; z.meet = NO_V; z.rq_tail = 0; z.edges = 0;
        Slot g2;
        cross_exchange<MULTI>(K, E, G, z, g2, deadline, true);
    }
    if (lane == 0 && G.phase == PH_DONE) write_result(K, C, G);
    DBG_STAMP2(2, G.phase);
    __syncwarp();
    return true;
}

// Query start, leader warp of each rank.  Works on the shared-memory state G / accumulators A.
template <bool MULTI>
__device__ __noinline__ bool leader_init(const KParams& K, const Env& E, LevelState& G, Acc& A, u64 deadline) {
    Ctrl* C = E.C;
    const u32 lane = threadIdx.x & 31;
    // the query: by value, or straight from the caller's pinned host memory (one PCIe read, no copy engine involved)
    u32 src = K.src, dst = K.dst;
    if (K.h_query) { src = *reinterpret_cast<const volatile u32*>(K.h_query); dst = *reinterpret_cast<const volatile u32*>(K.h_query + 1); }
    if (lane == 0) { C->q_src = src; C->q_dst = dst; }
    if (MULTI && K.async_peers && E.me != 0 && lane == 0) ptx::st_relaxed_sys(&ctrl_of(K, 0)->arrive[E.me], K.qn);
    // (the resets are dealt over the lanes: thirty-odd stores issued by one lane were a visible part of the start-up chain)
    if (lane < BAR_WAYS) C->bar_way[lane].count = 0;
    if (lane >= 16 && lane < 16 + MAXR) { C->sent[0][lane - 16] = 0; C->sent[1][lane - 16] = 0; }
    if (lane < MAXR) { G.rq_lo[0][lane] = 0; G.rq_hi[0][lane] = 0; G.rq_lo[1][lane] = 0; G.rq_hi[1][lane] = 0; }
    __syncwarp();
    if (lane == 0) {
        A.q_tail[0] = 0; A.q_tail[1] = 0;
        A.new_count = 0; A.meet_local = NO_V; A.edges = 0;
        C->hub_packed = 0;
        C->trace_count = 0;
        G.edges_total = 0; G.visited_total = 2;
        C->t_begin_ns = ptx::globaltimer_ns();
        G.phase = PH_TD_SOLO; G.side = 0; G.levels = 0; G.levels_bu = 0;
        G.depth[0] = G.depth[1] = 0;
        G.nf[0] = G.nf[1] = 1;
        for (u32 s = 0; s < 2; ++s) {
            G.q_lo[s] = G.q_hi[s] = 0;
            G.fbm_cur[s] = 0; G.fbm_valid[s] = 0;
        }
        G.length = -1; G.meet = NO_V; G.status = 0; G.wake = 0; G.local_front = 0; G.pad = 0; G.rounds_paired = 0; G.levels_grid = 0;
        G.hub_entries = 0; G.hub_chunks = 0; G.prev_phase = PH_TD_SOLO;
        G.rq_tag[0] = G.rq_tag[1] = 0;
        G.xepoch = C->xepoch;
        const bool bad = src >= K.n || dst >= K.n;
        if (bad) { G.status = 2; G.phase = PH_DONE; }
        else if (src == dst) { G.length = 0; G.meet = src; G.phase = PH_DONE; G.visited_total = 1; }
        else {
            const u32 hi = min(K.n, E.lo + K.B);
            if (src >= E.lo && src < hi) {
                atomicOr(reinterpret_cast<unsigned long long*>(&E.vis[src >> 5]), 1ull << (src & 31));
                E.q[0][0] = src; E.parent[0][src - E.lo] = NO_V;
                A.q_tail[0] = 1; G.q_hi[0] = 1;
            }
            if (dst >= E.lo && dst < hi) {
                atomicOr(reinterpret_cast<unsigned long long*>(&E.vis[dst >> 5]), 1ull << ((dst & 31) + 32));
                E.q[1][0] = dst; E.parent[1][dst - E.lo] = NO_V;
                A.q_tail[1] = 1; G.q_hi[1] = 1;
            }
        }
    }
    __syncwarp();
    bool ok = true;
    if (MULTI) {
        // start barrier: every rank has finished the previous query's cleanup and seeded this one
        // before anybody's remote traffic can arrive
        Slot z; z.count = 0; z.meet = NO_V; z.rq_tail = 0; z.edges = 0;
        Slot got;
        ok = cross_exchange<MULTI>(K, E, G, z, got, deadline, false);
    }
    if (lane == 0) {
        if (!ok) { atomicExch(&C->abort_flag, 1u); G.status = 1; G.phase = PH_DONE; }
        if (G.phase != PH_DONE) decide_next(K, G);
        if (G.phase == PH_DONE) write_result(K, C, G);
    }
    __syncwarp();
    return ok;
}

// Query continuation (KParams::resume): the front kernel (front.cuh) started the query and handed it over.  Rank 0 finds
// its level state in Ctrl::st, the other ranks theirs in the verdict rank 0 sent them.  Leader warp of each rank.
__device__ bool wait_verdict(Ctrl* C, u64 qn, u32* out, u64 deadline);
__device__ void state_from_verdict(const KParams& K, LevelState& G, const u32* v, u32 tag);

template <bool MULTI>
__device__ __noinline__ bool leader_resume(const KParams& K, const Env& E, LevelState& G, Acc& A, u64 deadline) {
    Ctrl* C = E.C;
    const u32 lane = threadIdx.x & 31;
    if (lane < BAR_WAYS) C->bar_way[lane].count = 0;
    bool ok = true;
    if (E.me == 0) {
        const u32* g = reinterpret_cast<const u32*>(&C->st);
        u32* s = reinterpret_cast<u32*>(&G);
        for (u32 i = lane; i < sizeof(LevelState) / 4; i += 32) s[i] = __ldcg(g + i);
        if (lane == 0) { pull_acc(A, &C->acc); C->hub_packed = 0; }
    } else if (MULTI) {
        u32 v[VD_WORDS];
        ok = wait_verdict(C, K.qn, v, deadline);
        if (lane == 0) ptx::fence_sys();        // the hand-over's bits and parents were fenced before the verdict was sent
        if (lane < MAXR) { C->sent[0][lane] = 0; C->sent[1][lane] = 0; }
        if (lane == 0) {
            const u64 xe = C->xepoch + 1;
            state_from_verdict(K, G, v, (u32)xe);
            G.xepoch = xe;
            A.q_tail[0] = 0; A.q_tail[1] = 0; A.new_count = 0; A.meet_local = NO_V; A.edges = 0;
            C->hub_packed = 0; C->trace_count = 0;
            if (!K.async_peers) C->t_begin_ns = ptx::globaltimer_ns();
            C->q_src = NO_V; C->q_dst = NO_V;
        }
    }
    __syncwarp();
    if (lane == 0) {
        if (!ok) { atomicExch(&C->abort_flag, 1u); G.status = 1; G.length = -1; G.phase = PH_DONE; }
        else decide_next(K, G);
        if (G.phase == PH_DONE) write_result(K, C, G);
    }
    __syncwarp();
    return ok;
}

// ------------------------------------------------------------------------------------------
// frontier access: flat index -> vertex, across the local window and the per-sender windows
// ------------------------------------------------------------------------------------------
template <bool MULTI>
__device__ __forceinline__ u32 fetch_frontier(const KParams& K, const Env& E, const LevelState& st, u32 side, u32 i,
                                              const u32* smem_front = nullptr) {
    const u32 nq = st.q_hi[side] - st.q_lo[side];
    if (!MULTI || i < nq) return smem_front ? smem_front[i] : __ldcg(E.q[side] + st.q_lo[side] + i);
    i -= nq;
    u32 v = 0;
#pragma unroll 1
    for (u32 p = 0; p < K.P; ++p) {
        const u32 lo = st.rq_lo[side][p], c = st.rq_hi[side][p] - lo;
        if (i < c) {
            // the sender's count may overtake its entries on the wire: wait for the tagged word
            const u64* e = rq_of(K, E.me, side, p) + lo + i;
            const u32 tag = st.rq_tag[side];
            u64 w = ptx::ld_relaxed_sys(e);
            for (u32 spins = 0; (u32)(w >> 32) != tag && spins < (1u << 24); ++spins) w = ptx::ld_relaxed_sys(e);
            if ((u32)(w >> 32) != tag) atomicExch(&E.C->abort_flag, 1u);
            v = (u32)w;
            break;
        }
        i -= c;
    }
    return v;
}

// address of entry i of the concatenated sender-segment windows (i counts past the local window)
__device__ __forceinline__ const u64* rq_entry(const KParams& K, const Env& E, const LevelState& st, u32 side, u32 i) {
#pragma unroll 1
    for (u32 p = 0; p < K.P; ++p) {
        const u32 lo = st.rq_lo[side][p], c = st.rq_hi[side][p] - lo;
        if (i < c) return rq_of(K, E.me, side, p) + lo + i;
        i -= c;
    }
    return nullptr;
}
// whole warp: lanes with e != nullptr read their tagged entry, polling until it carries `tag` (the sender's relaxed
// stores may land after the exchange words that announced them); warp-collective like every polling loop
__device__ __forceinline__ u32 fetch_rq(const Env& E, const u64* e, u32 tag) {
    u64 w = 0;
    bool done = e == nullptr;
    u32 spins = 0;
    while (true) {
        if (!done) { w = ptx::ld_relaxed_sys(e); done = (u32)(w >> 32) == tag; }
        if (__all_sync(FULL, done)) break;
        if (++spins > (1u << 22)) { if (!done) atomicExch(&E.C->abort_flag, 1u); break; }
    }
    return (u32)w;
}

// ------------------------------------------------------------------------------------------
// discovery, split in two so that several edges per lane can be in flight:
//   claim    test-and-set the visited bit of v for `side` (the atomic also returns the other
//            side's bit: the meet test costs nothing extra, even across NVLink)
//   enqueue  record the parent and append v to its owner's queue, warp-aggregated
// ------------------------------------------------------------------------------------------
struct Claim {
    u32 dest;
    bool isnew, meet;
};

template <bool MULTI>
__device__ __forceinline__ Claim claim(const KParams& K, const Env& E, u32 side, bool valid, u32 v, u64 seen_word) {
    Claim c;
    c.dest = E.me; c.isnew = false; c.meet = false;
    if (valid) {
        const u32 w = v >> 5;
        const u64 bit = 1ull << ((v & 31u) + 32u * side);
        const u64 obit = 1ull << ((v & 31u) + 32u * (1u - side));
        if (!(seen_word & bit)) {
            u64 old = ptx::atom_or_gpu(&E.vis[w], bit);
            if (!(old & bit)) {
                c.isnew = true;
                if (MULTI) {
                    c.dest = owner_of(K, v);
                    if (c.dest != E.me) {
                        // first time THIS rank sees v: ask the owner.  One round trip returns both
                        // "new for this side" and "already reached by the other side".
                        old = ptx::atom_or_sys(vis_of(K, c.dest) + w, bit);
                        c.isnew = !(old & bit);
                    }
                }
                c.meet = c.isnew && (old & obit);
            }
        }
    }
    return c;
}

template <bool MULTI>
__device__ __forceinline__ void enqueue(const KParams& K, const Env& E, u32 side, const Claim c, u32 u, u32 v, u32& cnt) {
    const u32 lane = threadIdx.x & 31;
    if (!MULTI) {
        const u32 m = __ballot_sync(FULL, c.isnew);
        if (c.isnew) {
            q_append(E, side, m, v);
            E.parent[side][v - E.lo] = u;
            ptx::prefetch_l2(E.row_ptr + (v - E.lo));
        }
    } else {
        const u32 key = c.isnew ? c.dest : 0xFFu;
        const u32 grp = __match_any_sync(FULL, key);
        if (c.isnew) {
            const u32 leader = __ffs(grp) - 1;
            const u32 rank_in = __popc(grp & lanemask_lt());
            u32 base = 0;
            if (c.dest == E.me) {
                q_append(E, side, grp, v);
                E.parent[side][v - E.lo] = u;
                ptx::prefetch_l2(E.row_ptr + (v - E.lo));
            } else {
                if (lane == leader) {
                    if (E.solo) { base = ptx::atoms_add(&E.s_sent[side * MAXR + c.dest], (u32)__popc(grp)); ptx::reds_add(E.s_rnew, (u32)__popc(grp)); }
                    else base = atomicAdd(&E.C->sent[side][c.dest], (u32)__popc(grp));
                }
                base = __shfl_sync(grp, base, leader);
                ptx::st_relaxed_sys(rq_of(K, c.dest, side, E.me) + base + rank_in, (u64)v | ((u64)E.tag << 32));
                ptx::st_peer(parent_of(K, c.dest, side) + (v - c.dest * K.B), u);
            }
        }
    }
    cnt += c.isnew ? 1u : 0u;
    if (c.meet) acc_note_meet(E, v);
}

// UNR edges per lane: all column loads first, then all filter loads, then all claims, then the
// enqueues — three dependent memory round trips per UNR*32 edges instead of per 32.  Deliberately
// NOT inlined: the level loop of a small search is executed once per level by a single warp, so
// its cost is instruction fetch; one shared copy of this body keeps the hot path inside the
// instruction cache (the fully inlined kernel was 124 KB of SASS and ran 2x slower).
static_assert(UNR == 4, "visit_batch takes four edges per lane");
template <bool MULTI>
__device__ __noinline__ u32 visit_batch(const KParams& K, const Env& E, u32 side, u32 okmask, u32 u0, u32 u1, u32 u2, u32 u3,
                                        const u32* p0, const u32* p1, const u32* p2, const u32* p3) {
    const bool ok[UNR] = {(okmask & 1u) != 0, (okmask & 2u) != 0, (okmask & 4u) != 0, (okmask & 8u) != 0};
    const u32 uu[UNR] = {u0, u1, u2, u3};
    const u32* const vp[UNR] = {p0, p1, p2, p3};
    u32 cnt = 0;
    u32 v[UNR];
    u64 seen[UNR];
    Claim c[UNR];
#pragma unroll
    for (int k = 0; k < UNR; ++k) v[k] = ok[k] ? (E.smem_src ? *vp[k] : __ldg(vp[k])) : 0u;
    DBG_STAMP(4, v[0] + v[1] + v[2] + v[3]);
#pragma unroll
    for (int k = 0; k < UNR; ++k) seen[k] = ok[k] ? (E.filter ? E.vis[v[k] >> 5] : 0ull) : ~0ull;   // filter only; may be stale
    {
        // Branch-free issue (measured: the branchy form serialises the round trips).  Slots with nothing to
        // claim OR a zero into a private scratch word, so all UNR local atomics, and then all UNR owner-side
        // atomics of the multi-rank case, are in flight together.
        u64* const dummy = dummy_of(K, E.me);
        const u32 sh = 32u * side, osh = 32u - sh;
        u64 old[UNR];
        bool act[UNR];
#pragma unroll
        for (int k = 0; k < UNR; ++k) {
            act[k] = ok[k] && !((seen[k] >> ((v[k] & 31u) + sh)) & 1ull);
            old[k] = ptx::atom_or_gpu(act[k] ? &E.vis[v[k] >> 5] : dummy, act[k] ? (1ull << ((v[k] & 31u) + sh)) : 0ull);
        }
        const u32 shz = sh + ptx::zero_after(old[UNR - 1]);     // == sh; keeps the tests below behind the last atomic
#pragma unroll
        for (int k = 0; k < UNR; ++k) {
            c[k].dest = E.me;
            c[k].isnew = act[k] && !((old[k] >> ((v[k] & 31u) + shz)) & 1ull);
        }
        if (MULTI) {
            bool rem[UNR];
            u64 rold[UNR];
#pragma unroll
            for (int k = 0; k < UNR; ++k) {
                const u32 dest = c[k].isnew ? owner_of(K, v[k]) : E.me;
                rem[k] = dest != E.me;
                c[k].dest = dest;
                rold[k] = ptx::atom_or_sys(rem[k] ? (vis_of(K, dest) + (v[k] >> 5)) : dummy, rem[k] ? (1ull << ((v[k] & 31u) + sh)) : 0ull);
            }
            const u32 shr = sh + ptx::zero_after(rold[UNR - 1]);
#pragma unroll
            for (int k = 0; k < UNR; ++k)
                if (rem[k]) { old[k] = rold[k]; c[k].isnew = !((rold[k] >> ((v[k] & 31u) + shr)) & 1ull); }
        }
#pragma unroll
        for (int k = 0; k < UNR; ++k) c[k].meet = c[k].isnew && ((old[k] >> ((v[k] & 31u) + osh)) & 1ull);
    }
    DBG_STAMP(5, (u32)c[0].isnew + (u32)c[1].isnew + (u32)c[2].isnew + (u32)c[3].isnew);
    if (MULTI) {
#pragma unroll
        for (int k = 0; k < UNR; ++k) enqueue<MULTI>(K, E, side, c[k], uu[k], v[k], cnt);
    } else {
#pragma unroll
        for (int k = 0; k < UNR; ++k) {
            const u32 m = __ballot_sync(FULL, c[k].isnew);
            if (c[k].isnew) {
                q_append(E, side, m, v[k]);
                E.parent[side][v[k] - E.lo] = uu[k];
                ptx::prefetch_l2(E.row_ptr + (v[k] - E.lo));
                ++cnt;
                if (c[k].meet) acc_note_meet(E, v[k]);
            }
        }
    }
    return cnt;
}

// a contiguous range of one row, walked by the whole warp
template <bool MULTI>
__device__ __forceinline__ void expand_range(const KParams& K, const Env& E, u32 side, u32 uu, u64 b, u64 e, u32& cnt) {
    const u32 lane = threadIdx.x & 31;
    for (u64 p = b; p < e; p += 32 * UNR) {
        const u64 i0 = p + lane;
        u32 okmask = 0;
#pragma unroll
        for (int k = 0; k < UNR; ++k) okmask |= (i0 + (u64)k * 32 < e) ? (1u << k) : 0u;
        const u32* c0 = E.col + i0;
        cnt += visit_batch<MULTI>(K, E, side, okmask, uu, uu, uu, uu, c0, c0 + 32, c0 + 64, c0 + 96);
    }
}

// ------------------------------------------------------------------------------------------
// top-down level.  Warps take `vpw` frontier vertices at a time (32 when the frontier is large,
// fewer when it is small so that the work spreads over the machine).
// ------------------------------------------------------------------------------------------
template <bool MULTI>
__device__ __noinline__ void expand_td(const KParams& K, const Env& E, const LevelState& st, u32 warp_gid, u32 nwarps,
                          const u32* smem_front = nullptr) {
    const u32 lane = threadIdx.x & 31;
    const u32 side = st.side;
    const u32 total = st.local_front;
    u32 vpw = 32;
    while (vpw > 1 && (u64)(vpw >> 1) * nwarps >= total) vpw >>= 1;
    u32 cnt = 0;
    u64 edges = 0;
    const u32 nbatches = (total + vpw - 1) / vpw;
    for (u32 b = warp_gid; b < nbatches; b += nwarps) {
        if (acc_meet_seen(E)) break;               // the search is over: any hit of this level is a shortest path
        const u32 i = b * vpw + lane;
        u32 u = 0;
        u64 r0 = 0, r1 = 0;
        DBG_STAMP(0, i);
        if (lane < vpw && i < total) {
            u = fetch_frontier<MULTI>(K, E, st, side, i, smem_front);
            DBG_STAMP(1, u);
            const u32 ul = u - E.lo;
            r0 = __ldg(E.row_ptr + ul);
            r1 = __ldg(E.row_ptr + ul + 1);
        }
        u64 deg = r1 - r0;
        DBG_STAMP(2, deg);
        edges += deg;
        // hub rows: hand them to the whole grid (hub phase) unless the list is full
        if (deg >= HUB_ROW) {
            const u32 nch = hub_nchunks(r0, r1);
            const unsigned long long inc = (1ull << 40) | (unsigned long long)nch;
            const unsigned long long old = atomicAdd(&E.C->hub_packed, inc);
            const u32 slot = (u32)(old >> 40);
            if (slot < HUB_CAP) {
                HubEntry h;
                h.u = u; h.nchunks = nch; h.r0 = r0; h.r1 = r1; h.chunk_base = old & 0xFFFFFFFFFFull;
                hub_of(K, E.me)[slot] = h;
                if (E.s_hub) *E.s_hub = 1u;
                deg = 0; r1 = r0;
            } else {
                atomicAdd(&E.C->hub_packed, 0ull - inc);
            }
        }
        // long rows: the whole warp walks one row, coalesced
        u32 big = __ballot_sync(FULL, deg >= BIG_ROW);
        while (big) {
            const u32 srcl = __ffs(big) - 1;
            big &= big - 1;
            const u32 uu = __shfl_sync(FULL, u, srcl);
            const u64 rb = __shfl_sync(FULL, r0, srcl), re = __shfl_sync(FULL, r1, srcl);
            expand_range<MULTI>(K, E, side, uu, rb, re, cnt);
        }
        // tiny rows (<= UNR entries): every lane walks its own row, no cross-lane bookkeeping
        if (__any_sync(FULL, deg != 0 && deg <= (u64)UNR)) {
            const u32 okmask = deg <= (u64)UNR ? ((1u << (u32)deg) - 1u) : 0u;
            const u32* c0 = E.col + r0;
            DBG_STAMP(3, c0);
            cnt += visit_batch<MULTI>(K, E, side, okmask, u, u, u, u, c0, c0 + 1, c0 + 2, c0 + 3);
            DBG_STAMP(6, cnt);
        }
        // short rows: flatten the rows of the batch into one edge index space (scan + search by shuffles)
        const u32 sdeg = (deg >= BIG_ROW || deg <= (u64)UNR) ? 0u : (u32)deg;
        u32 incl = sdeg;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const u32 t = __shfl_up_sync(FULL, incl, d);
            if ((int)lane >= d) incl += t;
        }
        const u32 excl = incl - sdeg;
        const u32 tot = __shfl_sync(FULL, incl, 31);
        for (u32 t = 0; t < tot; t += 32 * UNR) {
            u32 okmask = 0;
            u32 us[UNR];
            const u32* vp[UNR];
#pragma unroll
            for (int k = 0; k < UNR; ++k) {
                const u32 idx = t + k * 32 + lane;
                okmask |= (idx < tot) ? (1u << k) : 0u;
                u32 j = 0;
#pragma unroll
                for (int step = 16; step > 0; step >>= 1) {
                    const u32 cand = j + step;
                    const u32 ec = __shfl_sync(FULL, excl, cand & 31u);
                    if (cand < 32u && ec <= idx) j = cand;
                }
                us[k] = __shfl_sync(FULL, u, j);
                const u64 rb = __shfl_sync(FULL, r0, j);
                const u32 eb = __shfl_sync(FULL, excl, j);
                vp[k] = E.col + rb + (idx - eb);
            }
            cnt += visit_batch<MULTI>(K, E, side, okmask, us[0], us[1], us[2], us[3], vp[0], vp[1], vp[2], vp[3]);
        }
    }
    // flush this warp's counters
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        cnt += __shfl_xor_sync(FULL, cnt, d);
        edges += __shfl_xor_sync(FULL, edges, d);
    }
    if (lane == 0) acc_flush(E, cnt, edges);
    DBG_STAMP(7, cnt);
#ifdef BIBFS_FINE_TRACE
    if (blockIdx.x == 0 && threadIdx.x == 0) g_dbg_level++;
#endif
}

// hub phase: every warp of the rank takes 1 KB chunks of the deferred rows.  The chunk is fetched
// by the TMA engine (cp.async.bulk global -> shared, completion on an mbarrier; SASS UBLKCP) into a
// per-warp double buffer, so the next chunk streams in while the current one is being claimed.
struct HubChunk {
    u32 uu;            // frontier vertex the row belongs to
    u32 lo, hi;        // valid entries of the staged chunk: stage[lo..hi)
    u32 bytes;         // size of the bulk copy
    const u32* src;    // 16-byte aligned start of the chunk in col
    bool valid;
    u64 A;             // index of src in col
    u32 nch_word;      // the entry's nchunks word (wide rounds of the front kernel keep the side in bit 31)
};

// whole warp, same c in every lane.  The entries are ordered by chunk_base (slot and base come from one atomic): a
// 32-ary search, one load per step (18 entries: ONE step where a binary search takes five dependent L2 round trips),
// then the 32-byte entry is fetched by eight lanes at once.
__device__ __forceinline__ HubChunk hub_lookup(const HubEntry* H, u32 ne, const u32* col, u32 c) {
    const u32 lane = threadIdx.x & 31;
    HubChunk h;
    u32 lo = 0, n = ne;
    while (n > 1) {
        const u32 step = (n + 31u) / 32u;
        const u32 idx = lo + lane * step;
        const u64 b = (lane * step < n) ? __ldcg(&H[idx].chunk_base) : ~0ull;
        const u32 m = __ballot_sync(FULL, b <= (u64)c) | 1u;
        const u32 k = 31u - (u32)__clz(m);
        const u32 rest = n - k * step;
        lo += k * step;
        n = rest < step ? rest : step;
    }
    static_assert(sizeof(HubEntry) == 32, "eight words");
    const u32 w = (lane < 8 && ne) ? __ldcg(reinterpret_cast<const u32*>(&H[lo]) + lane) : 0u;
    h.uu = __shfl_sync(FULL, w, 0);
    const u32 nch = __shfl_sync(FULL, w, 1);
    const u64 r0 = (u64)__shfl_sync(FULL, w, 2) | ((u64)__shfl_sync(FULL, w, 3) << 32);
    const u64 r1 = (u64)__shfl_sync(FULL, w, 4) | ((u64)__shfl_sync(FULL, w, 5) << 32);
    const u64 base = (u64)__shfl_sync(FULL, w, 6) | ((u64)__shfl_sync(FULL, w, 7) << 32);
    h.valid = ne != 0 && (c - base) < nch;          // false: chunk of an entry that did not fit in the list
    const u64 A = (r0 & ~3ull) + (u64)(c - base) * HUB_CHUNK;     // 16-byte aligned
    const u64 Aend = min(A + HUB_CHUNK, r1);
    h.lo = (u32)(max(A, r0) - A);
    h.hi = (u32)(Aend - A);
    h.bytes = ((h.hi * 4u) + 15u) & ~15u;
    h.src = col + A;
    h.A = A;
    h.nch_word = nch;
    return h;
}

template <bool MULTI>
__device__ __noinline__ void expand_hubs(const KParams& K, const Env& E0, const LevelState& st, u32 warp_gid, u32 nwarps,
                                         u32* stage /* this warp's 2 x HUB_CHUNK words */, u64* bars /* this warp's 2 mbarriers */, u32& parity) {
    const u32 lane = threadIdx.x & 31;
    const u32 side = st.side;
    const HubEntry* H = hub_of(K, E0.me);
    const u32 ne = min(st.hub_entries, HUB_CAP);
    Env E = E0;
    E.smem_src = true;
    u32 cnt = 0;
    u32 c = warp_gid;
    HubChunk cur;
    cur.valid = false;
    if (c < st.hub_chunks) cur = hub_lookup(H, ne, E.col, c);
    if (lane == 0 && cur.valid) { ptx::mbar_expect_tx(&bars[0], cur.bytes); ptx::bulk_g2s(stage, cur.src, cur.bytes, &bars[0]); }
    for (u32 i = 0; c < st.hub_chunks; c += nwarps, ++i) {
        const u32 buf = i & 1u;
        if (acc_meet_seen(E)) {
            // stop early, but drain the copy already in flight so that the barrier phase stays in step
            if (cur.valid) { ptx::mbar_wait(&bars[buf], (parity >> buf) & 1u); parity ^= 1u << buf; }
            break;
        }
        // start the copy of this warp's next chunk into the other buffer
        const u32 cn = c + nwarps;
        HubChunk nxt;
        nxt.valid = false;
        if (cn < st.hub_chunks) nxt = hub_lookup(H, ne, E.col, cn);
        if (lane == 0 && nxt.valid) {
            ptx::fence_proxy_async();       // this buffer was read through the generic proxy two iterations ago
            ptx::mbar_expect_tx(&bars[buf ^ 1u], nxt.bytes);
            ptx::bulk_g2s(stage + (buf ^ 1u) * HUB_CHUNK, nxt.src, nxt.bytes, &bars[buf ^ 1u]);
        }
        if (cur.valid) {
            ptx::mbar_wait(&bars[buf], (parity >> buf) & 1u);
            parity ^= 1u << buf;
            const u32* sbuf = stage + buf * HUB_CHUNK;
            for (u32 p = cur.lo; p < cur.hi; p += 32 * UNR) {
                const u32 i0 = p + lane;
                u32 okmask = 0;
#pragma unroll
                for (int k = 0; k < UNR; ++k) okmask |= (i0 + k * 32 < cur.hi) ? (1u << k) : 0u;
                const u32* s0 = sbuf + i0;
                cnt += visit_batch<MULTI>(K, E, side, okmask, cur.uu, cur.uu, cur.uu, cur.uu, s0, s0 + 32, s0 + 64, s0 + 96);
            }
        }
        __syncwarp();
        cur = nxt;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) cnt += __shfl_xor_sync(FULL, cnt, d);
    if (lane == 0) acc_flush(E, cnt, 0);
}

// ------------------------------------------------------------------------------------------
// bottom-up level: every unvisited owned vertex looks for a parent in the frontier bitmap.
// Short rows are scanned by their lane, long rows by the whole warp (ballot = early exit).
// ------------------------------------------------------------------------------------------
constexpr int BU_W = 8;      // adjacency entries a lane has in flight in the bottom-up scans
constexpr u32 BU_HUB_ROW = 4096;      // bottom-up: rows at least this long are scanned by the whole grid (expand_bu_hubs)
template <bool MULTI>
__device__ __noinline__ void expand_bu(const KParams& K, const Env& E, const LevelState& st, u32 warp_gid, u32 nwarps) {
    const u32 lane = threadIdx.x & 31;
    const u32 side = st.side;
    const u32* fcur = fbm_of(K, E.me, side, st.fbm_cur[side]);
    u32* fnext = fbm_of(K, E.me, side, st.fbm_cur[side] ^ 1u);
    const u32 words = K.B / 32;
    const u32 w0 = E.lo / 32;
    u32 cnt = 0;
    u64 edges = 0;
    for (u32 wl = warp_gid; wl < words; wl += nwarps) {
        const u32 wg = w0 + wl;
        const u32 v = E.lo + wl * 32u + lane;
        const u64 word = E.vis[wg];
        const bool unvis = v < K.n && !((word >> (32u * side + lane)) & 1ull);
        bool found = false;
        u32 par = NO_V;
        u64 r0 = 0, r1 = 0;
        if (unvis) { r0 = __ldg(E.row_ptr + (v - E.lo)); r1 = __ldg(E.row_ptr + (v - E.lo) + 1); }
        const u64 deg = r1 - r0;
        if (unvis && deg < BIG_ROW) {
            // BU_W entries per step: the column loads and then the bitmap loads of a step are independent, so BU_W
            // L2 / DRAM round trips overlap where the one-entry loop had them back to back (the first hit still wins)
            for (u64 e = r0; e < r1 && !found; e += BU_W) {
                u32 u[BU_W], w[BU_W];
#pragma unroll
                for (int k = 0; k < BU_W; ++k) u[k] = __ldg(E.col + (e + k < r1 ? e + k : r0));
#pragma unroll
                for (int k = 0; k < BU_W; ++k) w[k] = fcur[u[k] >> 5];
#pragma unroll
                for (int k = 0; k < BU_W; ++k) {
                    if (!found && e + k < r1) {
                        ++edges;
                        if ((w[k] >> (u[k] & 31u)) & 1u) { found = true; par = u[k]; }
                    }
                }
            }
        }
        // rows of >= BU_HUB_ROW entries go to the hub list: the whole grid scans them chunk by chunk afterwards
        // (expand_bu_hubs).  One warp walking a 300,000-entry row alone was 1.6 ms of a 2.5 ms sweep (R-MAT scale 22).
        bool hub = unvis && deg >= BU_HUB_ROW;
        if (hub) {
            const u32 nch = hub_nchunks(r0, r1);
            const unsigned long long inc = (1ull << 40) | (unsigned long long)nch;
            const unsigned long long old = atomicAdd(&E.C->hub_packed, inc);
            const u32 slot = (u32)(old >> 40);
            if (slot < HUB_CAP) {
                HubEntry h;
                h.u = v; h.nchunks = nch; h.r0 = r0; h.r1 = r1; h.chunk_base = old & 0xFFFFFFFFFFull;
                hub_of(K, E.me)[slot] = h;
            } else {
                atomicAdd(&E.C->hub_packed, 0ull - inc);      // list full: the warp walks the row
                hub = false;
            }
        }
        u32 big = __ballot_sync(FULL, unvis && deg >= BIG_ROW && !hub);
        while (big) {
            const u32 sl = __ffs(big) - 1;
            big &= big - 1;
            const u64 b = __shfl_sync(FULL, r0, sl), e = __shfl_sync(FULL, r1, sl);
            u32 hitpar = NO_V;
            // 32 x BU_W entries per step (BU_W per lane, loads of a step independent of one another)
            for (u64 p = b; p < e && hitpar == NO_V; p += 32u * BU_W) {
                u32 u[BU_W], w[BU_W];
#pragma unroll
                for (int k = 0; k < BU_W; ++k) { const u64 idx = p + 32u * k + lane; u[k] = __ldg(E.col + (idx < e ? idx : b)); }
#pragma unroll
                for (int k = 0; k < BU_W; ++k) w[k] = fcur[u[k] >> 5];
#pragma unroll
                for (int k = 0; k < BU_W; ++k) {
                    const bool in = p + 32u * k + lane < e;
                    if (in && hitpar == NO_V) ++edges;
                    const u32 hm = __ballot_sync(FULL, in && ((w[k] >> (u[k] & 31u)) & 1u));
                    if (hm && hitpar == NO_V) hitpar = __shfl_sync(FULL, u[k], __ffs(hm) - 1);
                }
            }
            if (lane == sl && hitpar != NO_V) { found = true; par = hitpar; }
        }
        const u32 newmask = __ballot_sync(FULL, found);
        if (newmask) {
            if (lane == 0) {
                E.vis[wg] = word | ((u64)newmask << (32u * side));
                const u32 meetmask = newmask & (u32)(word >> (32u * (1u - side)));
                if (meetmask) atomicMin(&E.A->meet_local, E.lo + wl * 32u + (u32)(__ffs(meetmask) - 1));
            }
            if (found) {
                q_append(E, side, newmask, v);
                E.parent[side][v - E.lo] = par;
                ++cnt;
            }
        }
        if (lane == 0) fnext[wg] = newmask;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        cnt += __shfl_xor_sync(FULL, cnt, d);
        edges += __shfl_xor_sync(FULL, edges, d);
    }
    if (lane == 0) {
        if (cnt) atomicAdd(&E.A->new_count, cnt);
        if (edges) atomicAdd(&E.A->edges, (unsigned long long)edges);
    }
}

// second phase of a bottom-up level: the rows expand_bu deferred, one 256-entry chunk per warp and step, all warps of
// the rank.  The first chunk that finds a parent claims the vertex (atomic on its visited word); chunks that come later
// see the mark and skip their scan.
template <bool MULTI>
__device__ __noinline__ void expand_bu_hubs(const KParams& K, const Env& E, const LevelState& st, u32 warp_gid, u32 nwarps) {
    const u32 lane = threadIdx.x & 31;
    const u32 side = st.side;
    const u64 hp = ptx::ld_relaxed_gpu(reinterpret_cast<const u64*>(&E.C->hub_packed));
    const u32 ne = min((u32)(hp >> 40), HUB_CAP), nch = (u32)(hp & 0xFFFFFFFFFFull);
    if (!nch) return;
    const HubEntry* H = hub_of(K, E.me);
    const u32* fcur = fbm_of(K, E.me, side, st.fbm_cur[side]);
    u32* fnext = fbm_of(K, E.me, side, st.fbm_cur[side] ^ 1u);
    u32 cnt = 0;
    u64 edges = 0;
    static_assert(HUB_CHUNK == 32 * BU_W, "one step of the warp per chunk");
    for (u32 c = warp_gid; c < nch; c += nwarps) {
        const HubChunk h = hub_lookup(H, ne, E.col, c);
        if (!h.valid) continue;
        const u32 v = h.uu;
        const u64 bit = 1ull << (32u * side + (v & 31u));
        if (ptx::ld_relaxed_gpu(&E.vis[v >> 5]) & bit) continue;
        const u32* cc = E.col + h.A;
        u32 u[BU_W], w[BU_W];
#pragma unroll
        for (int k = 0; k < BU_W; ++k) { const u32 idx = h.lo + 32u * k + lane; u[k] = __ldg(cc + (idx < h.hi ? idx : h.lo)); }
#pragma unroll
        for (int k = 0; k < BU_W; ++k) w[k] = fcur[u[k] >> 5];
        u32 hitpar = NO_V;
#pragma unroll
        for (int k = 0; k < BU_W; ++k) {
            const bool in = h.lo + 32u * k + lane < h.hi;
            if (in) ++edges;
            const u32 hm = __ballot_sync(FULL, in && ((w[k] >> (u[k] & 31u)) & 1u));
            if (hm && hitpar == NO_V) hitpar = __shfl_sync(FULL, u[k], __ffs(hm) - 1);
        }
        if (hitpar != NO_V && lane == 0) {
            const u64 old = atomicOr(reinterpret_cast<unsigned long long*>(&E.vis[v >> 5]), (unsigned long long)bit);
            if (!(old & bit)) {
                q_append(E, side, 1u, v);
                E.parent[side][v - E.lo] = hitpar;
                atomicOr(&fnext[v >> 5], 1u << (v & 31u));
                if ((old >> (32u * (1u - side) + (v & 31u))) & 1ull) atomicMin(&E.A->meet_local, v);
                ++cnt;
            }
        }
        __syncwarp();
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) edges += __shfl_xor_sync(FULL, edges, d);
    if (lane == 0) {
        if (cnt) atomicAdd(&E.A->new_count, cnt);
        if (edges) atomicAdd(&E.A->edges, (unsigned long long)edges);
    }
}

// copy this rank's slice of a frontier bitmap into every peer's replica (8-byte stores over NVLink)
__device__ __noinline__ void push_fbm_slice(const KParams& K, const Env& E, u32 side, u32 buf, u32 gtid, u32 gthreads) {
    const u32 words2 = K.B / 64;                 // 8-byte units in the slice
    const u32 w0 = E.lo / 32;
    const uint2* src = reinterpret_cast<const uint2*>(fbm_of(K, E.me, side, buf) + w0);
    for (u32 p = 0; p < K.P; ++p) {
        if (p == E.me) continue;
        u32* dst = fbm_of(K, p, side, buf) + w0;
        for (u32 i = gtid; i < words2; i += gthreads) {
            const uint2 x = src[i];
            ptx::st_peer_v2(dst + 2 * i, x.x, x.y);
        }
    }
}

// ------------------------------------------------------------------------------------------
// shared memory of the search kernel
// ------------------------------------------------------------------------------------------
struct SoloLevel {          // written by thread 0 at the end of a level, read by everybody
    u32 side_flags;         // bit 0 side, bit 1 frontier is in the shared mirror, bit 2 which mirror buffer, bit 3 stop
    u32 total;              // frontier vertices of this rank (local window + windows of the sender segments)
    u32 q_lo;               // local window start in the global queue (used when not mirrored)
    u32 base_tail;          // local queue cursor at the start of the level (== local window end)
    u32 nq;                 // size of the local window
    u32 tag;                // epoch tag for entries sent to other ranks in this level
    u32 rq_tag;             // tag carried by the entries of the sender-segment windows
    u32 slice;              // cluster tier: frontier vertices per CTA of the cluster (0: the leader CTA takes them all)
};
static_assert(sizeof(SoloLevel) == 32, "two LDS.128");

enum : u32 { CL_HELP = 1, CL_EXIT = 2 };   // commands of the leader CTA to the other CTAs of its cluster

struct Smem {
    LevelState st;
    Acc acc;
    u32 fail, hub;
    u32 qcur[2], qvalid[2];
    unsigned long long pend_edges, pend_visited;   // solo levels: totals not yet added to Ctrl
    u64 hub_bar[WPC][2];                           // hub phase: one mbarrier per warp and buffer
    u32 stage_cnt, stage_base, cta_new, pad0;      // grid mode: per-CTA staging of discoveries / counters
    alignas(16) SoloLevel lv;                      // solo mode: descriptor of the current level (read over DSMEM by cluster helpers)
    u32 cl_seq, cl_cmd, cl_n, pad1;                // cluster tier: sequence number / command for the helper CTAs, cluster size
    unsigned long long cta_edges;
    u32 stage[STAGE_CAP];
    alignas(128) u32 q[2][2][SOLO_CAP];            // solo: frontier mirrors; hub phase: per-warp TMA staging (WPC x 2 KB)
};
static_assert(sizeof(u32) * 2 * 2 * SOLO_CAP >= WPC * 2 * HUB_CHUNK * sizeof(u32), "staging area too small");

__device__ __noinline__ void trace_level(const KParams& K, const Env& E, u32 side, u32 mode, u32 expanded, u32 discovered, u64 edges,
                                         u64 t_expand) {
    Ctrl* C = E.C;
    if (C->trace_count >= TRACE_CAP) return;
    bibfs_level_trace& T = trace_of(K, E.me)[C->trace_count++];
    T.side = side; T.mode = mode; T.frontier = expanded; T.discovered = discovered;
    T.t_ns = ptx::globaltimer_ns(); T.edges = edges; T.t_expand_ns = t_expand;
}

__device__ __noinline__ void solo_push_hub(const KParams& K, const Env& E, Smem& sm, u32 u, u64 r0, u64 r1) {
    const u32 nch = hub_nchunks(r0, r1);
    const unsigned long long inc = (1ull << 40) | (unsigned long long)nch;
    const unsigned long long old = atomicAdd(&E.C->hub_packed, inc);
    const u32 slot = (u32)(old >> 40);
    HubEntry h;
    h.u = u; h.nchunks = nch; h.r0 = r0; h.r1 = r1; h.chunk_base = old & 0xFFFFFFFFFFull;
    if (slot < HUB_CAP) hub_of(K, E.me)[slot] = h;   // solo frontiers hold <= 1024 vertices: never overflows
    sm.hub = 1u;
}

// Consecutive solo levels of a single-rank search, one frontier vertex per thread.  This is the
// latency path (reference-sized graphs finish here; a 4-digit-path grid spends thousands of levels
// here).  A single level is a chain of dependent memory operations executed by a handful of
// warps, so what counts is the length of that chain and the number of instructions on it:
//   one 16-byte shared load (level descriptor) -> frontier vertex (shared mirror) -> row offsets
//   (L2, prefetched at discovery) -> up to 8 column loads in flight -> up to 8 claims in flight
//   -> one shared-memory cursor atomic per discovering lane -> stores.
// The level bookkeeping runs in registers of thread 0; the other threads only ever read the
// 16-byte descriptor.  Nothing is published to the grid until it is needed.
constexpr int SW = 8;       // adjacency entries per lane per round
constexpr u32 SOLO_HUB_ROW = 128;   // solo mode packs the frontier into few warps: longer rows are shared by the CTA or the grid
constexpr u32 CL_ENLIST_MIN = 192;   // cluster tier: smallest frontier that is dealt over the cluster
constexpr u32 CL_CAP = 64;          // rows the CTA keeps for itself per level ...
constexpr u32 CL_MAX_EDGES = 16384; // ... as long as they add up to no more than this many entries

template <bool MULTI>
__device__ __noinline__ void solo_run(const KParams& K, Env E, Smem& sm, u64 deadline) {
    const u32 tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    LevelState& st = sm.st;
    SoloLevel& s_lv = sm.lv;
    __shared__ u32 s_sent[2][MAXR];      // MULTI: entries appended to each peer's segment (cumulative, per side)
    __shared__ u32 s_rnew;               // MULTI: vertices discovered at other ranks in this level
    __shared__ u64 s_mail[2];
    // rows of 128..: first choice is the CTA itself (all 16 warps share them, no grid wake-up needed)
    __shared__ u32 s_cl_n, s_cl_edges, s_cl_u[CL_CAP];
    __shared__ u64 s_cl_r0[CL_CAP], s_cl_r1[CL_CAP];            // MULTI: reduced (edges, meet<<32 | discovered) handed from the leader warp to thread 0
    E.solo = true; E.A = &sm.acc; E.sA = &sm.acc; E.filter = false; E.s_hub = &sm.hub;
    E.s_sent = &s_sent[0][0]; E.s_rnew = &s_rnew;
    // the pointers of the claim loop, named so that they stay in registers (E itself is a stack copy)
    const u32* const g_col = E.col;
    const u64* const g_row = E.row_ptr;
    u64* const g_vis = E.vis;
    u64* const dummy = dummy_of(K, E.me);
    const u32 g_lo = E.lo;
    // thread 0 keeps the level state in registers across levels
    u32 nf0 = st.nf[0], nf1 = st.nf[1], d0 = st.depth[0], d1 = st.depth[1], levels = st.levels;
    u32 qlo0 = st.q_lo[0], qlo1 = st.q_lo[1], qhi0 = st.q_hi[0], qhi1 = st.q_hi[1];
    u32 cur0 = sm.qcur[0], cur1 = sm.qcur[1], val0 = sm.qvalid[0], val1 = sm.qvalid[1];
    u64 pend_edges = 0, pend_vis = 0;
    auto make_level = [&](u32 sd) {
        SoloLevel lv;
        lv.total = st.local_front;
        // bit 4: the other CTAs of the cluster take part.  One SM's load/store path is the limit of a wide level
        // (measured: 2.6 us for <= 32 vertices, 3.7 us for 256, 5.5 us for 512 in ONE CTA), so the frontier is dealt
        // evenly over the cluster as soon as that beats the ~1 us a cluster round costs
        const bool enlist = !MULTI && sm.cl_n > 1 && lv.total > CL_ENLIST_MIN;
        lv.side_flags = sd | ((sd ? val1 : val0) ? 2u : 0u) | ((sd ? cur1 : cur0) ? 4u : 0u) | (enlist ? 16u : 0u);
        lv.q_lo = sd ? qlo1 : qlo0;
        lv.base_tail = sd ? qhi1 : qhi0;
        lv.nq = lv.base_tail - lv.q_lo;
        lv.tag = (u32)(st.xepoch + 1);
        lv.rq_tag = st.rq_tag[sd];
        lv.slice = enlist ? min((u32)BFS_THREADS, max(64u, ((lv.total + sm.cl_n - 1u) / sm.cl_n + 31u) & ~31u)) : 0u;
        s_lv = lv;
        // a frontier larger than this CTA: the other CTAs of the cluster take the slices beyond the first
        if (enlist) { sm.cl_cmd = CL_HELP; ptx::dsmem_st_release(ptx::smem_addr(&sm.cl_seq), sm.cl_seq + 1u); }
    };
    // MULTI: per-peer base pointers, looked up with one shared-memory load per remote claim / entry (computing them from
    // the kernel parameters costs ~25 instructions per remote discovery)
    __shared__ u64* s_pvis[MAXR];            // peer's visited words
    __shared__ u64* s_prq[2][MAXR];          // this rank's segment in the peer's arena, per side
    __shared__ u32* s_ppar[2][MAXR];         // peer's parent array per side, biased so that it is indexed by the global vertex id
    if (MULTI && tid < K.P) {
        s_pvis[tid] = vis_of(K, tid);
        for (u32 sd = 0; sd < 2; ++sd) {
            s_prq[sd][tid] = rq_of(K, tid, sd, E.me);
            s_ppar[sd][tid] = parent_of(K, tid, sd) - (size_t)tid * K.B;
        }
    }
    if (MULTI && tid < 2 * MAXR) s_sent[tid / MAXR][tid % MAXR] = ptx::ld_relaxed_gpu(&E.C->sent[tid / MAXR][tid % MAXR]);
    if (tid == 0) { make_level(st.side); s_rnew = 0; s_cl_n = 0; s_cl_edges = 0; }
    __syncthreads();
    while (true) {
        DBG_STAMP(0, tid);
        const uint4 raw = *reinterpret_cast<const uint4*>(&s_lv);
        const u32 side = raw.x & 1u, total = raw.y, base_tail = raw.w;
        if (raw.x & 8u) break;
        const bool cluster_level = !MULTI && (raw.x & 16u);
        const u32 cur = (raw.x >> 2) & 1u;
        u32* const out = sm.q[side][cur ^ 1u];
        const u32 sh = 32u * side, osh = 32u - sh;
        u32 nq = total, tag = 0, rq_tag = 0;
        u32 mine = total;                       // frontier vertices of THIS CTA (the first slice in a cluster level)
        if (MULTI) {
            const uint4 raw2 = *(reinterpret_cast<const uint4*>(&s_lv) + 1);
            nq = raw2.x; tag = raw2.y; rq_tag = raw2.z;
            E.tag = tag;
        } else if (cluster_level) {
            mine = min(total, s_lv.slice);
        }
        const bool wide = !MULTI && total > 64u;  // aggregate the cursor reservations per warp
        u32 u = 0;
        u64 r0 = 0, r1 = 0;
        if (MULTI && (warp + 1u) * 32u > nq && warp * 32u < total) {
            // this warp reads entries of sender segments (tagged, polled by the whole warp together)
            const u64* e = (tid >= nq && tid < total) ? rq_entry(K, E, st, side, tid - nq) : nullptr;
            u = fetch_rq(E, e, rq_tag);
        }
        if (tid < mine) {
            if (!MULTI || tid < nq) u = (raw.x & 2u) ? sm.q[side][cur][tid] : __ldcg(E.q[side] + raw.z + tid);
            r0 = __ldg(g_row + (u - g_lo));
            r1 = __ldg(g_row + (u - g_lo) + 1);
        }
        u32 deg = (u32)min(r1 - r0, (u64)0xFFFFFFFFu);
        DBG_STAMP(1, deg);
        const u32 nedges = deg;                      // statistics: hub rows count as inspected here
        // a small frontier sits in a warp or two: there even a 32-entry row is better dealt over all 16 warps (CTA list)
        // than walked by its warp, row after row at ~0.6 us each
        if (deg >= (total <= 64u ? 32u : SOLO_HUB_ROW)) {
            const u32 slot = deg < HUB_ROW * 8u ? ptx::atoms_add(&s_cl_n, 1u) : CL_CAP;
            if (slot < CL_CAP) { s_cl_u[slot] = u; s_cl_r0[slot] = r0; s_cl_r1[slot] = r1; ptx::reds_add(&s_cl_edges, deg); }
            else solo_push_hub(K, E, sm, u, r0, r1);          // very long row, or the CTA list is full: whole grid
            deg = 0;
        }
        u32 sdeg = deg < 32u ? deg : 0u;
        DBG_STAMP(2, sdeg);
        // rows shorter than a warp: each lane walks its own row, SW entries per round
        // one round = W adjacency entries per lane; W = 4 when no row of the warp has more than 4 entries left (grids,
        // most Erdos-Renyi rows): half the loads and claims of the 8-wide round, and one SM's load/store path is what
        // limits a wide level
        auto round = [&](auto Wc, const u32 e0) {
            constexpr int W = decltype(Wc)::value;
            u32 v[W];
            u64 old[W];
            // branch-free issue: idle slots re-read the row's first entry and OR a zero into a private
            // scratch word, so that all W loads and then all W claims are in flight together
#pragma unroll
            for (int k = 0; k < W; ++k) v[k] = ptx::ld_relaxed_gpu(g_col + r0 + ((e0 + k < sdeg) ? e0 + k : 0u));
#pragma unroll
            for (int k = 0; k < W; ++k) {
                const bool act = e0 + k < sdeg;
                old[k] = ptx::atom_or_gpu(act ? &g_vis[v[k] >> 5] : dummy, act ? (1ull << ((v[k] & 31u) + sh)) : 0ull);
            }
#pragma unroll
            for (int k = 0; k < W; ++k)
                if (!(e0 + k < sdeg)) old[k] = ~0ull;
            DBG_STAMP(3, v[0] + v[W - 1]);
            u32 newmask = 0, meetmask = 0, remmask = 0;
            const u32 shz = sh + ptx::zero_after(old[W - 1]);      // == sh; pins the tests behind the whole batch
#pragma unroll
            for (int k = 0; k < W; ++k) {
                const u32 isnew = ((u32)(old[k] >> ((v[k] & 31u) + shz)) & 1u) ^ 1u;
                newmask |= isnew << k;
            }
            if (MULTI && __any_sync(FULL, newmask != 0)) {
                // vertices of other ranks: the local bit was only a "not forwarded yet" filter — claim at the
                // owner.  Branch-free again so that the NVLink round trips (1.6 us each) overlap.
                u64 rold[W];
#pragma unroll
                for (int k = 0; k < W; ++k) {
                    const bool rem = ((newmask >> k) & 1u) && (v[k] - g_lo) >= K.B;
                    remmask |= rem ? (1u << k) : 0u;
                    u64* const addr = rem ? (s_pvis[owner_of(K, v[k])] + (v[k] >> 5)) : dummy;
                    rold[k] = ptx::atom_or_sys(addr, rem ? (1ull << ((v[k] & 31u) + sh)) : 0ull);
                }
                const u32 shr = sh + ptx::zero_after(rold[W - 1]);
#pragma unroll
                for (int k = 0; k < W; ++k) {
                    if ((remmask >> k) & 1u) {
                        old[k] = rold[k];
                        if ((rold[k] >> ((v[k] & 31u) + shr)) & 1ull) { newmask &= ~(1u << k); remmask &= ~(1u << k); }
                    }
                }
            }
#pragma unroll
            for (int k = 0; k < W; ++k)
                meetmask |= (((newmask >> k) & 1u) & (u32)(old[k] >> ((v[k] & 31u) + osh))) << k;
            DBG_STAMP(4, newmask);
            const u32 locmask = newmask & ~remmask;
            u32 wslot = 0;
            if (wide) wslot = warp_reserve((u32)__popc(locmask), [&](u32 n) { return ptx::atoms_add(&sm.acc.q_tail[side], n); });
            if (locmask) {
                u32 slot = wide ? wslot : ptx::atoms_add(&sm.acc.q_tail[side], (u32)__popc(locmask));
                u32* const qs = E.q[side];
                u32* const ps = E.parent[side];
#pragma unroll
                for (int k = 0; k < W; ++k) {
                    if ((locmask >> k) & 1u) {
                        qs[slot] = v[k];
                        ps[v[k] - g_lo] = u;
                        if (slot - base_tail < SOLO_CAP) out[slot - base_tail] = v[k];
                        ptx::prefetch_l2(g_row + (v[k] - g_lo));
                        ++slot;
                    }
                }
            }
            if (MULTI && remmask) {
#pragma unroll
                for (int k = 0; k < W; ++k) {
                    if ((remmask >> k) & 1u) {
                        const u32 dest = owner_of(K, v[k]);
                        const u32 slot = ptx::atoms_add(&s_sent[side][dest], 1u);
                        ptx::st_relaxed_sys(s_prq[side][dest] + slot, (u64)v[k] | ((u64)tag << 32));
                        ptx::st_peer(s_ppar[side][dest] + v[k], u);
                    }
                }
                ptx::reds_add(&s_rnew, (u32)__popc(remmask));
            }
            if (meetmask) {
#pragma unroll
                for (int k = 0; k < W; ++k)
                    if ((meetmask >> k) & 1u) ptx::reds_min(&sm.acc.meet_local, v[k]);
            }
                };
        for (u32 e0 = 0;;) {
            const u32 left = __reduce_max_sync(FULL, sdeg > e0 ? sdeg - e0 : 0u);
            if (left == 0) break;
            if (left <= 4u) { round(std::integral_constant<int, 4>{}, e0); e0 += 4; }
            else { round(std::integral_constant<int, SW>{}, e0); e0 += SW; }
        }
        DBG_STAMP(5, sdeg);
        // rows of a warp's length or more (below the hub threshold): the warp walks them together
        u32 big = __ballot_sync(FULL, deg >= 32u);
        if (big) {
            E.sq = out; E.sq_base = base_tail;
            u32 cnt = 0;     // (discoveries are derived from the queue cursor and s_rnew)
            while (big) {
                const u32 sl = __ffs(big) - 1;
                big &= big - 1;
                const u32 uu = __shfl_sync(FULL, u, sl);
                const u64 rb = __shfl_sync(FULL, r0, sl), re = __shfl_sync(FULL, r1, sl);
                expand_range<MULTI>(K, E, side, uu, rb, re, cnt);
            }
            (void)cnt;
        }
        // per-warp edge count (statistics only, off the critical path)
        {
            const u32 wsum = __reduce_add_sync(FULL, nedges);
            if (lane == 0 && wsum) ptx::reds_add64(&sm.acc.edges, wsum);
        }
        if (cluster_level) ptx::cluster_sync_all();      // the helpers' slices are done, their DSMEM updates visible
        else __syncthreads();
        if (s_cl_n) {
            // long rows collected by this level: shared by all warps of the CTA in 128-entry pieces, unless there
            // is so much of it that the grid should take over (hub phase)
            const u32 ncl = min(s_cl_n, CL_CAP);
            if (s_cl_edges <= CL_MAX_EDGES && !sm.hub) {
                E.sq = out; E.sq_base = base_tail;
                u32 cnt = 0, piece = 0;
                for (u32 i = 0; i < ncl; ++i) {
                    const u64 rb = s_cl_r0[i], re = s_cl_r1[i];
                    const u32 uu = s_cl_u[i];
                    for (u64 p = rb; p < re; p += 32 * UNR, ++piece)
                        if (piece % WPC == warp) expand_range<MULTI>(K, E, side, uu, p, min(re, p + 32 * UNR), cnt);
                }
                (void)cnt;
            } else if (tid < ncl) {
                solo_push_hub(K, E, sm, s_cl_u[tid], s_cl_r0[tid], s_cl_r1[tid]);
            }
            __syncthreads();
            if (tid == 0) { s_cl_n = 0; s_cl_edges = 0; }
        }
        DBG_STAMP(6, tid);
        // ---- level end
        u32 found = 0, meet = NO_V;
        if (MULTI) {
            // leader warp: exchange this rank's counts with every peer (tagged words over NVLink).
            // Not when a hub row was deferred: then the level is not over yet.
            if (warp == 0 && !sm.hub) {
                Slot mine, got;
                mine.count = (sm.acc.q_tail[side] - base_tail) + s_rnew;      // local cursor advance + claims at other ranks
                mine.meet = sm.acc.meet_local;
                mine.edges = (u32)sm.acc.edges;
                mine.rq_tail = lane < K.P ? s_sent[side][lane] : 0u;
                const bool ok = cross_exchange<MULTI>(K, E, st, mine, got, deadline, false);
                u32 total_new = got.count, meetmin = got.meet;
                u64 etot = got.edges;
#pragma unroll
                for (int d = 16; d > 0; d >>= 1) {
                    total_new += __shfl_xor_sync(FULL, total_new, d);
                    meetmin = min(meetmin, __shfl_xor_sync(FULL, meetmin, d));
                    etot += __shfl_xor_sync(FULL, etot, d);
                }
                if (lane < K.P) { st.rq_lo[side][lane] = st.rq_hi[side][lane]; st.rq_hi[side][lane] = got.rq_tail; }
                if (lane == 0) {
                    st.rq_tag[side] = (u32)st.xepoch;
                    s_mail[0] = etot;
                    s_mail[1] = ((u64)meetmin << 32) | total_new;
                    if (!ok) { sm.fail = 1; atomicExch(&E.C->abort_flag, 1u); }
                }
                __syncwarp();
            }
            if (tid == 0) {
                if (sm.hub) { found = (sm.acc.q_tail[side] - base_tail) + s_rnew; }
                else { found = (u32)s_mail[1]; meet = (u32)(s_mail[1] >> 32); }
            }
        } else if (tid == 0) {
            found = sm.acc.q_tail[side] - base_tail;
            meet = sm.acc.meet_local;
        }
        if (tid == 0) {
            Ctrl* C = E.C;
            const u32 tail = sm.acc.q_tail[side];
            u32 stop = 0;
            if (MULTI && sm.fail) {
                st.status = 1; st.length = -1; st.phase = PH_DONE; stop = 1;
            } else if (sm.hub) {
                // the level is not over: publish and let the grid expand the deferred rows
                const u64 hp = ptx::ld_relaxed_gpu(reinterpret_cast<const u64*>(&C->hub_packed));
                st.hub_entries = (u32)(hp >> 40);
                st.hub_chunks = (u32)(hp & 0xFFFFFFFFFFull);
                st.prev_phase = PH_TD_SOLO;
                st.phase = PH_TD_HUB;
                C->hub_packed = 0;
                sm.acc.new_count = found;
                sm.hub = 0;
                if (side) { cur1 ^= 1u; val1 = 0; } else { cur0 ^= 1u; val0 = 0; }
                stop = 1;
            } else {
                const u64 ed_local = sm.acc.edges;
                const u64 ed = MULTI ? s_mail[0] : ed_local;
                const u32 expanded = side ? nf1 : nf0;
                const u32 lfound = tail - base_tail;
                if (side) { nf1 = found; d1 += 1; qlo1 = base_tail; qhi1 = tail; cur1 ^= 1u; val1 = lfound <= SOLO_CAP; }
                else      { nf0 = found; d0 += 1; qlo0 = base_tail; qhi0 = tail; cur0 ^= 1u; val0 = lfound <= SOLO_CAP; }
                levels += 1;
                pend_edges += ed; pend_vis += found;
                sm.acc.edges = 0;
                if (MULTI) { s_rnew = 0; sm.acc.new_count = 0; }
                if (K.flags & BIBFS_F_TRACE) trace_level(K, E, side, PH_TD_SOLO, expanded, found, ed_local, ptx::globaltimer_ns());
                if (meet != NO_V) {
                    sm.acc.meet_local = NO_V;
                    st.length = (i32)(d0 + d1);
                    st.meet = meet;
                    st.phase = PH_DONE;
                    stop = 1;
                } else {
                    const u32 ns = nf0 <= nf1 ? 0u : 1u;           // smaller frontier first, ties -> source (v1)
                    const u32 nfn = ns ? nf1 : nf0;
                    u32 lf = ns ? qhi1 - qlo1 : qhi0 - qlo0;
                    if (MULTI)
                        for (u32 p = 0; p < K.P; ++p) lf += st.rq_hi[ns][p] - st.rq_lo[ns][p];
                    st.side = ns;
                    st.local_front = lf;
                    if (nfn == 0) { st.length = -1; st.phase = PH_DONE; stop = 1; }
                    else if (lf > K.solo_threshold || (K.flags & (BIBFS_F_FORCE_BU | BIBFS_F_FORCE_GRID)) ||
                             (!(K.flags & BIBFS_F_FORCE_TD) && K.bu_div != 0xFFFFFFFFu && nfn > K.n / K.bu_div)) stop = 2;
                    else make_level(ns);
                }
            }
            if (stop) {
                // write the register state back before anyone else looks at it
                st.nf[0] = nf0; st.nf[1] = nf1; st.depth[0] = d0; st.depth[1] = d1; st.levels = levels;
                st.q_lo[0] = qlo0; st.q_lo[1] = qlo1; st.q_hi[0] = qhi0; st.q_hi[1] = qhi1;
                st.fbm_valid[0] = 0; st.fbm_valid[1] = 0;
                sm.qcur[0] = cur0; sm.qcur[1] = cur1; sm.qvalid[0] = val0; sm.qvalid[1] = val1;
                sm.acc.new_count = (st.phase == PH_TD_HUB) ? sm.acc.new_count : 0;
                if (stop == 2) decide_next(K, st);
                st.edges_total += pend_edges;
                st.visited_total += pend_vis;
                if (MULTI)
                    for (u32 p = 0; p < MAXR; ++p) { C->sent[0][p] = s_sent[0][p]; C->sent[1][p] = s_sent[1][p]; }
                s_lv.side_flags = 8u;
            }
        }
        DBG_STAMP(7, st.phase);
#ifdef BIBFS_FINE_TRACE
        if (blockIdx.x == 0 && threadIdx.x == 0) g_dbg_level++;
#endif
        __syncthreads();
    }
    // leaving solo mode
    if (MULTI && st.phase == PH_DONE && warp == 0 && !sm.fail) {
        // every rank reaches the same verdict: one fenced exchange makes all remote parent stores visible
        Slot z; z.count = 0; z.meet = NO_V; z.rq_tail = 0; z.edges = 0;
        Slot g2;
        cross_exchange<MULTI>(K, E, st, z, g2, deadline, true);
    }
    if (tid == 0 && st.phase == PH_DONE) write_result(K, E.C, st);
    __syncthreads();
}

// ------------------------------------------------------------------------------------------
// cluster tier: the CTAs that share a thread-block cluster with the leader CTA help it with
// frontiers of 513..cl_n*512 vertices.  They find the level descriptor, the frontier mirror and
// the accumulators in the LEADER's shared memory through DSMEM (mapa + ld/st/atom.shared::cluster)
// and meet the leader at a hardware cluster barrier, which costs ~0.2 us where waking the whole
// grid through global memory costs ~7 us per level.  Single rank only.
// ------------------------------------------------------------------------------------------
__device__ __noinline__ void cluster_helper(const KParams& K, const Env& E, Smem& sm, u32 cl_rank, u32& seen_seq, u64 deadline) {
    const u32 tid = threadIdx.x, lane = tid & 31;
    __shared__ u32 s_cmd, s_seq;
    __shared__ __align__(16) SoloLevel s_lvc;
    // leader-side addresses (same layout in every CTA of the kernel)
    const u32 a_hub = ptx::dsmem_map(ptx::smem_addr(&sm.hub), 0);
    const u32 a_meet = ptx::dsmem_map(ptx::smem_addr(&sm.acc.meet_local), 0);
    const u32 a_edges = ptx::dsmem_map(ptx::smem_addr(&sm.acc.edges), 0);
    while (true) {
        if (tid < 32) {
            // warp-collective poll (see wait_gen): lane 0 reads the leader's sequence word over DSMEM, all lanes leave together.
            // (Measured alternative: the leader PUSHING descriptor + command into an inbox in each helper's own shared memory,
            //  polled locally - 9 % slower on the 4096 grid and 5 % on er100k; the helpers' tight local polling and the extra
            //  stores in the leader's loop cost more than the three DSMEM round trips they save.)
            u32 v = seen_seq, spins = 0;
            bool dead = false;
            const u32 a_seq = ptx::dsmem_map(ptx::smem_addr(&sm.cl_seq), 0), a_cmd = ptx::dsmem_map(ptx::smem_addr(&sm.cl_cmd), 0);
            const u32 a_lv = ptx::dsmem_map(ptx::smem_addr(&sm.lv), 0);
            while (true) {
                if (lane == 0) v = ptx::dsmem_ld(a_seq);
                v = __shfl_sync(FULL, v, 0);
                if (v != seen_seq) break;
                ptx::nanosleep(96);
                if (++spins > 4096u && (spins & 1023u) == 0 && expired_warp(E.C, deadline)) { dead = true; break; }
            }
            if (lane == 0) { v = ptx::dsmem_ld_acquire(a_seq); s_seq = v; s_cmd = dead ? (u32)CL_EXIT : ptx::dsmem_ld(a_cmd); }
            __syncwarp();
            if (!dead && s_cmd == CL_HELP && lane < sizeof(SoloLevel) / 4) reinterpret_cast<u32*>(&s_lvc)[lane] = ptx::dsmem_ld(a_lv + 4 * lane);
        }
        __syncthreads();
        seen_seq = s_seq;
        if (s_cmd != CL_HELP) return;
        const u32 side = s_lvc.side_flags & 1u, cur = (s_lvc.side_flags >> 2) & 1u;
        const u32 total = s_lvc.total, base_tail = s_lvc.base_tail;
        const u32 sh = 32u * side, osh = 32u - sh;
        const u32 slice = s_lvc.slice;
        const u32 i = tid < slice ? cl_rank * slice + tid : 0xFFFFFFFFu;
        const u32 a_tail = ptx::dsmem_map(ptx::smem_addr(&sm.acc.q_tail[side]), 0);
        const u32 a_out = ptx::dsmem_map(ptx::smem_addr(&sm.q[side][cur ^ 1u][0]), 0);
        u32 u = 0;
        u64 r0 = 0, r1 = 0;
        if (i < total) {
            u = (s_lvc.side_flags & 2u) ? ptx::dsmem_ld(ptx::dsmem_map(ptx::smem_addr(&sm.q[side][cur][i]), 0))
                                        : __ldcg(E.q[side] + s_lvc.q_lo + i);
            r0 = __ldg(E.row_ptr + (u - E.lo));
            r1 = __ldg(E.row_ptr + (u - E.lo) + 1);
        }
        u32 deg = (u32)min(r1 - r0, (u64)0xFFFFFFFFu);
        const u32 nedges = deg;
        if (deg >= 32u) {
            // long rows go to the grid-wide hub phase (the leader sees the flag after the cluster barrier)
            const u32 nch = hub_nchunks(r0, r1);
            const unsigned long long inc = (1ull << 40) | (unsigned long long)nch;
            const unsigned long long old = atomicAdd(&E.C->hub_packed, inc);
            const u32 slot = (u32)(old >> 40);
            HubEntry h;
            h.u = u; h.nchunks = nch; h.r0 = r0; h.r1 = r1; h.chunk_base = old & 0xFFFFFFFFFFull;
            if (slot < HUB_CAP) hub_of(K, E.me)[slot] = h;
            ptx::dsmem_st(a_hub, 1u);
            deg = 0;
        }
        u64* const dummy = dummy_of(K, E.me);
        auto round = [&](auto Wc, const u32 e0) {
            constexpr int W = decltype(Wc)::value;
            u32 v[W];
            u64 old[W];
#pragma unroll
            for (int k = 0; k < W; ++k) v[k] = ptx::ld_relaxed_gpu(E.col + r0 + ((e0 + k < deg) ? e0 + k : 0u));
#pragma unroll
            for (int k = 0; k < W; ++k) {
                const bool act = e0 + k < deg;
                old[k] = ptx::atom_or_gpu(act ? &E.vis[v[k] >> 5] : dummy, act ? (1ull << ((v[k] & 31u) + sh)) : 0ull);
            }
            u32 newmask = 0, meetmask = 0;
            const u32 shz = sh + ptx::zero_after(old[W - 1]);
#pragma unroll
            for (int k = 0; k < W; ++k) {
                const u32 isnew = (e0 + k < deg) ? (((u32)(old[k] >> ((v[k] & 31u) + shz)) & 1u) ^ 1u) : 0u;
                newmask |= isnew << k;
                meetmask |= (isnew & (u32)(old[k] >> ((v[k] & 31u) + osh))) << k;
            }
            // the leader's queue cursor: one DSMEM atomic per warp
            u32 slot = warp_reserve((u32)__popc(newmask), [&](u32 n) { return ptx::dsmem_atom_add(a_tail, n); });
            if (newmask) {
#pragma unroll
                for (int k = 0; k < W; ++k) {
                    if ((newmask >> k) & 1u) {
                        E.q[side][slot] = v[k];
                        E.parent[side][v[k] - E.lo] = u;
                        if (slot - base_tail < SOLO_CAP) ptx::dsmem_st(a_out + 4u * (slot - base_tail), v[k]);
                        ptx::prefetch_l2(E.row_ptr + (v[k] - E.lo));
                        ++slot;
                    }
                }
                if (meetmask) {
#pragma unroll
                    for (int k = 0; k < W; ++k)
                        if ((meetmask >> k) & 1u) ptx::dsmem_red_min(a_meet, v[k]);
                }
            }
                };
        for (u32 e0 = 0;;) {
            const u32 left = __reduce_max_sync(FULL, deg > e0 ? deg - e0 : 0u);
            if (left == 0) break;
            if (left <= 4u) { round(std::integral_constant<int, 4>{}, e0); e0 += 4; }
            else { round(std::integral_constant<int, SW>{}, e0); e0 += SW; }
        }
        {
            const u32 wsum = __reduce_add_sync(FULL, nedges);
            if (lane == 0 && wsum) ptx::dsmem_red_add64(a_edges, wsum);
        }
        ptx::cluster_sync_all();
    }
}

#include "pair_tier.cuh"

constexpr size_t SMEM_BYTES = sizeof(Smem) + sizeof(PairSmem);
static_assert(sizeof(Smem) % 16 == 0, "PairSmem follows Smem in dynamic shared memory");
static_assert(SMEM_BYTES <= 227 * 1024, "shared memory of the search kernel");

}  // namespace

// ------------------------------------------------------------------------------------------
// the kernel
// ------------------------------------------------------------------------------------------
template <bool MULTI>
__global__ void __launch_bounds__(BFS_THREADS, 1) k_bibfs(const __grid_constant__ KParams Kp) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    // The parameters are copied into shared memory and every function below works on THAT copy.  The device functions
    // of this file take `const KParams&`; a reference to a __grid_constant__ parameter is a generic pointer into the
    // parameter window, every K.x behind it is a generic load (~0.5 us, and reloaded after each asm memory clobber),
    // and those loads sat on the dependent chain of the levels (measured in the pair tier: +2 us per round from
    // owner_of's reciprocal alone).  Shared memory answers in ~30 cycles.
    __shared__ KParams sK;
    {
        const u32* src = reinterpret_cast<const u32*>(&Kp);
        u32* dst = reinterpret_cast<u32*>(&sK);
        for (u32 i = threadIdx.x; i < sizeof(KParams) / 4; i += BFS_THREADS) dst[i] = src[i];
    }
    __syncthreads();
    const KParams& K = sK;
    Smem& sm = *reinterpret_cast<Smem*>(smem_raw);
    PairSmem& ps = *reinterpret_cast<PairSmem*>(smem_raw + sizeof(Smem));
    KT_STAMP(0);
    LevelState& st = sm.st;        // this CTA's copy of the level state (the leader CTA's is authoritative)
    Acc& s_acc = sm.acc;           // leader CTA: accumulators of solo levels / staging at sync points
    u32& s_fail = sm.fail;
    u32& s_hub = sm.hub;           // solo level deferred at least one hub row
    auto& s_q = sm.q;              // solo mode: entries each side appended in its last level (double buffered)
    u32 (&s_qcur)[2] = sm.qcur;    // which buffer holds the side's current frontier
    u32 (&s_qvalid)[2] = sm.qvalid;  // ... and whether that mirror is complete
    const u32 tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const u32 vr = MULTI ? blockIdx.x / K.cpr : 0u;
    Env E;
    E.cta = blockIdx.x - vr * K.cpr;
    E.cpr = K.cpr;
    E.me = K.rank_lo + vr;
    E.lo = E.me * K.B;
    E.P = K.P; E.B = K.B; E.n = K.n; E.nw = K.L.nw;
    E.C = ctrl_of(K, E.me);
    E.A = &E.C->acc;
    E.sq = nullptr; E.sq_base = 0; E.s_hub = nullptr; E.filter = true;
    E.sA = &s_acc; E.solo = false; E.tag = 0; E.smem_src = false;
    E.stage = sm.stage; E.stage_cnt = &sm.stage_cnt; E.cta_new = &sm.cta_new; E.cta_edges = &sm.cta_edges; E.s_sent = nullptr; E.s_rnew = nullptr;
    E.vis = vis_of(K, E.me);
    E.parent[0] = parent_of(K, E.me, 0); E.parent[1] = parent_of(K, E.me, 1);
    E.q[0] = q_of(K, E.me, 0); E.q[1] = q_of(K, E.me, 1);
    E.row_ptr = K.row_ptr[E.me];
    E.col = K.col[E.me];
    Ctrl* C = E.C;
    const u64 deadline = ptx::globaltimer_ns() + (u64)K.watchdog_ms * 1000000ull;
    const u64 genbase = (u64)K.seq << 32;
    const bool is_leader_cta = (E.cta == 0);
    const u32 cl_n = MULTI ? 1u : ptx::cluster_nctarank(), cl_rank = ptx::cluster_ctarank();
    const bool in_leader_cluster = cl_n > 1 && blockIdx.x < cl_n;       // shares a cluster (and DSMEM) with the leader CTA
    u32 cl_seen = 0;
    if (tid == 0) { s_fail = 0; s_hub = 0; s_qvalid[0] = 0; s_qvalid[1] = 0; s_qcur[0] = 0; s_qcur[1] = 0; sm.pend_edges = 0; sm.pend_visited = 0; sm.stage_cnt = 0; sm.cta_new = 0; sm.cta_edges = 0; }
    if (lane == 0) { ptx::mbar_init(&sm.hub_bar[warp][0], 1); ptx::mbar_init(&sm.hub_bar[warp][1], 1); ptx::mbar_fence_init(); }
    u32 hub_parity = 0;            // phase bits of this warp's two hub barriers
    if (tid == 0) { sm.cl_seq = 0; sm.cl_cmd = 0; sm.cl_n = cl_n; }
    __syncthreads();
    // the helpers may poll the leader's command word once the leader CTA has started.  The leader itself only ARRIVES
    // here and collects the barrier when it first needs its cluster: the CTAs of a cluster start up to ~2 us apart
    bool cl_pending = false;
    if (cl_n > 1) {
        if (is_leader_cta) { ptx::cluster_arrive(); cl_pending = true; }
        else ptx::cluster_sync_all();
    }

    // non-leader CTAs: copy the published level state global -> shared after a wake-up
    auto load_state = [&]() {
        const u32* g = reinterpret_cast<const u32*>(&C->st);
        u32* s = reinterpret_cast<u32*>(&st);
        for (u32 i = tid; i < sizeof(LevelState) / 4; i += BFS_THREADS) s[i] = __ldcg(g + i);
        __syncthreads();
    };
    // leader warp: publish the shared state + accumulators and wake the other CTAs
    bool helpers_listening = cl_n > 1;      // leader CTA: the cluster helpers are waiting for commands (initially, and during solo phases)
    auto publish_and_wake = [&]() {
        if (helpers_listening) {
            // the helpers of this cluster leave their command loop before the grid-wide release goes out
            if (lane == 0) { sm.cl_cmd = CL_EXIT; ptx::dsmem_st_release(ptx::smem_addr(&sm.cl_seq), sm.cl_seq + 1u); }
            __syncwarp();
        }
        helpers_listening = cl_n > 1 && st.phase == PH_TD_SOLO;
        if (lane == 0) st.wake += 1;
        __syncwarp();
        u32* g = reinterpret_cast<u32*>(&C->st);
        const u32* s = reinterpret_cast<const u32*>(&st);
        for (u32 i = lane; i < sizeof(LevelState) / 4; i += 32) g[i] = s[i];
        __syncwarp();
        if (lane == 0) {
            C->acc = s_acc;
            ptx::fence_gpu();
            ptx::st_release_gpu(&C->bar_gen, genbase | (u64)st.wake);
        }
        __syncwarp();
    };

    // ---- query start.  The leader CTA always works on the shared-memory state; it publishes it
    // whenever the next phase needs the other CTAs.
    if (is_leader_cta) {
        if (warp == 0) {
            KT_STAMP(1);
            const bool ok = K.resume ? leader_resume<MULTI>(K, E, st, s_acc, deadline) : leader_init<MULTI>(K, E, st, s_acc, deadline);
            KT_STAMP(2);
            if (!ok && lane == 0) s_fail = 1;
            if (lane == 0 && !K.resume) {   // the seeds are the first mirrored entries
                const u32 src = C->q_src, dst = C->q_dst;     // as resolved by leader_init
                if (s_acc.q_tail[0] == 1) { s_q[0][0][0] = src; s_qvalid[0] = 1; }
                if (s_acc.q_tail[1] == 1) { s_q[1][0][0] = dst; s_qvalid[1] = 1; }
            }
            __syncwarp();
            if (st.phase != PH_TD_SOLO) publish_and_wake();
        }
        __syncthreads();
    } else {
        if (!MULTI && K.warm_lines && E.cpr > 1) {
            // the graph fits L2 with room to spare: the CTAs that would only wait stream the vertex records and the
            // visited words into L2 while the leader runs the first rounds (they were evicted since the last query
            // whenever anything else used the GPU; a cold record costs ~700 ns on the leader's chain, a hit 154 ns)
            const char* rec = reinterpret_cast<const char*>(K.vrec[E.me]);
            const char* vw = reinterpret_cast<const char*>(E.vis);
            const u32 vis_lines = (E.nw * 8u + 127u) / 128u;
            const u32 stride = (E.cpr - 1u) * BFS_THREADS;
            for (u32 i = (E.cta - 1u) * BFS_THREADS + tid; i < K.warm_lines + vis_lines; i += stride)
                ptx::prefetch_l2(i < K.warm_lines ? rec + 128ull * i : vw + 128ull * (i - K.warm_lines));
        }
        if (in_leader_cluster) cluster_helper(K, E, sm, cl_rank, cl_seen, deadline);   // may be needed before the first release
        if (warp == 0 && !wait_gen(C, genbase | 1u, deadline) && lane == 0) s_fail = 1;
        __syncthreads();
        if (!s_fail) load_state();
    }
    bool alive = (s_fail == 0);
    // latency tier first (leader CTA): paired rounds on vertex records, while both frontiers are small
    bool pair_first = !MULTI && !K.resume && K.vrec[E.me] != nullptr && !(K.flags & (BIBFS_F_FORCE_GRID | BIBFS_F_FORCE_BU));

    // Synchronisation point of grid-mode phases.  Returns false if the query was aborted.
    auto sync_grid = [&](int kind, u32 mode) -> bool {
        __syncthreads();
        // ---- CTA flush: one queue reservation and one counter update per CTA (skipped by CTAs that found nothing)
        if (sm.stage_cnt | sm.cta_new | (u32)(sm.cta_edges != 0)) {
            if (tid == 0) {
                const u32 cnt = min(sm.stage_cnt, STAGE_CAP);
                sm.stage_base = cnt ? atomicAdd(&C->acc.q_tail[st.side], cnt) : 0u;
                if (sm.cta_new) atomicAdd(&C->acc.new_count, sm.cta_new);
                if (sm.cta_edges) atomicAdd(&C->acc.edges, sm.cta_edges);
            }
            __syncthreads();
            {
                const u32 cnt = min(sm.stage_cnt, STAGE_CAP);
                u32* qd = E.q[st.side] + sm.stage_base;
                for (u32 i = tid; i < cnt; i += BFS_THREADS) qd[i] = sm.stage[i];
            }
            __syncthreads();
            if (tid == 0) { sm.stage_cnt = 0; sm.cta_new = 0; sm.cta_edges = 0; }
        }
        if (!is_leader_cta) {
            if (warp == 0) {
                if (lane == 0) {
                    if (MULTI && mode == PH_BU) ptx::fence_sys();   // bitmap slices pushed with plain stores
                    ptx::red_release_gpu_add(&C->bar_way[E.cta % BAR_WAYS].count, 1u);
                }
                __syncwarp();
                DBG_STAMP3(4, tid);      // arrived
                if (!wait_gen(C, genbase | (u64)(st.wake + 1u), deadline, false) && lane == 0) s_fail = 1;
                DBG_STAMP3(5, tid);      // released
            }
            __syncthreads();
            if (s_fail) return false;
            load_state();
            DBG_STAMP3(6, st.phase);     // state loaded
            return true;
        }
        if (warp == 0) {
            bool ok = wait_count(C, E.cpr, deadline);
            if (lane == 0) pull_acc(s_acc, &C->acc);    // the grid's accumulators -> shared memory
            if (lane == 0 && kind == SYNC_LEVEL) s_qvalid[st.side] = 0;   // grid levels do not feed the solo mirror
            __syncwarp();
            if (ok) ok = leader_sync<MULTI>(K, E, st, s_acc, kind, mode, deadline, true);
            else if (lane == 0) { atomicExch(&C->abort_flag, 1u); st.status = 1; st.phase = PH_DONE; write_result(K, C, st); }
            if (!ok && lane == 0) s_fail = 1;
            __syncwarp();
            publish_and_wake();     // grid mode: every sync point releases the other CTAs
        }
        __syncthreads();
        return s_fail == 0;
    };

    while (alive) {
        const u32 phase = st.phase;
        if (phase == PH_DONE) break;
        if (phase == PH_TD_SOLO) {
            if (!is_leader_cta) {   // parked until the leader wakes the grid
                if (in_leader_cluster) cluster_helper(K, E, sm, cl_rank, cl_seen, deadline);   // ... helping it meanwhile
                if (warp == 0 && !wait_gen(C, genbase | (u64)(st.wake + 1u), deadline) && lane == 0) s_fail = 1;
                __syncthreads();
                if (s_fail) { alive = false; break; }
                load_state();
                continue;
            }
            if (pair_first) {
                pair_first = false;
                KT_STAMP(3);
                {
                    PairIO io;
                    io.log[0] = E.q[0]; io.log[1] = E.q[1]; io.par[0] = E.parent[0]; io.par[1] = E.parent[1];
                    pair_run<false, PT_CAP_SMALL>(K, E, ps, io, C->q_src, C->q_dst, false);
                    if (tid == 0) pair_finish<false>(K, E, st, s_acc, s_qvalid, ps.out, ps.edges);
                    __syncthreads();
                }
                KT_STAMP(4);
                if (st.phase != PH_TD_SOLO) { if (warp == 0) publish_and_wake(); __syncthreads(); continue; }
            }
            if (cl_pending) { ptx::cluster_wait(); cl_pending = false; }
            solo_run<MULTI>(K, E, sm, deadline);
            if (s_fail) { if (warp == 0) publish_and_wake(); __syncthreads(); alive = false; break; }
            if (warp == 0) publish_and_wake();      // left solo mode: the grid is needed (or the query is done)
            __syncthreads();
        } else if (phase == PH_TD_GRID) {
            E.A = &C->acc;
            E.tag = (u32)(st.xepoch + 1);
            E.filter = st.nf[st.side] >= FILTER_MIN;
            expand_td<MULTI>(K, E, st, E.cta * WPC + warp, E.cpr * WPC);
            alive = sync_grid(SYNC_LEVEL, PH_TD_GRID);
        } else if (phase == PH_TD_HUB) {
            E.A = &C->acc;
            E.tag = (u32)(st.xepoch + 1);
            if (is_leader_cta && tid == 0) { s_qvalid[0] = 0; s_qvalid[1] = 0; }    // the mirrors double as TMA staging
            expand_hubs<MULTI>(K, E, st, E.cta * WPC + warp, E.cpr * WPC, &sm.q[0][0][0] + warp * 2 * HUB_CHUNK, sm.hub_bar[warp], hub_parity);
            alive = sync_grid(SYNC_LEVEL, PH_TD_HUB);
        } else {  // PH_BU
            E.A = &C->acc;
            const u32 side = st.side;
            const u32 gtid = E.cta * BFS_THREADS + tid, gthreads = E.cpr * BFS_THREADS;
            if (!st.fbm_valid[side]) {
                // queue windows -> bitmap slice, then replicate the slice
                u32* f = fbm_of(K, E.me, side, st.fbm_cur[side]);
                const u32 w0 = E.lo / 32, words = K.B / 32;
                for (u32 i = gtid; i < words; i += gthreads) f[w0 + i] = 0;
                alive = sync_grid(SYNC_LOCAL, PH_BU);
                if (!alive) break;
                for (u32 i = gtid; i < st.local_front; i += gthreads) {
                    const u32 u = fetch_frontier<MULTI>(K, E, st, side, i);
                    atomicOr(&f[u >> 5], 1u << (u & 31u));
                }
                if (MULTI) {
                    alive = sync_grid(SYNC_LOCAL, PH_BU);
                    if (!alive) break;
                    push_fbm_slice(K, E, side, st.fbm_cur[side], gtid, gthreads);
                    alive = sync_grid(SYNC_CROSS, PH_BU);
                } else {
                    alive = sync_grid(SYNC_LOCAL, PH_BU);
                }
                if (!alive) break;
            }
            expand_bu<MULTI>(K, E, st, E.cta * WPC + warp, E.cpr * WPC);
            if (K.long_rows) {       // the graph may have rows for the hub list: they are scanned by all warps of the rank
                alive = sync_grid(SYNC_LOCAL, PH_BU);
                if (!alive) break;
                expand_bu_hubs<MULTI>(K, E, st, E.cta * WPC + warp, E.cpr * WPC);
            }
            if (MULTI) {
                alive = sync_grid(SYNC_LOCAL, PH_BU);
                if (!alive) break;
                push_fbm_slice(K, E, side, st.fbm_cur[side] ^ 1u, gtid, gthreads);
            }
            alive = sync_grid(SYNC_LEVEL, PH_BU);
        }
    }
    if (cl_pending) { ptx::cluster_wait(); cl_pending = false; }
    if (!alive) {         // aborted: the host must reset the context (it still gets the status word)
        if (K.fuse_path && is_leader_cta && E.me == 0) { __syncthreads(); result_to_host(K); }
        return;
    }

    // ---- rank 0's leader CTA reconstructs the path right here (one launch per query instead of two); with several ranks it
    // walks owner to owner through the peer mappings — every rank has passed the final fenced exchange, and no rank writes
    // parents again before rank 0's NEXT query hands over, which comes after this kernel
    KT_STAMP(5);
    if (K.fuse_path && is_leader_cta && E.me == 0) {
        __syncthreads();
        if (K.fuse_path == 2u) path_body(K);
        else result_to_host(K);
    }
    KT_STAMP(6);

    // ---- cleanup for the next query: clear the visited bits this query set.  Every vertex a
    // single-rank search marked sits in a queue, so a small search clears exactly those words;
    // otherwise (and always with several ranks, whose filter bits are not logged) sweep the bitmap.
    {
        const u32 gtid = E.cta * BFS_THREADS + tid, gthreads = E.cpr * BFS_THREADS;
        const u32 t0 = __ldcg(&C->acc.q_tail[0]), t1 = __ldcg(&C->acc.q_tail[1]);
        if (!MULTI && (u64)(t0 + t1) * 8ull < (u64)E.nw) {
            for (u32 i = gtid; i < t0; i += gthreads) E.vis[__ldcg(E.q[0] + i) >> 5] = 0ull;
            for (u32 i = gtid; i < t1; i += gthreads) E.vis[__ldcg(E.q[1] + i) >> 5] = 0ull;
        } else {
            ulonglong2* p = reinterpret_cast<ulonglong2*>(E.vis);
            const u32 n2 = E.nw / 2;
            for (u32 i = gtid; i < n2; i += gthreads) p[i] = make_ulonglong2(0ull, 0ull);
        }
    }
    KT_STAMP(7);
}

// ------------------------------------------------------------------------------------------
// path kernel: walk both parent chains from the meeting vertex, owner to owner
// (reference: v1/main-v1.cpp:86-97 walks on the host; v2/v4 re-run a whole serial BFS,
//  v2/second_try.cpp:137-162, v4/mpi_bas.cpp:141-164)
// ------------------------------------------------------------------------------------------
// whole CTA (any size >= 64 threads).  par0 / par1: parent arrays indexed by the vertex id (front kernel), or null:
// look the parent up at the vertex's owner
static __device__ __forceinline__ void path_body_impl(const KParams& K, const u32* par0, const u32* par1) {
    Ctrl* C = ctrl_of(K, K.rank_lo);
    Result& R = C->res;
    u32* hdr = reinterpret_cast<u32*>(K.arena[K.rank_lo] + K.L.off_path);   // [copy of Result | path]
    u32* path = hdr + PATH_HDR_WORDS;
    const u64 t0 = ptx::globaltimer_ns();
    const bool ok = R.status == 0 && R.length >= 0 && (u32)R.length + 1u <= K.max_path;
    const u32 L = ok ? (u32)R.length : 0u;
    const u32 ds = R.depth_s;
    if (ok && threadIdx.x == 0) {
        u32 cur = R.meet;
        for (u32 pos = ds;; --pos) {
            path[pos] = cur;
            if (pos == 0 || cur >= K.n) break;
            if (par0) cur = __ldcg(par0 + cur);
            else { const u32 o = owner_of(K, cur); cur = __ldcg(parent_of(K, o, 0) + (cur - o * K.B)); }
        }
    } else if (ok && threadIdx.x == 32) {
        u32 cur = R.meet;
        for (u32 pos = ds; pos < L && cur < K.n; ++pos) {
            if (par1) cur = __ldcg(par1 + cur);
            else { const u32 o = owner_of(K, cur); cur = __ldcg(parent_of(K, o, 1) + (cur - o * K.B)); }
            path[pos + 1] = cur;
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        R.path_len = ok ? L + 1u : 0u;
        const u64 dt = ptx::globaltimer_ns() - t0;
        R.path_ns = dt > 0xFFFFFFFFull ? 0xFFFFFFFFu : (u32)dt;
    }
    __syncthreads();
    static_assert(sizeof(Result) == PATH_HDR_WORDS * 4, "result header size");
    // [result | path] goes to the device buffer and, when the path is short, straight to the caller's pinned buffer
    // (posted PCIe writes).  Only the threads that wrote fence, once; the completion word follows behind that fence and
    // needs none of its own (the host polls it, and the end of the kernel flushes it anyway).
    const u32 plen = ok ? L + 1u : 0u;
    const u32 words = PATH_HDR_WORDS + ((K.h_out && plen <= K.h_out_words) ? plen : 0u);
    for (u32 i = threadIdx.x; i < words; i += blockDim.x) {
        const u32 w = i < PATH_HDR_WORDS ? reinterpret_cast<const u32*>(&R)[i] : __ldcg(hdr + i);
        if (i < PATH_HDR_WORDS) hdr[i] = w;
        if (K.h_out) { K.h_out[i] = w; if (!(K.flags & KF_HOST_SYNCS)) __threadfence_system(); }
    }
    if (K.h_out && K.h_seq) {
        __syncthreads();
        if (threadIdx.x == 0) *reinterpret_cast<volatile u32*>(K.h_seq) = K.seq;
    }
}

__device__ __noinline__ void path_body(const KParams& K) { path_body_impl(K, nullptr, nullptr); }
__device__ __noinline__ void path_body_local(const KParams& K, const u32* par0, const u32* par1) { path_body_impl(K, par0, par1); }

// no path wanted: only the 64-byte result goes to the host (whole CTA)
__device__ __noinline__ void result_to_host(const KParams& K) {
    const Result& R = ctrl_of(K, K.rank_lo)->res;
    if (!K.h_out) return;
    if (threadIdx.x < PATH_HDR_WORDS) { K.h_out[threadIdx.x] = reinterpret_cast<const u32*>(&R)[threadIdx.x]; if (!(K.flags & KF_HOST_SYNCS)) __threadfence_system(); }
    __syncthreads();
    if (threadIdx.x == 0 && K.h_seq) *reinterpret_cast<volatile u32*>(K.h_seq) = K.seq;
}

#include "front.cuh"

// ------------------------------------------------------------------------------------------
// launch helpers used by ctx.cu
// ------------------------------------------------------------------------------------------
// opt in to the kernel's dynamic shared memory; the attribute is per device, so contexts call this at creation
static int bfs_prepare(bool multi) {
    if (multi) BIBFS_CUDA_TRY(cudaFuncSetAttribute(k_bibfs<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
    else BIBFS_CUDA_TRY(cudaFuncSetAttribute(k_bibfs<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
    return 0;
}

int bfs_max_ctas_per_device(bool multi, int* out) {
    int dev = 0, sms = 0, per = 0;
    BIBFS_CUDA_TRY(cudaGetDevice(&dev));
    BIBFS_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    if (bfs_prepare(multi)) return 1;
    if (multi) BIBFS_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, k_bibfs<true>, BFS_THREADS, SMEM_BYTES));
    else BIBFS_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, k_bibfs<false>, BFS_THREADS, SMEM_BYTES));
    if (per < 1) return fail("search kernel does not fit on an SM");
    *out = sms * per;
    return 0;
}

int bfs_launch(const KParams& K, bool multi, cudaStream_t st) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(K.cpr * K.vranks);
    cfg.blockDim = dim3(BFS_THREADS);
    cfg.dynamicSmemBytes = SMEM_BYTES;
    cfg.stream = st;
    cudaLaunchAttribute attrs[2];
    attrs[0].id = cudaLaunchAttributeCooperative;
    attrs[0].val.cooperative = 1;
    cfg.numAttrs = 1;
    if (!multi && K.cluster > 1) {
        attrs[1].id = cudaLaunchAttributeClusterDimension;
        attrs[1].val.clusterDim.x = K.cluster; attrs[1].val.clusterDim.y = 1; attrs[1].val.clusterDim.z = 1;
        cfg.numAttrs = 2;
    }
    cfg.attrs = attrs;
    if (multi) BIBFS_CUDA_TRY(cudaLaunchKernelEx(&cfg, k_bibfs<true>, K));
    else BIBFS_CUDA_TRY(cudaLaunchKernelEx(&cfg, k_bibfs<false>, K));
    count_launch();
    return 0;
}

// how many CTAs can be co-resident when the single-rank kernel is launched in clusters of `cluster`
int bfs_max_ctas_clustered(u32 cluster, int* out) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(cluster * 64);
    cfg.blockDim = dim3(BFS_THREADS);
    cfg.dynamicSmemBytes = SMEM_BYTES;
    cudaLaunchAttribute attr;
    attr.id = cudaLaunchAttributeClusterDimension;
    attr.val.clusterDim.x = cluster; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = 1;
    cfg.attrs = &attr;
    cfg.numAttrs = 1;
    int n = 0;
    if (cluster > 8)      // beyond the portable limit: opt in (B200 schedules 16-CTA clusters inside one GPC)
        BIBFS_CUDA_TRY(cudaFuncSetAttribute(k_bibfs<false>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    BIBFS_CUDA_TRY(cudaOccupancyMaxActiveClusters(&n, k_bibfs<false>, &cfg));
    *out = n * (int)cluster;
    return 0;
}

#ifdef BIBFS_FINE_TRACE
extern "C" int bibfs_debug_read(uint64_t* out, uint32_t* levels) {
    cudaMemcpyFromSymbol(out, g_dbg, sizeof(u64) * 256 * 8);
    cudaMemcpyFromSymbol(out + 256 * 8, g_dbg2, sizeof(u64) * 256 * 8);
    cudaMemcpyFromSymbol(levels, g_dbg_level, 4);
    u32 z = 0;
    cudaMemcpyToSymbol(g_dbg_level, &z, 4);
    return 0;
}
extern "C" int bibfs_debug_read_kt(uint64_t* out32) {
    cudaMemcpyFromSymbol(out32, g_kt, sizeof(u64) * 32);
    return 0;
}
#endif

constexpr unsigned FRONT_WARM_CTAS = 48;

template <bool MULTI>
static int front_cluster_fit(u32 cluster, int* fit) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(cluster);
    cfg.blockDim = dim3(BFS_THREADS);
    cfg.dynamicSmemBytes = sizeof(FrontSmem);
    cudaLaunchAttribute attr;
    attr.id = cudaLaunchAttributeClusterDimension;
    attr.val.clusterDim.x = cluster; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = 1;
    cfg.attrs = &attr;
    cfg.numAttrs = 1;
    *fit = 0;
    return cudaOccupancyMaxActiveClusters(fit, k_front<MULTI>, &cfg) == cudaSuccess ? 0 : 1;
}

// opt in to the front kernel's shared memory and pick the size of its cluster: 16 (beyond the portable 8; B200 takes it),
// else 8, 4, 2, 1.  BIBFS_FRONT_CLUSTER=n overrides.
int front_prepare(bool multi, u32* cluster_out) {
    if (multi) BIBFS_CUDA_TRY(cudaFuncSetAttribute(k_front<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FrontSmem)));
    else BIBFS_CUDA_TRY(cudaFuncSetAttribute(k_front<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FrontSmem)));
    if (multi) cudaFuncSetAttribute(k_front<true>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    else cudaFuncSetAttribute(k_front<false>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    cudaGetLastError();
    const char* want = getenv("BIBFS_FRONT_CLUSTER");
    u32 cl = want ? (u32)atoi(want) : 16u;
    if (cl < 1 || cl > 16 || (cl & (cl - 1))) cl = 16;
    for (; cl > 1; cl /= 2) {
        int fit = 0;
        const int rc = multi ? front_cluster_fit<true>(cl, &fit) : front_cluster_fit<false>(cl, &fit);
        cudaGetLastError();
        if (rc == 0 && fit >= 1) break;
    }
    *cluster_out = cl;
    return 0;
}

int front_launch(const KParams& K, bool multi, u32 cluster, cudaStream_t st) {
    // the first cluster runs the query; the clusters behind it only warm L2 (graphs that fit it with room to spare)
    const unsigned warm = K.warm_lines ? (FRONT_WARM_CTAS + cluster - 1) / cluster * cluster : 0u;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(cluster + warm);
    cfg.blockDim = dim3(BFS_THREADS);
    cfg.dynamicSmemBytes = sizeof(FrontSmem);
    cfg.stream = st;
    cudaLaunchAttribute attr;
    attr.id = cudaLaunchAttributeClusterDimension;
    attr.val.clusterDim.x = cluster; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = 1;
    cfg.attrs = &attr;
    cfg.numAttrs = cluster > 1 ? 1 : 0;
    if (multi) BIBFS_CUDA_TRY(cudaLaunchKernelEx(&cfg, k_front<true>, K));
    else BIBFS_CUDA_TRY(cudaLaunchKernelEx(&cfg, k_front<false>, K));
    count_launch();
    return 0;
}

// one launch for a batch of queries: `lanes` clusters of `cluster` CTAs, lane i works with d_lane_K[i]
int front_batch_prepare(bool multi) {
    if (multi) BIBFS_CUDA_TRY(cudaFuncSetAttribute(k_front_batch<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FrontSmem)));
    else BIBFS_CUDA_TRY(cudaFuncSetAttribute(k_front_batch<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FrontSmem)));
    if (multi) cudaFuncSetAttribute(k_front_batch<true>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    else cudaFuncSetAttribute(k_front_batch<false>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    cudaGetLastError();
    return 0;
}

int front_batch_launch(const KParams* d_lane_K, unsigned long long* d_queries, Result* h_res, u32 nq, u32 lanes, bool multi, u32 cluster, cudaStream_t st) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(lanes * cluster);
    cfg.blockDim = dim3(BFS_THREADS);
    cfg.dynamicSmemBytes = sizeof(FrontSmem);
    cfg.stream = st;
    cudaLaunchAttribute attr;
    attr.id = cudaLaunchAttributeClusterDimension;
    attr.val.clusterDim.x = cluster; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = 1;
    cfg.attrs = &attr;
    cfg.numAttrs = cluster > 1 ? 1 : 0;
    if (multi) BIBFS_CUDA_TRY(cudaLaunchKernelEx(&cfg, k_front_batch<true>, d_lane_K, d_queries, h_res, nq, (Mailbox*)nullptr, 0u, 0u));
    else BIBFS_CUDA_TRY(cudaLaunchKernelEx(&cfg, k_front_batch<false>, d_lane_K, d_queries, h_res, nq, (Mailbox*)nullptr, 0u, 0u));
    count_launch();
    return 0;
}

// the resident front kernel: one lane that takes its queries from the mailbox until it is told to leave or idles out
int front_resident_launch(const KParams* d_lane_K, Mailbox* mb, u32 start_seq, u32 idle_us, u32 cluster, cudaStream_t st) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(cluster);
    cfg.blockDim = dim3(BFS_THREADS);
    cfg.dynamicSmemBytes = sizeof(FrontSmem);
    cfg.stream = st;
    cudaLaunchAttribute attr;
    attr.id = cudaLaunchAttributeClusterDimension;
    attr.val.clusterDim.x = cluster; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = 1;
    cfg.attrs = &attr;
    cfg.numAttrs = cluster > 1 ? 1 : 0;
    BIBFS_CUDA_TRY(cudaLaunchKernelEx(&cfg, k_front_batch<false>, d_lane_K, (unsigned long long*)nullptr, (Result*)nullptr, 0u, mb, start_seq, idle_us));
    count_launch();
    return 0;
}

int wait_launch(const KParams& K, cudaStream_t st) {
    k_wait<<<1, 32, 0, st>>>(K);
    BIBFS_KERNEL_CHECK();
    return 0;
}

}  // namespace bibfs
