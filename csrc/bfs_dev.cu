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
//        -> expand_td: frontier-proportional work, warp-cooperative neighbour gathering,
//           expand_bu: bottom-up sweep over unvisited owned vertices against a frontier bitmap
//   check_intersect(_kernel) + host O(N) scan  (v3:45-62, comp.cu:40-47, mpi_bas.cpp:81-87)
//        -> fused: the visited bitmap word carries both sides' bits, so the atomic OR that claims
//           a vertex also returns "was it reached from the other side" (discovery-time test)
//   MPI_Allreduce(BOR) of the next frontier  (mpi_bas.cpp:107; second_try.cpp:82-104)
//        -> a remote atom.or on the OWNER's bitmap word + remote stores of the parent and of the
//           vertex into the owner's per-sender queue segment, straight over NVLink (peer pointers)
//   MPI_Allreduce(LOR/SUM) for found / changed / frontier sizes  (mpi_bas.cpp:85,124; second_try:115-124)
//        -> one 16-byte slot store per peer + a release flag; every rank sums the slots itself
//   the blocking collectives' implicit barrier
//        -> device-side barrier: local arrive counter + generation word, cross-rank epoch flags
//           written with st.release.sys and polled with ld.acquire.sys.  A watchdog turns a lost
//           rank into an error instead of a hang.
//
// Scheduling: the side with the smaller (global) frontier is expanded, ties to the source side
// (v1/main-v1.cpp:51, mpi_bas.cpp:90-92).  Small local frontiers are expanded by CTA 0 alone with
// __syncthreads between levels ("solo"); the other CTAs sleep on the generation word.  Large
// ones wake the whole grid; very large ones switch to bottom-up (Beamer et al., SC'12).
//
// Exactness: levels are always completed; the first level in which any discovered vertex is
// already marked by the other side ends the search and length = depth_s + depth_t (all hits of
// that level have the same length — SURVEY 7.4 item 6, the argument that makes v1 correct).
//
// With P "virtual ranks" hosted by one launch (blockIdx.x / cpr = rank) the very same protocol
// runs on a single GPU, which is how the multi-rank path is tested without 8 GPUs.
#include "bfs.h"
#include "ptx.cuh"

namespace bibfs {

namespace {

constexpr u32 FULL = 0xffffffffu;
constexpr u32 BIG_ROW = 64;          // rows at least this long are strip-mined by the whole warp

struct Env {
    u32 me, lo, P, B, n, nw;
    u32 cta, cpr;
    Ctrl* C;
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
__device__ __forceinline__ u32* rq_of(const KParams& K, u32 r, u32 side, u32 sender) {
    return reinterpret_cast<u32*>(K.arena[r] + K.L.off_rq) + ((size_t)side * K.P + sender) * K.B;
}
__device__ __forceinline__ bibfs_level_trace* trace_of(const KParams& K, u32 r) {
    return reinterpret_cast<bibfs_level_trace*>(K.arena[r] + K.L.off_trace);
}

__device__ __forceinline__ u32 lanemask_lt() {
    u32 m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

// ------------------------------------------------------------------------------------------
// waiting with a watchdog
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ bool expired(Ctrl* C, u64 deadline) {
    return ptx::globaltimer_ns() > deadline || ptx::ld_relaxed_gpu(&C->abort_flag) != 0;
}

__device__ bool wait_gen(Ctrl* C, u64 want, u64 deadline) {
    u32 spins = 0;
    while (ptx::ld_acquire_gpu(&C->bar_gen) < want) {
        if (++spins > 32) {
            ptx::nanosleep(40);
            if ((spins & 255u) == 0 && expired(C, deadline)) return false;
        }
    }
    return true;
}

__device__ bool wait_count(Ctrl* C, u32 want, u64 deadline) {
    u32 spins = 0;
    while (ptx::ld_acquire_gpu(&C->bar_count) < want) {
        if (++spins > 64 && (spins & 255u) == 0 && expired(C, deadline)) return false;
    }
    return true;
}

__device__ bool wait_arrive(Ctrl* C, u32 p, u64 want, u64 deadline) {
    u32 spins = 0;
    while (ptx::ld_acquire_sys(&C->arrive[p]) < want) {
        if (++spins > 64 && (spins & 255u) == 0 && expired(C, deadline)) return false;
    }
    return true;
}

// ------------------------------------------------------------------------------------------
// leader side: decisions
// ------------------------------------------------------------------------------------------
__device__ void decide_next(const KParams& K, LevelState& G) {
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

__device__ void write_result(const KParams& K, Ctrl* C, const LevelState& G) {
    Result& R = C->res;
    R.length = G.length;
    R.meet = G.meet;
    R.depth_s = G.depth[0];
    R.depth_t = G.depth[1];
    R.levels = G.levels;
    R.levels_bu = G.levels_bu;
    R.status = (i32)G.status;
    R.path_len = 0;
    R.edges_scanned = C->edges_total;
    R.t_end_ns = ptx::globaltimer_ns();
    C->xepoch = G.xepoch;
}

// publish a new generation: everything the leader warp wrote becomes visible to the other CTAs
__device__ __forceinline__ void bump_gen(const KParams& K, Ctrl* C, LevelState& G) {
    G.wake += 1;
    __threadfence();
    ptx::st_release_gpu(&C->bar_gen, ((u64)K.seq << 32) | G.wake);
}

// Cross-rank handshake executed by the leader warp (lanes 0..P-1 each talk to one peer).
// Writes `mine` into slot[parity][me] of every peer, raises arrive[me] there, waits until every
// peer has done the same here and returns the peer's slot in `got`.  Returns false on timeout.
template <bool MULTI>
__device__ bool cross_exchange(const KParams& K, const Env& E, LevelState& G, const Slot mine_for_lane, Slot& got, u64 deadline) {
    const u32 lane = threadIdx.x & 31;
    got = mine_for_lane;
    if (!MULTI) return true;
    const u64 epoch = G.xepoch + 1;
    const u32 par = (u32)(epoch & 1);
    bool ok = true;
    if (lane < K.P) {
        Ctrl* PC = ctrl_of(K, lane);
        ptx::st_peer_v4(reinterpret_cast<u32*>(&PC->slot[par][E.me]), mine_for_lane.count, mine_for_lane.meet,
                        mine_for_lane.rq_tail, mine_for_lane.edges);
        ptx::fence_sys();
        ptx::st_release_sys(&PC->arrive[E.me], epoch);
        ok = wait_arrive(E.C, lane, epoch, deadline);
        if (ok) {
            const volatile Slot* s = &E.C->slot[par][lane];
            got.count = s->count; got.meet = s->meet; got.rq_tail = s->rq_tail; got.edges = s->edges;
        }
    } else {
        got.count = 0; got.meet = NO_V; got.rq_tail = 0; got.edges = 0;
    }
    ok = __all_sync(FULL, ok);
    if (threadIdx.x == 0) G.xepoch = epoch;
    __syncwarp();
    return ok;
}

enum SyncKind : int { SYNC_LEVEL = 0, SYNC_LOCAL = 1, SYNC_CROSS = 2 };

// Work of the leader warp (warp 0 of CTA 0 of the rank) at a synchronisation point, after every
// participating local CTA has arrived.  G is the rank's LevelState in global memory.
template <bool MULTI>
__device__ bool leader_sync(const KParams& K, const Env& E, int kind, bool solo, u32 level_mode, u64 deadline) {
    Ctrl* C = E.C;
    LevelState& G = C->st;
    const u32 lane = threadIdx.x & 31;
    if (lane == 0) ptx::st_relaxed_gpu(&C->bar_count, 0u);
    if (kind == SYNC_LOCAL) {
        if (lane == 0 && !solo) bump_gen(K, C, G);
        __syncwarp();
        return true;
    }
    const u32 side = G.side;
    Slot mine;
    mine.count = ptx::ld_relaxed_gpu(&C->new_count);
    mine.meet = ptx::ld_relaxed_gpu(&C->meet_local);
    mine.rq_tail = (MULTI && lane < K.P) ? ptx::ld_relaxed_gpu(&C->sent[side][lane]) : 0u;
    const u64 my_edges = ptx::ld_relaxed_gpu(reinterpret_cast<const u64*>(&C->edges_acc));
    mine.edges = (u32)my_edges;
    Slot got;
    const bool ok = cross_exchange<MULTI>(K, E, G, mine, got, deadline);
    if (!MULTI && lane != 0) { got.count = 0; got.meet = NO_V; got.edges = 0; }   // single rank: lane 0 holds the totals
    if (!ok) {
        if (lane == 0) {
            atomicExch(&C->abort_flag, 1u);
            G.status = 1; G.length = -1; G.phase = PH_DONE;
            write_result(K, C, G);
            bump_gen(K, C, G);
        }
        __syncwarp();
        return false;
    }
    if (kind == SYNC_CROSS) {
        if (lane == 0 && !solo) bump_gen(K, C, G);
        __syncwarp();
        return true;
    }
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
    __syncwarp();
    if (lane == 0) {
        const u32 expanded = G.nf[side];
        G.nf[side] = total;
        G.depth[side] += 1;
        G.levels += 1;
        if (level_mode == PH_BU) { G.levels_bu += 1; G.fbm_cur[side] ^= 1u; G.fbm_valid[side] = 1; }
        else G.fbm_valid[side] = 0;
        G.q_lo[side] = G.q_hi[side];
        G.q_hi[side] = ptx::ld_relaxed_gpu(&C->q_tail[side]);
        C->edges_total += etot;   // all ranks' adjacency entries of this level
        C->res.vertices_visited += total;
        if ((K.flags & BIBFS_F_TRACE) && C->trace_count < TRACE_CAP) {
            bibfs_level_trace& T = trace_of(K, E.me)[C->trace_count++];
            T.side = side; T.mode = level_mode; T.frontier = expanded; T.discovered = total;
            T.t_ns = ptx::globaltimer_ns(); T.edges = my_edges;
        }
        // reset the per-level accumulators
        C->new_count = 0;
        C->meet_local = NO_V;
        C->edges_acc = 0;
        if (meetmin != NO_V) {
            G.length = (i32)(G.depth[0] + G.depth[1]);
            G.meet = meetmin;
            G.phase = PH_DONE;
        } else {
            decide_next(K, G);
        }
        if (G.phase == PH_DONE) write_result(K, C, G);
        if (!solo || G.phase != PH_TD_SOLO) bump_gen(K, C, G);
        else __threadfence();
    }
    __syncwarp();
    return true;
}

// Query start, leader warp of each rank.
template <bool MULTI>
__device__ bool leader_init(const KParams& K, const Env& E, u64 deadline) {
    Ctrl* C = E.C;
    LevelState& G = C->st;
    const u32 lane = threadIdx.x & 31;
    const u32 src = K.query_in_ctrl ? C->q_src : K.src, dst = K.query_in_ctrl ? C->q_dst : K.dst;
    if (lane == 0) {
        C->bar_count = 0;
        C->q_tail[0] = 0; C->q_tail[1] = 0;
        for (u32 p = 0; p < MAXR; ++p) { C->sent[0][p] = 0; C->sent[1][p] = 0; }
        C->new_count = 0; C->meet_local = NO_V; C->edges_acc = 0; C->edges_total = 0;
        C->trace_count = 0;
        C->res.vertices_visited = 2;
        C->res.t_begin_ns = ptx::globaltimer_ns();
        G.phase = PH_TD_SOLO; G.side = 0; G.levels = 0; G.levels_bu = 0;
        G.depth[0] = G.depth[1] = 0;
        G.nf[0] = G.nf[1] = 1;
        for (u32 s = 0; s < 2; ++s) {
            G.q_lo[s] = G.q_hi[s] = 0;
            G.fbm_cur[s] = 0; G.fbm_valid[s] = 0;
            for (u32 p = 0; p < MAXR; ++p) { G.rq_lo[s][p] = 0; G.rq_hi[s][p] = 0; }
        }
        G.length = -1; G.meet = NO_V; G.status = 0; G.wake = 0; G.local_front = 0;
        G.xepoch = C->xepoch;
        const bool bad = src >= K.n || dst >= K.n;
        if (bad) { G.status = 2; G.phase = PH_DONE; }
        else if (src == dst) { G.length = 0; G.meet = src; G.phase = PH_DONE; C->res.vertices_visited = 1; }
        else {
            const u32 hi = min(K.n, E.lo + K.B);
            if (src >= E.lo && src < hi) {
                atomicOr(reinterpret_cast<unsigned long long*>(&E.vis[src >> 5]), 1ull << (src & 31));
                E.q[0][0] = src; E.parent[0][src - E.lo] = NO_V;
                C->q_tail[0] = 1; G.q_hi[0] = 1;
            }
            if (dst >= E.lo && dst < hi) {
                atomicOr(reinterpret_cast<unsigned long long*>(&E.vis[dst >> 5]), 1ull << ((dst & 31) + 32));
                E.q[1][0] = dst; E.parent[1][dst - E.lo] = NO_V;
                C->q_tail[1] = 1; G.q_hi[1] = 1;
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
        ok = cross_exchange<MULTI>(K, E, G, z, got, deadline);
    }
    if (lane == 0) {
        if (!ok) { atomicExch(&C->abort_flag, 1u); G.status = 1; G.phase = PH_DONE; }
        if (G.phase != PH_DONE) decide_next(K, G);
        if (G.phase == PH_DONE) write_result(K, C, G);
        if (G.phase != PH_TD_SOLO) bump_gen(K, C, G);
        else __threadfence();
    }
    __syncwarp();
    return ok;
}

// ------------------------------------------------------------------------------------------
// frontier access: flat index -> vertex, across the local window and the per-sender windows
// ------------------------------------------------------------------------------------------
template <bool MULTI>
__device__ __forceinline__ u32 fetch_frontier(const KParams& K, const Env& E, const LevelState& st, u32 side, u32 i) {
    const u32 nq = st.q_hi[side] - st.q_lo[side];
    if (!MULTI || i < nq) return __ldcg(E.q[side] + st.q_lo[side] + i);
    i -= nq;
    u32 v = 0;
#pragma unroll 1
    for (u32 p = 0; p < K.P; ++p) {
        const u32 lo = st.rq_lo[side][p], c = st.rq_hi[side][p] - lo;
        if (i < c) { v = __ldcg(rq_of(K, E.me, side, p) + lo + i); break; }
        i -= c;
    }
    return v;
}

// ------------------------------------------------------------------------------------------
// discovery: claim v for `side`, record its parent, enqueue it at its owner, test the meet
// ------------------------------------------------------------------------------------------
template <bool MULTI>
__device__ __forceinline__ void visit(const KParams& K, const Env& E, u32 side, bool valid, u32 u, u32 v, u32& cnt) {
    const u32 lane = threadIdx.x & 31;
    bool isnew = false, meet = false;
    u32 dest = E.me;
    if (valid) {
        const u32 w = v >> 5;
        const u64 bit = 1ull << ((v & 31u) + 32u * side);
        const u64 obit = 1ull << ((v & 31u) + 32u * (1u - side));
        const u64 cur = E.vis[w];                       // filter only; may be stale
        if (!(cur & bit)) {
            u64 old = ptx::atom_or_gpu(&E.vis[w], bit);
            if (!(old & bit)) {
                isnew = true;
                if (MULTI) {
                    dest = v / K.B;
                    if (dest != E.me) {
                        // first time THIS rank sees v: ask the owner.  One round trip returns both
                        // "new for this side" and "already reached by the other side".
                        old = ptx::atom_or_sys(vis_of(K, dest) + w, bit);
                        isnew = !(old & bit);
                    }
                }
                meet = isnew && (old & obit);
            }
        }
    }
    if (!MULTI) {
        const u32 m = __ballot_sync(FULL, isnew);
        if (m) {
            const u32 leader = __ffs(m) - 1;
            u32 base = 0;
            if (lane == leader) base = atomicAdd(&E.C->q_tail[side], (u32)__popc(m));
            base = __shfl_sync(FULL, base, leader);
            if (isnew) {
                E.q[side][base + __popc(m & lanemask_lt())] = v;
                E.parent[side][v - E.lo] = u;
            }
        }
    } else {
        const u32 key = isnew ? dest : 0xFFu;
        const u32 grp = __match_any_sync(FULL, key);
        if (isnew) {
            const u32 leader = __ffs(grp) - 1;
            const u32 rank_in = __popc(grp & lanemask_lt());
            u32 base = 0;
            if (dest == E.me) {
                if (lane == leader) base = atomicAdd(&E.C->q_tail[side], (u32)__popc(grp));
                base = __shfl_sync(grp, base, leader);
                E.q[side][base + rank_in] = v;
                E.parent[side][v - E.lo] = u;
            } else {
                if (lane == leader) base = atomicAdd(&E.C->sent[side][dest], (u32)__popc(grp));
                base = __shfl_sync(grp, base, leader);
                ptx::st_peer(rq_of(K, dest, side, E.me) + base + rank_in, v);
                ptx::st_peer(parent_of(K, dest, side) + (v - dest * K.B), u);
            }
        }
    }
    cnt += isnew ? 1u : 0u;
    if (meet) atomicMin(&E.C->meet_local, v);
}

// ------------------------------------------------------------------------------------------
// top-down level: warps take 32 frontier vertices at a time
// ------------------------------------------------------------------------------------------
template <bool MULTI>
__device__ void expand_td(const KParams& K, const Env& E, const LevelState& st, u32 warp_gid, u32 nwarps) {
    const u32 lane = threadIdx.x & 31;
    const u32 side = st.side;
    const u32 total = st.local_front;
    u32 cnt = 0;
    u64 edges = 0;
    for (u32 base = warp_gid * 32u; base < total; base += nwarps * 32u) {
        const u32 i = base + lane;
        u32 u = 0;
        u64 r0 = 0, r1 = 0;
        if (i < total) {
            u = fetch_frontier<MULTI>(K, E, st, side, i);
            const u32 ul = u - E.lo;
            r0 = __ldg(E.row_ptr + ul);
            r1 = __ldg(E.row_ptr + ul + 1);
        }
        const u64 deg = r1 - r0;
        edges += deg;
        // long rows: the whole warp walks one row, coalesced
        u32 big = __ballot_sync(FULL, deg >= BIG_ROW);
        while (big) {
            const u32 srcl = __ffs(big) - 1;
            big &= big - 1;
            const u32 uu = __shfl_sync(FULL, u, srcl);
            const u64 b = __shfl_sync(FULL, r0, srcl), e = __shfl_sync(FULL, r1, srcl);
            for (u64 p = b; p < e; p += 32) {
                const u64 idx = p + lane;
                const bool ok = idx < e;
                const u32 v = ok ? __ldg(E.col + idx) : 0u;
                visit<MULTI>(K, E, side, ok, uu, v, cnt);
            }
        }
        // short rows: flatten the 32 rows into one edge index space (scan + search by shuffles)
        const u32 sdeg = deg >= BIG_ROW ? 0u : (u32)deg;
        u32 incl = sdeg;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const u32 t = __shfl_up_sync(FULL, incl, d);
            if ((int)lane >= d) incl += t;
        }
        const u32 excl = incl - sdeg;
        const u32 tot = __shfl_sync(FULL, incl, 31);
        for (u32 t = 0; t < tot; t += 32) {
            const u32 idx = t + lane;
            const bool ok = idx < tot;
            u32 j = 0;
#pragma unroll
            for (int step = 16; step > 0; step >>= 1) {
                const u32 cand = j + step;
                const u32 ec = __shfl_sync(FULL, excl, cand & 31u);
                if (cand < 32u && ec <= idx) j = cand;
            }
            const u32 uu = __shfl_sync(FULL, u, j);
            const u64 rb = __shfl_sync(FULL, r0, j);
            const u32 eb = __shfl_sync(FULL, excl, j);
            const u32 v = ok ? __ldg(E.col + rb + (idx - eb)) : 0u;
            visit<MULTI>(K, E, side, ok, uu, v, cnt);
        }
    }
    // flush this warp's counters
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        cnt += __shfl_xor_sync(FULL, cnt, d);
        edges += __shfl_xor_sync(FULL, edges, d);
    }
    if (lane == 0) {
        if (cnt) atomicAdd(&E.C->new_count, cnt);
        if (edges) atomicAdd(&E.C->edges_acc, (unsigned long long)edges);
    }
}

// ------------------------------------------------------------------------------------------
// bottom-up level: every unvisited owned vertex looks for a parent in the frontier bitmap
// ------------------------------------------------------------------------------------------
template <bool MULTI>
__device__ void expand_bu(const KParams& K, const Env& E, const LevelState& st, u32 warp_gid, u32 nwarps) {
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
        if (unvis) {
            const u64 r0 = __ldg(E.row_ptr + (v - E.lo)), r1 = __ldg(E.row_ptr + (v - E.lo) + 1);
            for (u64 e = r0; e < r1; ++e) {
                const u32 u = __ldg(E.col + e);
                ++edges;
                if ((fcur[u >> 5] >> (u & 31u)) & 1u) { found = true; par = u; break; }
            }
        }
        const u32 newmask = __ballot_sync(FULL, found);
        if (newmask) {
            u32 base = 0;
            if (lane == 0) {
                E.vis[wg] = word | ((u64)newmask << (32u * side));
                const u32 meetmask = newmask & (u32)(word >> (32u * (1u - side)));
                if (meetmask) atomicMin(&E.C->meet_local, E.lo + wl * 32u + (u32)(__ffs(meetmask) - 1));
                base = atomicAdd(&E.C->q_tail[side], (u32)__popc(newmask));
            }
            base = __shfl_sync(FULL, base, 0);
            if (found) {
                E.q[side][base + __popc(newmask & lanemask_lt())] = v;
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
        if (cnt) atomicAdd(&E.C->new_count, cnt);
        if (edges) atomicAdd(&E.C->edges_acc, (unsigned long long)edges);
    }
}

// copy this rank's slice of a frontier bitmap into every peer's replica (8-byte stores over NVLink)
__device__ void push_fbm_slice(const KParams& K, const Env& E, u32 side, u32 buf, u32 gtid, u32 gthreads) {
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

}  // namespace

// ------------------------------------------------------------------------------------------
// the kernel
// ------------------------------------------------------------------------------------------
template <bool MULTI>
__global__ void __launch_bounds__(BFS_THREADS, 1) k_bibfs(const __grid_constant__ KParams K) {
    __shared__ LevelState st;
    __shared__ u32 s_fail;
    const u32 tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr u32 WPC = BFS_THREADS / 32;
    const u32 vr = MULTI ? blockIdx.x / K.cpr : 0u;
    Env E;
    E.cta = blockIdx.x - vr * K.cpr;
    E.cpr = K.cpr;
    E.me = K.rank_lo + vr;
    E.lo = E.me * K.B;
    E.P = K.P; E.B = K.B; E.n = K.n; E.nw = K.L.nw;
    E.C = ctrl_of(K, E.me);
    E.vis = vis_of(K, E.me);
    E.parent[0] = parent_of(K, E.me, 0); E.parent[1] = parent_of(K, E.me, 1);
    E.q[0] = q_of(K, E.me, 0); E.q[1] = q_of(K, E.me, 1);
    E.row_ptr = K.row_ptr[E.me];
    E.col = K.col[E.me];
    Ctrl* C = E.C;
    const u64 deadline = ptx::globaltimer_ns() + (u64)K.watchdog_ms * 1000000ull;
    const u64 genbase = (u64)K.seq << 32;
    if (tid == 0) s_fail = 0;
    __syncthreads();

    // ---- query start
    if (E.cta == 0) {
        if (warp == 0) {
            const bool ok = leader_init<MULTI>(K, E, deadline);
            if (!ok && lane == 0) s_fail = 1;
        }
    } else if (tid == 0) {
        if (!wait_gen(C, genbase | 1u, deadline)) s_fail = 1;
    }
    __syncthreads();

    // every CTA keeps a shared-memory copy of the level state
    auto load_state = [&]() {
        const u32* g = reinterpret_cast<const u32*>(&C->st);
        u32* s = reinterpret_cast<u32*>(&st);
        for (u32 i = tid; i < sizeof(LevelState) / 4; i += BFS_THREADS) s[i] = __ldcg(g + i);
        __syncthreads();
    };

    // synchronisation point shared by all phases.  Returns false if the query was aborted.
    auto sync_point = [&](int kind, bool solo, u32 mode) -> bool {
        __syncthreads();
        if (solo) {
            if (warp == 0 && !leader_sync<MULTI>(K, E, kind, true, mode, deadline) && lane == 0) s_fail = 1;
        } else if (E.cta != 0) {
            if (tid == 0) {
                if (MULTI) ptx::fence_sys();
                ptx::red_release_gpu_add(&C->bar_count, 1u);
                if (!wait_gen(C, genbase | (u64)(st.wake + 1u), deadline)) s_fail = 1;
            }
        } else if (warp == 0) {
            bool ok = true;
            if (lane == 0) ok = wait_count(C, E.cpr - 1u, deadline);
            ok = __shfl_sync(FULL, ok, 0);
            if (ok) ok = leader_sync<MULTI>(K, E, kind, false, mode, deadline);
            else if (lane == 0) { atomicExch(&C->abort_flag, 1u); C->st.status = 1; C->st.phase = PH_DONE; write_result(K, C, C->st); }
            if (!ok && lane == 0) s_fail = 1;
        }
        __syncthreads();
        if (s_fail) return false;
        load_state();
        return true;
    };

    bool alive = (s_fail == 0);
    if (alive) load_state();

    while (alive) {
        const u32 phase = st.phase;
        if (phase == PH_DONE) break;
        if (phase == PH_TD_SOLO) {
            if (E.cta != 0) {   // parked until the leader wakes the grid
                if (tid == 0 && !wait_gen(C, genbase | (u64)(st.wake + 1u), deadline)) s_fail = 1;
                __syncthreads();
                if (s_fail) { alive = false; break; }
                load_state();
                continue;
            }
            expand_td<MULTI>(K, E, st, warp, WPC);
            alive = sync_point(SYNC_LEVEL, true, PH_TD_SOLO);
        } else if (phase == PH_TD_GRID) {
            expand_td<MULTI>(K, E, st, E.cta * WPC + warp, E.cpr * WPC);
            alive = sync_point(SYNC_LEVEL, false, PH_TD_GRID);
        } else {  // PH_BU
            const u32 side = st.side;
            const u32 gtid = E.cta * BFS_THREADS + tid, gthreads = E.cpr * BFS_THREADS;
            if (!st.fbm_valid[side]) {
                // queue windows -> bitmap slice, then replicate the slice
                u32* f = fbm_of(K, E.me, side, st.fbm_cur[side]);
                const u32 w0 = E.lo / 32, words = K.B / 32;
                for (u32 i = gtid; i < words; i += gthreads) f[w0 + i] = 0;
                alive = sync_point(SYNC_LOCAL, false, PH_BU);
                if (!alive) break;
                for (u32 i = gtid; i < st.local_front; i += gthreads) {
                    const u32 u = fetch_frontier<MULTI>(K, E, st, side, i);
                    atomicOr(&f[u >> 5], 1u << (u & 31u));
                }
                if (MULTI) {
                    alive = sync_point(SYNC_LOCAL, false, PH_BU);
                    if (!alive) break;
                    push_fbm_slice(K, E, side, st.fbm_cur[side], gtid, gthreads);
                    alive = sync_point(SYNC_CROSS, false, PH_BU);
                } else {
                    alive = sync_point(SYNC_LOCAL, false, PH_BU);
                }
                if (!alive) break;
            }
            expand_bu<MULTI>(K, E, st, E.cta * WPC + warp, E.cpr * WPC);
            if (MULTI) {
                alive = sync_point(SYNC_LOCAL, false, PH_BU);
                if (!alive) break;
                push_fbm_slice(K, E, side, st.fbm_cur[side] ^ 1u, gtid, gthreads);
            }
            alive = sync_point(SYNC_LEVEL, false, PH_BU);
        }
    }
    if (!alive) return;   // aborted: the host must reset the context

    // ---- cleanup for the next query: clear the visited bits this query set.  Every vertex a
    // single-rank search marked sits in a queue, so a small search clears exactly those words;
    // otherwise (and always with several ranks, whose filter bits are not logged) sweep the bitmap.
    {
        const u32 gtid = E.cta * BFS_THREADS + tid, gthreads = E.cpr * BFS_THREADS;
        const u32 t0 = __ldcg(&C->q_tail[0]), t1 = __ldcg(&C->q_tail[1]);
        if (!MULTI && (u64)(t0 + t1) * 8ull < (u64)E.nw) {
            for (u32 i = gtid; i < t0; i += gthreads) E.vis[__ldcg(E.q[0] + i) >> 5] = 0ull;
            for (u32 i = gtid; i < t1; i += gthreads) E.vis[__ldcg(E.q[1] + i) >> 5] = 0ull;
        } else {
            ulonglong2* p = reinterpret_cast<ulonglong2*>(E.vis);
            const u32 n2 = E.nw / 2;
            for (u32 i = gtid; i < n2; i += gthreads) p[i] = make_ulonglong2(0ull, 0ull);
        }
    }
}

// ------------------------------------------------------------------------------------------
// path kernel: walk both parent chains from the meeting vertex, owner to owner
// (reference: v1/main-v1.cpp:86-97 walks on the host; v2/v4 re-run a whole serial BFS,
//  v2/second_try.cpp:137-162, v4/mpi_bas.cpp:141-164)
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(64) k_path(const __grid_constant__ KParams K) {
    Ctrl* C = ctrl_of(K, K.rank_lo);
    Result& R = C->res;
    u32* path = reinterpret_cast<u32*>(K.arena[K.rank_lo] + K.L.off_path);
    if (R.status != 0 || R.length < 0) return;
    const u32 L = (u32)R.length;
    if (L + 1u > K.max_path) return;
    const u32 ds = R.depth_s;
    if (threadIdx.x == 0) {
        u32 cur = R.meet;
        for (u32 pos = ds;; --pos) {
            path[pos] = cur;
            if (pos == 0 || cur >= K.n) break;
            const u32 o = cur / K.B;
            cur = __ldcg(parent_of(K, o, 0) + (cur - o * K.B));
        }
    } else if (threadIdx.x == 32) {
        u32 cur = R.meet;
        for (u32 pos = ds; pos < L && cur < K.n; ++pos) {
            const u32 o = cur / K.B;
            cur = __ldcg(parent_of(K, o, 1) + (cur - o * K.B));
            path[pos + 1] = cur;
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) R.path_len = L + 1u;
}

// ------------------------------------------------------------------------------------------
// launch helpers used by ctx.cu
// ------------------------------------------------------------------------------------------
int bfs_max_ctas_per_device(bool multi, int* out) {
    int dev = 0, sms = 0, per = 0;
    BIBFS_CUDA_TRY(cudaGetDevice(&dev));
    BIBFS_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    if (multi) BIBFS_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, k_bibfs<true>, BFS_THREADS, 0));
    else BIBFS_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, k_bibfs<false>, BFS_THREADS, 0));
    if (per < 1) return fail("search kernel does not fit on an SM");
    *out = sms * per;
    return 0;
}

int bfs_launch(const KParams& K, bool multi, cudaStream_t st) {
    void* args[] = {(void*)&K};
    dim3 grid(K.cpr * K.vranks), block(BFS_THREADS);
    if (multi) BIBFS_CUDA_TRY(cudaLaunchCooperativeKernel((void*)k_bibfs<true>, grid, block, args, 0, st));
    else BIBFS_CUDA_TRY(cudaLaunchCooperativeKernel((void*)k_bibfs<false>, grid, block, args, 0, st));
    count_launch();
    return 0;
}

int path_launch(const KParams& K, cudaStream_t st) {
    k_path<<<1, 64, 0, st>>>(K);
    BIBFS_KERNEL_CHECK();
    return 0;
}

}  // namespace bibfs
