// pair_tier.cuh — the latency tier of the search kernel: PAIRED rounds on inline-adjacency records.
// (included by bfs_dev.cu inside its anonymous namespace, after Env / Smem)
//
// A reference-sized query (ER 100k: ~3,000 vertices touched in 14 levels) is a chain of dependent
// memory round trips; the work is nothing.  The plain solo level has three on its chain
// (row offsets -> adjacency entries -> visited word) and runs one side per level.  This tier cuts
// the chain to ONE round trip per round, and halves the number of rounds:
//
//   * vertex records ("vrec", built once per shard by k_vrec_build): two 32-byte sectors per vertex, the first =
//     {degree, row start, first six neighbours}, the second = neighbours 7..14 (read only by rows that long).  When v is a CANDIDATE (read from its parent's
//     record) its own record is fetched in the same instruction slot as the atomic that claims it:
//     the claim (285 ns) and the fetch (154 ns, L2) overlap, and a newly discovered vertex enters
//     the next frontier WITH its neighbours already in shared memory.  A round is
//         shared memory -> { claim || record fetch } -> shared memory -> one __syncthreads.
//     Rows longer than six entries read the second sector, rows longer than 14 the rest from col[] (further trips,
//     for those rows only).
//   * both directions expand in the SAME round on disjoint warps (the reference's v2/v3 expand S
//     and T in one iteration too, v2/second_try.cpp:70-105 — and print ceil(d/2)).
//     Exact length rule used here: balls are disjoint when a round starts, so every vertex that
//     ends the round with both marks was marked by at least one side in this round.  If it sits in
//     one of the two frontiers that were expanded (its mark is OLD, depth d_x) the path through it
//     has d_s + d_t + 1 hops; if both marks are new, d_s + d_t + 2.  After a round that saw a hit,
//     every frontier thread re-reads the visited word of ITS OWN frontier vertex: any other-side
//     mark there is an old hit.  Cost: one extra load, in the last round only.
//   * level state lives in registers of EVERY thread (each derives the next round from two
//     counters and a flag word in shared memory), so a round has one barrier, not two, and no
//     thread-0 bookkeeping between them (that was 40 % of a solo level).  The counters rotate
//     through three sets so that nobody resets a word somebody else may still read.
//
// Inside the general kernel the tier runs at the start of a query while both frontiers fit one CTA (<= 512 entries
// per side) and no row of >= PT_MAXDEG entries has been discovered; then it hands the search to the general machinery
// (solo / cluster / grid / hub / bottom-up) through the same LevelState those use.  The front kernel (front.cuh) runs
// it with 768 entries per side, keeps the first PT_LOG_CAP log entries per side in shared memory for the path walk, and
// continues with WIDE rounds on its whole cluster (pair_wide below) where the narrow rounds stop.

// PT_MARK: points of a round where the fine-trace build takes time stamps.  The regular build keeps a NEVER-TAKEN branch
// at each of them (query flag bit 31 is rejected by bibfs_ctx_query): measured on B200, the same round loop runs 36.3 us
// per er100k query (device clock, L2-hot) without the branches, 36.3 us with empty asm statements in their place, and
// 24.7 us with them — ptxas schedules the batch of record loads and claims differently when the basic blocks end there
// (profiles/r2/scheduling_marks.md).  The fine-trace build (27.3 us including its stamps) gave the hint.
#ifdef BIBFS_FINE_TRACE
#define PT_MARK(slot, dep) DBG_STAMP(slot, dep)
#else
#define PT_MARK(slot, dep)                                                                                   \
    do {                                                                                                     \
        if (threadIdx.x == 0 && mark_on) {                                                                   \
            u64 _t;                                                                                          \
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(_t) : "l"((u64)(dep)) : "memory");             \
            *reinterpret_cast<volatile u64*>(E.vis) = _t;                                                    \
        }                                                                                                    \
    } while (0)
#endif

constexpr u32 PT_CAP_SMALL = 512;    // frontier entries per side: inside the general kernel (shares its shared memory) ...
constexpr u32 PT_LOG_CAP = 1280;     // front kernel: (vertex, log index of its parent) pairs per side kept in shared memory for the path walk
constexpr u32 PT_CAP_FRONT = 768;    // ... and in the front kernel (frontiers beyond the CTA's 512 threads are walked in passes)
constexpr u32 PT_MAXDEG = 32;        // a discovered row at least this long ends the tier
constexpr int PT_NB = 6;             // neighbours inlined in a vertex record
constexpr int PT_W = 4;              // adjacency entries per lane per batch beyond the inlined ones
enum : u32 { PT_HIT = 1u, PT_BIG = 2u, PT_OVF = 4u };

// State of the tier when its rounds stop: written by thread 0 of pair_run, pushed to the other CTAs of the front kernel's
// cluster when the rounds go WIDE (pair_wide), and finally turned into the LevelState by pair_finish.
struct PairOut {
    u32 n[2], base[2], d[2];         // frontier sizes, log cursors, depths per side
    u32 levels, visited;
    i32 length;                      // -2: still searching (hand over / go wide)
    u32 meet, depth_s, depth_t;
    u32 go_wide;                     // the rounds continue on the whole cluster
    u32 went_wide;                   // ... and did
    u32 pad[2];
};
static_assert(sizeof(PairOut) == 64, "pushed over DSMEM as 16 words");
constexpr u32 PT_REC_ENTRIES = PT_NB + 2 * PT_W;      // neighbours held by a vertex record (both sectors): 14
constexpr u32 PT_WIDE_MIN = 160;     // frontier positions of a round from which the whole cluster takes it
constexpr u32 PT_WIDE_CAP = 1u << 16;  // ... and the frontier size per side at which even the cluster hands over

template <u32 CAP>
struct PairSmemT {
    u32 cnt[3][2];                   // vertices discovered in the round, per side (three rotating sets)
    u32 flag[3];
    u32 meet_new[3];                 // smallest vertex a claim of the round found marked by the other side
    u32 tr_edges[3][2];              // (tracing only) adjacency entries per side
    u32 meet_old[2];                 // old-hit test: smallest frontier vertex of side x that carries the other side's mark
    unsigned long long edges, visited;
    PairOut out;                     // where the rounds stopped (pair_run / pair_wide -> pair_finish)
    alignas(16) u32 fv[2][2][CAP];               // frontier vertices [side][buffer]
    alignas(16) uint4 fr[2][2][CAP][2];          // ... and their records
    // front kernel only: the first PT_LOG_CAP entries of each side's log, vertex | (log index of its parent) << 32.  A
    // search that ends inside the narrow rounds walks its path and clears its marks from here instead of chasing
    // par[] through L2 (a dependent load of a partially written sector per hop: ~550 ns each, measured)
    static constexpr bool HAS_LOG = CAP == PT_CAP_FRONT;
    alignas(16) unsigned long long slog[2][HAS_LOG ? PT_LOG_CAP : 1];
};
using PairSmem = PairSmemT<PT_CAP_SMALL>;

__device__ __forceinline__ void reds_or(u32* p, u32 v) {
    asm volatile("red.shared.or.b32 [%0], %1;" ::"r"(ptx::smem_addr(p)), "r"(v) : "memory");
}

// positions of a paired round (both sides when they are of comparable size, else the smaller one)
__device__ __forceinline__ u32 both_span(u32 n0, u32 n1) {
    const u32 mx = max(n0, n1), mn = min(n0, n1);
    return (mx <= 4u * mn || mx <= 32u) ? n0 + n1 : mn;
}

// What the tier works on.  Single rank: the rank's own queue / parent arrays (so that the general tiers continue on
// them).  Several ranks (MULTI): rank 0 runs the tier ALONE over the whole graph — records and long rows of other
// ranks' vertices are read through the peer mappings, claims go to rank 0's own all-vertex visited words, and the
// vertices it reaches are logged locally (log / pall regions of rank 0's arena); the other ranks are not touched
// unless the search has to be handed over to them (front_handover below).
struct PairIO {
    u32* log[2];        // vertices reached per side, in discovery order (entry 0 = the seed)
    u32* par[2];        // parent of a vertex per side, indexed by the vertex id
};

// Narrow rounds: one CTA.  Leaves its state in ps.out (thread 0) and the adjacency entries it inspected in ps.edges;
// pair_finish turns that into the LevelState.  allow_wide: the CTA leads a cluster (front kernel) and stops with
// out.go_wide = 1 when a round is large enough for all of it (pair_wide).
template <bool MULTI, u32 PT_CAP>
__device__ __noinline__ void pair_run(const KParams& K, Env E, PairSmemT<PT_CAP>& ps, const PairIO io,
                                      const u32 src, const u32 dst, const bool allow_wide) {
    const u32 tid = threadIdx.x, lane = tid & 31;
    u64* const g_vis = E.vis;
    const bool tracing = (K.flags & BIBFS_F_TRACE) != 0;
    // Everything the round loop needs from the kernel parameters is copied into registers here: K is reached through a
    // pointer, and every asm statement of the loop is a memory clobber, so each K.x inside it would be a fresh load on
    // the chain of the round (owner_of's reciprocal alone cost two of them per round: +2 us per round, measured).
    const u64 div_b = K.div_B;
    const bool mark_on = (K.flags & 0x80000000u) != 0; (void)mark_on;      // never set (see PT_MARK)
    const u32 wide_min = K.wide_min;
    const u32* const e_col = E.col;
    u32* const log0 = io.log[0]; u32* const log1 = io.log[1];
    u32* const par0 = io.par[0]; u32* const par1 = io.par[1];
    auto owner = [div_b](u32 v) -> u32 { return (u32)__umul64hi(div_b, (u64)v); };
    // record / adjacency bases indexed by the GLOBAL vertex id resp. the owner's row start
    __shared__ const uint4* s_vrecb[MAXR];
    __shared__ const u32* s_colb[MAXR];
    if (MULTI && tid < K.P) { s_vrecb[tid] = K.vrec[tid] - 4 * (size_t)tid * K.B; s_colb[tid] = K.col[tid]; }
    const uint4* const vrec0 = MULTI ? nullptr : K.vrec[E.me] - 4 * (size_t)E.lo;
    auto rec_of = [&](u32 v) -> const uint4* { return (MULTI ? s_vrecb[owner(v)] : vrec0) + 4 * (size_t)v; };
    // (records and adjacency never change: gpu-scope loads are enough also when they come from a peer's memory, and
    //  system-scope loads were measured 2 us per round slower even on local memory)
    auto ld_rec_if = [&](bool on, const uint4* p) -> uint4 { return ptx::ld_v4_gpu_if(on, p); };
    if (MULTI) __syncthreads();
    // level state, identical in every thread
    u32 n0 = 1, n1 = 1, base0 = 1, base1 = 1, d0 = 0, d1 = 0, levels = 0, buf0 = 0, buf1 = 0, par = 0;
    u64 my_edges = 0;
    if (tid < 2) {
        const u32 v = tid ? dst : src;
        const uint4* p = rec_of(v);
        const uint4 a = ld_rec_if(true, p), b = ld_rec_if(true, p + 1);
        ps.fv[tid][0][0] = v;
        if (PairSmemT<PT_CAP>::HAS_LOG) ps.slog[tid][0] = (unsigned long long)v;
        ps.fr[tid][0][0][0] = a;
        ps.fr[tid][0][0][1] = b;
        ps.meet_old[tid] = a.x >= PT_MAXDEG ? 0u : NO_V;       // (borrowed until the first barrier: "seed row too long")
    }
    if (tid < 3) { ps.cnt[tid][0] = 0; ps.cnt[tid][1] = 0; ps.flag[tid] = 0; ps.meet_new[tid] = NO_V; ps.tr_edges[tid][0] = 0; ps.tr_edges[tid][1] = 0; }
    if (tid == 32) { ps.edges = 0; ps.visited = 0; }
    __syncthreads();
    bool leave = (ps.meet_old[0] & ps.meet_old[1]) != NO_V;
    __syncthreads();
    if (tid < 2) ps.meet_old[tid] = NO_V;
    i32 length = -2;                 // -2: still searching
    u32 meet = NO_V, depth_s = 0, depth_t = 0;
    u32 visited = 0, go_wide = 0;
    if (leave && allow_wide) go_wide = 1;          // a seed with a long row: straight to the cluster

    while (!leave) {
        // ---- which sides expand: both, unless one frontier dwarfs the other (then the smaller, ties -> source, like v1)
        const u32 mx = max(n0, n1), mn = min(n0, n1);
        if (mn > PT_CAP) break;
        if (allow_wide && both_span(n0, n1) > wide_min) { go_wide = 1; break; }
        const u32 pad0 = (n0 + 31u) & ~31u;
        const bool both = (mx <= 4u * mn || mx <= 32u) && mx <= PT_CAP;
        const bool e0 = both || n0 <= n1, e1 = both || n1 < n0;
        const u32 off1 = e0 ? pad0 : 0u;                        // the target side starts on a warp boundary
        const u32 span = off1 + (e1 ? n1 : 0u);                 // frontier positions of the round; the CTA walks them in passes of 512
        const bool wide = (e0 ? n0 : 0u) + (e1 ? n1 : 0u) > 64u;  // aggregate the cursor reservations per warp
        const u32 pb0 = buf0, pb1 = buf1;
        u32 fl = 0;
        for (u32 vt = tid; (vt & ~31u) < span; vt += BFS_THREADS) {
        const u32 side = (e0 && (!e1 || vt < off1)) ? 0u : 1u;      // warp-uniform
        const u32 idx = side ? vt - off1 : vt;
        const bool have = side ? (e1 && vt >= off1 && idx < n1) : (e0 && vt < n0);
        const u32 sh = 32u * side, osh = 32u - sh;
        const u32 buf = side ? buf1 : buf0;
        const u32 base = side ? base1 : base0;
        const unsigned long long plog = (unsigned long long)(base - (side ? n1 : n0) + idx) << 32;      // where u sits in the log
        unsigned long long* const slog = ps.slog[side];
        u32 u = 0, deg = 0, r0 = 0;
        uint4 ra = make_uint4(0, 0, 0, 0), rb = ra;
        if (have) {
            u = ps.fv[side][buf][idx];
            ra = ps.fr[side][buf][idx][0];
            rb = ps.fr[side][buf][idx][1];
            deg = ra.x; r0 = ra.y;
        }
        PT_MARK(0, u + deg);
        my_edges += deg;
        if (tracing && deg) ptx::reds_add(&ps.tr_edges[par][side], deg);
        u32* const out_v = ps.fv[side][buf ^ 1u];
        uint4(*const out_r)[2] = ps.fr[side][buf ^ 1u];
        u32* const cursor = &ps.cnt[par][side];
        u32* const qs = side ? log1 : log0;
        u32* const pp = side ? par1 : par0;
        const u32* const g_col = MULTI ? s_colb[owner(u)] : e_col;

        // one batch of W candidates: claims and record fetches in flight together (branch-free: idle slots OR a zero
        // into the thread's scratch word and re-read the thread's own record)
        auto batch = [&](auto Wc, const u32* v, const u32 nact) {
            constexpr int W = decltype(Wc)::value;
            uint4 qa[W], qb[W];
            u64 old[W];
#pragma unroll
            for (int k = 0; k < W; ++k) {
                const bool act = (u32)k < nact;
                const uint4* p = rec_of(act ? v[k] : u);
                qa[k] = ld_rec_if(act, p);
                qb[k] = ld_rec_if(act, p + 1);
            }
#pragma unroll
            for (int k = 0; k < W; ++k) {
                const bool act = (u32)k < nact;
                old[k] = ptx::atom_or_gpu_if(act, &g_vis[(act ? v[k] : u) >> 5], 1ull << ((v[k] & 31u) + sh));
            }
            PT_MARK(1, qb[W - 1].w);
            PT_MARK(2, old[W - 1]);
            const u32 shz = sh + ptx::zero_after(old[W - 1]);      // == sh; pins the tests behind the whole batch
            u32 newmask = 0, meetmask = 0;
#pragma unroll
            for (int k = 0; k < W; ++k) {
                const u32 isnew = (u32)k < nact ? (((u32)(old[k] >> ((v[k] & 31u) + shz)) & 1u) ^ 1u) : 0u;
                newmask |= isnew << k;
                meetmask |= (isnew & (u32)(old[k] >> ((v[k] & 31u) + osh))) << k;
            }
            const u32 nnew = (u32)__popc(newmask);
            u32 slot = 0;
            if (wide) slot = warp_reserve(nnew, [&](u32 n) { return ptx::atoms_add(cursor, n); });
            else if (nnew) slot = ptx::atoms_add(cursor, nnew);
            PT_MARK(3, slot);
            if (newmask) {
#pragma unroll
                for (int k = 0; k < W; ++k) {
                    if ((newmask >> k) & 1u) {
                        qs[base + slot] = v[k];
                        pp[v[k]] = u;
                        if (PairSmemT<PT_CAP>::HAS_LOG && base + slot < PT_LOG_CAP) slog[base + slot] = plog | v[k];
                        if (slot < PT_CAP) { out_v[slot] = v[k]; out_r[slot][0] = qa[k]; out_r[slot][1] = qb[k]; }
                        else fl |= PT_OVF;
                        if (qa[k].x >= PT_MAXDEG) fl |= PT_BIG;
                        ++slot;
                    }
                }
                if (meetmask) {
                    fl |= PT_HIT;
#pragma unroll
                    for (int k = 0; k < W; ++k)
                        if ((meetmask >> k) & 1u) ptx::reds_min(&ps.meet_new[par], v[k]);
                }
            }
        };
        if (__any_sync(FULL, have)) {
            {
                const u32 c[PT_NB] = {ra.z, ra.w, rb.x, rb.y, rb.z, rb.w};
                batch(std::integral_constant<int, PT_NB>{}, c, min(deg, (u32)PT_NB));
            }
            // entries 7..14 come from the record's second sector, the rest (rare) from col[]
            if (__any_sync(FULL, deg > (u32)PT_NB)) {
                const bool ext = deg > (u32)PT_NB;
                const uint4* xp = rec_of(u) + 2;
                const uint4 xa = ld_rec_if(ext, xp), xb = ld_rec_if(ext, xp + 1);
                {
                    const u32 c[PT_W] = {xa.x, xa.y, xa.z, xa.w};
                    batch(std::integral_constant<int, PT_W>{}, c, min(ext ? deg - PT_NB : 0u, (u32)PT_W));
                }
                if (__any_sync(FULL, deg > (u32)PT_NB + PT_W)) {
                    const u32 c[PT_W] = {xb.x, xb.y, xb.z, xb.w};
                    batch(std::integral_constant<int, PT_W>{}, c, min(deg > (u32)PT_NB + PT_W ? deg - PT_NB - PT_W : 0u, (u32)PT_W));
                }
            }
            for (u32 e = PT_NB + 2 * PT_W; __any_sync(FULL, e < deg); e += PT_W) {
                u32 c[PT_W];
                const u32 left = deg > e ? deg - e : 0u;
#pragma unroll
                for (int k = 0; k < PT_W; ++k) {
                    const u32* a = g_col + r0 + ((u32)k < left ? e + k : 0u);
                    c[k] = ptx::ld_relaxed_gpu(a);
                }
                batch(std::integral_constant<int, PT_W>{}, c, min(left, (u32)PT_W));
            }
        }
        }       // passes
        if (fl) reds_or(&ps.flag[par], fl);
        PT_MARK(4, fl);
        __syncthreads();
        PT_MARK(5, tid);
        // ---- round end: every thread derives the next state from the round's counters
        const u32 c0 = ps.cnt[par][0], c1 = ps.cnt[par][1], f = ps.flag[par], mnew = ps.meet_new[par];
        if (tid == 0) {
            const u32 z = (par + 2u) % 3u;       // last read before this barrier, next used after the next one
            ps.cnt[z][0] = 0; ps.cnt[z][1] = 0; ps.flag[z] = 0; ps.meet_new[z] = NO_V;
            if (tracing) {
                const u64 now = ptx::globaltimer_ns();
                if (e0) trace_level(K, E, 0, PH_TD_SOLO | 16u, n0, c0, ps.tr_edges[par][0], now);
                if (e1) trace_level(K, E, 1, PH_TD_SOLO | 16u, n1, c1, ps.tr_edges[par][1], now);
                ps.tr_edges[z][0] = 0; ps.tr_edges[z][1] = 0;
            }
        }
        const u32 c0_prev = n0, c1_prev = n1;        // sizes of the frontiers this round expanded
        if (e0) { d0 += 1; levels += 1; base0 += c0; n0 = c0; buf0 ^= 1u; visited += c0; }
        if (e1) { d1 += 1; levels += 1; base1 += c1; n1 = c1; buf1 ^= 1u; visited += c1; }
        par = (par + 1u) % 3u;
        PT_MARK(6, n0 + n1);
#ifdef BIBFS_FINE_TRACE
        if (blockIdx.x == 0 && threadIdx.x == 0) { if (g_dbg_level < 256) g_dbg[g_dbg_level][7] = ((u64)(e0 ? 1 : 0) | (e1 ? 2 : 0)) | ((u64)n0 << 8) | ((u64)n1 << 32); g_dbg_level++; }
#endif
        if (f & PT_HIT) {
            length = (i32)(d0 + d1); meet = mnew; depth_s = d0;
            if (e0 && e1) {
                // old-hit test (see the header): is a vertex of an expanded frontier marked by the other side now?
                for (u32 vt = tid; vt < span; vt += BFS_THREADS) {
                    const u32 side = vt < off1 ? 0u : 1u;
                    const u32 idx = side ? vt - off1 : vt;
                    if (idx < (side ? c1_prev : c0_prev)) {
                        const u32 u = ps.fv[side][side ? pb1 : pb0][idx];
                        const u64 w = ptx::ld_relaxed_gpu(&g_vis[u >> 5]);
                        if ((w >> ((u & 31u) + 32u - 32u * side)) & 1ull) ptx::reds_min(&ps.meet_old[side], u);
                    }
                }
                __syncthreads();
                const u32 m0 = ps.meet_old[0], m1 = ps.meet_old[1];
                if (m0 != NO_V) { length = (i32)(d0 + d1 - 1u); meet = m0; depth_s = d0 - 1u; }        // in the source frontier, reached by the target side
                else if (m1 != NO_V) { length = (i32)(d0 + d1 - 1u); meet = m1; depth_s = d0; }
            }
            depth_t = (u32)length - depth_s;
            break;
        }
        if (n0 == 0 || n1 == 0) { length = -1; depth_s = d0; depth_t = d1; break; }      // a side is exhausted: no path
        if (allow_wide && (f & (PT_OVF | PT_BIG))) { go_wide = 1; break; }     // the cluster takes long rows and large frontiers (the log holds what the mirror dropped)
        if (f & (PT_BIG | PT_OVF)) break;
    }
    // ---- leave the tier: totals, then either the verdict or the hand-over to the general machinery
    {
        u64 e = my_edges;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) e += __shfl_xor_sync(FULL, e, d);
        if (lane == 0 && e) ptx::reds_add64(&ps.edges, e);
    }
    __syncthreads();
    if (tid == 0) {
        PairOut& o = ps.out;
        o.n[0] = n0; o.n[1] = n1; o.base[0] = base0; o.base[1] = base1; o.d[0] = d0; o.d[1] = d1;
        o.levels = levels; o.visited = visited; o.length = length; o.meet = meet; o.depth_s = depth_s; o.depth_t = depth_t;
        o.go_wide = go_wide; o.went_wide = 0;
    }
    __syncthreads();
}

// Turn the tier's final state into the LevelState the rest of the machinery uses (thread 0 of the leader CTA; callers
// synchronise).  `edges`: adjacency entries the tier inspected.
template <bool MULTI>
__device__ __noinline__ void pair_finish(const KParams& K, const Env& E, LevelState& st, Acc& acc, u32* qvalid, const PairOut& o, u64 edges) {
    st.nf[0] = o.n[0]; st.nf[1] = o.n[1];
    st.depth[0] = o.d[0]; st.depth[1] = o.d[1];
    st.levels = o.levels;
    st.q_lo[0] = o.base[0] - o.n[0]; st.q_hi[0] = o.base[0];
    st.q_lo[1] = o.base[1] - o.n[1]; st.q_hi[1] = o.base[1];
    acc.q_tail[0] = o.base[0]; acc.q_tail[1] = o.base[1];
    acc.new_count = 0; acc.meet_local = NO_V; acc.edges = 0;
    qvalid[0] = 0; qvalid[1] = 0;                   // the general tiers read the frontier from the queue
    st.edges_total += edges;
    st.visited_total += o.visited;
    st.rounds_paired = o.levels | (o.went_wide ? 0x80000000u : 0u);      // bit 31: wide rounds were used (launch policy of ctx.cu)
    if (o.length != -2) {
        st.length = o.length; st.meet = o.length >= 0 ? o.meet : NO_V;
        st.depth[0] = o.depth_s; st.depth[1] = o.depth_t;
        st.phase = PH_DONE;
        write_result(K, E.C, st);
    } else if (!MULTI) {
        decide_next(K, st);
        if (st.phase == PH_DONE) write_result(K, E.C, st);
    }
}

// ------------------------------------------------------------------------------------------
// WIDE rounds: the whole thread-block cluster of the front kernel.
//
// One SM's load/store path is what limits a round of more than a couple of hundred frontier
// vertices (a 512-thread narrow round takes ~5 us), so from PT_WIDE_MIN positions on the round is
// dealt over all CTAs of the cluster.  Differences to the narrow round:
//   * the frontier is read from the LOG in global memory (every reached vertex is appended there
//     anyway) and the record of a frontier vertex is fetched by the thread that expands it — two
//     more L2 trips per round, but no record fetch per CANDIDATE and no shared-memory mirror, so
//     the frontier size is not bounded by one CTA's shared memory (a 4096 x 4096 grid stays here
//     for all of its 8,190 levels);
//   * cursors, flags and the meeting vertex live in Ctrl::wide (L2 atomics, one reservation per
//     warp), the round ends at barrier.cluster, and every thread of every CTA derives the next
//     round from those counters, exactly as in the narrow round;
//   * rows of >= PT_MAXDEG entries (hubs of power-law graphs) are cut into 256-entry chunks that ALL warps of the
//     cluster take in a second phase of the round, behind one more cluster barrier; graphs without such rows
//     (KParams::long_rows, from the records' degree maximum) skip that barrier.
// All CTAs call this together (the leader after pair_run, the others from the start-up barrier);
// `o` is each CTA's own copy of the state and is updated in place.  Returns the adjacency entries
// inspected by the whole cluster (valid in the leader CTA).
// ------------------------------------------------------------------------------------------
//   * discoveries are STAGED per CTA in shared memory ((vertex, parent) pairs in the space the narrow rounds use for
//     their frontier mirror) and flushed with ONE reservation per CTA and side on the global cursor: 256 warps
//     bumping the same L2 word serialise at ~40 ns apiece (10 us per round, measured on the 4096 x 4096 grid).
template <bool MULTI>
__device__ __noinline__ u64 pair_wide(const KParams& K, Env E, const PairIO io, PairOut& o, const u32 cl_n, const u32 cl_rank,
                                      u64* const stage /* 2 x stage_cap entries of this CTA's shared memory */, const u32 stage_cap) {
    const u32 tid = threadIdx.x, lane = tid & 31;
    Ctrl::Wide& W = E.C->wide;
    u64* const g_vis = E.vis;
    const bool tracing = (K.flags & BIBFS_F_TRACE) != 0;
    const u64 div_b = K.div_B;
    const u32* const e_col = E.col;
    u32* const log0 = io.log[0]; u32* const log1 = io.log[1];
    u32* const par0 = io.par[0]; u32* const par1 = io.par[1];
    auto owner = [div_b](u32 v) -> u32 { return (u32)__umul64hi(div_b, (u64)v); };
    __shared__ const uint4* s_vrecb[MAXR];
    __shared__ const u32* s_colb[MAXR];
    if (MULTI && tid < K.P) { s_vrecb[tid] = K.vrec[tid] - 4 * (size_t)tid * K.B; s_colb[tid] = K.col[tid]; }
    const uint4* const vrec0 = MULTI ? nullptr : K.vrec[E.me] - 4 * (size_t)E.lo;
    auto rec_of = [&](u32 v) -> const uint4* { return (MULTI ? s_vrecb[owner(v)] : vrec0) + 4 * (size_t)v; };
    __shared__ u32 s_cnt[2], s_gbase[2];
    if (tid < 2) s_cnt[tid] = 0;
    __syncthreads();
    u32 n0 = o.n[0], n1 = o.n[1], base0 = o.base[0], base1 = o.base[1], d0 = o.d[0], d1 = o.d[1], levels = o.levels, visited = o.visited;
    u32 par = 0;
    u64 my_edges = 0;
    i32 length = -2;
    u32 meet = NO_V, depth_s = 0, depth_t = 0;
    const u32 gthreads = cl_n * BFS_THREADS, gt0 = cl_rank * BFS_THREADS + tid;

    const bool long_rows = K.long_rows != 0;
    HubEntry* const H = hub_of(K, E.me);
    const u32 gwarp = gt0 >> 5, gwarps = gthreads >> 5;

    while (true) {
        const u32 mx = max(n0, n1), mn = min(n0, n1);
        if (mn > PT_WIDE_CAP) break;
        const u32 pad0 = (n0 + 31u) & ~31u;
        // both sides while they are of comparable size AND fit one pass of the cluster; beyond that only the smaller one
        // (v1's policy: fewer entries to inspect), and such a round may stop as soon as any claim has hit the other side:
        // with one side expanding every hit of the round gives the same length (what v1's `break` does, main-v1.cpp:61)
        // (graphs with long rows: always one side - a hub row multiplies the work of a round by 10^5, and only a
        //  single-side round may stop early)
        const bool both = !long_rows && (mx <= 4u * mn || mx <= 32u) && ((n0 + 31u) & ~31u) + n1 <= gthreads;
        const bool e0 = both || n0 <= n1, e1 = both || n1 < n0;
        auto hit_seen = [&]() -> bool {
            u32 f = 0;
            if (lane == 0) f = ptx::ld_relaxed_gpu(&W.flag[par]);
            return (__shfl_sync(FULL, f, 0) & PT_HIT) != 0;
        };
        const u32 off1 = e0 ? pad0 : 0u;
        const u32 span = off1 + (e1 ? n1 : 0u);
        const u32 flo0 = base0 - n0, flo1 = base1 - n1;          // the frontiers are the last n entries of the logs
        u32 fl = 0;
        // claims of one batch of W candidates of parent `pu` (side `sd`); discoveries go to the CTA's stage
        auto batch = [&](auto Wc, const u32* v, const u32 nact, const u32 pu, const u32 sd) {
            constexpr int Wn = decltype(Wc)::value;
            const u32 sh = 32u * sd, osh = 32u - sh;
            u64 old[Wn];
#pragma unroll
            for (int k = 0; k < Wn; ++k) {
                const bool act = (u32)k < nact;
                old[k] = ptx::atom_or_gpu_if(act, &g_vis[(act ? v[k] : pu) >> 5], 1ull << ((v[k] & 31u) + sh));
            }
            const u32 shz = sh + ptx::zero_after(old[Wn - 1]);
            u32 newmask = 0, meetmask = 0;
#pragma unroll
            for (int k = 0; k < Wn; ++k) {
                const u32 isnew = (u32)k < nact ? (((u32)(old[k] >> ((v[k] & 31u) + shz)) & 1u) ^ 1u) : 0u;
                newmask |= isnew << k;
                meetmask |= (isnew & (u32)(old[k] >> ((v[k] & 31u) + osh))) << k;
            }
            u32 slot = warp_reserve((u32)__popc(newmask), [&](u32 n) { return ptx::atoms_add(&s_cnt[sd], n); });
            if (newmask) {
                u64* const stg = stage + (size_t)sd * stage_cap;
#pragma unroll
                for (int k = 0; k < Wn; ++k) {
                    if ((newmask >> k) & 1u) {
                        if (slot < stage_cap) stg[slot] = (u64)v[k] | ((u64)pu << 32);
                        else {      // stage full: straight to the log
                            (sd ? log1 : log0)[(sd ? base1 : base0) + atomicAdd(&W.cnt[par][sd], 1u)] = v[k];
                            (sd ? par1 : par0)[v[k]] = pu;
                        }
                        ++slot;
                    }
                }
                if (meetmask) {
#pragma unroll
                    for (int k = 0; k < Wn; ++k)
                        if ((meetmask >> k) & 1u) atomicMin(&W.meet[par], v[k]);
                    if (!(fl & PT_HIT)) atomicOr(&W.flag[par], PT_HIT);      // at once: single-side rounds stop on it
                    fl |= PT_HIT;
                }
            }
        };
        // ---- phase 1: every frontier position; rows shorter than PT_MAXDEG are walked by their lane, longer ones are
        // cut into chunks for the whole cluster (phase 2)
        for (u32 vt = gt0; (vt & ~31u) < span; vt += gthreads) {
            if (!both && vt >= gthreads && hit_seen()) break;
            const u32 side = (e0 && (!e1 || vt < off1)) ? 0u : 1u;      // warp-uniform
            const u32 idx = side ? vt - off1 : vt;
            const bool have = side ? (e1 && vt >= off1 && idx < n1) : (e0 && vt < n0);
            u32 u = 0;
            if (have) u = ptx::ld_relaxed_gpu((side ? log1 + flo1 : log0 + flo0) + idx);
            const uint4* rp = rec_of(u);
            const uint4 ra = ptx::ld_v4_gpu_if(have, rp), rb = ptx::ld_v4_gpu_if(have, rp + 1);
            const u32 deg = ra.x, r0 = ra.y;
            my_edges += deg;
            if (tracing && deg) atomicAdd(&W.tr_edges[par][side], deg);
            if (!__any_sync(FULL, have)) continue;
            const u32* const g_col = MULTI ? s_colb[owner(u)] : e_col;
            bool lng = deg >= PT_MAXDEG;
            if (lng) {
                const u32 nch = hub_nchunks(r0, (u64)r0 + deg);
                const unsigned long long inc = (1ull << 40) | (unsigned long long)nch;
                const unsigned long long was = atomicAdd(&W.long_packed[par], inc);
                const u32 hs = (u32)(was >> 40);
                if (hs < HUB_CAP) {
                    HubEntry h;
                    h.u = u; h.nchunks = nch | (side << 31); h.r0 = r0; h.r1 = (u64)r0 + deg; h.chunk_base = was & 0xFFFFFFFFFFull;
                    H[hs] = h;
                    lng = false;
                } else {
                    atomicAdd(&W.long_packed[par], 0ull - inc);       // list full: the warp walks the row below
                }
            }
            const u32 sdeg = deg < PT_MAXDEG ? deg : 0u;
            {
                const u32 c[PT_NB] = {ra.z, ra.w, rb.x, rb.y, rb.z, rb.w};
                batch(std::integral_constant<int, PT_NB>{}, c, min(sdeg, (u32)PT_NB), u, side);
            }
            if (__any_sync(FULL, sdeg > (u32)PT_NB)) {          // entries 7..14: the record's second sector
                const bool ext = sdeg > (u32)PT_NB;
                const uint4 xa = ptx::ld_v4_gpu_if(ext, rp + 2), xb = ptx::ld_v4_gpu_if(ext, rp + 3);
                {
                    const u32 c[PT_W] = {xa.x, xa.y, xa.z, xa.w};
                    batch(std::integral_constant<int, PT_W>{}, c, min(ext ? sdeg - PT_NB : 0u, (u32)PT_W), u, side);
                }
                if (__any_sync(FULL, sdeg > (u32)PT_NB + PT_W)) {
                    const u32 c[PT_W] = {xb.x, xb.y, xb.z, xb.w};
                    batch(std::integral_constant<int, PT_W>{}, c, min(sdeg > (u32)PT_NB + PT_W ? sdeg - PT_NB - PT_W : 0u, (u32)PT_W), u, side);
                }
            }
            // entries 15..31 live in col[] (another GPU's memory when rank 0 works alone on a partitioned graph): ALL of a
            // lane's loads are issued before the first claim, so that the row costs one trip to col[] instead of five
            if (__any_sync(FULL, sdeg > PT_REC_ENTRIES)) {
                constexpr int TAIL = (int)(PT_MAXDEG - 1u - PT_REC_ENTRIES);       // 17
                const u32 rem = sdeg > PT_REC_ENTRIES ? sdeg - PT_REC_ENTRIES : 0u;
                const u32* const tp = g_col + r0 + PT_REC_ENTRIES;
                u32 c[TAIL + 3] = {};
#pragma unroll
                for (int k = 0; k < TAIL; ++k) c[k] = ptx::ld_u32_gpu_if((u32)k < rem, tp + k);
#pragma unroll
                for (int j = 0; j < TAIL; j += PT_W)
                    if (__any_sync(FULL, rem > (u32)j)) batch(std::integral_constant<int, PT_W>{}, c + j, min(rem > (u32)j ? rem - j : 0u, (u32)PT_W), u, side);
            }
            u32 big = __ballot_sync(FULL, lng);          // (only when the chunk list overflowed)
            while (big) {
                const u32 sl = __ffs(big) - 1;
                big &= big - 1;
                const u32 uu = __shfl_sync(FULL, u, sl), rr = __shfl_sync(FULL, r0, sl), dd = __shfl_sync(FULL, deg, sl);
                const u32* const cc = (MULTI ? s_colb[owner(uu)] : e_col) + rr;
                for (u32 e = 0; e < dd; e += 32u * PT_W) {
                    u32 c[PT_W];
                    u32 nact = 0;
#pragma unroll
                    for (int k = 0; k < PT_W; ++k) {
                        const u32 i = e + 32u * k + lane;
                        c[k] = ptx::ld_relaxed_gpu(cc + (i < dd ? i : 0u));
                        nact += i < dd ? 1u : 0u;
                    }
                    batch(std::integral_constant<int, PT_W>{}, c, nact, uu, side);
                }
            }
        }
        // ---- phase 2 (graphs with long rows): the chunks of the long rows, dealt over all warps of the cluster
        if (long_rows) {
            ptx::cluster_sync_all();
            const u64 hp = ptx::ld_relaxed_gpu(reinterpret_cast<const u64*>(&W.long_packed[par]));
            const u32 ne = min((u32)(hp >> 40), HUB_CAP), nchunks = (u32)(hp & 0xFFFFFFFFFFull);
            for (u32 c = gwarp; c < nchunks; c += gwarps) {
                if (!both && hit_seen()) break;
                const HubChunk h = hub_lookup(H, ne, (const u32*)nullptr, c);
                const u32 sd = h.nch_word >> 31;
                const u32* const cc = (MULTI ? s_colb[owner(h.uu)] : e_col) + h.A;
                // a chunk (256 entries) is fetched in ONE go: 8 loads per lane in flight, then the claims
                for (u32 e = h.lo; e < h.hi; e += 64u * PT_W) {
                    u32 cv[2 * PT_W];
#pragma unroll
                    for (int k = 0; k < 2 * PT_W; ++k) {
                        const u32 i = e + 32u * k + lane;
                        cv[k] = ptx::ld_u32_gpu_if(i < h.hi, cc + i);
                    }
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        u32 nact = 0;
#pragma unroll
                        for (int k = 0; k < PT_W; ++k) nact += e + 32u * (PT_W * j + k) + lane < h.hi ? 1u : 0u;
                        if (__any_sync(FULL, nact != 0)) batch(std::integral_constant<int, PT_W>{}, cv + PT_W * j, nact, h.uu, sd);
                    }
                }
            }
        }
        if (fl & ~PT_HIT) atomicOr(&W.flag[par], fl & ~PT_HIT);
        // ---- flush the CTA's stage: one reservation per side
        __syncthreads();
        if (tid < 2) { const u32 c = min(s_cnt[tid], stage_cap); s_gbase[tid] = c ? atomicAdd(&W.cnt[par][tid], c) : 0u; }
        __syncthreads();
        for (u32 side = 0; side < 2; ++side) {
            const u32 c = min(s_cnt[side], stage_cap);
            u32* const qs = side ? log1 : log0;
            u32* const pp = side ? par1 : par0;
            const u32 gb = (side ? base1 : base0) + s_gbase[side];
            const u64* const stg = stage + (size_t)side * stage_cap;
            for (u32 i = tid; i < c; i += BFS_THREADS) {
                const u64 e = stg[i];
                qs[gb + i] = (u32)e;
                pp[(u32)e] = (u32)(e >> 32);
            }
        }
        __syncthreads();
        if (tid < 2) s_cnt[tid] = 0;
        ptx::cluster_sync_all();
        // ---- round end
        const u32 c0 = ptx::ld_relaxed_gpu(&W.cnt[par][0]), c1 = ptx::ld_relaxed_gpu(&W.cnt[par][1]);
        const u32 f = ptx::ld_relaxed_gpu(&W.flag[par]), mnew = ptx::ld_relaxed_gpu(&W.meet[par]);
        if (cl_rank == 0 && tid == 0) {
            const u32 z = (par + 2u) % 3u;       // last read before this barrier, next used after the next one
            if (tracing) {
                const u64 now = ptx::globaltimer_ns();
                if (e0) trace_level(K, E, 0, PH_TD_SOLO | 16u | 32u, n0, c0, W.tr_edges[par][0], now);
                if (e1) trace_level(K, E, 1, PH_TD_SOLO | 16u | 32u, n1, c1, W.tr_edges[par][1], now);
            }
            W.cnt[z][0] = 0; W.cnt[z][1] = 0; W.flag[z] = 0; W.meet[z] = NO_V; W.tr_edges[z][0] = 0; W.tr_edges[z][1] = 0;
            W.long_packed[z] = 0;
        }
        const u32 p0 = n0, p1 = n1;
        if (e0) { d0 += 1; levels += 1; base0 += c0; n0 = c0; visited += c0; }
        if (e1) { d1 += 1; levels += 1; base1 += c1; n1 = c1; visited += c1; }
        par = (par + 1u) % 3u;
        if (f & PT_HIT) {
            length = (i32)(d0 + d1); meet = mnew; depth_s = d0;
            if (e0 && e1) {
                // old-hit test: a vertex of an expanded frontier that carries the other side's mark now
                for (u32 vt = gt0; vt < span; vt += gthreads) {
                    const u32 side = vt < off1 ? 0u : 1u;
                    const u32 idx = side ? vt - off1 : vt;
                    if (idx < (side ? p1 : p0)) {
                        const u32 u = ptx::ld_relaxed_gpu((side ? log1 + flo1 : log0 + flo0) + idx);
                        const u64 w = ptx::ld_relaxed_gpu(&g_vis[u >> 5]);
                        if ((w >> ((u & 31u) + 32u - 32u * side)) & 1ull) atomicMin(&W.meet_old[side], u);
                    }
                }
                ptx::cluster_sync_all();
                const u32 m0 = ptx::ld_relaxed_gpu(&W.meet_old[0]), m1 = ptx::ld_relaxed_gpu(&W.meet_old[1]);
                if (m0 != NO_V) { length = (i32)(d0 + d1 - 1u); meet = m0; depth_s = d0 - 1u; }
                else if (m1 != NO_V) { length = (i32)(d0 + d1 - 1u); meet = m1; depth_s = d0; }
            }
            depth_t = (u32)length - depth_s;
            break;
        }
        if (n0 == 0 || n1 == 0) { length = -1; depth_s = d0; depth_t = d1; break; }
    }
    // ---- totals; every CTA leaves with the same state
    {
        u64 e = my_edges;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) e += __shfl_xor_sync(FULL, e, d);
        if (lane == 0 && e) atomicAdd(&W.edges, (unsigned long long)e);
    }
    ptx::cluster_sync_all();
    if (tid == 0) {
        o.n[0] = n0; o.n[1] = n1; o.base[0] = base0; o.base[1] = base1; o.d[0] = d0; o.d[1] = d1;
        o.levels = levels; o.visited = visited; o.length = length; o.meet = meet; o.depth_s = depth_s; o.depth_t = depth_t;
        o.go_wide = 0; o.went_wide = 1;
    }
    __syncthreads();
    return ptx::ld_relaxed_gpu(reinterpret_cast<const u64*>(&W.edges));
}
