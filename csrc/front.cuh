// front.cuh — query flow with a FRONT kernel (included by bfs_dev.cu after the search kernel).
//
//   k_front   one CTA — or one thread-block cluster of up to 16 — on the GPU that hosts rank 0.  Starts the query, runs the
//             pair tier (pair_tier.cuh: narrow rounds in the leader CTA, wide rounds on the whole cluster) and, when the
//             search ends there — every reference-sized query does, and so do grids and power-law graphs up to frontiers
//             of 65536 per side — walks the path (from the shared-memory log when the rounds stayed narrow), delivers
//             [result | path] to the caller's pinned buffer and clears the visited words it set (all CTAs that took
//             part; as a sweep once the search has covered a sizeable part of the graph).  It needs no cooperative
//             launch and no grid barrier: the fixed cost of a query (start-up + teardown of the general kernel: ~25 us
//             of a 47 us query) shrinks to the launch of a small kernel (~10 us).  When the tier has to stop it leaves
//             the search where the general kernel k_bibfs can pick it up (KParams::resume): the level state in Ctrl::st,
//             and with several ranks the HAND-OVER below.  The host launches k_bibfs only then.
//   k_front_batch   the same query flow as a loop: every cluster of the grid is a lane with its own arena that draws
//             queries from a counter (query pool), or ONE lane that stays resident and takes its queries from a mailbox
//             in pinned host memory (resident solver).
//   several ranks: rank 0's front kernel works alone on the whole graph through the peer mappings, so a small
//             query costs the other GPUs nothing: no per-level handshake, no remote claim.  The reference pays an
//             MPI_Allreduce over int[N] per level whatever the frontier size (v4/mpi_bas.cpp:107).
//             verdict   rank 0 -> every peer, VD_WORDS tagged words (payload | query number << 32) in
//                       Ctrl::verdict[query & 1]: DONE + the result, or HANDOVER + the level state of that peer.
//             hand-over rank 0 scatters what it has reached to the owners: visited bits (red.or.sys), parents,
//                       and the current frontiers into the owners' per-sender queue segments (tagged entries, the
//                       same format the fused levels use), fences, then sends the verdict.  From there on all
//                       ranks run the fused multi-GPU levels of k_bibfs.
//             k_wait    (one process per GPU) a one-warp kernel on each peer: posts "I have entered query q" to
//                       rank 0 (arrival words: rank 0 never writes into a peer that is still busy with the previous
//                       query, and never gets more than one verdict ahead), waits for the verdict and hands it to its
//                       host, which launches k_bibfs(resume) on HANDOVER.

namespace {

// wide rounds stage their discoveries in the record mirror of the narrow rounds: (vertex, parent) pairs, per side
constexpr u32 WIDE_STAGE_CAP = sizeof(PairSmemT<PT_CAP_FRONT>::fr) / 16;

struct FrontSmem {
    LevelState st;
    Acc acc;
    u32 qvalid[2];
    u32 sent[2][MAXR];     // hand-over: frontier entries written into each peer's segment, per side
    u32 own[2];            // hand-over: frontier entries of rank 0's own vertices
    u32 fail, pad;
    alignas(16) PairSmemT<PT_CAP_FRONT> ps;
};
// The kernel must stay within the 132 KB shared-memory configuration of the SM (dynamic + ~1 KB static + 1 KB reserved):
// what shared memory does not take is L1, and with the 228 KB configuration (28 KB of L1 left) every round of the front
// kernel was 15 % (narrow) to 45 % (wide, grid 4096 x 4096: 13.9 -> 20.3 ms) slower, measured; 164 KB and below are alike.
static_assert(sizeof(FrontSmem) + 2048 <= 132 * 1024, "front kernel: shared memory beyond the 132 KB configuration");

// leader warp: lane p waits until word `w` of rank 0 holds a value >= want
__device__ bool wait_arrivals(const KParams& K, Ctrl* C, u64 want, u64 deadline) {
    const u32 lane = threadIdx.x & 31;
    bool done = !(lane > 0 && lane < K.P);
    u32 spins = 0;
    while (true) {
        if (!done) done = ptx::ld_relaxed_sys(&C->arrive[lane]) >= want;
        if (__all_sync(FULL, done)) return true;
        if (++spins > 64 && (spins & 255u) == 0 && expired_warp(C, deadline)) return false;
    }
}

// leader warp of rank 0: lane p sends the verdict to peer p
__device__ void send_verdict(const KParams& K, const LevelState& st, u32 kind, u64 qn, const u32 (*sent)[MAXR]) {
    const u32 lane = threadIdx.x & 31;
    if (lane == 0 || lane >= K.P) return;
    u64* dst = ctrl_of(K, lane)->verdict[qn & 1];
    const u64 tag = (u64)(u32)qn << 32;
    u32 w[VD_WORDS];
    w[VD_KIND] = kind; w[VD_LENGTH] = (u32)st.length; w[VD_MEET] = st.meet;
    w[VD_D0] = st.depth[0]; w[VD_D1] = st.depth[1]; w[VD_LEVELS] = st.levels;
    w[VD_NF0] = st.nf[0]; w[VD_NF1] = st.nf[1];
    w[VD_SENT0] = sent ? sent[0][lane] : 0u; w[VD_SENT1] = sent ? sent[1][lane] : 0u;
    w[VD_EDGES_LO] = (u32)st.edges_total; w[VD_EDGES_HI] = (u32)(st.edges_total >> 32);
    w[VD_VIS_LO] = (u32)st.visited_total; w[VD_VIS_HI] = (u32)(st.visited_total >> 32);
    w[VD_PAIRED] = st.rounds_paired; w[VD_STATUS] = st.status;
#pragma unroll
    for (u32 i = 0; i < VD_WORDS; ++i) ptx::st_relaxed_sys(dst + i, tag | w[i]);
}

// whole warp of a peer: poll this rank's verdict words until all carry the tag of query qn
__device__ bool wait_verdict(Ctrl* C, u64 qn, u32* out /* VD_WORDS */, u64 deadline) {
    const u32 lane = threadIdx.x & 31;
    const u64* src = C->verdict[qn & 1];
    bool done = lane >= VD_WORDS;
    u64 w = 0;
    u32 spins = 0;
    bool ok = true;
    while (true) {
        if (!done) { w = ptx::ld_relaxed_sys(src + lane); done = (u32)(w >> 32) == (u32)qn; }
        if (__all_sync(FULL, done)) break;
        if (++spins > 64 && (spins & 255u) == 0 && expired_warp(C, deadline)) { ok = false; break; }
    }
#pragma unroll
    for (u32 i = 0; i < VD_WORDS; ++i) out[i] = (u32)__shfl_sync(FULL, w, i);
    return ok;
}

// level state of a peer (rank > 0) after a hand-over, from its verdict
__device__ void state_from_verdict(const KParams& K, LevelState& G, const u32* v, u32 tag) {
    G.phase = PH_TD_SOLO; G.side = 0; G.levels = v[VD_LEVELS]; G.levels_bu = 0;
    G.depth[0] = v[VD_D0]; G.depth[1] = v[VD_D1];
    G.nf[0] = v[VD_NF0]; G.nf[1] = v[VD_NF1];
    for (u32 s = 0; s < 2; ++s) {
        G.q_lo[s] = G.q_hi[s] = 0;
        G.fbm_cur[s] = 0; G.fbm_valid[s] = 0;
        for (u32 p = 0; p < MAXR; ++p) { G.rq_lo[s][p] = 0; G.rq_hi[s][p] = 0; }
        G.rq_tag[s] = tag;
    }
    G.rq_hi[0][0] = v[VD_SENT0]; G.rq_hi[1][0] = v[VD_SENT1];
    G.local_front = 0; G.hub_entries = 0; G.hub_chunks = 0; G.prev_phase = PH_TD_SOLO;
    G.length = -1; G.meet = NO_V; G.status = 0; G.wake = 0; G.pad = 0;
    G.rounds_paired = v[VD_PAIRED]; G.levels_grid = 0;
    G.edges_total = (u64)v[VD_EDGES_LO] | ((u64)v[VD_EDGES_HI] << 32);
    G.visited_total = (u64)v[VD_VIS_LO] | ((u64)v[VD_VIS_HI] << 32);
}

// Hand-over (whole CTA of rank 0's front kernel): everything the tier reached goes to its owner.
__device__ void front_handover(const KParams& K, const Env& E, FrontSmem& fs, const PairIO io, u32 tag) {
    const u32 tid = threadIdx.x;
    LevelState& st = fs.st;
    if (tid < 2 * MAXR) fs.sent[tid / MAXR][tid % MAXR] = 0;
    if (tid < 2) fs.own[tid] = 0;
    __syncthreads();
    for (u32 side = 0; side < 2; ++side) {
        const u32 total = st.q_hi[side], front_lo = st.q_lo[side];      // log window of the current frontier
        u32* const own_parent = E.parent[side];
        for (u32 i = tid; i < total; i += BFS_THREADS) {
            const u32 v = io.log[side][i];
            const u32 par = io.par[side][v];
            const u32 o = owner_of(K, v);
            if (o == E.me) {
                own_parent[v - E.lo] = par;
                if (i >= front_lo) E.q[side][ptx::atoms_add(&fs.own[side], 1u)] = v;
            } else {
                ptx::red_or_sys(reinterpret_cast<u32*>(vis_of(K, o) + (v >> 5)) + side, 1u << (v & 31u));
                ptx::st_peer(parent_of(K, o, side) + (v - o * K.B), par);
                if (i >= front_lo)
                    ptx::st_relaxed_sys(rq_of(K, o, side, E.me) + ptx::atoms_add(&fs.sent[side][o], 1u), (u64)v | ((u64)tag << 32));
            }
        }
    }
    __threadfence_system();        // bits and parents were plain / relaxed stores: order them before the verdict
    __syncthreads();
    if (tid == 0) {
        for (u32 s = 0; s < 2; ++s) {
            st.q_lo[s] = 0; st.q_hi[s] = fs.own[s];
            fs.acc.q_tail[s] = fs.own[s];
            for (u32 p = 0; p < MAXR; ++p) { st.rq_lo[s][p] = 0; st.rq_hi[s][p] = 0; E.C->sent[s][p] = fs.sent[s][p]; }
            st.rq_tag[s] = tag;
        }
    }
    __syncthreads();
}

// ---- pieces shared by the front kernel and its loop form (k_front_batch) ----------------------------------------------------

// rank 0's view of the search: its own arena, the leader CTA's shared-memory accumulators
__device__ __forceinline__ Env front_env(const KParams& K, FrontSmem& fs) {
    Env E;
    E.cta = 0; E.cpr = 1; E.me = 0; E.lo = 0;
    E.P = K.P; E.B = K.B; E.n = K.n; E.nw = K.L.nw;
    E.C = ctrl_of(K, 0);
    E.A = &fs.acc; E.sA = &fs.acc;
    E.sq = nullptr; E.sq_base = 0; E.s_hub = nullptr; E.filter = false; E.solo = true; E.tag = 0; E.smem_src = false;
    E.stage = nullptr; E.stage_cnt = nullptr; E.cta_new = nullptr; E.cta_edges = nullptr; E.s_sent = nullptr; E.s_rnew = nullptr;
    E.vis = vis_of(K, 0);
    E.parent[0] = parent_of(K, 0, 0); E.parent[1] = parent_of(K, 0, 1);
    E.q[0] = q_of(K, 0, 0); E.q[1] = q_of(K, 0, 1);
    E.row_ptr = K.row_ptr[0];
    E.col = K.col[0];
    return E;
}

// where the tier logs what it reaches: the rank's own queues / parents (single rank), the all-vertex log / parent regions
// of rank 0's arena (several ranks)
template <bool MULTI>
__device__ __forceinline__ PairIO front_io(const KParams& K, const Env& E) {
    PairIO io;
    if (MULTI) {
        u32* lg = reinterpret_cast<u32*>(K.arena[0] + K.L.off_log);
        u32* pa = reinterpret_cast<u32*>(K.arena[0] + K.L.off_pall);
        const size_t nv = (size_t)K.B * K.P;
        io.log[0] = lg; io.log[1] = lg + nv; io.par[0] = pa; io.par[1] = pa + nv;
    } else {
        io.log[0] = E.q[0]; io.log[1] = E.q[1]; io.par[0] = E.parent[0]; io.par[1] = E.parent[1];
    }
    return io;
}

// thread 0 of the leader CTA: level state and seeds of a new query (or its immediate answer: bad arguments, src == dst)
__device__ __forceinline__ void front_begin_query(const KParams& K, const Env& E, FrontSmem& fs, const PairIO& io, const u32 src, const u32 dst,
                                                  const bool bad) {
    Ctrl* const C = E.C;
    LevelState& G = fs.st;
    C->q_src = src; C->q_dst = dst;
    C->trace_count = 0;
    C->t_begin_ns = ptx::globaltimer_ns();
    fs.acc.q_tail[0] = 1; fs.acc.q_tail[1] = 1; fs.acc.new_count = 0; fs.acc.meet_local = NO_V; fs.acc.edges = 0;
    fs.fail = 0;
    G.phase = PH_TD_SOLO; G.side = 0; G.levels = 0; G.levels_bu = 0;
    G.depth[0] = G.depth[1] = 0; G.nf[0] = G.nf[1] = 1;
    for (u32 s = 0; s < 2; ++s) {
        G.q_lo[s] = 0; G.q_hi[s] = 1; G.fbm_cur[s] = 0; G.fbm_valid[s] = 0; G.rq_tag[s] = 0;
        for (u32 p = 0; p < MAXR; ++p) { G.rq_lo[s][p] = 0; G.rq_hi[s][p] = 0; }
    }
    G.length = -1; G.meet = NO_V; G.status = 0; G.wake = 0; G.local_front = 1; G.pad = 0; G.rounds_paired = 0; G.levels_grid = 0;
    G.hub_entries = 0; G.hub_chunks = 0; G.prev_phase = PH_TD_SOLO;
    G.edges_total = 0; G.visited_total = 2;
    G.xepoch = C->xepoch;
    if (bad) { G.status = 2; G.phase = PH_DONE; fs.acc.q_tail[0] = 0; fs.acc.q_tail[1] = 0; G.q_hi[0] = G.q_hi[1] = 0; }
    else if (src == dst) { G.length = 0; G.meet = src; G.phase = PH_DONE; G.visited_total = 1; fs.acc.q_tail[0] = 0; fs.acc.q_tail[1] = 0; G.q_hi[0] = G.q_hi[1] = 0; }
    else {
        // seeds: marked in rank 0's visited words whoever owns them; the owners learn about them at a hand-over
        atomicOr(reinterpret_cast<unsigned long long*>(&E.vis[src >> 5]), 1ull << (src & 31));
        atomicOr(reinterpret_cast<unsigned long long*>(&E.vis[dst >> 5]), 1ull << ((dst & 31) + 32));
        io.log[0][0] = src; io.par[0][src] = NO_V;
        io.log[1][0] = dst; io.par[1][dst] = NO_V;
    }
    if (G.phase == PH_DONE) write_result(K, C, G);
}

// counters of the wide rounds (thread 0 of the leader CTA, before the helpers are released into them)
__device__ __forceinline__ void wide_reset(Ctrl::Wide& W) {
    for (u32 z = 0; z < 3; ++z) { W.cnt[z][0] = 0; W.cnt[z][1] = 0; W.flag[z] = 0; W.meet[z] = NO_V; W.tr_edges[z][0] = 0; W.tr_edges[z][1] = 0; }
    W.meet_old[0] = NO_V; W.meet_old[1] = NO_V; W.edges = 0; W.long_packed[0] = 0; W.long_packed[1] = 0; W.long_packed[2] = 0;
}

// "not finished here": what the host (or the lane's caller) sees when the front kernel leaves a search to somebody else
__device__ __forceinline__ void escalated_result(Result& R, const LevelState& st) {
    R.length = -1; R.meet = NO_V; R.status = ST_ESCALATED; R.path_len = 0; R.levels = st.levels; R.rounds_paired = st.rounds_paired;
    R.depth_s = st.depth[0]; R.depth_t = st.depth[1]; R.levels_bu = 0; R.levels_grid = 0;
    R.edges_scanned = st.edges_total; R.vertices_visited = st.visited_total; R.search_ns = 0; R.path_ns = 0;
}

// Undo the marks of a finished search: share `part` of `parts` (the CTAs of the cluster after wide rounds; 0 of 1: the
// leader CTA alone).  A search that reached a sizeable part of the graph (a grid: millions of vertices) sweeps the visited
// words instead of looking every logged vertex up: one CTA clearing 6 M entries one by one took 5 of the 13.6 ms of a
// 4096 x 4096 grid query (measured: the peers' wait kernels saw the verdict that much earlier).
__device__ __forceinline__ void clear_marks(const Env& E, const PairIO& io, const u32 t0, const u32 t1, const u32 part, const u32 parts) {
    const u32 gt = part * BFS_THREADS + threadIdx.x, gn = parts * BFS_THREADS;
    if ((u64)(t0 + t1) * 8ull >= (u64)E.nw) {
        ulonglong2* p = reinterpret_cast<ulonglong2*>(E.vis);
        const u32 n2 = E.nw / 2;
        for (u32 i = gt; i < n2; i += gn) p[i] = make_ulonglong2(0ull, 0ull);
        if ((E.nw & 1u) && gt == 0) E.vis[E.nw - 1] = 0ull;
    } else {
        for (u32 i = gt; i < t0; i += gn) E.vis[io.log[0][i] >> 5] = 0ull;
        for (u32 i = gt; i < t1; i += gn) E.vis[io.log[1][i] >> 5] = 0ull;
    }
}

// Path of a search that ended inside the narrow rounds (whole CTA): both parent chains are walked in the shared-memory
// log (a few cycles per hop), the path goes to the arena and, when it fits, with the result straight into the caller's
// pinned buffer.  t0 / t1: log entries per side (<= PT_LOG_CAP); s_path: PT_CAP_FRONT words of scratch; s_idx: 2 words.
__device__ __noinline__ void path_from_log(const KParams& K, const unsigned long long (*slog)[PT_LOG_CAP], const u32 t0, const u32 t1,
                                           u32* s_path, u32* s_idx) {
    Ctrl* C = ctrl_of(K, K.rank_lo);
    Result& R = C->res;
    u32* hdr = reinterpret_cast<u32*>(K.arena[K.rank_lo] + K.L.off_path);   // [copy of Result | path]
    u32* path = hdr + PATH_HDR_WORDS;
    const u32 tid = threadIdx.x;
    const u64 tb = ptx::globaltimer_ns();
    const u32 L = (u32)R.length, ds = R.depth_s, meet = R.meet;
    const bool ok = L + 1u <= K.max_path;
    if (tid < 2) s_idx[tid] = 0;
    __syncthreads();
    for (u32 i = tid; i < t0; i += BFS_THREADS) if ((u32)slog[0][i] == meet) s_idx[0] = i;
    for (u32 i = tid; i < t1; i += BFS_THREADS) if ((u32)slog[1][i] == meet) s_idx[1] = i;
    __syncthreads();
    // (indices are masked into the log: a corrupted entry may give a wrong path, never a stray access)
    if (ok && tid == 0) {
        u32 idx = s_idx[0];
        for (u32 pos = ds;; --pos) {
            const unsigned long long e = slog[0][min(idx, PT_LOG_CAP - 1u)];
            path[pos] = (u32)e;
            if (pos < PT_CAP_FRONT) s_path[pos] = (u32)e;
            if (pos == 0) break;
            idx = (u32)(e >> 32);
        }
    } else if (ok && tid == 32) {
        unsigned long long e = slog[1][min(s_idx[1], PT_LOG_CAP - 1u)];
        for (u32 pos = ds; pos < L; ++pos) {
            e = slog[1][min((u32)(e >> 32), PT_LOG_CAP - 1u)];
            path[pos + 1] = (u32)e;
            if (pos + 1 < PT_CAP_FRONT) s_path[pos + 1] = (u32)e;
        }
    }
    __syncthreads();
    if (tid == 0) {
        R.path_len = ok ? L + 1u : 0u;
        const u64 dt = ptx::globaltimer_ns() - tb;
        R.path_ns = dt > 0xFFFFFFFFull ? 0xFFFFFFFFu : (u32)dt;
    }
    __syncthreads();
    const u32 plen = ok ? L + 1u : 0u;
    const u32 words = PATH_HDR_WORDS + ((K.h_out && plen <= K.h_out_words && plen <= PT_CAP_FRONT) ? plen : 0u);
    for (u32 i = tid; i < words; i += BFS_THREADS) {
        const u32 w = i < PATH_HDR_WORDS ? reinterpret_cast<const u32*>(&R)[i] : s_path[i - PATH_HDR_WORDS];
        if (i < PATH_HDR_WORDS) hdr[i] = w;
        if (K.h_out) { K.h_out[i] = w; if (!(K.flags & KF_HOST_SYNCS)) __threadfence_system(); }
    }
    if (K.h_out && K.h_seq) {
        __syncthreads();
        if (tid == 0) *reinterpret_cast<volatile u32*>(K.h_seq) = K.seq;
    }
}

}  // namespace

// ------------------------------------------------------------------------------------------
// the front kernel: one CTA on rank 0's GPU
// ------------------------------------------------------------------------------------------
template <bool MULTI>
__global__ void __launch_bounds__(BFS_THREADS, 1) k_front(const __grid_constant__ KParams Kp) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ KParams sK;          // (see k_bibfs: the callees read the parameters from shared memory, not through the parameter window)
    {
        const u32* src = reinterpret_cast<const u32*>(&Kp);
        u32* dst = reinterpret_cast<u32*>(&sK);
        for (u32 i = threadIdx.x; i < sizeof(KParams) / 4; i += BFS_THREADS) dst[i] = src[i];
    }
    __syncthreads();
    const KParams& K = sK;
    FrontSmem& fs = *reinterpret_cast<FrontSmem*>(smem_raw);
    LevelState& st = fs.st;
    const u32 tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const u32 cl_n = ptx::cluster_nctarank(), cl_rank = ptx::cluster_ctarank();
    if (blockIdx.x >= cl_n) {
        // warm-up CTAs, the clusters behind the first one (launched only for graphs that fit L2 with room to spare): stream rank 0's vertex records and its
        // visited words into L2 and leave — a cold record costs ~700 ns on the chain of a round, a hit 154 ns, and one
        // CTA could not issue the ~25,000 prefetches of a 100k-vertex graph without stalling its own rounds
        const char* rec = reinterpret_cast<const char*>(K.vrec[0]);
        const char* vw = reinterpret_cast<const char*>(vis_of(K, 0));
        const u32 vis_lines = (K.L.nw * 8u + 127u) / 128u;
        const u32 stride = (gridDim.x - cl_n) * BFS_THREADS;
        for (u32 i = (blockIdx.x - cl_n) * BFS_THREADS + tid; i < K.warm_lines + vis_lines; i += stride)
            ptx::prefetch_l2(i < K.warm_lines ? rec + 128ull * i : vw + 128ull * (i - K.warm_lines));
        return;
    }
    const Env E = front_env(K, fs);
    Ctrl* const C = E.C;
    const u64 deadline = ptx::globaltimer_ns() + (u64)K.watchdog_ms * 1000000ull;
    KT_STAMP(0);
    const PairIO io = front_io<MULTI>(K, E);
    if (cl_rank > 0) {
        // the other CTAs of the working cluster: wait until the leader CTA has decided whether its rounds go wide (it pushes
        // its state into this CTA's shared memory before it arrives at the barrier), take part if so, and leave
        ptx::cluster_sync_all();
        if (fs.ps.out.go_wide) {
            pair_wide<MULTI>(K, E, io, fs.ps.out, cl_n, cl_rank, reinterpret_cast<u64*>(&fs.ps.fr[0][0][0][0]), WIDE_STAGE_CAP);
            // the search ended in the wide rounds (or the lane gives up): every CTA clears its share of the marks
            if (fs.ps.out.length != -2 || K.lone) clear_marks(E, io, fs.ps.out.base[0], fs.ps.out.base[1], cl_rank, cl_n);
        }
        return;
    }
    bool cl_pending = cl_n > 1;       // the helpers are waiting at the start-up barrier
    // leader CTA: hand `fs.ps.out` to the helpers and release them (all threads)
    auto release_helpers = [&]() {
        if (!cl_pending) return;
        cl_pending = false;
        if (fs.ps.out.go_wide && tid == 0) {
            wide_reset(C->wide);
        }
        if (tid < 16u * (cl_n - 1u)) {
            const u32 c = 1u + tid / 16u, i = tid % 16u;
            ptx::dsmem_st(ptx::dsmem_map(ptx::smem_addr(&fs.ps.out), c) + 4u * i, reinterpret_cast<const u32*>(&fs.ps.out)[i]);
        }
        ptx::cluster_sync_all();
    };
    // ---- query start
    u32 src = K.src, dst = K.dst;
    if (K.h_query) {
        // the query sits in the caller's pinned host memory: ONE 8-byte read over PCIe by one thread, handed on through
        // shared memory (sixteen warps reading it themselves queued up behind one another: +28 us per query, measured)
        if (tid == 0) {
            const unsigned long long q = *reinterpret_cast<const volatile unsigned long long*>(K.h_query);
            fs.own[0] = (u32)q; fs.own[1] = (u32)(q >> 32);
        }
        __syncthreads();
        src = fs.own[0]; dst = fs.own[1];
        __syncthreads();
    }
    const u64 qn = K.qn;
    const bool bad = src >= K.n || dst >= K.n;
    bool tier = !bad && src != dst && !(K.flags & (BIBFS_F_FORCE_GRID | BIBFS_F_FORCE_BU));
    if (MULTI) for (u32 p = 0; p < K.P; ++p) tier = tier && K.vrec[p] != nullptr;
    else tier = tier && K.vrec[0] != nullptr;
    if (tid == 0) {
        front_begin_query(K, E, fs, io, src, dst, bad);
    }
    if (MULTI && K.async_peers && !K.lone && warp == 1) {
        // never more than one verdict ahead of a peer (the verdict slots alternate by query parity)
        if (!wait_arrivals(K, C, qn - 1, deadline) && lane == 0) fs.fail = 1;
    }
    __syncthreads();
    KT_STAMP(2);
    const bool ran_tier = st.phase != PH_DONE && !fs.fail && tier;
    if (ran_tier) {
        KT_STAMP(3);
        pair_run<MULTI, PT_CAP_FRONT>(K, E, fs.ps, io, src, dst, cl_n > 1);
        u64 edges = fs.ps.edges;
        if (fs.ps.out.go_wide) {
            release_helpers();
            edges += pair_wide<MULTI>(K, E, io, fs.ps.out, cl_n, 0u, reinterpret_cast<u64*>(&fs.ps.fr[0][0][0][0]), WIDE_STAGE_CAP);
        }
        if (tid == 0) pair_finish<MULTI>(K, E, st, fs.acc, fs.qvalid, fs.ps.out, edges);
        __syncthreads();
        KT_STAMP(4);
    }
    if (cl_pending) {                 // the rounds never went wide (or there were none): let the helpers go
        if (tid == 0) fs.ps.out.go_wide = 0;
        __syncthreads();
        release_helpers();
    }
    if (fs.fail) {
        if (tid == 0) { st.status = 1; st.length = -1; st.phase = PH_DONE; write_result(K, C, st); atomicExch(&C->abort_flag, 1u); }
        __syncthreads();
    }
    if (st.phase == PH_DONE) {
        // ---- the query ended here: verdict to the peers, path, result, cleanup
        if (MULTI && !K.lone && warp == 0 && !fs.fail) send_verdict(K, st, VK_DONE, qn, nullptr);
        KT_STAMP(5);
        const u32 t0 = fs.acc.q_tail[0], t1 = fs.acc.q_tail[1];
        // a search that never left the narrow rounds has its whole log in shared memory: walk the path and clear the
        // marks from there (no L2 round trip per hop / per entry)
        const bool in_smem = ran_tier && !fs.ps.out.went_wide && t0 <= PT_LOG_CAP && t1 <= PT_LOG_CAP;
        if (K.fuse_path == 2u) {
            if (in_smem && st.status == 0 && st.length > 0) path_from_log(K, fs.ps.slog, t0, t1, &fs.ps.fv[0][0][0], fs.own);
            else path_body_local(K, io.par[0], io.par[1]);
        } else if (K.fuse_path) result_to_host(K);
        KT_STAMP(6);
        if (in_smem) {
            for (u32 i = tid; i < t0; i += BFS_THREADS) E.vis[(u32)fs.ps.slog[0][i] >> 5] = 0ull;
            for (u32 i = tid; i < t1; i += BFS_THREADS) E.vis[(u32)fs.ps.slog[1][i] >> 5] = 0ull;
        } else {
            const bool shared = ran_tier && fs.ps.out.went_wide && fs.ps.out.length != -2;      // the helpers take part (see there)
            clear_marks(E, io, t0, t1, 0u, shared ? cl_n : 1u);
        }
        KT_STAMP(7);
        return;
    }
    // ---- the search goes on in the general kernel
    if (K.lone) {
        // pool lane: nobody to hand over to.  Undo the marks, report "handed over"; the caller answers the query elsewhere
        clear_marks(E, io, fs.acc.q_tail[0], fs.acc.q_tail[1], 0u, (ran_tier && fs.ps.out.went_wide) ? cl_n : 1u);
        if (tid == 0) {
            escalated_result(C->res, st);
        }
        __syncthreads();
        if (K.fuse_path) result_to_host(K);
        return;
    }
    if (MULTI) {
        if (K.async_peers && warp == 0 && !wait_arrivals(K, C, qn, deadline) && lane == 0) fs.fail = 1;
        __syncthreads();
        if (fs.fail) {
            if (tid == 0) { st.status = 1; st.length = -1; st.phase = PH_DONE; write_result(K, C, st); atomicExch(&C->abort_flag, 1u); }
            __syncthreads();
            if (K.fuse_path) result_to_host(K);
            return;
        }
        const u32 tag = (u32)(st.xepoch + 1);
        front_handover(K, E, fs, io, tag);
        if (warp == 0) send_verdict(K, st, VK_HANDOVER, qn, fs.sent);
        if (tid == 0) { st.xepoch += 1; C->xepoch = st.xepoch; }
        // rank 0's filter bits of other ranks' vertices stay set: "already forwarded to the owner"
    }
    __syncthreads();
    {
        u32* g = reinterpret_cast<u32*>(&C->st);
        const u32* s = reinterpret_cast<const u32*>(&st);
        for (u32 i = tid; i < sizeof(LevelState) / 4; i += BFS_THREADS) g[i] = s[i];
        if (tid == 0) {
            C->acc = fs.acc;
            escalated_result(C->res, st);
        }
    }
    __syncthreads();
    if (K.fuse_path) result_to_host(K);
}

// ------------------------------------------------------------------------------------------
// k_front_batch: MANY queries in one launch (the query pool of ctx.cu).
//
// 64 front kernels on 64 streams are bound by the launch rate (~5 us apiece: 0.32 ms of a 0.35 ms batch, measured).
// Here every cluster of the grid is a LANE: it owns one search arena (lane_K[lane], uploaded once per pool), draws the
// next query from a counter behind the query list (queries[nq], zero at launch; {src, dst} packed into 64 bits per
// query) until the list is used up, and delivers each 64-byte result into h_res[q] (pinned host memory).  A lane runs the query the way a `lone` front kernel does: pair tier, wide rounds on
// its own cluster, no peers; a search it cannot finish is reported as ST_ESCALATED and left to the caller.  Lanes never
// talk to one another, so the grid may be larger than what is resident at once.
// RESIDENT mode (mb != null, one lane): the queries come one by one through a mailbox in pinned host memory and the
// kernel stays until it is told to leave or has been idle for idle_us; [result | path] go back to the mailbox
// (bibfs_resident_* in ctx.cu).
// ------------------------------------------------------------------------------------------
template <bool MULTI>
__global__ void __launch_bounds__(BFS_THREADS, 1) k_front_batch(const KParams* __restrict__ lane_K, unsigned long long* queries,
                                                                Result* h_res, const u32 nq,
                                                                Mailbox* const mb, const u32 start_seq, const u32 idle_us) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ KParams sK;
    __shared__ u32 s_query[4];
    const u32 tid = threadIdx.x;
    const u32 cl_n = ptx::cluster_nctarank(), cl_rank = ptx::cluster_ctarank();
    constexpr u32 LANE_EXIT = 0xFFFFFFFFu;      // PairOut::go_wide value that sends the helpers home
    const u32 lane_id = blockIdx.x / cl_n;
    {
        const u32* src = reinterpret_cast<const u32*>(lane_K + lane_id);
        u32* dst = reinterpret_cast<u32*>(&sK);
        for (u32 i = tid; i < sizeof(KParams) / 4; i += BFS_THREADS) dst[i] = src[i];
    }
    __syncthreads();
    const KParams& K = sK;
    FrontSmem& fs = *reinterpret_cast<FrontSmem*>(smem_raw);
    LevelState& st = fs.st;
    const Env E = front_env(K, fs);
    Ctrl* const C = E.C;
    const PairIO io = front_io<MULTI>(K, E);
    u64* const wide_stage = reinterpret_cast<u64*>(&fs.ps.fr[0][0][0][0]);
    bool tier_ok = true;
    if (MULTI) for (u32 p = 0; p < K.P; ++p) tier_ok = tier_ok && K.vrec[p] != nullptr;
    else tier_ok = K.vrec[0] != nullptr;

    u32 last_seq = start_seq;        // (thread 0 of the leader CTA) sequence number of the last mailbox query taken
    auto push_out = [&]() {          // leader CTA: fs.ps.out into every helper's shared memory, then the barrier they wait at
        if (tid < 16u * (cl_n - 1u)) {
            const u32 c = 1u + tid / 16u, i = tid % 16u;
            ptx::dsmem_st(ptx::dsmem_map(ptx::smem_addr(&fs.ps.out), c) + 4u * i, reinterpret_cast<const u32*>(&fs.ps.out)[i]);
        }
        ptx::cluster_sync_all();
    };
    while (true) {
        if (cl_rank > 0) {
            // helpers: the leader pushes its state into fs.ps.out before it arrives at this barrier
            ptx::cluster_sync_all();
            const u32 go = fs.ps.out.go_wide;
            if (go == LANE_EXIT) break;
            if (go) {
                pair_wide<MULTI>(K, E, io, fs.ps.out, cl_n, cl_rank, wide_stage, WIDE_STAGE_CAP);
                const u32 t0 = fs.ps.out.base[0], t1 = fs.ps.out.base[1];
                ptx::cluster_sync_all();        // nobody still reads its copy of the state when the leader pushes the next one
                clear_marks(E, io, t0, t1, cl_rank, cl_n);
                ptx::cluster_sync_all();        // ... and no stale clear lands in the marks of the lane's next query
            }
            continue;
        }
        if (tid == 0 && mb) {
            // resident mode: poll the mailbox in the host's pinned memory (one 16-byte load per PCIe round trip) until a new
            // sequence number shows up; leave when told to, or after idle_us without a query
            const u64 t_idle = ptx::globaltimer_ns();
            uint4 w;
            u32 spins = 0;
            bool leave = false;
            while (true) {
                w = ptx::ld_v4_sys(reinterpret_cast<const uint4*>(mb));
                if (w.z != last_seq) { leave = (w.w & MB_EXIT) != 0; break; }
                if ((++spins & 15u) == 0 && ptx::globaltimer_ns() - t_idle > (u64)idle_us * 1000ull) { leave = true; break; }
            }
            s_query[2] = leave ? 1u : 0u;
            if (!leave) {
                last_seq = w.z;
                s_query[0] = w.x; s_query[1] = w.y;
                sK.seq = w.z;
                sK.flags = w.w & BIBFS_F_NO_PATH;
                sK.fuse_path = (w.w & BIBFS_F_NO_PATH) ? 1u : 2u;
            }
        } else if (tid == 0) {
            const u32 q = (u32)atomicAdd(queries + nq, 1ull);
            s_query[2] = q < nq ? 0u : 1u;
            if (q < nq) {
                const u64 w = ptx::ld_relaxed_gpu(reinterpret_cast<const u64*>(queries + q));
                s_query[0] = (u32)w; s_query[1] = (u32)(w >> 32);
                sK.h_out = reinterpret_cast<u32*>(h_res + q);
            }
        }
        __syncthreads();
        if (s_query[2]) {
            if (mb && tid == 0) { __threadfence_system(); *reinterpret_cast<volatile u32*>(&mb->alive) = 0u; }
            if (cl_n > 1) {
                if (tid == 0) fs.ps.out.go_wide = LANE_EXIT;
                __syncthreads();
                push_out();
            }
            break;
        }
        const u32 src = s_query[0], dst = s_query[1];
        const bool bad = src >= K.n || dst >= K.n;
        const bool tier = tier_ok && !bad && src != dst;
        if (tid == 0) {
            front_begin_query(K, E, fs, io, src, dst, bad);
            fs.ps.out.go_wide = 0;
        }
        __syncthreads();
        bool wide = false;
        if (tier) {
            pair_run<MULTI, PT_CAP_FRONT>(K, E, fs.ps, io, src, dst, cl_n > 1);
            wide = fs.ps.out.go_wide != 0;
        }
        if (cl_n > 1) {
            // hand fs.ps.out to the helpers and release them
            if (wide && tid == 0) {
                wide_reset(C->wide);
            }
            __syncthreads();
            push_out();
        }
        if (tier) {
            u64 edges = fs.ps.edges;
            if (wide) {
                edges += pair_wide<MULTI>(K, E, io, fs.ps.out, cl_n, 0u, wide_stage, WIDE_STAGE_CAP);
                ptx::cluster_sync_all();
            }
            if (tid == 0) pair_finish<MULTI>(K, E, st, fs.acc, fs.qvalid, fs.ps.out, edges);
            __syncthreads();
        }
        // ---- result (or "not finished here") and, in resident mode, the path to the host; then undo the marks (from the
        // shared-memory log when the search never left the narrow rounds)
        if (st.phase != PH_DONE && tid == 0) {
            escalated_result(C->res, st);
        }
        __syncthreads();
        const u32 t0 = fs.acc.q_tail[0], t1 = fs.acc.q_tail[1];
        const bool in_smem = tier && !fs.ps.out.went_wide && t0 <= PT_LOG_CAP && t1 <= PT_LOG_CAP;
        if (K.fuse_path == 2u && st.phase == PH_DONE) {
            if (in_smem && st.status == 0 && st.length > 0) path_from_log(K, fs.ps.slog, t0, t1, &fs.ps.fv[0][0][0], fs.own);
            else path_body_local(K, io.par[0], io.par[1]);
        } else result_to_host(K);
        if (in_smem) {
            for (u32 i = tid; i < t0; i += BFS_THREADS) E.vis[(u32)fs.ps.slog[0][i] >> 5] = 0ull;
            for (u32 i = tid; i < t1; i += BFS_THREADS) E.vis[(u32)fs.ps.slog[1][i] >> 5] = 0ull;
        } else clear_marks(E, io, t0, t1, 0u, wide ? cl_n : 1u);
        if (wide) ptx::cluster_sync_all();
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------
// k_wait: a peer (rank > 0) in the one-process-per-GPU deployment; one warp
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(32) k_wait(const __grid_constant__ KParams K) {
    const u32 lane = threadIdx.x;
    const u32 me = K.rank_lo;
    Ctrl* const C = ctrl_of(K, me);
    const u64 deadline = ptx::globaltimer_ns() + (u64)K.watchdog_ms * 1000000ull;
    const u64 qn = K.qn;
    if (lane == 0) { C->t_begin_ns = ptx::globaltimer_ns(); ptx::st_relaxed_sys(&ctrl_of(K, 0)->arrive[me], qn); }
    u32 v[VD_WORDS];
    const bool ok = wait_verdict(C, qn, v, deadline);
    if (lane == 0) {
        Result& R = C->res;
        R.length = -1; R.meet = NO_V; R.path_len = 0; R.levels_bu = 0; R.levels_grid = 0; R.path_ns = 0;
        if (!ok) { R.status = 1; atomicExch(&C->abort_flag, 1u); }
        else {
            R.depth_s = v[VD_D0]; R.depth_t = v[VD_D1]; R.levels = v[VD_LEVELS]; R.rounds_paired = v[VD_PAIRED];
            R.edges_scanned = (u64)v[VD_EDGES_LO] | ((u64)v[VD_EDGES_HI] << 32);
            R.vertices_visited = (u64)v[VD_VIS_LO] | ((u64)v[VD_VIS_HI] << 32);
            if (v[VD_KIND] == VK_DONE) { R.length = (i32)v[VD_LENGTH]; R.meet = v[VD_MEET]; R.status = (i32)v[VD_STATUS]; }
            else R.status = ST_ESCALATED;
        }
        const u64 dt = ptx::globaltimer_ns() - C->t_begin_ns;
        R.search_ns = dt > 0xFFFFFFFFull ? 0xFFFFFFFFu : (u32)dt;
    }
    __syncwarp();
    if (K.h_out) {      // the verdict goes straight to the host's pinned buffer, then the completion word it may be polling
        if (lane < PATH_HDR_WORDS) { K.h_out[lane] = reinterpret_cast<const u32*>(&C->res)[lane]; if (!(K.flags & KF_HOST_SYNCS)) __threadfence_system(); }
        __syncwarp();
        if (lane == 0 && K.h_seq) *reinterpret_cast<volatile u32*>(K.h_seq) = K.seq;
    }
}
