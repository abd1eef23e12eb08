// bfs.h — device-resident data model of a search context (shared by bfs_dev.cu and ctx.cu).
//
// One ARENA per rank holds every piece of per-query state.  All arenas have the same layout, so
// a rank reaches a peer's bitmap / parents / queues by adding the same offsets to the peer's
// base pointer (mapped through CUDA IPC or cudaDeviceEnablePeerAccess).
//
//   Ctrl                      control block: query, barriers, exchange slots, level state, result
//   vis   u64[NW]             word w = vertices 32w..32w+31: low 32 bits "reached from source",
//                             high 32 bits "reached from target".  NW covers ALL vertices: the
//                             slice of the owned vertices is authoritative, the rest is a local
//                             "already forwarded to its owner" filter.
//   fbm   u32[2][2][NW]       frontier bitmaps for bottom-up levels (side x double buffer), replicated
//   parent u32[2][B]          BFS-tree parent of owned vertices per side (global ids)
//   q     u32[2][B]           append-only queue of owned vertices discovered locally, per side
//   rq    u64[2][P][B]        (P>1) per-sender segments: owned vertices discovered by rank p, each entry
//                             = vertex | epoch tag << 32 so the owner can tell a landed entry from a stale one
//   path  u32[max_path]       path output (rank 0 of the context group writes it)
//   trace LevelTrace[TRACE_CAP]
// The frontier of a side is a WINDOW [lo,hi) of its queue plus one window per sender segment;
// nothing is copied or compacted between levels.
#pragma once
#include "common.h"

namespace bibfs {

constexpr u32 MAXR = BIBFS_MAX_RANKS;
constexpr u32 NO_V = 0xFFFFFFFFu;
constexpr u32 TRACE_CAP = 32768;
constexpr int BFS_THREADS = 512;     // 512 threads x 1 CTA per SM: up to 128 registers, so eight claims per lane stay in flight

enum Phase : u32 { PH_TD_SOLO = 0, PH_TD_GRID = 1, PH_BU = 2, PH_DONE = 3, PH_TD_HUB = 4 };

struct Slot {          // what a rank tells every peer at the end of a level
    u32 count;         // vertices this rank discovered in the level (local + remote)
    u32 meet;          // smallest meeting vertex it saw, NO_V if none
    u32 rq_tail;       // cumulative entries it has appended to the receiver's segment (this side)
    u32 edges;         // adjacency entries it inspected (low 32 bits, statistics)
};
// On the wire every 32-bit field travels in its own 8-byte word together with the 32-bit epoch
// tag of the exchange ("LL" style): the receiver polls the words until all tags match, so no
// release fence / acquire load pair is needed (fence.acq_rel.sys costs ~1.35 us on B200 and an
// acquire/release ping-pong 5.3 us round trip, against 2.4 us for relaxed stores and polls).
struct LSlot {
    u64 w[4];
};

struct LevelState {
    u32 phase;
    u32 side;
    u32 levels, levels_bu;
    u32 depth[2];
    u32 nf[2];                 // global frontier size per side
    u32 q_lo[2], q_hi[2];      // local queue window per side
    u32 rq_lo[2][MAXR], rq_hi[2][MAXR];
    u32 rq_tag[2];             // epoch tag carried by the entries of the current rq windows, per side
    u32 fbm_cur[2], fbm_valid[2];
    u32 local_front;           // vertices this rank expands in the coming level
    u32 hub_entries, hub_chunks; // hub phase work (PH_TD_HUB)
    u32 prev_phase;            // phase that produced the hub list (solo or grid)
    i32 length;
    u32 meet;
    u32 status;
    u32 wake;                  // number of local wake-ups issued so far (gen low word)
    u32 rounds_paired;         // levels run by the pair tier (pair_tier.cuh)
    u32 levels_grid, pad;      // levels that needed the whole grid (grid top-down, hub phase, bottom-up)
    u64 xepoch;                // cross-rank barrier epoch
    u64 edges_total;           // adjacency entries inspected so far (all ranks)
    u64 visited_total;         // vertices reached so far (all ranks, both sides)
};

struct alignas(16) Result {   // 64 bytes (PATH_HDR_WORDS words)
    i32 length;
    u32 meet;
    u32 depth_s, depth_t;
    u32 levels, levels_bu;
    u32 path_len;
    i32 status;
    u64 edges_scanned;
    u64 vertices_visited;
    u32 levels_grid;           // levels that ran on the whole grid: feeds the launch-geometry policy of ctx.cu
    u32 rounds_paired;         // levels run by the pair tier
    u32 search_ns, path_ns;    // device-measured duration of the search and of the path walk (saturating)
};

// level accumulators: one copy in Ctrl (grid mode), one in shared memory (solo mode)
struct Acc {
    u32 q_tail[2];     // append cursor of the local queue per side (persists through the query)
    u32 new_count;     // vertices discovered in the current level
    u32 meet_local;    // smallest meeting vertex seen in the current level
    unsigned long long edges;   // adjacency entries inspected in the current level
};

struct HubEntry {      // a frontier vertex whose row is expanded by the whole grid
    u32 u;
    u32 nchunks;
    u64 r0, r1;
    u64 chunk_base;    // first global chunk id of this row
};
constexpr u32 HUB_ROW = 1024;     // rows at least this long are deferred to the hub phase
constexpr u32 HUB_CHUNK = 256;    // adjacency entries per hub work item (1 KB: one bulk copy, 16-byte aligned in col)
// chunk i of a row [r0,r1) covers col[A + i*HUB_CHUNK, A + (i+1)*HUB_CHUNK) with A = r0 rounded down to 4 entries,
// so that every chunk starts on a 16-byte boundary (cp.async.bulk requirement)
BIBFS_HD u32 hub_nchunks(u64 r0, u64 r1) { return (u32)((r1 - (r0 & ~3ull) + HUB_CHUNK - 1) / HUB_CHUNK); }
constexpr u32 BAR_WAYS = 16;
// verdict of the front kernel, one tagged word per field
enum : u32 { VD_KIND = 0, VD_LENGTH, VD_MEET, VD_D0, VD_D1, VD_LEVELS, VD_NF0, VD_NF1, VD_SENT0, VD_SENT1, VD_EDGES_LO, VD_EDGES_HI,
             VD_VIS_LO, VD_VIS_HI, VD_PAIRED, VD_STATUS, VD_WORDS };
enum : u32 { VK_DONE = 1, VK_HANDOVER = 2 };
constexpr u32 KF_HOST_SYNCS = 0x40000000u;   // KParams::flags (internal): the host waits for the STREAM before it reads the pinned result, so the
                                             // kernel's writes to it need no fence of their own (a fence.sys is a PCIe round trip on the query's chain)
constexpr i32 ST_ESCALATED = 100;    // Result::status of a front kernel that handed the search over (never reaches the caller)
constexpr u32 PATH_HDR_WORDS = 16;   // sizeof(Result) / 4
constexpr u32 HUB_CAP = 16384;          // hub entries per level per rank (overflow: expanded in place)
constexpr u32 DUMMY_WORDS = 512 * 512;   // one private scratch word per thread of up to 512 CTAs: target of no-op atomics

struct Ctrl {
    // written by the host (H2D copy from pinned memory) before the launch
    alignas(128) u32 q_src;
    u32 q_dst;
    // local grid barrier
    // arrive counters: CTA c bumps counter c % BAR_WAYS (one 128-byte line each); 148 arrivals on ONE word serialise
    // in the L2 atomic unit for ~2 us, sixteen words are bumped in parallel
    struct alignas(128) BarWay { u32 count; } bar_way[BAR_WAYS];
    alignas(128) u64 bar_gen;            // (launch seq << 32) | wake-ups
    // cross-rank exchange: slot[parity][p] is written by rank p (tagged words, relaxed stores)
    alignas(128) LSlot slot[2][MAXR];    // [epoch parity][sender]
    // accumulators, reset by the leader each level / query
    alignas(128) Acc acc;
    unsigned long long hub_packed;       // (entries << 40) | chunks, appended atomically
    u32 sent[2][MAXR];
    unsigned long long edges_total;
    // wide rounds of the front kernel's cluster (pair_tier.cuh): counters shared by its CTAs, three rotating sets
    struct alignas(128) Wide {
        u32 cnt[3][2];
        u32 flag[3];
        u32 meet[3];
        u32 tr_edges[3][2];
        u32 meet_old[2];
        unsigned long long edges;
        unsigned long long long_packed[3];   // rows of >= PT_MAXDEG entries met in the round: (entries << 40) | chunks
    } wide;
    // front kernel (k_front) <-> peers: see "query flow" in bfs_dev.cu
    alignas(128) u64 arrive[MAXR];       // at rank 0: arrive[p] = queries rank p has entered (one process per GPU only)
    alignas(128) u64 verdict[2][VD_WORDS];   // written by rank 0: payload | (u32)query number << 32, [query number & 1]
    // persisted across launches
    alignas(128) u64 xepoch;
    u32 abort_flag;
    u32 trace_count;
    u64 t_begin_ns;            // device clock (%globaltimer) when the search proper began
    alignas(128) LevelState st;
    alignas(128) Result res;
};

struct Layout {
    size_t off_vis, off_fbm, off_parent, off_q, off_rq, off_path, off_trace, off_hub, off_dummy, total;
    size_t off_log, off_pall, total0;   // P > 1, rank 0 only: log of the vertices its front kernel reached and their parents (all vertices)
    u32 nw;           // bitmap words covering P*B vertices
};

inline Layout make_layout(u32 B, u32 P, u32 max_path) {
    auto al = [](size_t x) { return (x + 511) / 512 * 512; };
    Layout L;
    L.nw = (u32)(((u64)B * P) / 32);
    size_t o = al(sizeof(Ctrl));
    L.off_vis = o;    o += al((size_t)L.nw * 8);
    L.off_fbm = o;    o += al((size_t)L.nw * 4 * 4);
    L.off_parent = o; o += al((size_t)B * 4 * 2);
    L.off_q = o;      o += al((size_t)B * 4 * 2);
    L.off_rq = o;     o += (P > 1) ? al((size_t)B * 8 * 2 * P) : 0;   // tagged 8-byte entries
    L.off_path = o;   o += al((size_t)max_path * 4 + PATH_HDR_WORDS * 4);   // [copy of Result | path]: one D2H fetches both
    L.off_trace = o;  o += al((size_t)TRACE_CAP * sizeof(bibfs_level_trace));
    L.off_hub = o;    o += al((size_t)HUB_CAP * sizeof(HubEntry));
    L.off_dummy = o;  o += al((size_t)DUMMY_WORDS * 8);
    L.total = o;
    L.off_log = o;    o += (P > 1) ? al((size_t)B * P * 4 * 2) : 0;
    L.off_pall = o;   o += (P > 1) ? al((size_t)B * P * 4 * 2) : 0;
    L.total0 = o;
    return L;
}

struct KParams {
    u32 n, B, P;
    u32 rank_lo, vranks, cpr;      // ranks hosted by this launch, CTAs per rank
    u32 cluster;                   // thread-block cluster size of the launch (single rank; 1 = no clusters)
    u32 solo_threshold, bu_div, flags, max_path;
    u32 seq;                       // launch sequence number (local barrier generation prefix)
    u32 watchdog_ms;
    u32 fuse_path;                 // single rank, work of the leader CTA at the end of the search kernel: 0 nothing, 1 copy the result to h_out,
                                   // 2 walk the parent chains first (no path kernel)
    u32 src, dst, query_in_ctrl;   // query by value, or (query_in_ctrl) read from Ctrl::q_src/q_dst
    u32 resume;                    // k_bibfs continues a search the front kernel handed over (no query start)
    u32 async_peers;               // the other ranks are driven by other processes: arrival words order the queries
    u32 lone;                      // front kernel of a query-pool lane: rank 0 works without its peers (no verdicts, no hand-over;
                                   // a search it cannot finish is reported back as handed over and answered by the group)
    u32 long_rows;                 // the graph has rows of >= 32 entries (or nobody knows): wide rounds deal them over the cluster
    u32 wide_min;                  // front kernel: frontier positions of a round from which its whole cluster takes the round
    u64 qn;                        // number of this query in the context group (host counter, identical on every rank)
    u64 div_B;                     // 2^64 / B rounded up: owner(v) = umulhi64(div_B, v), exact for 32-bit v
    const u32* h_query;            // pinned host memory holding {src,dst}: read by the kernel itself (zero copy), or null
    u32* h_out;                    // pinned host buffer [Result | first path entries] written by the path kernel, or null
    u32 h_out_words;               // path entries that fit in h_out
    u32* h_seq;                    // pinned word that receives K.seq after h_out is complete (host polls it), or null
    Layout L;
    char* arena[MAXR];             // base of every rank's arena as seen from this device
    const u64* row_ptr[MAXR];      // CSR shards of the hosted ranks (index = rank)
    const u32* col[MAXR];
    const uint4* vrec[MAXR];       // vertex records of the shards (pair tier), or null
    u32 warm_lines;                // > 0: the graph is small enough for L2 — idle CTAs prefetch this many 128-byte lines of vrec at query start
};

// Mailbox of a resident front kernel (pinned host memory, mapped into the device's address space).
struct alignas(64) Mailbox {
    // host -> device: one 16-byte word group, `seq` written last (x86 stores stay in order; the device reads all 16 bytes
    // with one load)
    u32 src, dst, seq, op;                 // op: MB_QUERY | flags, or MB_EXIT
    u32 pad0[12];
    // device -> host
    alignas(64) u32 alive;                 // 1 while the kernel polls; 0 once it has left (idle or told to)
    u32 out_seq;                           // == seq of the query whose [result | path] is complete in `out`
    u32 pad1[14];
    alignas(64) u32 out[PATH_HDR_WORDS + 240];
};
constexpr u32 MB_EXIT = 0x80000000u;
static_assert(offsetof(Mailbox, seq) == 8 && offsetof(Mailbox, alive) == 64 && offsetof(Mailbox, out_seq) == 68 && offsetof(Mailbox, out) == 128,
              "mailbox layout: the query is one 16-byte load, host-written and device-written words sit in different 64-byte lines");

// implemented in bfs_dev.cu
int bfs_max_ctas_per_device(bool multi, int* out);
int bfs_max_ctas_clustered(u32 cluster, int* out);
#if defined(__CUDACC__) || defined(BIBFS_WITH_CUDA_RUNTIME)
int bfs_launch(const KParams& K, bool multi, cudaStream_t st);
int front_launch(const KParams& K, bool multi, u32 cluster, cudaStream_t st);
int wait_launch(const KParams& K, cudaStream_t st);
int front_batch_prepare(bool multi);
int front_batch_launch(const KParams* d_lane_K, unsigned long long* d_queries, Result* h_res, u32 nq, u32 lanes, bool multi, u32 cluster, cudaStream_t st);
int front_resident_launch(const KParams* d_lane_K, Mailbox* mb, u32 start_seq, u32 idle_us, u32 cluster, cudaStream_t st);
int front_prepare(bool multi, u32* cluster_out);
#endif

}  // namespace bibfs
