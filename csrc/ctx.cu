// ctx.cu — search-context management: arenas, peer mapping (CUDA IPC or in-process peer access),
// query launch, result / timing / trace read-back.
//
// This is the counterpart of the reference's v4 C API (v4/comp.h:8-36, v4/comp.cu:49-133:
// cudaInitGraph / cudaInitFrontiers / cudaExpandFrontier / cudaCheckIntersect / cudaFreeGraph)
// with three differences: state lives in a handle instead of file-static device pointers
// (comp.cu:13-17) so several graphs / ranks can coexist in one process; a device is selected
// explicitly (v4 never calls cudaSetDevice); and the per-level host calls are gone —
// bibfs_ctx_query enqueues the whole search.
#include "bfs.h"
#include <atomic>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <vector>

using namespace bibfs;

struct bibfs_ctx {
    bibfs_ctx_desc d;
    int device = 0;
    u32 B = 0;
    Layout L;
    char* arena_local[MAXR] = {};   // arenas allocated here (index = hosted slot)
    char* arena_all[MAXR] = {};     // every rank's arena as seen from this device (index = rank)
    bool ipc_opened[MAXR] = {};
    u32 seq = 0;
    u32 cpr = 0;
    u32 cluster = 1;
    u32 cpr_small = 0;           // BIBFS_GEOMETRY=small: launch the general kernel as one cluster (measured: no gain, kept as a switch)
    int geometry = 0;
    // ---- query flow (front.cuh): front kernel first, general kernel only when the front kernel hands the search over
    const u32* col_all[MAXR] = {};       // every rank's adjacency / vertex records as seen from this device
    const void* vrec_all[MAXR] = {};
    bool hosts0 = false;         // this context hosts rank 0 (runs the front kernel, walks the path, delivers the result)
    bool hosts_all = false;      // ... and every other rank too (single rank, or virtual ranks)
    bool lone = false;           // lane of a query pool on a partitioned graph: rank 0 without its peers (front kernel only)
    bool async_peers = false;    // peers live in other processes (attached through IPC): each rank resumes on its own verdict
    u32 front_cluster = 1;       // thread-block cluster of the front kernel (wide rounds)
    // the cluster launch costs ~1.5 us more than a single CTA: contexts whose queries never go wide drop it, and take it
    // back when a single-CTA front kernel has to hand a search over
    bool front_narrow = true;        // start with the single CTA: the first search that outgrows it brings the cluster in
    int no_wide_streak = 0;
    int front_mode = 0;          // BIBFS_FRONT: 0 policy, 1 always, 2 never
    int front_score = 0;         // policy: rises when the front kernel hands over after a few levels (hub graphs), falls otherwise
    u64 qn = 0;                  // queries issued (identical on every rank of the group)
    int stage = 0;               // what the last bibfs_ctx_query launched: 0 general kernel, 1 front kernel / wait kernel, 2 nothing yet (sync peer)
    u32 pend_src = 0, pend_dst = 0, pend_flags = 0;
    const u32* pend_hq = nullptr;
    bool last_fused = false;     // the last query wrote its result (and path) to h_stage from inside the search kernel
    bool last_lean = false;      // ... and was launched without events (BIBFS_F_DEVICE_CLOCK)
    u32* h_seq = nullptr;        // completion word inside h_stage
    bool multi = false;
    cudaEvent_t ev[3] = {};
    char* h_stage = nullptr;        // pinned: Result + path
    size_t h_stage_bytes = 0;
    u32 last_flags = 0;
};

namespace {
constexpr u32 HOST_PATH_WORDS = 240;   // path entries the path kernel writes straight into the pinned result buffer
struct DeviceGuard {
    int prev = -1;
    bool active = false;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) == cudaSuccess && prev != dev) { cudaSetDevice(dev); active = true; }
    }
    ~DeviceGuard() { if (active) cudaSetDevice(prev); }
};
size_t zero_prefix(const Layout& L) { return L.off_parent; }   // Ctrl + vis + fbm
}  // namespace

extern "C" size_t bibfs_ctx_arena_bytes(const bibfs_ctx* ctx) { return ctx ? (ctx->hosts0 ? ctx->L.total0 : ctx->L.total) : 0; }

extern "C" int bibfs_ctx_create(bibfs_ctx** out, const bibfs_ctx_desc* desc) {
    if (!out || !desc) return fail("NULL argument");
    if (desc->n == 0 || desc->n > 0xFFFFFF00u) return fail("n out of range");
    if (desc->nranks < 1 || desc->nranks > MAXR) return fail("nranks must be 1..%u", MAXR);
    if (desc->vranks < 1 || desc->rank_lo + desc->vranks > desc->nranks) return fail("bad hosted rank range");
    for (u32 i = 0; i < desc->vranks; ++i)
        if (!desc->d_row_ptr[i]) return fail("hosted rank %u has no CSR", desc->rank_lo + i);
    for (u32 i = 0; i < desc->vranks; ++i)
        if (reinterpret_cast<uintptr_t>(desc->d_col[i]) & 15u)
            return fail("d_col of hosted rank %u is not 16-byte aligned (hub rows are fetched with cp.async.bulk)", desc->rank_lo + i);
    int coop = 0, dev = 0;
    BIBFS_CUDA_TRY(cudaGetDevice(&dev));
    BIBFS_CUDA_TRY(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
    if (!coop) return fail("device %d does not support cooperative launches", dev);

    bibfs_ctx* c = new bibfs_ctx();
    c->d = *desc;
    c->device = dev;
    c->multi = desc->nranks > 1;
    c->B = bibfs_block_size(desc->n, desc->nranks);
    if (c->d.max_path == 0) c->d.max_path = 1u << 16;
    if (c->d.solo_threshold == 0 || c->d.solo_threshold > (u32)BFS_THREADS) c->d.solo_threshold = BFS_THREADS;   // one frontier vertex per thread
    if (c->d.bottom_up_div == 0) c->d.bottom_up_div = 24;
    if (c->d.watchdog_ms == 0) c->d.watchdog_ms = 4000;
    c->L = make_layout(c->B, desc->nranks, c->d.max_path);

    int maxc = 0, sms = 0;
    if (bfs_max_ctas_per_device(c->multi, &maxc)) { delete c; return 1; }
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    u32 per_rank = (u32)(sms < maxc ? sms : maxc) / desc->vranks;   // one CTA per SM keeps the barrier short
    if (desc->max_ctas && desc->max_ctas < per_rank) per_rank = desc->max_ctas;
    if (per_rank < 1) { delete c; return fail("not enough SMs for %u hosted ranks", desc->vranks); }
    c->cpr = per_rank;
    // single rank: launch in thread-block clusters (16 - beyond the portable 8, B200 takes it - else 8, 4, 2) so that the leader CTA can enlist its neighbours over
    // DSMEM: frontiers up to cluster x 512 vertices stay off the grid-wide barrier.  BIBFS_CLUSTER=n overrides.
    if (!c->multi && desc->vranks == 1 && per_rank >= 8 && !getenv("BIBFS_NO_CLUSTER")) {
        const char* want = getenv("BIBFS_CLUSTER");
        const u32 first = want ? (u32)atoi(want) : 16u;
        for (u32 cl = first; cl >= 2; cl /= 2) {
            if (cl > 16 || (cl & (cl - 1))) continue;
            int fit = 0;
            if (bfs_max_ctas_clustered(cl, &fit) == 0 && fit >= (int)(2 * cl)) {
                c->cluster = cl;
                c->cpr = (per_rank < (u32)fit ? per_rank : (u32)fit) / cl * cl;
                c->cpr_small = cl;
                if (c->d.solo_threshold == (u32)BFS_THREADS) c->d.solo_threshold = cl * BFS_THREADS;
                break;
            }
            cudaGetLastError();
        }
        cudaGetLastError();
    }

    if (const char* g = getenv("BIBFS_GEOMETRY")) c->geometry = !strcmp(g, "small") ? 1 : !strcmp(g, "full") ? 2 : 0;
    if (const char* f = getenv("BIBFS_FRONT")) c->front_mode = !strcmp(f, "1") ? 1 : !strcmp(f, "0") ? 2 : 0;
    c->hosts0 = desc->rank_lo == 0;
    c->hosts_all = desc->vranks == desc->nranks;
    for (u32 i = 0; i < desc->vranks; ++i) { c->col_all[desc->rank_lo + i] = desc->d_col[i]; c->vrec_all[desc->rank_lo + i] = desc->d_vrec[i]; }
    if (c->hosts0 && front_prepare(c->multi, &c->front_cluster)) { delete c; return 1; }
    if (getenv("BIBFS_FRONT_CLUSTER") || c->front_cluster == 1) c->front_narrow = false;
    for (u32 i = 0; i < desc->vranks; ++i)
        if (reinterpret_cast<uintptr_t>(desc->d_vrec[i]) & 31u) { delete c; return fail("d_vrec of hosted rank %u is not 32-byte aligned", desc->rank_lo + i); }
    for (u32 i = 0; i < desc->vranks; ++i) {
        const size_t bytes = (desc->rank_lo + i == 0) ? c->L.total0 : c->L.total;     // rank 0 also holds the front kernel's log
        cudaError_t e = cudaMalloc((void**)&c->arena_local[i], bytes);
        if (e == cudaSuccess) e = cudaMemset(c->arena_local[i], 0, zero_prefix(c->L));
        if (e != cudaSuccess) {
            for (u32 j = 0; j <= i; ++j) if (c->arena_local[j]) cudaFree(c->arena_local[j]);
            delete c;
            return fail("arena allocation of %zu bytes failed: %s", c->L.total, cudaGetErrorString(e));
        }
        c->arena_all[desc->rank_lo + i] = c->arena_local[i];
    }
    for (int k = 0; k < 3; ++k) cudaEventCreate(&c->ev[k]);
    c->h_stage_bytes = 256 + sizeof(Result) + (size_t)c->d.max_path * 4;
    if (cudaHostAlloc((void**)&c->h_stage, c->h_stage_bytes, cudaHostAllocDefault) != cudaSuccess) {
        bibfs_ctx_destroy(c);
        return fail("pinned staging allocation failed");
    }
    memset(c->h_stage, 0, c->h_stage_bytes);
    c->h_seq = reinterpret_cast<u32*>(c->h_stage + sizeof(Result) + (size_t)c->d.max_path * 4);   // in the 256-byte tail
    BIBFS_CUDA_TRY(cudaDeviceSynchronize());
    *out = c;
    return 0;
}

extern "C" int bibfs_ctx_destroy(bibfs_ctx* c) {
    if (!c) return 0;
    DeviceGuard g(c->device);
    cudaDeviceSynchronize();
    for (u32 r = 0; r < MAXR; ++r)
        if (c->ipc_opened[r] && c->arena_all[r]) cudaIpcCloseMemHandle(c->arena_all[r]);
    for (u32 i = 0; i < MAXR; ++i)
        if (c->arena_local[i]) cudaFree(c->arena_local[i]);
    for (int k = 0; k < 3; ++k) if (c->ev[k]) cudaEventDestroy(c->ev[k]);
    if (c->h_stage) cudaFreeHost(c->h_stage);
    delete c;
    return 0;
}

extern "C" int bibfs_ctx_ipc_handle(bibfs_ctx* c, uint32_t i, void* handle64) {
    if (!c || i >= c->d.vranks) return fail("bad hosted slot");
    DeviceGuard g(c->device);
    cudaIpcMemHandle_t h;
    BIBFS_CUDA_TRY(cudaIpcGetMemHandle(&h, c->arena_local[i]));
    static_assert(sizeof(h) == 64, "IPC handle size");
    memcpy(handle64, &h, 64);
    return 0;
}

extern "C" int bibfs_ctx_attach_ipc(bibfs_ctx* c, uint32_t rank, const void* handle64) {
    if (!c || rank >= c->d.nranks) return fail("bad rank");
    if (c->arena_all[rank]) return fail("rank %u is already attached", rank);
    DeviceGuard g(c->device);
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    void* p = nullptr;
    BIBFS_CUDA_TRY(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    c->arena_all[rank] = (char*)p;
    c->ipc_opened[rank] = true;
    c->async_peers = true;
    return 0;
}

extern "C" int bibfs_ctx_attach_local(bibfs_ctx* c, uint32_t rank, bibfs_ctx* owner) {
    if (!c || !owner || rank >= c->d.nranks) return fail("bad argument");
    if (rank < owner->d.rank_lo || rank >= owner->d.rank_lo + owner->d.vranks) return fail("owner does not host rank %u", rank);
    if (owner->device != c->device) {
        DeviceGuard g(c->device);
        int can = 0;
        BIBFS_CUDA_TRY(cudaDeviceCanAccessPeer(&can, c->device, owner->device));
        if (!can) return fail("device %d cannot access device %d", c->device, owner->device);
        cudaError_t e = cudaDeviceEnablePeerAccess(owner->device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return fail("enable peer access: %s", cudaGetErrorString(e));
        cudaGetLastError();
    }
    c->arena_all[rank] = owner->arena_local[rank - owner->d.rank_lo];
    c->col_all[rank] = owner->d.d_col[rank - owner->d.rank_lo];
    c->vrec_all[rank] = owner->d.d_vrec[rank - owner->d.rank_lo];
    return 0;
}

extern "C" int bibfs_ctx_set_peer_graph(bibfs_ctx* c, uint32_t rank, const uint32_t* d_col, const void* d_vrec) {
    if (!c || rank >= c->d.nranks) return fail("bad rank");
    if (reinterpret_cast<uintptr_t>(d_vrec) & 31u) return fail("d_vrec must be 32-byte aligned");
    c->col_all[rank] = d_col;
    c->vrec_all[rank] = d_vrec;
    return 0;
}

extern "C" int bibfs_ctx_reset(bibfs_ctx* c, void* stream) {
    if (!c) return fail("NULL ctx");
    DeviceGuard g(c->device);
    c->qn = 0;
    for (u32 i = 0; i < c->d.vranks; ++i)
        BIBFS_CUDA_TRY(cudaMemsetAsync(c->arena_local[i], 0, zero_prefix(c->L), (cudaStream_t)stream));
    return 0;
}

static void fill_params(bibfs_ctx* c, KParams& K, u32 src, u32 dst, bool in_ctrl, u32 flags) {
    memset(&K, 0, sizeof(K));
    K.n = c->d.n; K.B = c->B; K.P = c->d.nranks;
    K.rank_lo = c->d.rank_lo; K.vranks = c->d.vranks; K.cpr = c->cpr;
    K.cluster = c->cluster;
    if (c->cpr_small && c->cpr_small < c->cpr && c->geometry == 1) K.cpr = c->cpr_small;
    K.solo_threshold = c->d.solo_threshold;
    K.bu_div = c->d.bottom_up_div;
    K.flags = flags;
    K.max_path = c->d.max_path;
    K.seq = ++c->seq;
    K.watchdog_ms = c->d.watchdog_ms;
    K.src = src; K.dst = dst; K.query_in_ctrl = 0u;
    K.div_B = ~0ull / c->B + 1ull;
    K.h_query = nullptr; K.h_out = nullptr; K.h_out_words = 0;
    K.L = c->L;
    for (u32 r = 0; r < MAXR; ++r) K.arena[r] = c->arena_all[r];
    for (u32 i = 0; i < c->d.vranks; ++i) {
        K.row_ptr[c->d.rank_lo + i] = c->d.d_row_ptr[i];
        K.col[c->d.rank_lo + i] = c->d.d_col[i];
    }
    for (u32 r = 0; r < c->d.nranks; ++r) {
        K.col[r] = c->col_all[r];
        K.vrec[r] = reinterpret_cast<const uint4*>(c->vrec_all[r]);
    }
    K.qn = c->qn;
    K.long_rows = (c->d.max_degree == 0 || c->d.max_degree >= 32) ? 1u : 0u;
    { static const u32 wm = getenv("BIBFS_WIDE_MIN") ? (u32)atoi(getenv("BIBFS_WIDE_MIN")) : 512u; K.wide_min = wm; }
    K.async_peers = c->async_peers ? 1u : 0u;
    K.lone = c->lone ? 1u : 0u;
    // records + visited words of a small graph are streamed into L2 by the idle CTAs at query start (<= 24 MB: a fifth of L2)
    if (c->vrec_all[0] && c->hosts0) {
        // how many of the records are local and contiguous behind rank 0's (its own shard; every shard with a record replica)
        u32 local = 1;
        while (local < c->d.nranks && c->vrec_all[local] == (const char*)c->vrec_all[0] + (size_t)local * c->B * 64) ++local;
        const u64 bytes = local == c->d.nranks ? (u64)c->d.n * 64ull : (u64)c->B * 64ull;
        if (bytes <= (24ull << 20)) K.warm_lines = (u32)((bytes + 127ull) / 128ull);
    }
}

// Does this query start with the front kernel?  Needs every rank's vertex records on this device.  Policy (contexts that
// host every rank or whose peers decide alike — they see the same verdicts): hand-overs after fewer than eight levels
// (graphs with hub rows next to almost every vertex) push the score up, everything else pulls it down; at 4 the context
// goes straight to the general kernel and looks at the front kernel again every 32nd query.  Contexts of a single
// process spanning several devices (peers attached locally) leave the choice to the caller (BIBFS_F_NO_FRONT).
static bool use_front(const bibfs_ctx* c, u32 flags) {
    if (flags & (BIBFS_F_NO_FRONT | BIBFS_F_FORCE_GRID | BIBFS_F_FORCE_BU)) return false;
    for (u32 r = 0; r < c->d.nranks; ++r)
        if (!c->vrec_all[r] || !c->col_all[r]) return false;
    if (c->front_mode) return c->front_mode == 1;
    if (!c->hosts_all && !c->async_peers) return true;
    return c->front_score < 4 || (c->qn & 31u) == 0;
}

static int launch_general(bibfs_ctx* c, KParams& K, bool want_path, bool resume, cudaStream_t st) {
    K.resume = resume ? 1u : 0u;
    const bool lean = c->last_lean;
    c->last_fused = c->hosts0 && (want_path || lean || c->multi || resume);
    K.fuse_path = c->last_fused ? (want_path ? 2u : 1u) : 0u;
    if (c->last_fused) { K.h_out = reinterpret_cast<u32*>(c->h_stage); K.h_out_words = HOST_PATH_WORDS; K.h_seq = c->h_seq; }
    else { K.h_out = nullptr; K.h_out_words = 0; K.h_seq = nullptr; }
    if (bfs_launch(K, c->multi, st)) return 1;
    if (!lean) BIBFS_CUDA_TRY(cudaEventRecord(c->ev[1], st));
    return 0;
}

extern "C" int bibfs_ctx_query(bibfs_ctx* c, uint32_t src, uint32_t dst, const uint32_t* h_query_pinned,
                               uint32_t flags, void* stream) {
    if (!c) return fail("NULL ctx");
    if (flags & ~BIBFS_F_ALL) return fail("unknown query flag bits 0x%x", flags & ~BIBFS_F_ALL);
    if (reinterpret_cast<uintptr_t>(h_query_pinned) & 7u) return fail("h_query_pinned must be 8-byte aligned");
    for (u32 r = 0; r < c->d.nranks; ++r)
        if (!c->arena_all[r]) return fail("rank %u is not attached", r);
    DeviceGuard g(c->device);
    cudaStream_t st = (cudaStream_t)stream;
    c->qn += 1;
    KParams K;
    fill_params(c, K, src, dst, false, flags);
    K.h_query = h_query_pinned;          // read by the kernel itself: pinned memory is device-visible under UVA
    const bool want_path = !(flags & BIBFS_F_NO_PATH);
    c->last_lean = (flags & BIBFS_F_DEVICE_CLOCK) != 0;
    if (!c->last_lean) K.flags |= KF_HOST_SYNCS;
    c->last_flags = flags;
    c->pend_src = src; c->pend_dst = dst; c->pend_flags = flags; c->pend_hq = h_query_pinned;
    if (!c->last_lean) BIBFS_CUDA_TRY(cudaEventRecord(c->ev[0], st));
    if (use_front(c, flags)) {
        // ---- front kernel on the context that hosts rank 0; a wait kernel on peers in other processes; nothing on peers of
        // the same process (their caller resumes them with bibfs_ctx_resume when rank 0's context reports a hand-over)
        c->stage = 1;
        if (c->hosts0) {
            c->last_fused = true;
            K.fuse_path = want_path ? 2u : 1u;
            K.h_out = reinterpret_cast<u32*>(c->h_stage); K.h_out_words = HOST_PATH_WORDS; K.h_seq = c->h_seq;
            if (front_launch(K, c->multi, c->front_narrow ? 1u : c->front_cluster, st)) return 1;
        } else if (c->async_peers) {
            c->last_fused = true;         // the wait kernel delivers the verdict to the pinned buffer itself
            K.fuse_path = 1u;
            K.h_out = reinterpret_cast<u32*>(c->h_stage); K.h_out_words = 0; K.h_seq = c->h_seq;
            if (wait_launch(K, st)) return 1;
        } else {
            c->last_fused = false;
            c->stage = 2;
        }
        if (!c->last_lean) BIBFS_CUDA_TRY(cudaEventRecord(c->ev[1], st));
        return 0;
    }
    c->stage = 0;
    return launch_general(c, K, want_path, false, st);
}

// Launch the general kernel for a query that the front kernel handed over.  Called by bibfs_ctx_result itself where the
// context can decide alone; by the caller on EVERY context of a single-process multi-device group after rank 0's
// context returned status BIBFS_STATUS_HANDED_OVER.
extern "C" int bibfs_ctx_resume(bibfs_ctx* c, void* stream) {
    if (!c) return fail("NULL ctx");
    if (c->stage == 0) return fail("no handed-over query to resume");
    DeviceGuard g(c->device);
    cudaStream_t st = (cudaStream_t)stream;
    KParams K;
    fill_params(c, K, c->pend_src, c->pend_dst, false, c->pend_flags);
    if (!c->last_lean) K.flags |= KF_HOST_SYNCS;
    c->stage = 0;
    return launch_general(c, K, !(c->pend_flags & BIBFS_F_NO_PATH), true, st);
}

// wait for what the last launch delivers into h_stage (fused) or copy the 64-byte result from the arena
static int wait_stage(bibfs_ctx* c, cudaStream_t st) {
    Result* hr = reinterpret_cast<Result*>(c->h_stage);
    if (c->last_fused) {
        // the kernel itself wrote h_stage and then the completion word: poll that word (posted PCIe writes arrive in
        // order) instead of waiting for the kernel to retire; look at the stream now and then in case it failed
        volatile u32* seq = c->h_seq;
        bool done = false;
        if (c->last_lean) {
            for (u32 spins = 0;; ++spins) {
                if (*seq == c->seq) { done = true; break; }
#if defined(__x86_64__) || defined(__i386__)
                __builtin_ia32_pause();
#endif
                if ((spins & 4095u) == 4095u && cudaStreamQuery(st) != cudaErrorNotReady) break;
            }
            std::atomic_thread_fence(std::memory_order_acquire);
        }
        if (!done) {
            BIBFS_CUDA_TRY(cudaStreamSynchronize(st));
            if (*seq != c->seq) return fail("the search kernel finished without delivering a result (aborted launch?)");
        }
    } else {
        BIBFS_CUDA_TRY(cudaMemcpyAsync(hr, c->arena_local[0] + offsetof(Ctrl, res), sizeof(Result), cudaMemcpyDeviceToHost, st));
        BIBFS_CUDA_TRY(cudaStreamSynchronize(st));
    }
    return 0;
}

extern "C" int bibfs_ctx_result(bibfs_ctx* c, bibfs_result* res, uint32_t* path, uint32_t path_cap, void* stream) {
    if (!c || !res) return fail("NULL argument");
    DeviceGuard g(c->device);
    cudaStream_t st = (cudaStream_t)stream;
    Result* hr = reinterpret_cast<Result*>(c->h_stage);
    u32* hp = reinterpret_cast<u32*>(c->h_stage + sizeof(Result));
    const bool has_path = !(c->last_flags & BIBFS_F_NO_PATH) && c->hosts0;
    const u32 first = has_path ? HOST_PATH_WORDS : 0u;
    *res = bibfs_result();
    if (c->stage == 2) {            // peer of a single-process group, front stage: nothing was launched here
        res->length = -1; res->meet = BIBFS_NO_VERTEX; res->status = BIBFS_STATUS_HANDED_OVER;
        return 0;
    }
    const bool was_front = c->stage == 1;
    if (wait_stage(c, st)) return 1;
    if (was_front && c->hosts0) {
        const bool wide = (hr->rounds_paired >> 31) != 0, handed0 = hr->status == ST_ESCALATED;
        if (c->front_narrow) { if (handed0) { c->front_narrow = false; c->no_wide_streak = 0; } }
        else if (wide) c->no_wide_streak = 0;
        else if (++c->no_wide_streak >= 8 && !getenv("BIBFS_FRONT_CLUSTER")) c->front_narrow = true;
    }
    hr->rounds_paired &= 0x7FFFFFFFu;
    if (was_front) {
        const bool handed = hr->status == ST_ESCALATED;
        // policy input: a hand-over after a handful of levels means the front kernel only added a launch
        if (handed && hr->rounds_paired < 8) c->front_score = c->front_score + 2 > 8 ? 8 : c->front_score + 2;
        else if (c->front_score > 0) c->front_score -= 1;
        if (handed) {
            if (!c->hosts_all && !c->async_peers) {      // the caller resumes every context of the group
                res->length = -1; res->meet = BIBFS_NO_VERTEX; res->status = BIBFS_STATUS_HANDED_OVER;
                res->levels = hr->levels; res->levels_paired = hr->rounds_paired;
                return 0;
            }
            if (bibfs_ctx_resume(c, stream)) return 1;
            if (wait_stage(c, st)) return 1;
        }
    }
    res->length = hr->length;
    res->meet = hr->meet;
    res->depth_s = hr->depth_s;
    res->depth_t = hr->depth_t;
    res->levels = hr->levels;
    res->levels_bottom_up = hr->levels_bu;
    res->status = hr->status;
    res->edges_scanned = hr->edges_scanned;
    res->vertices_visited = hr->vertices_visited;
    res->levels_grid = hr->levels_grid;
    res->levels_paired = hr->rounds_paired & 0x7FFFFFFFu;
    res->path_len = 0;
    float a = 0;
    if (c->last_lean) {
        res->search_ms = hr->search_ns * 1e-6;        // the device's own clock, launch latency not included
        res->path_ms = hr->path_ns * 1e-6;
    } else if (cudaEventElapsedTime(&a, c->ev[0], c->ev[1]) == cudaSuccess) {
        // the path walk runs inside the last kernel: split the event time with the device's own clock
        res->path_ms = c->last_fused ? hr->path_ns * 1e-6 : 0.0;
        res->search_ms = a > res->path_ms ? a - res->path_ms : a;
    }
    cudaGetLastError();
    if (hr->status == 1) return fail("search aborted by the device watchdog (a rank did not arrive); reset the context");
    if (hr->status == 2) return fail("src/dst out of range (n=%u)", c->d.n);
    if (has_path && hr->status == 0 && hr->length >= 0 && path && (hr->path_len == 0 || hr->path_len > path_cap))
        res->status = 3;     // a path exists but did not fit the device buffer (max_path) or the caller's (path_cap)
    if (path && has_path && hr->path_len && hr->path_len <= path_cap) {
        if (hr->path_len > first) {      // long path: fetch all of it from the device buffer
            BIBFS_CUDA_TRY(cudaMemcpyAsync(hp, c->arena_local[0] + c->L.off_path + sizeof(Result), (size_t)hr->path_len * 4,
                                           cudaMemcpyDeviceToHost, st));
            BIBFS_CUDA_TRY(cudaStreamSynchronize(st));
        }
        memcpy(path, hp, (size_t)hr->path_len * 4);
        res->path_len = hr->path_len;
    }
    return 0;
}

// query + result in one call (one FFI crossing per query for interpreted callers)
extern "C" int bibfs_ctx_query_result(bibfs_ctx* c, uint32_t src, uint32_t dst, const uint32_t* h_query_pinned, uint32_t flags,
                                      bibfs_result* res, uint32_t* path, uint32_t path_cap, void* stream) {
    if (bibfs_ctx_query(c, src, dst, h_query_pinned, flags, stream)) return 1;
    return bibfs_ctx_result(c, res, path, path_cap, stream);
}

extern "C" int bibfs_ctx_timing(bibfs_ctx* c, float* search_ms, float* path_ms) {
    if (!c) return fail("NULL ctx");
    DeviceGuard g(c->device);
    const Result* hr = reinterpret_cast<const Result*>(c->h_stage);
    float a = 0, b = c->last_fused ? hr->path_ns * 1e-6f : 0.0f;
    if (c->last_lean) a = hr->search_ns * 1e-6f;
    else {
        BIBFS_CUDA_TRY(cudaEventElapsedTime(&a, c->ev[0], c->ev[1]));
        a = a > b ? a - b : a;
    }
    if (search_ms) *search_ms = a;
    if (path_ms) *path_ms = b;
    return 0;
}

extern "C" int bibfs_ctx_trace(bibfs_ctx* c, uint32_t i, bibfs_level_trace* out, uint32_t cap, uint32_t* count) {
    if (!c || i >= c->d.vranks || !count) return fail("bad argument");
    DeviceGuard g(c->device);
    u32 k = 0;
    BIBFS_CUDA_TRY(cudaMemcpy(&k, c->arena_local[i] + offsetof(Ctrl, trace_count), 4, cudaMemcpyDeviceToHost));
    if (k > TRACE_CAP) k = TRACE_CAP;
    *count = k;
    if (out && cap) {
        const u32 m = k < cap ? k : cap;
        if (m) BIBFS_CUDA_TRY(cudaMemcpy(out, c->arena_local[i] + c->L.off_trace, (size_t)m * sizeof(bibfs_level_trace), cudaMemcpyDeviceToHost));
    }
    return 0;
}

// ------------------------------------------------------------------------------------------
// query pool: many independent queries in flight at once
//
// A bidirectional search touches a sliver of the graph and is bound by the latency of its
// level chain, not by bandwidth, so one query cannot use more than a few SMs — but 148 SMs can
// chase 148 chains at once.  The pool owns `lanes` single-rank contexts (one search arena each); a batch is ONE
// launch in which every thread-block cluster is a lane that takes every lanes-th query (k_front_batch, front.cuh)
// and writes its 64-byte results straight into pinned memory.  (The bit-parallel MS-BFS engine
// of msbfs_dev.cu sweeps all N vertices per level instead: better only when the searches are
// heavy and overlap.)  No counterpart in the reference (one query per process).
// ------------------------------------------------------------------------------------------
struct bibfs_pool {
    u32 lane_cluster = 1;        // cluster size of a lane's front kernel
    bool multi = false;
    std::vector<bibfs_ctx*> ctx;
    std::vector<cudaStream_t> st;
    std::vector<cudaEvent_t> done;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    Result* h_res = nullptr;
    size_t h_cap = 0;
    // one launch per batch (k_front_batch): the lanes' kernel parameters on the device, uploaded when they change, and
    // the queries of the batch behind a pinned staging copy
    KParams* h_lane_K = nullptr;
    KParams* d_lane_K = nullptr;
    bool lanes_dirty = true;
    unsigned long long* h_q = nullptr;
    unsigned long long* d_q = nullptr;
    int sms = 0;
    int device = 0;
};

extern "C" int bibfs_pool_destroy(bibfs_pool* p) {
    if (!p) return 0;
    DeviceGuard g(p->device);
    cudaDeviceSynchronize();
    for (auto c : p->ctx) bibfs_ctx_destroy(c);
    for (auto s : p->st) cudaStreamDestroy(s);
    for (auto e : p->done) cudaEventDestroy(e);
    if (p->ev0) cudaEventDestroy(p->ev0);
    if (p->ev1) cudaEventDestroy(p->ev1);
    if (p->h_res) cudaFreeHost(p->h_res);
    if (p->h_lane_K) cudaFreeHost(p->h_lane_K);
    if (p->d_lane_K) cudaFree(p->d_lane_K);
    if (p->h_q) cudaFreeHost(p->h_q);
    if (p->d_q) cudaFree(p->d_q);
    delete p;
    return 0;
}

extern "C" int bibfs_pool_create(bibfs_pool** out, const bibfs_ctx_desc* desc, uint32_t lanes, uint32_t ctas_per_lane) {
    if (!out || !desc) return fail("NULL argument");
    if (desc->vranks != 1 || desc->rank_lo != 0) return fail("a query pool is created on the context that hosts rank 0 (vranks = 1)");
    if (lanes < 1 || lanes > 1024) return fail("lanes must be 1..1024");
    bibfs_pool* p = new bibfs_pool();
    cudaGetDevice(&p->device);
    bibfs_ctx_desc d = *desc;
    d.max_ctas = ctas_per_lane ? ctas_per_lane : 1;
    d.max_path = 16;                       // the pool returns lengths and meeting vertices, not paths
    for (uint32_t i = 0; i < lanes; ++i) {
        bibfs_ctx* c = nullptr;
        if (bibfs_ctx_create(&c, &d)) { std::string why = bibfs_last_error(); bibfs_pool_destroy(p); return fail("lane %u: %s", i, why.c_str()); }
        c->lone = desc->nranks > 1;
        p->ctx.push_back(c);
        cudaStream_t s;
        cudaEvent_t e;
        if (cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) != cudaSuccess || cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) {
            bibfs_pool_destroy(p);
            return fail("stream/event creation failed");
        }
        p->st.push_back(s);
        p->done.push_back(e);
    }
    cudaEventCreate(&p->ev0);
    cudaEventCreate(&p->ev1);
    p->multi = desc->nranks > 1;
    // lanes run their front kernels side by side: small clusters (4 CTAs: long rows and wide rounds still work)
    p->lane_cluster = p->ctx[0]->front_cluster >= 4 ? 4u : p->ctx[0]->front_cluster;
    if (const char* e = getenv("BIBFS_POOL_CLUSTER")) { const u32 v = (u32)atoi(e); if (v >= 1 && v <= p->ctx[0]->front_cluster && !(v & (v - 1))) p->lane_cluster = v; }
    cudaDeviceGetAttribute(&p->sms, cudaDevAttrMultiProcessorCount, p->device);
    if (cudaHostAlloc((void**)&p->h_lane_K, lanes * sizeof(KParams), cudaHostAllocDefault) != cudaSuccess ||
        cudaMalloc((void**)&p->d_lane_K, lanes * sizeof(KParams)) != cudaSuccess || front_batch_prepare(p->multi)) {
        bibfs_pool_destroy(p);
        return fail("query pool: allocation of the lane parameters failed");
    }
    *out = p;
    return 0;
}

extern "C" int bibfs_pool_set_peer_graph(bibfs_pool* p, uint32_t rank, const uint32_t* d_col, const void* d_vrec) {
    if (!p) return fail("NULL pool");
    p->lanes_dirty = true;
    for (auto c : p->ctx)
        if (bibfs_ctx_set_peer_graph(c, rank, d_col, d_vrec)) return 1;
    return 0;
}

extern "C" int bibfs_pool_query(bibfs_pool* p, const uint32_t* h_src, const uint32_t* h_dst, uint32_t nq,
                                int32_t* h_len, uint32_t* h_meet, uint32_t* h_levels, uint64_t* h_edges, float* ms) {
    if (!p || !h_src || !h_dst) return fail("NULL argument");
    DeviceGuard g(p->device);
    if (nq > p->h_cap) {
        if (p->h_res) cudaFreeHost(p->h_res);
        if (p->h_q) cudaFreeHost(p->h_q);
        if (p->d_q) cudaFree(p->d_q);
        p->h_res = nullptr; p->h_q = nullptr; p->d_q = nullptr;
        p->h_cap = 0;
        const size_t cap = nq < 1024 ? 1024 : nq;
        BIBFS_CUDA_TRY(cudaHostAlloc((void**)&p->h_res, cap * sizeof(Result), cudaHostAllocDefault));
        BIBFS_CUDA_TRY(cudaHostAlloc((void**)&p->h_q, (cap + 1) * sizeof(unsigned long long), cudaHostAllocDefault));      // + the lanes' query counter
        BIBFS_CUDA_TRY(cudaMalloc((void**)&p->d_q, (cap + 1) * sizeof(unsigned long long)));
        p->h_cap = cap;
    }
    const size_t L = p->ctx.size();
    // with vertex records for every rank the whole batch is ONE launch of k_front_batch: every cluster of its grid is a
    // lane (pair tier, wide rounds, long rows) that takes every L-th query and delivers the 64-byte results straight
    // into the pinned array; without records the general kernel of the single-rank lanes answers, one launch per query
    bool front = true;
    for (u32 r = 0; r < p->ctx[0]->d.nranks; ++r) front = front && p->ctx[0]->vrec_all[r] && p->ctx[0]->col_all[r];
    if (p->multi && !front) return fail("a pool on a partitioned graph needs every rank's vertex records (bibfs_pool_set_peer_graph)");
    if (nq == 0) { if (ms) *ms = 0.f; return 0; }
    if (front) {
        cudaStream_t st = p->st[0];
        if (p->lanes_dirty) {
            for (size_t i = 0; i < L; ++i) {
                KParams& K = p->h_lane_K[i];
                fill_params(p->ctx[i], K, 0, 0, false, BIBFS_F_NO_PATH);
                K.lone = 1u; K.warm_lines = 0; K.async_peers = 0;
                K.fuse_path = 1u; K.h_out = nullptr; K.h_out_words = 0; K.h_seq = nullptr;
            }
            BIBFS_CUDA_TRY(cudaMemcpyAsync(p->d_lane_K, p->h_lane_K, L * sizeof(KParams), cudaMemcpyHostToDevice, st));
            p->lanes_dirty = false;
        }
        for (uint32_t q = 0; q < nq; ++q) p->h_q[q] = (unsigned long long)h_src[q] | ((unsigned long long)h_dst[q] << 32);
        p->h_q[nq] = 0ull;         // the counter the lanes draw their queries from
        const u32 lanes = (u32)(nq < L ? nq : L);
        // cluster per lane: the pool's size while every lane is resident at once, smaller (down to 2) when that buys residency
        u32 cl = p->lane_cluster;
        if (!getenv("BIBFS_POOL_CLUSTER")) while (cl > 2 && lanes * cl > (u32)p->sms) cl /= 2;
        BIBFS_CUDA_TRY(cudaEventRecord(p->ev0, st));
        BIBFS_CUDA_TRY(cudaMemcpyAsync(p->d_q, p->h_q, ((size_t)nq + 1) * sizeof(unsigned long long), cudaMemcpyHostToDevice, st));
        if (front_batch_launch(p->d_lane_K, p->d_q, p->h_res, nq, lanes, p->multi, cl, st)) return 1;
        BIBFS_CUDA_TRY(cudaEventRecord(p->ev1, st));
    } else {
        BIBFS_CUDA_TRY(cudaEventRecord(p->ev0, p->st[0]));
        for (size_t i = 1; i < L; ++i) BIBFS_CUDA_TRY(cudaStreamWaitEvent(p->st[i], p->ev0, 0));
        for (uint32_t q = 0; q < nq; ++q) {
            const size_t lane = q % L;
            bibfs_ctx* c = p->ctx[lane];
            KParams K;
            c->qn += 1;
            fill_params(c, K, h_src[q], h_dst[q], false, BIBFS_F_NO_PATH);
            if (bfs_launch(K, false, p->st[lane])) return 1;
            BIBFS_CUDA_TRY(cudaMemcpyAsync(&p->h_res[q], c->arena_local[0] + offsetof(Ctrl, res), sizeof(Result), cudaMemcpyDeviceToHost, p->st[lane]));
        }
        for (size_t i = 1; i < L; ++i) {
            BIBFS_CUDA_TRY(cudaEventRecord(p->done[i], p->st[i]));
            BIBFS_CUDA_TRY(cudaStreamWaitEvent(p->st[0], p->done[i], 0));
        }
        BIBFS_CUDA_TRY(cudaEventRecord(p->ev1, p->st[0]));
    }
    BIBFS_CUDA_TRY(cudaEventSynchronize(p->ev1));
    if (ms) BIBFS_CUDA_TRY(cudaEventElapsedTime(ms, p->ev0, p->ev1));
    int bad = 0;
    for (uint32_t q = 0; q < nq; ++q) {
        if (front && p->h_res[q].status == ST_ESCALATED && !p->multi) {
            // a search beyond the front kernel (frontier > 65536 per side): the general kernel, one at a time
            bibfs_ctx* c = p->ctx[0];
            KParams K;
            fill_params(c, K, h_src[q], h_dst[q], false, BIBFS_F_NO_PATH);
            if (bfs_launch(K, false, p->st[0])) return 1;
            BIBFS_CUDA_TRY(cudaMemcpyAsync(&p->h_res[q], c->arena_local[0] + offsetof(Ctrl, res), sizeof(Result), cudaMemcpyDeviceToHost, p->st[0]));
            BIBFS_CUDA_TRY(cudaStreamSynchronize(p->st[0]));
        }
        const Result& r = p->h_res[q];
        if (r.status == 1) bad = 1;
        if (h_len) h_len[q] = r.status == 0 ? r.length : (r.status == ST_ESCALATED ? -2 : -1);      // -2: left to the caller (partitioned graph)
        if (h_meet) h_meet[q] = r.meet;
        if (h_levels) h_levels[q] = r.levels;
        if (h_edges) h_edges[q] = r.edges_scanned;
    }
    if (bad) return fail("a pooled query was aborted by the device watchdog");
    return 0;
}

// ------------------------------------------------------------------------------------------
// resident solver: one front kernel that stays on the GPU and is fed through a mailbox in pinned host memory
// (k_front_batch in resident mode, front.cuh).  The host side is three stores and a poll.
// ------------------------------------------------------------------------------------------
struct bibfs_resident {
    bibfs_ctx* ctx = nullptr;          // owns the arena; also answers what the front kernel cannot finish
    Mailbox* mb = nullptr;             // pinned, device-visible (UVA)
    KParams* d_K = nullptr;
    cudaStream_t st = nullptr, st2 = nullptr;      // st: the resident kernel; st2: copies beside it
    u32 cluster = 1, idle_us = 50000;
    u32 seq = 0;                       // sequence number of the last query posted
    bool running = false;              // a kernel has been launched and has not been seen leaving
    uint64_t launches = 0;
    int device = 0;
};

static int resident_launch(bibfs_resident* r) {
    // the new kernel treats everything up to the last COMPLETED query as old: a query posted while the previous kernel was
    // leaving is still in the mailbox and is taken at once
    const u32 done = *reinterpret_cast<volatile u32*>(&r->mb->out_seq);
    *reinterpret_cast<volatile u32*>(&r->mb->alive) = 1u;
    std::atomic_thread_fence(std::memory_order_seq_cst);
    if (front_resident_launch(r->d_K, r->mb, done, r->idle_us, r->cluster, r->st)) return 1;
    r->running = true;
    r->launches += 1;
    return 0;
}

// tell the kernel to leave (if it is there) and wait until it has
static int resident_send_home(bibfs_resident* r) {
    if (!r->running) return 0;
    Mailbox* mb = r->mb;
    if (*reinterpret_cast<volatile u32*>(&mb->alive)) {
        mb->op = MB_EXIT;
        std::atomic_thread_fence(std::memory_order_release);
        *reinterpret_cast<volatile u32*>(&mb->seq) = ++r->seq;
    }
    BIBFS_CUDA_TRY(cudaStreamSynchronize(r->st));
    // (an EXIT is never answered: bring out_seq up to date so that the next kernel does not take it for a query)
    mb->out_seq = r->seq;
    r->running = false;
    return 0;
}

extern "C" int bibfs_resident_stop(bibfs_resident* r) {
    if (!r) return 0;
    DeviceGuard g(r->device);
    int rc = resident_send_home(r);
    if (r->st) { cudaStreamSynchronize(r->st); cudaStreamDestroy(r->st); }
    if (r->st2) cudaStreamDestroy(r->st2);
    if (r->ctx) bibfs_ctx_destroy(r->ctx);
    if (r->d_K) cudaFree(r->d_K);
    if (r->mb) cudaFreeHost(r->mb);
    delete r;
    return rc;
}

extern "C" int bibfs_resident_start(bibfs_resident** out, const bibfs_ctx_desc* desc, uint32_t cluster, uint32_t idle_ms) {
    if (!out || !desc) return fail("NULL argument");
    if (desc->nranks != 1 || desc->vranks != 1 || desc->rank_lo != 0) return fail("the resident solver works on a single-rank graph");
    if (!desc->d_vrec[0]) return fail("the resident solver needs the graph's vertex records (bibfs_vrec_build_device)");
    if (cluster == 0) cluster = 1;
    if (cluster > 16 || (cluster & (cluster - 1))) return fail("cluster must be 1, 2, 4, 8 or 16");
    bibfs_resident* r = new bibfs_resident();
    cudaGetDevice(&r->device);
    r->idle_us = (idle_ms ? idle_ms : 50u) * 1000u;
    if (bibfs_ctx_create(&r->ctx, desc)) { delete r; return 1; }
    if (cluster > r->ctx->front_cluster) cluster = r->ctx->front_cluster;
    r->cluster = cluster;
    if (cudaHostAlloc((void**)&r->mb, sizeof(Mailbox), cudaHostAllocDefault) != cudaSuccess ||
        cudaMalloc((void**)&r->d_K, sizeof(KParams)) != cudaSuccess ||
        cudaStreamCreateWithFlags(&r->st, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&r->st2, cudaStreamNonBlocking) != cudaSuccess || front_batch_prepare(false)) {
        bibfs_resident_stop(r);
        return fail("resident solver: allocation failed");
    }
    memset(r->mb, 0, sizeof(Mailbox));
    KParams K;
    fill_params(r->ctx, K, 0, 0, false, 0);
    K.lone = 1u; K.warm_lines = 0; K.async_peers = 0;
    K.fuse_path = 2u;
    K.h_out = r->mb->out; K.h_out_words = 240; K.h_seq = &r->mb->out_seq;
    if (cudaMemcpy(r->d_K, &K, sizeof(K), cudaMemcpyHostToDevice) != cudaSuccess) { bibfs_resident_stop(r); return fail("resident solver: upload failed"); }
    *out = r;
    return 0;
}

extern "C" int bibfs_resident_launches(const bibfs_resident* r, uint64_t* launches) {
    if (!r || !launches) return fail("NULL argument");
    *launches = r->launches;
    return 0;
}

extern "C" int bibfs_resident_query(bibfs_resident* r, uint32_t src, uint32_t dst, uint32_t flags,
                                    bibfs_result* res, uint32_t* path, uint32_t path_cap) {
    if (!r || !res) return fail("NULL argument");
    if (flags & ~(uint32_t)BIBFS_F_NO_PATH) return fail("the resident solver takes flags 0 or BIBFS_F_NO_PATH");
    DeviceGuard g(r->device);
    Mailbox* mb = r->mb;
    volatile u32* alive = &mb->alive;
    volatile u32* out_seq = &mb->out_seq;
    if (!r->running || !*alive) {
        if (r->running) { BIBFS_CUDA_TRY(cudaStreamSynchronize(r->st)); r->running = false; }
        if (resident_launch(r)) return 1;
    }
    // post the query: payload first, sequence number last
    const u32 seq = ++r->seq;
    mb->src = src; mb->dst = dst; mb->op = flags & BIBFS_F_NO_PATH;
    std::atomic_thread_fence(std::memory_order_release);
    *reinterpret_cast<volatile u32*>(&mb->seq) = seq;
    const auto t0 = std::chrono::steady_clock::now();
    const double limit_s = 2.0 * r->ctx->d.watchdog_ms * 1e-3 + 1.0;
    for (u32 spins = 0;; ++spins) {
        if (*out_seq == seq) break;
#if defined(__x86_64__) || defined(__i386__)
        __builtin_ia32_pause();
#endif
        if ((spins & 1023u) == 1023u) {
            if (!*alive) {
                // the kernel left (idle) as the query arrived: the next one finds it in the mailbox
                BIBFS_CUDA_TRY(cudaStreamSynchronize(r->st));
                r->running = false;
                if (*out_seq == seq) break;
                if (resident_launch(r)) return 1;
            }
            if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > limit_s)
                return fail("the resident kernel did not answer within %.1f s", limit_s);
        }
    }
    std::atomic_thread_fence(std::memory_order_acquire);
    const Result* hr = reinterpret_cast<const Result*>(mb->out);
    if (hr->status == ST_ESCALATED) {
        // beyond the front kernel: send it home and let the general kernel answer on the same arena
        if (resident_send_home(r)) return 1;
        return bibfs_ctx_query_result(r->ctx, src, dst, nullptr, flags, res, path, path_cap, r->st);
    }
    memset(res, 0, sizeof(*res));
    res->length = hr->length; res->meet = hr->meet; res->depth_s = hr->depth_s; res->depth_t = hr->depth_t;
    res->levels = hr->levels; res->levels_bottom_up = hr->levels_bu; res->status = hr->status;
    res->edges_scanned = hr->edges_scanned; res->vertices_visited = hr->vertices_visited;
    res->levels_grid = hr->levels_grid; res->levels_paired = hr->rounds_paired & 0x7FFFFFFFu;
    res->search_ms = hr->search_ns * 1e-6; res->path_ms = hr->path_ns * 1e-6;
    res->path_len = 0;
    if (hr->status == 1) return fail("search aborted by the device watchdog");
    if (hr->status == 2) return fail("src/dst out of range (n=%u)", r->ctx->d.n);
    const bool want_path = !(flags & BIBFS_F_NO_PATH);
    if (want_path && hr->status == 0 && hr->length >= 0 && path) {
        const u32 plen = hr->path_len;
        if (plen == 0 || plen > path_cap) res->status = 3;
        else if (plen <= 240) { memcpy(path, mb->out + PATH_HDR_WORDS, (size_t)plen * 4); res->path_len = plen; }
        else {
            // long path: it sits in the arena (the kernel is idle now: the copy runs beside it on another stream)
            BIBFS_CUDA_TRY(cudaMemcpyAsync(path, r->ctx->arena_local[0] + r->ctx->L.off_path + sizeof(Result), (size_t)plen * 4,
                                           cudaMemcpyDeviceToHost, r->st2));
            BIBFS_CUDA_TRY(cudaStreamSynchronize(r->st2));
            res->path_len = plen;
        }
    }
    return 0;
}

// ------------------------------------------------------------------------------------------
// sharing device buffers that this library did not allocate (the graph tensors of the PyTorch
// layer) with the other processes of a one-process-per-GPU group.  A CUDA IPC handle names a
// whole ALLOCATION; a tensor usually sits somewhere inside a block of the framework's caching
// allocator, so the exporter also reports the tensor's offset from the allocation base
// (cuMemGetAddressRange, taken from the driver library that is already loaded).  The importer maps
// each distinct allocation once per process and device.
// ------------------------------------------------------------------------------------------
#include <dlfcn.h>
#include <map>
#include <mutex>

namespace {
typedef int (*cuMemGetAddressRange_t)(unsigned long long*, size_t*, unsigned long long);
cuMemGetAddressRange_t address_range_fn() {
    static cuMemGetAddressRange_t fn = [] {
        void* h = dlopen("libcuda.so.1", RTLD_NOW | RTLD_GLOBAL);
        return h ? (cuMemGetAddressRange_t)dlsym(h, "cuMemGetAddressRange_v2") : (cuMemGetAddressRange_t) nullptr;
    }();
    return fn;
}
std::mutex g_ipc_mu;
std::map<std::pair<int, std::string>, char*> g_ipc_open;
}  // namespace

extern "C" int bibfs_copy_device(void* d_dst, const void* d_src, size_t bytes, void* stream) {
    if (!bytes) return 0;
    if (!d_dst || !d_src) return fail("NULL argument");
    BIBFS_CUDA_TRY(cudaMemcpyAsync(d_dst, d_src, bytes, cudaMemcpyDefault, (cudaStream_t)stream));
    return 0;
}

extern "C" int bibfs_ipc_export(const void* d_ptr, void* handle64, uint64_t* offset) {
    if (!d_ptr || !handle64 || !offset) return fail("NULL argument");
    cuMemGetAddressRange_t fn = address_range_fn();
    if (!fn) return fail("the CUDA driver library is not available");
    unsigned long long base = 0;
    size_t size = 0;
    BIBFS_CUDA_TRY(cudaFree(0));      // make sure the runtime's context is current for the driver call
    if (fn(&base, &size, (unsigned long long)(uintptr_t)d_ptr) != 0) return fail("cuMemGetAddressRange failed for %p", d_ptr);
    cudaIpcMemHandle_t h;
    BIBFS_CUDA_TRY(cudaIpcGetMemHandle(&h, (void*)(uintptr_t)base));
    memcpy(handle64, &h, 64);
    *offset = (uint64_t)((uintptr_t)d_ptr - (uintptr_t)base);
    return 0;
}

extern "C" int bibfs_ipc_import(const void* handle64, uint64_t offset, void** d_ptr) {
    if (!handle64 || !d_ptr) return fail("NULL argument");
    int dev = 0;
    BIBFS_CUDA_TRY(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lock(g_ipc_mu);
    const std::pair<int, std::string> key(dev, std::string((const char*)handle64, 64));
    auto it = g_ipc_open.find(key);
    char* base = nullptr;
    if (it != g_ipc_open.end()) base = it->second;
    else {
        cudaIpcMemHandle_t h;
        memcpy(&h, handle64, 64);
        void* p = nullptr;
        BIBFS_CUDA_TRY(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        base = (char*)p;
        g_ipc_open[key] = base;
    }
    *d_ptr = base + offset;
    return 0;
}
