// cpu_ranks.cpp — multi-rank CPU bidirectional BFS with bitset frontiers ("cpu-ranks" variant).
//
// Equivalent of the reference's v2 (v2/second_try.cpp:53-129): six bitsets (frontier / next /
// visited x source / target), work strided over ranks, an OR-merge and a handful of reductions at
// a barrier each level.  MPI is not available on the target box, so the ranks are threads and the
// MPI_Allreduce(BOR) calls (second_try.cpp:82-85,101-104) become atomic fetch_or on shared
// bitsets; the LOR / SUM reductions (:115,:123-124) become per-rank slots summed after a barrier.
// Differences on purpose: the smaller side is expanded (one side per level, like v1/v4) so the
// reported length is the true hop count (v2 prints ceil(d/2)); parents are recorded so
// the path comes from the search itself instead of a second serial BFS (second_try.cpp:137-162);
// the search stops as soon as the side to expand is empty (Q9); small frontiers are expanded by rank 0 alone from
// vertex lists (solo_phase below), the bitset levels of all ranks take over when a frontier outgrows 2,048 vertices.
#include "common.h"
#include <atomic>
#include <chrono>
#include <cstdlib>
#include <thread>
#include <vector>

using namespace bibfs;

namespace {

struct SpinBarrier {
    std::atomic<u32> count{0};
    std::atomic<u32> gen{0};
    u32 n;
    explicit SpinBarrier(u32 n_) : n(n_) {}
    void wait() {
        u32 g = gen.load(std::memory_order_acquire);
        if (count.fetch_add(1, std::memory_order_acq_rel) + 1 == n) {
            count.store(0, std::memory_order_relaxed);
            gen.store(g + 1, std::memory_order_release);
        } else {
            // levels last microseconds: spin (with a pause hint) before giving the core away
            u32 spins = 0;
            while (gen.load(std::memory_order_acquire) == g) {
#if defined(__x86_64__) || defined(__i386__)
                __builtin_ia32_pause();
#endif
                if (++spins > 4000) std::this_thread::yield();   // oversubscribed box: let the other ranks run
            }
        }
    }
};

struct Shared {
    u32 n, L, R;
    const u64* row_ptr;
    const u32* col;
    std::vector<std::atomic<u64>> front[2], vis[2], next;
    std::vector<u32> par[2];
    // per-rank reduction slots
    std::vector<u64> slot_new, slot_scan;
    std::vector<u32> slot_meet;
    // level state (written by rank 0 between barriers)
    int side = 0;
    bool done = false;
    u64 nf[2] = {1, 1};
    u32 depth[2] = {0, 0};
    u32 levels = 0;
    u32 meet = BIBFS_NO_VERTEX;
    u64 scanned = 0, visited = 2;
    Shared(u32 n_, u32 R_) : n(n_), L((n_ + 63) / 64), R(R_), next(L) {
        for (int s = 0; s < 2; ++s) {
            front[s] = std::vector<std::atomic<u64>>(L);
            vis[s] = std::vector<std::atomic<u64>>(L);
            par[s].assign(n, BIBFS_NO_VERTEX);
        }
        slot_new.assign(R, 0);
        slot_scan.assign(R, 0);
        slot_meet.assign(R, BIBFS_NO_VERTEX);
    }
};

// Solo phase (rank 0 alone, the other ranks wait at one barrier): as long as the frontier of the side to expand is
// small, a level is a walk over a vertex LIST - no bitset sweep, no barrier, no atomics contention.  Same idea as the
// solo mode of the GPU kernel, and the "queue mode" of the reference's v2 prototype (v2/main-v2.cpp:57-69, frontier
// queues exchanged with Allgatherv).  The bitset frontiers are kept in step, so the parallel levels can take over at
// any level boundary.  Reference-sized graphs (average degree 2.2) finish here.
// (BIBFS_CPU_SOLO_MAX overrides the 2,048 of the hand-over: the tests force the parallel levels with a small value)
static size_t solo_max_frontier() {
    const char* e = getenv("BIBFS_CPU_SOLO_MAX");
    return e ? (size_t)strtoull(e, nullptr, 10) : 2048;
}

void solo_phase(Shared& S, u32 src, u32 dst) {
    std::vector<u32> fl[2] = {{src}, {dst}}, nxt;
    const size_t solo_max = solo_max_frontier();
    while (!S.done) {
        const int s = S.side, o = 1 - s;
        if (fl[s].size() > solo_max) return;                       // hand over to all ranks
        u32 meet = BIBFS_NO_VERTEX;
        u64 scanned = 0;
        nxt.clear();
        for (u32 u : fl[s]) {
            for (u64 e = S.row_ptr[u]; e < S.row_ptr[u + 1]; ++e) {
                const u32 v = S.col[e];
                ++scanned;
                const u64 mv = 1ull << (v & 63);
                const u64 w = S.vis[s][v >> 6].load(std::memory_order_relaxed);
                if (w & mv) continue;
                S.vis[s][v >> 6].store(w | mv, std::memory_order_relaxed);
                S.par[s][v] = u;
                nxt.push_back(v);
                if ((S.vis[o][v >> 6].load(std::memory_order_relaxed) & mv) && v < meet) meet = v;
            }
        }
        // keep the bitset frontier of this side in step with the list
        for (u32 u : fl[s]) S.front[s][u >> 6].store(0, std::memory_order_relaxed);
        for (u32 v : nxt) S.front[s][v >> 6].fetch_or(1ull << (v & 63), std::memory_order_relaxed);
        fl[s].swap(nxt);
        S.nf[s] = fl[s].size();
        S.visited += fl[s].size();
        S.scanned += scanned;
        ++S.depth[s];
        ++S.levels;
        if (meet != BIBFS_NO_VERTEX) { S.meet = meet; S.done = true; }
        else {
            S.side = S.nf[0] <= S.nf[1] ? 0 : 1;
            if (S.nf[S.side] == 0) S.done = true;
        }
    }
}

void worker(Shared& S, SpinBarrier& bar, u32 rank, std::chrono::steady_clock::time_point* t_start, u32 src, u32 dst) {
    // start line: the timed region begins when every rank is up (v2 starts MPI_Wtime after MPI_Init
    // and the graph broadcast, second_try.cpp:66)
    bar.wait();
    if (t_start) *t_start = std::chrono::steady_clock::now();
    if (rank == 0) solo_phase(S, src, dst);
    bar.wait();
    while (true) {
        const int s = S.side, o = 1 - s;
        if (S.done) return;
        u64 found_new = 0, scanned = 0;
        u32 meet = BIBFS_NO_VERTEX;
        // expansion: words strided by rank (v2 strides vertex ids, second_try.cpp:71)
        for (u32 w = rank; w < S.L; w += S.R) {
            u64 bits = S.front[s][w].load(std::memory_order_relaxed);
            while (bits) {
                const u32 u = w * 64 + (u32)__builtin_ctzll(bits);
                bits &= bits - 1;
                for (u64 e = S.row_ptr[u]; e < S.row_ptr[u + 1]; ++e) {
                    const u32 v = S.col[e];
                    ++scanned;
                    const u64 mv = 1ull << (v & 63);
                    if (S.vis[s][v >> 6].load(std::memory_order_relaxed) & mv) continue;
                    const u64 old = S.vis[s][v >> 6].fetch_or(mv, std::memory_order_relaxed);
                    if (old & mv) continue;
                    S.par[s][v] = u;
                    S.next[v >> 6].fetch_or(mv, std::memory_order_relaxed);
                    ++found_new;
                    if (S.vis[o][v >> 6].load(std::memory_order_relaxed) & mv)
                        if (v < meet) meet = v;
                }
            }
        }
        S.slot_new[rank] = found_new;
        S.slot_scan[rank] = scanned;
        S.slot_meet[rank] = meet;
        bar.wait();
        // swap: frontier[s] <- next, next <- 0 (each rank its own words)
        for (u32 w = rank; w < S.L; w += S.R) {
            S.front[s][w].store(S.next[w].load(std::memory_order_relaxed), std::memory_order_relaxed);
            S.next[w].store(0, std::memory_order_relaxed);
        }
        if (rank == 0) {
            u64 tot = 0;
            u32 mt = BIBFS_NO_VERTEX;
            for (u32 r = 0; r < S.R; ++r) {
                tot += S.slot_new[r];
                S.scanned += S.slot_scan[r];
                if (S.slot_meet[r] < mt) mt = S.slot_meet[r];
            }
            S.nf[s] = tot;
            S.visited += tot;
            ++S.depth[s];
            ++S.levels;
            if (mt != BIBFS_NO_VERTEX) { S.meet = mt; S.done = true; }
            else {
                S.side = S.nf[0] <= S.nf[1] ? 0 : 1;
                if (S.nf[S.side] == 0) S.done = true;
            }
        }
        bar.wait();
    }
}

}  // namespace

extern "C" int bibfs_cpu_ranks(uint32_t n, const uint64_t* row_ptr, const uint32_t* col,
                               uint32_t src, uint32_t dst, int ranks,
                               bibfs_result* res, uint32_t* path, uint32_t path_cap) {
    if (!res) return fail("res is NULL");
    *res = bibfs_result();
    res->meet = BIBFS_NO_VERTEX;
    res->length = -1;
    if (src >= n || dst >= n) { res->status = 2; return fail("src/dst out of range (n=%u)", n); }
    if (ranks < 1 || ranks > 256) return fail("ranks must be in 1..256");
    if (src == dst) {
        res->length = 0; res->meet = src; res->vertices_visited = 1;
        if (path && path_cap >= 1) { path[0] = src; res->path_len = 1; }
        return 0;
    }
    Shared S(n, (u32)ranks);
    S.row_ptr = row_ptr;
    S.col = col;
    S.front[0][src >> 6] = 1ull << (src & 63);
    S.vis[0][src >> 6] = 1ull << (src & 63);
    S.front[1][dst >> 6].fetch_or(1ull << (dst & 63));
    S.vis[1][dst >> 6].fetch_or(1ull << (dst & 63));
    SpinBarrier bar((u32)ranks);

    std::chrono::steady_clock::time_point t0;
    std::vector<std::thread> th;
    for (int r = 1; r < ranks; ++r) th.emplace_back(worker, std::ref(S), std::ref(bar), (u32)r, nullptr, src, dst);
    worker(S, bar, 0, &t0, src, dst);
    auto t1 = std::chrono::steady_clock::now();
    for (auto& t : th) t.join();

    res->search_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
    res->levels = S.levels;
    res->depth_s = S.depth[0];
    res->depth_t = S.depth[1];
    res->edges_scanned = S.scanned;
    res->vertices_visited = S.visited;
    if (S.meet == BIBFS_NO_VERTEX) return 0;
    auto p0 = std::chrono::steady_clock::now();
    std::vector<u32> a, b;
    for (u32 c = S.meet; c != BIBFS_NO_VERTEX; c = S.par[0][c]) a.push_back(c);
    for (u32 c = S.meet; c != BIBFS_NO_VERTEX; c = S.par[1][c]) b.push_back(c);
    const u32 hops = (u32)(a.size() - 1 + b.size() - 1);
    res->length = (i32)hops;
    res->meet = S.meet;
    if (path && path_cap >= hops + 1) {
        u32 k = 0;
        for (size_t i = a.size(); i-- > 0;) path[k++] = a[i];
        for (size_t i = 1; i < b.size(); ++i) path[k++] = b[i];
        res->path_len = k;
    }
    res->path_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - p0).count();
    return 0;
}
