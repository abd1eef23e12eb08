// cpu_serial.cpp — serial bidirectional BFS on CSR (the "serial" variant and the test oracle).
//
// Same algorithm as the reference's v1 (v1/main-v1.cpp:50-97): expand the side whose frontier
// has fewer vertices (ties -> source side), mark + record the parent at discovery, test the other
// side's visited flag at discovery time, rebuild the path from the two parent chains.
// Deliberate differences: CSR instead of vector<vector<int>>; src == dst answers 0 (v1 prints 2
// or "No path"); arguments are validated (Q9); the timed region is the level loop
// only, like v1 (v1/main-v1.cpp:49,82).  policy 0 replays v1's schedule exactly, including the
// early break in the middle of a level, so `edges_scanned` is the common GTEPS numerator of
// BASELINE.md; policy 1 finishes the level (the schedule of the GPU kernels).
#include "common.h"
#include <chrono>
#include <vector>

using namespace bibfs;

extern "C" int bibfs_serial(uint32_t n, const uint64_t* row_ptr, const uint32_t* col,
                            uint32_t src, uint32_t dst, int policy,
                            bibfs_result* res, uint32_t* path, uint32_t path_cap) {
    if (!res) return fail("res is NULL");
    *res = bibfs_result();
    res->meet = BIBFS_NO_VERTEX;
    res->length = -1;
    if (src >= n || dst >= n) { res->status = 2; return fail("src/dst out of range (n=%u)", n); }
    if (src == dst) {
        res->length = 0;
        res->meet = src;
        res->vertices_visited = 1;
        if (path && path_cap >= 1) { path[0] = src; res->path_len = 1; }
        return 0;
    }
    std::vector<unsigned char> vis[2] = {std::vector<unsigned char>(n, 0), std::vector<unsigned char>(n, 0)};
    std::vector<u32> par[2] = {std::vector<u32>(n, BIBFS_NO_VERTEX), std::vector<u32>(n, BIBFS_NO_VERTEX)};
    std::vector<u32> front[2], next;
    vis[0][src] = 1; front[0].push_back(src);
    vis[1][dst] = 1; front[1].push_back(dst);
    u32 depth[2] = {0, 0};
    u32 meet = BIBFS_NO_VERTEX;
    u64 scanned = 0, visited = 2;
    u32 levels = 0;

    auto t0 = std::chrono::steady_clock::now();
    while (meet == BIBFS_NO_VERTEX && !front[0].empty() && !front[1].empty()) {
        const int s = front[0].size() <= front[1].size() ? 0 : 1;
        const int o = 1 - s;
        next.clear();
        bool stop = false;
        for (size_t i = 0; i < front[s].size() && !stop; ++i) {
            const u32 u = front[s][i];
            for (u64 e = row_ptr[u]; e < row_ptr[u + 1]; ++e) {
                const u32 v = col[e];
                ++scanned;
                if (!vis[s][v]) {
                    vis[s][v] = 1;
                    par[s][v] = u;
                    next.push_back(v);
                    ++visited;
                    if (vis[o][v]) {
                        if (meet == BIBFS_NO_VERTEX || v < meet) meet = v;
                        if (policy == 0) { meet = v; stop = true; break; }
                    }
                }
            }
        }
        front[s].swap(next);
        ++depth[s];
        ++levels;
    }
    auto t1 = std::chrono::steady_clock::now();
    res->search_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
    res->levels = levels;
    res->depth_s = depth[0];
    res->depth_t = depth[1];
    res->edges_scanned = scanned;
    res->vertices_visited = visited;
    if (meet == BIBFS_NO_VERTEX) return 0;

    auto p0 = std::chrono::steady_clock::now();
    std::vector<u32> a, b;
    for (u32 c = meet; c != BIBFS_NO_VERTEX; c = par[0][c]) a.push_back(c);
    for (u32 c = meet; c != BIBFS_NO_VERTEX; c = par[1][c]) b.push_back(c);
    const u32 hops = (u32)(a.size() - 1 + b.size() - 1);
    res->length = (i32)hops;
    res->meet = meet;
    if (path && path_cap >= hops + 1) {
        u32 k = 0;
        for (size_t i = a.size(); i-- > 0;) path[k++] = a[i];
        for (size_t i = 1; i < b.size(); ++i) path[k++] = b[i];
        res->path_len = k;
    }
    auto p1 = std::chrono::steady_clock::now();
    res->path_ms = std::chrono::duration<double, std::milli>(p1 - p0).count();
    return 0;
}

extern "C" int bibfs_validate_path(uint32_t n, const uint64_t* row_ptr, const uint32_t* col,
                                   uint32_t src, uint32_t dst, const uint32_t* path, uint32_t path_len) {
    if (path_len == 0) return fail("empty path");
    if (path[0] != src) return fail("path starts at %u, not src %u", path[0], src);
    if (path[path_len - 1] != dst) return fail("path ends at %u, not dst %u", path[path_len - 1], dst);
    for (u32 i = 0; i + 1 < path_len; ++i) {
        const u32 u = path[i], v = path[i + 1];
        if (u >= n || v >= n) return fail("path vertex out of range at hop %u", i);
        bool ok = false;
        for (u64 e = row_ptr[u]; e < row_ptr[u + 1] && !ok; ++e) ok = (col[e] == v);
        if (!ok) return fail("hop %u: (%u,%u) is not an edge", i, u, v);
    }
    return 0;
}
