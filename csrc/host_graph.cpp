// host_graph.cpp — CPU generators and CPU CSR construction.
//
// The host CSR builder is the same count -> prefix -> fill scheme the reference uses
// (v3/bibfs_cuda_only.cu:89-99, v4/mpi_bas.cpp:45-57) but for an arbitrary row range (a 1-D
// partition shard), with 64-bit offsets, optional row sorting, and self loops dropped.  It runs
// the very same Gen::for_each stream as the sm_100a kernels in csr_dev.cu / gen_dev.cu.
#include "gen.h"
#include <algorithm>
#include <vector>

using namespace bibfs;

static int resolve_gen(u32 n, u64 m, const u32* edges, const bibfs_gen_spec* spec, Gen& g) {
    if (spec) {
        std::string why;
        if (!make_gen(*spec, g, why)) return fail("generator spec: %s", why.c_str());
        if (g.n != n) return fail("generator n=%llu != n=%u", (unsigned long long)g.n, n);
        return 0;
    }
    if (!edges && m) return fail("no edges and no generator spec");
    g = make_edge_gen(n, m, edges);
    return 0;
}

extern "C" int bibfs_gen_count_host(const bibfs_gen_spec* spec, uint64_t* m_out) {
    Gen g;
    std::string why;
    if (!make_gen(*spec, g, why)) return fail("generator spec: %s", why.c_str());
    u64 cnt = 0;
    for (u64 c = 0; c < g.nchunks; ++c) g.for_each(c, [&](u32, u32) { ++cnt; });
    *m_out = cnt;
    return 0;
}

extern "C" int bibfs_gen_edges_host(const bibfs_gen_spec* spec, uint32_t* edges, uint64_t cap, uint64_t* m_out) {
    Gen g;
    std::string why;
    if (!make_gen(*spec, g, why)) return fail("generator spec: %s", why.c_str());
    u64 cnt = 0;
    bool overflow = false;
    for (u64 c = 0; c < g.nchunks; ++c)
        g.for_each(c, [&](u32 u, u32 v) {
            if (cnt < cap) { edges[2 * cnt] = u; edges[2 * cnt + 1] = v; } else overflow = true;
            ++cnt;
        });
    *m_out = cnt;
    if (overflow) return fail("edge buffer too small: need %llu records, have %llu", (unsigned long long)cnt, (unsigned long long)cap);
    return 0;
}

extern "C" int bibfs_csr_count_host(uint32_t n, uint64_t m, const uint32_t* edges, const bibfs_gen_spec* spec,
                                    uint32_t row_lo, uint32_t row_hi, uint64_t* row_ptr) {
    if (row_lo > row_hi || row_hi > n) return fail("bad row range [%u,%u) for n=%u", row_lo, row_hi, n);
    Gen g;
    if (resolve_gen(n, m, edges, spec, g)) return 1;
    const u32 rows = row_hi - row_lo;
    std::vector<u64> deg(rows + 1, 0);
    bool bad = false;
    for (u64 c = 0; c < g.nchunks; ++c)
        g.for_each(c, [&](u32 u, u32 v) {
            if (u >= n || v >= n) { bad = true; return; }
            if (u == v) return;
            if (u >= row_lo && u < row_hi) ++deg[u - row_lo];
            if (v >= row_lo && v < row_hi) ++deg[v - row_lo];
        });
    if (bad) return fail("edge endpoint >= n");
    u64 acc = 0;
    for (u32 i = 0; i < rows; ++i) { row_ptr[i] = acc; acc += deg[i]; }
    row_ptr[rows] = acc;
    return 0;
}

extern "C" int bibfs_csr_fill_host(uint32_t n, uint64_t m, const uint32_t* edges, const bibfs_gen_spec* spec,
                                   uint32_t row_lo, uint32_t row_hi, const uint64_t* row_ptr, uint32_t* col,
                                   int sort_rows) {
    if (row_lo > row_hi || row_hi > n) return fail("bad row range [%u,%u) for n=%u", row_lo, row_hi, n);
    Gen g;
    if (resolve_gen(n, m, edges, spec, g)) return 1;
    const u32 rows = row_hi - row_lo;
    std::vector<u64> cur(row_ptr, row_ptr + rows);
    for (u64 c = 0; c < g.nchunks; ++c)
        g.for_each(c, [&](u32 u, u32 v) {
            if (u == v || u >= n || v >= n) return;
            if (u >= row_lo && u < row_hi) col[cur[u - row_lo]++] = v;
            if (v >= row_lo && v < row_hi) col[cur[v - row_lo]++] = u;
        });
    if (sort_rows)
        for (u32 i = 0; i < rows; ++i) std::sort(col + row_ptr[i], col + row_ptr[i + 1]);
    return 0;
}

extern "C" uint32_t bibfs_block_size(uint32_t n, uint32_t nranks) {
    if (nranks == 0) nranks = 1;
    u64 b = ((u64)n + nranks - 1) / nranks;
    b = (b + 63) / 64 * 64;   // bitmap words never straddle two owners
    if (b == 0) b = 64;
    return (u32)b;
}
