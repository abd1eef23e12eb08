// bench_prep.cpp — stand-alone helper of bench.py (no CUDA, no libbibfs_b200.so).
//
// Both arms of bench.py need the same synthetic graph, the same query list and the same GTEPS
// numerator, and the reference arm must get them without mapping this repository's library into
// its process.  This program
//   * generates the graph with the counter-based generator of csrc/gen.h (header only; identical
//     to what the sm_100a generator kernels produce, tests/test_gpu_build.py),
//   * optionally writes it as the reference's .bin edge list (graphs/generate_graph.py:35-39),
//   * builds a host CSR with sorted rows, and for every "src dst" pair read from stdin prints
//         src dst length v1_edges levels
//     where v1_edges = directed adjacency entries the reference's v1 schedule inspects
//     (v1/main-v1.cpp:50-81: smaller frontier first, ties to the source side, stop at the first
//     meeting vertex in the middle of a level) — the common numerator of BASELINE.md §3.
//
//   bench_prep er:<n>:<p>:<seed> | rmat:<scale>:<edgefactor>:<n>:<seed> | grid:<w>:<h>  [--bin FILE] [--need K] < pairs
#include "../../csrc/gen.h"
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

namespace bibfs {   // gen.h only declares these; nothing here reports through them
void set_error(const char*, ...) {}
int fail(const char*, ...) { return 1; }
void count_launch(u64) {}
}
using namespace bibfs;

static std::vector<std::string> split(const std::string& s, char c) {
    std::vector<std::string> out;
    size_t a = 0;
    for (size_t i = 0; i <= s.size(); ++i)
        if (i == s.size() || s[i] == c) { out.push_back(s.substr(a, i - a)); a = i + 1; }
    return out;
}

int main(int argc, char** argv) {
    if (argc < 2) { fprintf(stderr, "usage: bench_prep <spec> [--bin FILE] < pairs\n"); return 2; }
    const char* bin_path = nullptr;
    unsigned long long need = ~0ull;       // stop after this many usable pairs (connected, >= 2 hops; the first pair always counts)
    for (int i = 2; i + 1 < argc; ++i) {
        if (!strcmp(argv[i], "--bin")) bin_path = argv[i + 1];
        if (!strcmp(argv[i], "--need")) need = strtoull(argv[i + 1], 0, 10);
    }
    auto f = split(argv[1], ':');
    bibfs_gen_spec s;
    memset(&s, 0, sizeof(s));
    if (f[0] == "er" && f.size() == 4) {
        s.kind = GEN_ER; s.n = strtoull(f[1].c_str(), 0, 10); s.p = strtod(f[2].c_str(), 0); s.seed = strtoull(f[3].c_str(), 0, 10);
    } else if (f[0] == "rmat" && f.size() == 5) {
        s.kind = GEN_RMAT; s.scale = (uint32_t)atoi(f[1].c_str()); s.n = strtoull(f[3].c_str(), 0, 10); s.seed = strtoull(f[4].c_str(), 0, 10);
        s.m = strtoull(f[2].c_str(), 0, 10) * s.n;
        s.a16 = (uint32_t)(0.57 * 65536); s.ab16 = (uint32_t)((0.57 + 0.19) * 65536); s.abc16 = (uint32_t)((0.57 + 0.19 + 0.19) * 65536);
    } else if (f[0] == "grid" && f.size() == 3) {
        s.kind = GEN_GRID; s.width = (uint32_t)atoi(f[1].c_str()); s.height = (uint32_t)atoi(f[2].c_str()); s.n = (uint64_t)s.width * s.height;
    } else { fprintf(stderr, "bad spec %s\n", argv[1]); return 2; }
    Gen g;
    std::string why;
    if (!make_gen(s, g, why)) { fprintf(stderr, "generator: %s\n", why.c_str()); return 2; }
    const u32 n = (u32)g.n;

    // pass 1: degrees (and the .bin file).  Chunks are dealt over host threads; the generator is a pure function of
    // the chunk index, so the graph does not depend on the thread count.
    unsigned T = std::thread::hardware_concurrency();
    if (T < 1) T = 1;
    if (T > 32) T = 32;
    if (g.nchunks < 4096 || bin_path) T = 1;          // the .bin writer keeps the generator's edge order
    std::vector<u64> rp(n + 1, 0);
    u64 records = 0;
    FILE* out = bin_path ? fopen(bin_path, "wb") : nullptr;
    if (bin_path && !out) { perror(bin_path); return 1; }
    std::vector<u32> buf;
    if (out) { u32 hdr[2] = {n, 0}; fwrite(hdr, 4, 2, out); }
    auto parallel_chunks = [&](auto body) {
        std::vector<std::thread> th;
        const u64 per = (g.nchunks + T - 1) / T;
        for (unsigned t = 0; t < T; ++t)
            th.emplace_back([&, t] { const u64 hi = std::min(g.nchunks, (t + 1) * per); for (u64 c = t * per; c < hi; ++c) body(c); });
        for (auto& x : th) x.join();
    };
    parallel_chunks([&](u64 c) {
        u64 local = 0;
        g.for_each(c, [&](u32 u, u32 v) {
            ++local;
            if (out) { buf.push_back(u); buf.push_back(v); if (buf.size() >= (1u << 20)) { fwrite(buf.data(), 4, buf.size(), out); buf.clear(); } }
            if (u == v || u >= n || v >= n) return;
            __atomic_fetch_add(&rp[u + 1], 1, __ATOMIC_RELAXED);
            __atomic_fetch_add(&rp[v + 1], 1, __ATOMIC_RELAXED);
        });
        __atomic_fetch_add(&records, local, __ATOMIC_RELAXED);
    });
    if (out) {
        if (!buf.empty()) fwrite(buf.data(), 4, buf.size(), out);
        if (records > 0xFFFFFFFFull) { fprintf(stderr, "too many edges for the .bin header\n"); return 1; }
        u32 m32 = (u32)records;
        fseek(out, 4, SEEK_SET); fwrite(&m32, 4, 1, out); fclose(out);
    }
    for (u32 i = 0; i < n; ++i) rp[i + 1] += rp[i];
    std::vector<u32> col(rp[n]);
    {
        std::vector<u64> cur(rp.begin(), rp.end() - 1);
        parallel_chunks([&](u64 c) {
            g.for_each(c, [&](u32 u, u32 v) {
                if (u == v || u >= n || v >= n) return;
                col[__atomic_fetch_add(&cur[u], 1, __ATOMIC_RELAXED)] = v;
                col[__atomic_fetch_add(&cur[v], 1, __ATOMIC_RELAXED)] = u;
            });
        });
        std::vector<std::thread> th;          // rows sorted: the replay below does not depend on the fill order
        const u32 per = (n + T - 1) / T;
        for (unsigned t = 0; t < T; ++t)
            th.emplace_back([&, t] { const u32 hi = std::min<u64>(n, (u64)(t + 1) * per); for (u32 i = t * per; i < hi; ++i) std::sort(col.begin() + rp[i], col.begin() + rp[i + 1]); });
        for (auto& x : th) x.join();
    }
    printf("graph %u %llu %llu\n", n, (unsigned long long)records, (unsigned long long)rp[n]);

    // pass 2: replay the v1 schedule for every pair on stdin
    std::vector<unsigned char> mark(n, 0);          // bit 0: reached from src, bit 1: reached from dst
    std::vector<u32> touched, fr[2], nx;
    unsigned long long a, b, usable = 0, seen = 0;
    while (usable < need && scanf("%llu %llu", &a, &b) == 2) {
        ++seen;
        const u32 src = (u32)a, dst = (u32)b;
        long long length = -1;
        u64 edges = 0;
        u32 levels = 0;
        if (src >= n || dst >= n) { printf("%u %u -2 0 0\n", src, dst); continue; }
        if (src == dst) { printf("%u %u 0 0 0\n", src, dst); continue; }
        fr[0].assign(1, src); fr[1].assign(1, dst);
        mark[src] |= 1; mark[dst] |= 2; touched.assign({src, dst});
        u32 depth[2] = {0, 0};
        bool met = false;
        while (!met && !fr[0].empty() && !fr[1].empty()) {
            const int sd = fr[0].size() <= fr[1].size() ? 0 : 1;
            const unsigned char mine = (unsigned char)(1u << sd), other = (unsigned char)(2u >> sd);
            nx.clear();
            for (size_t i = 0; i < fr[sd].size() && !met; ++i) {
                const u32 u = fr[sd][i];
                for (u64 e = rp[u]; e < rp[u + 1]; ++e) {
                    const u32 v = col[e];
                    ++edges;
                    if (mark[v] & mine) continue;
                    if (!mark[v]) touched.push_back(v);
                    mark[v] |= mine;
                    nx.push_back(v);
                    if (mark[v] & other) { met = true; break; }
                }
            }
            fr[sd].swap(nx);
            ++depth[sd]; ++levels;
        }
        if (met) length = (long long)depth[0] + depth[1];
        for (u32 v : touched) mark[v] = 0;
        printf("%u %u %lld %llu %u\n", src, dst, length, (unsigned long long)edges, levels);
        if (length >= 2 || (seen == 1 && length >= 0)) ++usable;
    }
    return 0;
}
