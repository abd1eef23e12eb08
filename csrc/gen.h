// gen.h — graph generators as chunked, counter-based edge streams (host + device).
//
// A generator is a pure function  chunk index -> a few (u,v) records, so the same code feeds
//   * the edge-list writers (count -> scan -> emit)            gen_dev.cu / gen_host.cpp
//   * the fused CSR builder (degree histogram / fill straight from the generator, the edge
//     list is never materialised; each rank keeps only the rows it owns)   csr_dev.cu
// Kinds: Erdos-Renyi G(n,p) by geometric skipping inside fixed windows of the pair-index space
// (what the reference gets from networkx fast_gnp, graphs/generate_graph.py:31), R-MAT /
// Graph500 Kronecker, 2-D grid, and a plain edge array (for .bin files).
#pragma once
#include "rng.h"

namespace bibfs {

enum GenKind : int { GEN_ER = 0, GEN_RMAT = 1, GEN_GRID = 2, GEN_EDGES = 3 };

struct Gen {
    int kind;
    u32 scale;
    u64 n;
    u64 m;            // RMAT: edges; EDGES: records
    u64 nchunks;
    u64 chunk_len;    // ER: pair indices per chunk
    u64 npairs;       // ER: n(n-1)/2
    double inv_log1mp;  // ER: 1/ln(1-p)  (negative)
    int er_all;       // ER: p >= 1 -> complete graph
    u32 a16, ab16, abc16;
    u32 width, height;
    u32 k0, k1;       // seed halves
    u32 pm0, pm1, pa0, pa1;  // id scrambling constants
    const u32* edges; // EDGES

    // unrank pair index -> (u,v), u < v, idx = v(v-1)/2 + u
    BIBFS_HD static void unrank(u64 idx, u32& u, u32& v) {
        double d = sqrt(8.0 * (double)idx + 1.0);
        u64 vv = (u64)((1.0 + d) * 0.5);
        while (vv * (vv - 1) / 2 > idx) --vv;
        while (vv * (vv + 1) / 2 <= idx) ++vv;
        v = (u32)vv;
        u = (u32)(idx - vv * (vv - 1) / 2);
    }

    BIBFS_HD u32 scramble(u32 x) const {
        // bijection on [0,2^scale): odd multiply, add, xorshift — twice
        const u32 mask = (scale >= 32) ? 0xFFFFFFFFu : ((1u << scale) - 1u);
        const u32 sh = scale > 1 ? scale / 2 : 1;
        x = (x * pm0 + pa0) & mask;
        x ^= x >> sh;
        x = (x * pm1 + pa1) & mask;
        x ^= x >> sh;
        return x;
    }

    template <class F>
    BIBFS_HD void for_each(u64 chunk, F&& f) const {
        if (kind == GEN_EDGES) {
            f(edges[2 * chunk], edges[2 * chunk + 1]);
        } else if (kind == GEN_GRID) {
            u32 x = (u32)(chunk % width), y = (u32)(chunk / width);
            if (x + 1 < width) f((u32)chunk, (u32)chunk + 1);
            if (y + 1 < height) f((u32)chunk, (u32)chunk + width);
        } else if (kind == GEN_RMAT) {
            u32 u = 0, v = 0;
            u32 lvl = 0;
            for (u32 blk = 0; lvl < scale; ++blk) {
                Philox r = philox4x32((u32)chunk, (u32)(chunk >> 32), blk, 0x524D4154u, k0, k1);
#pragma unroll
                for (int w = 0; w < 4; ++w) {
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        if (lvl < scale) {
                            u32 t = (r.c[w] >> (16 * h)) & 0xFFFFu;
                            u32 ub = (t >= ab16) ? 1u : 0u;
                            u32 vb = (t >= a16 && t < ab16) || (t >= abc16) ? 1u : 0u;
                            u = (u << 1) | ub;
                            v = (v << 1) | vb;
                            ++lvl;
                        }
                    }
                }
            }
            u = scramble(u);
            v = scramble(v);
            if (n < (1ull << scale)) { u = (u32)(u % n); v = (u32)(v % n); }
            f(u, v);
        } else {  // GEN_ER
            u64 lo = chunk * chunk_len;
            u64 hi = lo + chunk_len;
            if (hi > npairs) hi = npairs;
            if (er_all) {
                for (u64 i = lo; i < hi; ++i) { u32 a, b; unrank(i, a, b); f(a, b); }
                return;
            }
            u64 pos = lo;            // next candidate pair index
            for (u32 it = 0;; ++it) {
                Philox r = philox4x32((u32)chunk, (u32)(chunk >> 32), it, 0x45524E59u, k0, k1);
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    double U = u01_open_closed(r.c[2 * h], r.c[2 * h + 1]);
                    double q = det_log(U) * inv_log1mp;      // >= 0
                    u64 skip = (q >= 9.0e18) ? (u64)9000000000000000000ull : (u64)q;
                    if (skip >= hi - pos) return;
                    pos += skip;
                    u32 a, b;
                    unrank(pos, a, b);
                    f(a, b);
                    ++pos;
                    if (pos >= hi) return;
                }
            }
        }
    }
};

// Build the derived generator from the public spec.  Returns false on bad parameters.
inline bool make_gen(const bibfs_gen_spec& s, Gen& g, std::string& why) {
    g = Gen();
    g.kind = s.kind;
    g.n = s.n;
    g.k0 = (u32)s.seed;
    g.k1 = (u32)(s.seed >> 32);
    g.edges = nullptr;
    if (s.n == 0 || s.n > 0xFFFFFFFEull) { why = "n out of range"; return false; }
    switch (s.kind) {
        case GEN_ER: {
            if (!(s.p >= 0.0) || s.p > 1.0) { why = "p must be in [0,1]"; return false; }
            if (s.n > (1ull << 31)) { why = "ER supports n <= 2^31"; return false; }
            g.npairs = s.n * (s.n - 1) / 2;
            g.er_all = s.p >= 1.0;
            if (s.p <= 0.0 || g.npairs == 0) { g.nchunks = 0; g.chunk_len = 1; g.inv_log1mp = -1.0; return true; }
            if (g.er_all) {
                g.chunk_len = 64;
            } else {
                double cl = 64.0 / s.p;
                g.chunk_len = cl >= 1.8e19 ? g.npairs : (u64)cl;
                if (g.chunk_len < 64) g.chunk_len = 64;
                g.inv_log1mp = 1.0 / det_log(1.0 - s.p);
            }
            if (g.chunk_len > g.npairs) g.chunk_len = g.npairs;
            g.nchunks = (g.npairs + g.chunk_len - 1) / g.chunk_len;
            return true;
        }
        case GEN_RMAT: {
            if (s.scale == 0 || s.scale > 32) { why = "scale must be 1..32"; return false; }
            if (s.n > (1ull << s.scale)) { why = "n exceeds 2^scale"; return false; }
            if (!(s.a16 <= s.ab16 && s.ab16 <= s.abc16 && s.abc16 <= 65536)) { why = "bad RMAT thresholds"; return false; }
            g.scale = s.scale;
            g.m = s.m;
            g.nchunks = s.m;
            g.a16 = s.a16; g.ab16 = s.ab16; g.abc16 = s.abc16;
            Philox r = philox4x32(0x5343524Du, 1, 2, 3, g.k0, g.k1);
            g.pm0 = r.c[0] | 1u; g.pm1 = r.c[1] | 1u; g.pa0 = r.c[2]; g.pa1 = r.c[3];
            return true;
        }
        case GEN_GRID: {
            if (s.width == 0 || s.height == 0 || (u64)s.width * s.height != s.n) { why = "grid needs width*height == n"; return false; }
            g.width = s.width; g.height = s.height;
            g.nchunks = s.n;
            return true;
        }
        default:
            why = "unknown generator kind";
            return false;
    }
}

inline Gen make_edge_gen(u32 n, u64 m, const u32* edges) {
    Gen g = Gen();
    g.kind = GEN_EDGES;
    g.n = n;
    g.m = m;
    g.nchunks = m;
    g.edges = edges;
    return g;
}

}  // namespace bibfs
