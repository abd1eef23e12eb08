// cli_main.cpp — native `bibfs` command line.
//
//   bibfs <graph.bin> <src> <dst> [--variant serial|cpu-ranks|gpu|multi-gpu|naive-gpu]
//                                 [--ranks R] [--gpus G] [--vranks V] [--repeat K] [--trace] [--quiet-path]
//   bibfs show <graph.bin> [--dense] [--limit K]      edge-list / CSR dump (dense print for tiny N)
//   bibfs show-dense <legacy.bin>                      the old N + N*N byte format of v2/read_in.cpp
//   bibfs gen er|rmat|grid ... --out file.bin          see usage()
//
// Same argv shape and the same stdout lines as the reference binaries:
//   serial     v1/main-v1.cpp:84-101   "Shortest path length = d" / "Path: ..." / "Serial bidirectional BFS took t seconds"
//   cpu-ranks  v2/second_try.cpp:134-135,159-161   "[Time] bidir-bitset BFS = t seconds" (U+2011 hyphen as in the source)
//   gpu        v3/bibfs_cuda_only.cu:208,224       "DEBUG: gpu bfs time = t ms", "GPU BFS: PATH FOUND|NO PATH" (+ the length and path v3 lacks)
//   multi-gpu  v4/mpi_bas.cpp:137-138,161-163      "[Time] MPI+CUDA BI-BFS = t s" (kept verbatim so benchmark_test.sh's awk still parses it)
// plus one "[phases] load=.. build=.. h2d=.. search=.. path=.. ms" line (per-phase timing is new).
#include "common.h"
#define BIBFS_WITH_CUDA_RUNTIME 1
#include <cuda_runtime.h>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

using namespace bibfs;
typedef std::chrono::steady_clock Clock;
static double ms_since(Clock::time_point t0) { return std::chrono::duration<double, std::milli>(Clock::now() - t0).count(); }

static int usage(const char* a0) {
    std::cerr << "Usage: " << a0 << " <graph_file> <src> <dst> [--variant serial|cpu-ranks|gpu|multi-gpu|naive-gpu]\n"
              << "            [--ranks R] [--gpus G] [--vranks V] [--repeat K] [--trace] [--quiet-path]\n"
              << "       " << a0 << " show <graph_file> [--dense] [--limit K]\n"
              << "       " << a0 << " show-dense <legacy_dense_file>\n"
              << "       " << a0 << " gen er --n N --p P [--seed S] --out F\n"
              << "       " << a0 << " gen rmat --scale S --edgefactor E [--n N] [--seed S] --out F\n"
              << "       " << a0 << " gen grid --width W --height H --out F\n";
    return 1;
}

#define CK(x) do { if (x) { std::cerr << "error: " << bibfs_last_error() << "\n"; return 1; } } while (0)
#define CU(x) do { cudaError_t _e = (x); if (_e != cudaSuccess) { std::cerr << "CUDA error: " << cudaGetErrorString(_e) << " at " #x "\n"; return 1; } } while (0)

struct HostGraph {
    u32 n = 0;
    u64 m = 0;
    std::vector<u32> edges;
    std::vector<u64> row_ptr;
    std::vector<u32> col;
};

static int load_graph(const char* path, HostGraph& g, bool build_csr) {
    CK(bibfs_bin_header(path, &g.n, &g.m));
    g.edges.resize(2 * g.m);
    CK(bibfs_bin_read(path, g.edges.data(), g.m));
    if (build_csr) {
        g.row_ptr.resize((size_t)g.n + 1);
        CK(bibfs_csr_count_host(g.n, g.m, g.edges.data(), nullptr, 0, g.n, g.row_ptr.data()));
        g.col.resize(g.row_ptr[g.n]);
        CK(bibfs_csr_fill_host(g.n, g.m, g.edges.data(), nullptr, 0, g.n, g.row_ptr.data(), g.col.data(), 0));
    }
    return 0;
}

static void print_path(const std::vector<u32>& path, u32 len, const char* prefix) {
    std::cout << prefix;
    for (u32 i = 0; i < len; ++i) std::cout << path[i] << (i + 1 < len ? ' ' : '\n');
}

static int cmd_show(int argc, char** argv) {
    if (argc < 3) return usage(argv[0]);
    bool dense = false;
    u64 limit = ~0ull;
    for (int i = 3; i < argc; ++i) {
        if (!strcmp(argv[i], "--dense")) dense = true;
        else if (!strcmp(argv[i], "--limit") && i + 1 < argc) limit = strtoull(argv[++i], nullptr, 10);
    }
    HostGraph g;
    if (load_graph(argv[2], g, true)) return 1;
    std::cout << "Loaded CSR with shape (" << g.n << ", " << g.n << ") and " << g.row_ptr[g.n] << " nonzeros (" << g.m << " undirected edges)\n";
    if (dense) {
        if (g.n > 64) { std::cerr << "--dense is for N <= 64\n"; return 1; }
        for (u32 u = 0; u < g.n; ++u) {
            std::vector<int> row(g.n, 0);
            for (u64 e = g.row_ptr[u]; e < g.row_ptr[u + 1]; ++e) row[g.col[e]] = 1;
            for (u32 v = 0; v < g.n; ++v) std::cout << row[v] << (v + 1 < g.n ? ' ' : '\n');
        }
        return 0;
    }
    u64 k = 0;
    for (u32 u = 0; u < g.n && k < limit; ++u)
        for (u64 e = g.row_ptr[u]; e < g.row_ptr[u + 1] && k < limit; ++e, ++k) std::cout << u << " -> " << g.col[e] << "\n";
    return 0;
}

static int cmd_show_dense(int argc, char** argv) {
    if (argc < 3) return usage(argv[0]);
    u32 n = 0;
    CK(bibfs_dense_header(argv[2], &n));
    std::vector<uint8_t> a((size_t)n * n);
    CK(bibfs_dense_read(argv[2], a.data(), n));
    std::cout << "Adjacency matrix (" << n << "x" << n << "):\n";
    for (u32 i = 0; i < n; ++i)
        for (u32 j = 0; j < n; ++j) std::cout << (int)a[(size_t)i * n + j] << (j + 1 < n ? ' ' : '\n');
    return 0;
}

static int cmd_gen(int argc, char** argv) {
    if (argc < 3) return usage(argv[0]);
    bibfs_gen_spec s;
    memset(&s, 0, sizeof(s));
    s.seed = 1;
    s.a16 = (u32)(0.57 * 65536); s.ab16 = (u32)(0.76 * 65536); s.abc16 = (u32)(0.95 * 65536);
    std::string kind = argv[2], out;
    u64 ef = 16;
    for (int i = 3; i < argc; ++i) {
        std::string a = argv[i];
        auto next = [&]() -> const char* { return i + 1 < argc ? argv[++i] : "0"; };
        if (a == "--n") s.n = strtoull(next(), nullptr, 10);
        else if (a == "--p") s.p = atof(next());
        else if (a == "--seed") s.seed = strtoull(next(), nullptr, 10);
        else if (a == "--scale") s.scale = (u32)atoi(next());
        else if (a == "--edgefactor") ef = strtoull(next(), nullptr, 10);
        else if (a == "--width") s.width = (u32)atoi(next());
        else if (a == "--height") s.height = (u32)atoi(next());
        else if (a == "--out") out = next();
        else return usage(argv[0]);
    }
    if (out.empty()) return usage(argv[0]);
    if (kind == "er") s.kind = 0;
    else if (kind == "rmat") { s.kind = 1; if (!s.n) s.n = 1ull << s.scale; s.m = ef * s.n; }
    else if (kind == "grid") { s.kind = 2; s.n = (u64)s.width * s.height; }
    else return usage(argv[0]);
    u64 m = 0;
    CK(bibfs_gen_count_host(&s, &m));
    std::vector<u32> e(2 * m);
    CK(bibfs_gen_edges_host(&s, e.data(), m, &m));
    CK(bibfs_bin_write(out.c_str(), (u32)s.n, m, e.data()));
    std::cout << "Wrote " << out << ": N=" << s.n << " M=" << m << "\n";
    return 0;
}

int main(int argc, char** argv) {
    if (argc >= 2 && !strcmp(argv[1], "show")) return cmd_show(argc, argv);
    if (argc >= 2 && !strcmp(argv[1], "show-dense")) return cmd_show_dense(argc, argv);
    if (argc >= 2 && !strcmp(argv[1], "gen")) return cmd_gen(argc, argv);
    if (argc < 4) return usage(argv[0]);
    const char* file = argv[1];
    char* endp = nullptr;
    const long long src_ll = strtoll(argv[2], &endp, 10);
    if (*endp) return usage(argv[0]);
    const long long dst_ll = strtoll(argv[3], &endp, 10);
    if (*endp) return usage(argv[0]);
    std::string variant = "serial";
    int ranks = 4, gpus = 1, vranks = 0, repeat = 1;
    bool trace = false, quiet_path = false;
    for (int i = 4; i < argc; ++i) {
        std::string a = argv[i];
        if (a == "--variant" && i + 1 < argc) variant = argv[++i];
        else if (a == "--ranks" && i + 1 < argc) ranks = atoi(argv[++i]);
        else if (a == "--gpus" && i + 1 < argc) gpus = atoi(argv[++i]);
        else if (a == "--vranks" && i + 1 < argc) vranks = atoi(argv[++i]);
        else if (a == "--repeat" && i + 1 < argc) repeat = atoi(argv[++i]);
        else if (a == "--trace") trace = true;
        else if (a == "--quiet-path") quiet_path = true;
        else return usage(argv[0]);
    }

    auto t_load = Clock::now();
    HostGraph g;
    const bool host_csr = true;   // every variant here builds the CSR on the host; the Python layer uses the device builder
    if (bibfs_bin_header(file, &g.n, &g.m)) { std::cerr << "Cannot open " << file << ": " << bibfs_last_error() << "\n"; return 1; }
    g.edges.resize(2 * g.m);
    if (bibfs_bin_read(file, g.edges.data(), g.m)) { std::cerr << bibfs_last_error() << "\n"; return 1; }
    const double load_ms = ms_since(t_load);
    if (src_ll < 0 || dst_ll < 0 || src_ll >= (long long)g.n || dst_ll >= (long long)g.n) {
        std::cerr << "src/dst must be in [0," << g.n << ")\n";
        return 1;
    }
    const u32 src = (u32)src_ll, dst = (u32)dst_ll;
    auto t_build = Clock::now();
    if (host_csr) {
        g.row_ptr.resize((size_t)g.n + 1);
        CK(bibfs_csr_count_host(g.n, g.m, g.edges.data(), nullptr, 0, g.n, g.row_ptr.data()));
        g.col.resize(g.row_ptr[g.n]);
        CK(bibfs_csr_fill_host(g.n, g.m, g.edges.data(), nullptr, 0, g.n, g.row_ptr.data(), g.col.data(), 0));
    }
    const double build_ms = ms_since(t_build);

    bibfs_result res;
    std::vector<u32> path((size_t)g.n + 1);
    double h2d_ms = 0;

    if (variant == "serial" || variant == "cpu-ranks") {
        for (int r = 0; r < repeat; ++r) {
            if (variant == "serial") CK(bibfs_serial(g.n, g.row_ptr.data(), g.col.data(), src, dst, 0, &res, path.data(), (u32)path.size()));
            else CK(bibfs_cpu_ranks(g.n, g.row_ptr.data(), g.col.data(), src, dst, ranks, &res, path.data(), (u32)path.size()));
        }
        if (res.length < 0) {
            if (variant == "serial") std::cout << "No path found between " << src << " and " << dst << "\n";
            else std::cout << "No path found\n";
        } else {
            std::cout << "Shortest path length = " << res.length << "\n";
            if (variant == "serial" && !quiet_path) print_path(path, res.path_len, "Path: ");
        }
        if (variant == "serial") std::cout << "Serial bidirectional BFS took " << res.search_ms * 1e-3 << " seconds\n";
        else {
            std::cout << "[Time] bidir\xE2\x80\x91" "bitset BFS = " << res.search_ms * 1e-3 << " seconds\n";
            if (res.length >= 0 && !quiet_path) print_path(path, res.path_len, "Path: ");
        }
    } else if (variant == "naive-gpu") {
        u64* d_rp; u32* d_col;
        auto t_h2d = Clock::now();
        CU(cudaMalloc(&d_rp, g.row_ptr.size() * 8)); CU(cudaMalloc(&d_col, g.col.size() * 4 + 4));
        CU(cudaMemcpy(d_rp, g.row_ptr.data(), g.row_ptr.size() * 8, cudaMemcpyHostToDevice));
        CU(cudaMemcpy(d_col, g.col.data(), g.col.size() * 4, cudaMemcpyHostToDevice));
        h2d_ms = ms_since(t_h2d);
        int32_t len = -1; float ms = 0;
        for (int r = 0; r < repeat; ++r) CK(bibfs_naive_levelsync_device(g.n, d_rp, d_col, src, dst, &len, &ms, nullptr));
        res = bibfs_result(); res.length = len; res.search_ms = ms;
        std::cout << "DEBUG: gpu bfs time = " << ms << " ms\n";
        if (len >= 0) std::cout << "Shortest path length = " << len << "\n";
        std::cout << "GPU BFS: " << (len >= 0 ? "PATH FOUND" : "NO PATH") << "\n";
    } else if (variant == "gpu" || variant == "multi-gpu") {
        int ndev = 0;
        CU(cudaGetDeviceCount(&ndev));
        if (ndev < 1) { std::cerr << "no CUDA device\n"; return 1; }
        u32 P = 1, ngroups = 1, vr = 1;
        if (variant == "multi-gpu") {
            if (vranks > 0) { P = (u32)vranks; ngroups = 1; vr = P; }         // P ranks emulated on one GPU
            else { P = (u32)gpus; ngroups = P; vr = 1; }
            if (P < 1 || P > BIBFS_MAX_RANKS) { std::cerr << "ranks must be 1.." << BIBFS_MAX_RANKS << "\n"; return 1; }
            if (vranks <= 0 && (int)P > ndev) { std::cerr << "only " << ndev << " GPUs visible; use --vranks to emulate\n"; return 1; }
        }
        const u32 B = bibfs_block_size(g.n, P);
        auto t_h2d = Clock::now();
        std::vector<bibfs_ctx*> ctx(ngroups, nullptr);
        std::vector<u64*> d_rp(P, nullptr);
        std::vector<u32*> d_col(P, nullptr);
        std::vector<void*> d_vrec(P, nullptr);        // 32-byte vertex records: the front kernel's index (csrc/vrec_dev.cu)
        u32 max_degree = 0;
        for (u32 v = 0; v < g.n; ++v) { const u64 dg = g.row_ptr[v + 1] - g.row_ptr[v]; if (dg > max_degree) max_degree = (u32)(dg > 0xFFFFFFFFull ? 0xFFFFFFFFull : dg); }
        for (u32 r = 0; r < P; ++r) {
            const int dev = (ngroups > 1) ? (int)r : 0;
            CU(cudaSetDevice(dev));
            const u32 lo = r * B < g.n ? r * B : g.n, hi = (u64)(r + 1) * B < g.n ? (r + 1) * B : g.n;
            // shard rows [lo,hi): offsets rebased to the shard, B+1 entries (rows past n are empty)
            std::vector<u64> rp((size_t)B + 1);
            const u64 base = g.row_ptr[lo];
            for (u32 i = 0; i <= B; ++i) { const u32 row = lo + i < hi ? lo + i : hi; rp[i] = g.row_ptr[row] - base; }
            const u64 nnz = g.row_ptr[hi] - base;
            CU(cudaMalloc(&d_rp[r], rp.size() * 8)); CU(cudaMalloc(&d_col[r], nnz * 4 + 16));
            CU(cudaMemcpy(d_rp[r], rp.data(), rp.size() * 8, cudaMemcpyHostToDevice));
            if (nnz) CU(cudaMemcpy(d_col[r], g.col.data() + base, nnz * 4, cudaMemcpyHostToDevice));
            if (hi > lo && nnz < (1ull << 32)) {
                CU(cudaMalloc(&d_vrec[r], (size_t)B * 64));
                CU(cudaMemset(d_vrec[r], 0, (size_t)B * 64));
                CK(bibfs_vrec_build_device(d_rp[r], d_col[r], hi - lo, nnz, d_vrec[r], nullptr, nullptr));
            }
        }
        for (u32 gi = 0; gi < ngroups; ++gi) {
            CU(cudaSetDevice((int)gi));
            bibfs_ctx_desc d;
            memset(&d, 0, sizeof(d));
            d.n = g.n; d.nranks = P; d.rank_lo = gi * vr; d.vranks = vr;
            for (u32 i = 0; i < vr; ++i) { d.d_row_ptr[i] = d_rp[gi * vr + i]; d.d_col[i] = d_col[gi * vr + i]; d.d_vrec[i] = d_vrec[gi * vr + i]; }
            d.max_degree = max_degree;
            d.max_path = g.n + 1 < (1u << 20) ? (1u << 20) : g.n + 1;
            CK(bibfs_ctx_create(&ctx[gi], &d));
        }
        for (u32 a = 0; a < ngroups; ++a)
            for (u32 b = 0; b < ngroups; ++b)
                if (a != b) CK(bibfs_ctx_attach_local(ctx[a], b * vr, ctx[b]));
        h2d_ms = ms_since(t_h2d);
        const u32 flags = trace ? BIBFS_F_TRACE : 0u;
        double worst_search = 0, worst_path = 0;
        for (int r = 0; r < repeat; ++r) {
            for (u32 gi = 0; gi < ngroups; ++gi) CK(bibfs_ctx_query(ctx[gi], src, dst, nullptr, flags, nullptr));
            worst_search = worst_path = 0;
            for (u32 gi = 0; gi < ngroups; ++gi) {
                bibfs_result rr;
                CK(bibfs_ctx_result(ctx[gi], &rr, gi == 0 ? path.data() : nullptr, (u32)path.size(), nullptr));
                if (gi == 0 && rr.status == BIBFS_STATUS_HANDED_OVER) {
                    // several devices in this process: rank 0's front kernel handed the search over, every context continues it
                    for (u32 gj = 0; gj < ngroups; ++gj) CK(bibfs_ctx_resume(ctx[gj], nullptr));
                    CK(bibfs_ctx_result(ctx[0], &rr, path.data(), (u32)path.size(), nullptr));
                } else if (gi > 0 && rr.status == BIBFS_STATUS_HANDED_OVER) {
                    continue;           // rank 0 answered alone
                }
                if (gi == 0) res = rr;
                if (rr.search_ms > worst_search) worst_search = rr.search_ms;   // max over ranks
                if (rr.path_ms > worst_path) worst_path = rr.path_ms;
            }
        }
        res.search_ms = worst_search; res.path_ms = worst_path;
        if (variant == "gpu") {
            std::cout << "DEBUG: gpu bfs time = " << res.search_ms << " ms\n";
            if (res.length >= 0) {
                std::cout << "Shortest path length = " << res.length << "\n";
                if (!quiet_path) print_path(path, res.path_len, "Path: ");
            }
            std::cout << "GPU BFS: " << (res.length >= 0 ? "PATH FOUND" : "NO PATH") << "\n";
        } else {
            if (res.length >= 0) std::cout << "Shortest path length = " << res.length << "\n";
            else std::cout << "No path found\n";
            std::cout << "[Time] MPI+CUDA BI\xE2\x80\x91" "BFS = " << res.search_ms * 1e-3 << " s\n";
            if (res.length >= 0 && !quiet_path) print_path(path, res.path_len, "Path: ");
        }
        if (trace) {
            std::vector<bibfs_level_trace> tr(32768);
            u32 cnt = 0;
            CK(bibfs_ctx_trace(ctx[0], 0, tr.data(), (u32)tr.size(), &cnt));
            u64 t_prev = cnt ? tr[0].t_ns : 0;
            for (u32 i = 0; i < cnt && i < tr.size(); ++i) {
                std::cout << "[level " << i << "] side=" << (tr[i].side ? 'T' : 'S') << " mode=" << tr[i].mode << " frontier=" << tr[i].frontier
                          << " discovered=" << tr[i].discovered << " edges=" << tr[i].edges << " dt_us=" << (tr[i].t_ns - t_prev) * 1e-3 << "\n";
                t_prev = tr[i].t_ns;
            }
        }
        for (auto c : ctx) bibfs_ctx_destroy(c);
    } else {
        return usage(argv[0]);
    }
    std::cout << "[phases] load=" << load_ms << " build=" << build_ms << " h2d=" << h2d_ms << " search=" << res.search_ms
              << " path=" << res.path_ms << " ms  levels=" << res.levels << " edges_scanned=" << res.edges_scanned << "\n";
    return 0;
}
