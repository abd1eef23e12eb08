/*
 * bibfs_b200.h — C ABI of libbibfs_b200.so
 *
 * Blackwell-native bidirectional-BFS shortest-path library.  Everything the Python layer
 * (bidirectional_bfs_b200/) and the native CLI (csrc/cli_main.cpp) do goes through the
 * functions below; they are plain C so any FFI can call them.
 *
 * Reference parity (Tomaszbrauntsch/Bidirectional-BFS): the reference's only library-like
 * surface is v4/comp.h:8-36 (cudaInitGraph / cudaInitFrontiers / cudaExpandFrontier /
 * cudaCheckIntersect / cudaFreeGraph).  Here a graph is a CSR shard, a search context owns
 * all per-query state, and ONE launch answers a whole query (no host round trip per level).
 *
 * Conventions
 *   - every function returns 0 on success, non-zero on failure; bibfs_last_error() gives the
 *     message for the calling thread.
 *   - vertex ids are uint32, edge offsets uint64 (the reference's int row_ptr overflows at
 *     2M >= 2^31, v4/comp.cu:49-55).
 *   - "d_" pointers are device pointers on the current CUDA device, others are host pointers.
 *   - cudaStream_t is passed as void* so that this header needs no CUDA include.
 */
#ifndef BIBFS_B200_H
#define BIBFS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BIBFS_MAX_RANKS 8
#define BIBFS_NO_VERTEX 0xFFFFFFFFu

/* ------------------------------------------------------------------------------------------
 * errors / build info
 * ---------------------------------------------------------------------------------------- */
const char* bibfs_last_error(void);
const char* bibfs_version(void);
/* number of kernel launches this library has issued in this process (bench.py reports it) */
uint64_t bibfs_launch_count(void);

/* ------------------------------------------------------------------------------------------
 * .bin edge-list files: <u32 N><u32 M> then M x <u32 u, u32 v>, little endian, each undirected
 * edge once (writer: reference graphs/generate_graph.py:35-39; readers v1/main-v1.cpp:26-40).
 * ---------------------------------------------------------------------------------------- */
int bibfs_bin_header(const char* path, uint32_t* n, uint64_t* m);
/* edges must hold 2*m uint32; validates ids < n (the reference does no validation, Q9). */
int bibfs_bin_read(const char* path, uint32_t* edges, uint64_t m);
int bibfs_bin_write(const char* path, uint32_t n, uint64_t m, const uint32_t* edges);
/* legacy dense format of v2/read_in.cpp:13-24: <u32 N> + N*N bytes; out holds N*N bytes. */
int bibfs_dense_header(const char* path, uint32_t* n);
int bibfs_dense_read(const char* path, uint8_t* out, uint32_t n);

/* ------------------------------------------------------------------------------------------
 * graph generators (counter-based, seeded; identical output on host and device)
 *   kind 0  Erdos-Renyi G(n,p)  (reference generate_graph.py:31 uses networkx fast_gnp, unseeded)
 *   kind 1  R-MAT / Graph500 Kronecker (m edges, 2^scale id space folded onto n, ids scrambled)
 *   kind 2  2-D grid  width x height, 4-neighbour
 * ---------------------------------------------------------------------------------------- */
typedef struct {
    int32_t  kind;
    uint32_t scale;          /* RMAT: id space 2^scale                                   */
    uint64_t n;              /* vertices                                                   */
    uint64_t m;              /* RMAT: edges to draw; ER/GRID: ignored on input             */
    double   p;              /* ER edge probability                                        */
    uint32_t a16, ab16, abc16; /* RMAT cumulative quadrant thresholds scaled by 65536     */
    uint32_t width, height;  /* GRID                                                       */
    uint64_t seed;
} bibfs_gen_spec;

/* exact number of (u,v) records the generator emits (ER needs a counting pass). */
int bibfs_gen_count_host(const bibfs_gen_spec* spec, uint64_t* m_out);
int bibfs_gen_edges_host(const bibfs_gen_spec* spec, uint32_t* edges, uint64_t cap, uint64_t* m_out);
/* device versions; d_ws is caller-provided scratch of at least bibfs_gen_workspace_bytes(). */
size_t bibfs_gen_workspace_bytes(const bibfs_gen_spec* spec);
int bibfs_gen_count_device(const bibfs_gen_spec* spec, void* d_ws, size_t ws_bytes,
                           uint64_t* m_out, void* stream);
int bibfs_gen_edges_device(const bibfs_gen_spec* spec, uint32_t* d_edges, uint64_t cap,
                           void* d_ws, size_t ws_bytes, uint64_t* m_out, void* stream);

/* ------------------------------------------------------------------------------------------
 * CSR construction for the row range [row_lo,row_hi) of the mirrored (undirected) graph.
 * Self loops are dropped.  Source of edges is either an edge array or a generator spec
 * (fused: the edge list is never materialised).  Reference: host loops in
 * v3/bibfs_cuda_only.cu:89-99 and v4/mpi_bas.cpp:45-57.
 * Two-step protocol: *_count fills row_ptr[0..rows] (uint64), then the caller allocates
 * col[row_ptr[rows]] and calls *_fill.
 * ---------------------------------------------------------------------------------------- */
int bibfs_csr_count_host(uint32_t n, uint64_t m, const uint32_t* edges, const bibfs_gen_spec* spec,
                         uint32_t row_lo, uint32_t row_hi, uint64_t* row_ptr);
int bibfs_csr_fill_host(uint32_t n, uint64_t m, const uint32_t* edges, const bibfs_gen_spec* spec,
                        uint32_t row_lo, uint32_t row_hi, const uint64_t* row_ptr, uint32_t* col,
                        int sort_rows);

size_t bibfs_csr_workspace_bytes(uint32_t rows);
int bibfs_csr_count_device(uint32_t n, uint64_t m, const uint32_t* d_edges, const bibfs_gen_spec* spec,
                           uint32_t row_lo, uint32_t row_hi, uint64_t* d_row_ptr,
                           void* d_ws, size_t ws_bytes, uint64_t* nnz_out, void* stream);
int bibfs_csr_fill_device(uint32_t n, uint64_t m, const uint32_t* d_edges, const bibfs_gen_spec* spec,
                          uint32_t row_lo, uint32_t row_hi, const uint64_t* d_row_ptr, uint32_t* d_col,
                          void* d_ws, size_t ws_bytes, int sort_rows, void* stream);
/* exclusive scan uint32 -> uint64 (out has count+1 entries; out[count] = total). */
size_t bibfs_scan_workspace_bytes(uint64_t count);
int bibfs_scan_u32_device(const uint32_t* d_in, uint64_t* d_out, uint64_t count,
                          void* d_ws, size_t ws_bytes, void* stream);
/* The same scan for callers that cannot allocate (no counterpart in the reference, which
 * allocates everything with cudaMalloc up front, v4/comp.cu:49-73).  Meant for code that is
 * handed only its inputs and outputs — a custom operator inside an inference server or another
 * framework's stream pipeline — and must neither exceed a memory budget fixed at start-up nor
 * call cudaMalloc/cudaFree, which can synchronise the device under the host application.
 * The scratch is a fixed array inside this library's device image instead:
 *   - every GPU that loads the library holds it (128 KB) from load until the process exits;
 *   - count must be <= bibfs_scan_noalloc_max_count() (2^26); larger inputs are rejected with an
 *     error, use bibfs_scan_u32_device with your own scratch for those;
 *   - the call performs no cudaMalloc / cudaFree and calls no allocating routine;
 *   - the FIRST call on each device looks up the array's device address (a one-off runtime
 *     query, with that device current); make that first call before starting any stream capture;
 *   - all calls on a device share the array and it keeps the previous call's contents (nothing
 *     in it needs to start at zero, each call overwrites what it reads); two calls that overlap
 *     on the GPU — different streams, or different host threads — race on it and give wrong
 *     results: serialise them on one stream.
 * d_in / d_out are the caller's, exactly as for bibfs_scan_u32_device (d_out has count+1 entries). */
uint64_t bibfs_scan_noalloc_max_count(void);
int bibfs_scan_u32_device_noalloc(const uint32_t* d_in, uint64_t* d_out, uint64_t count, void* stream);

/* ------------------------------------------------------------------------------------------
 * results
 * ---------------------------------------------------------------------------------------- */
typedef struct {
    int32_t  length;         /* hops; -1 = no path                                          */
    uint32_t meet;           /* meeting vertex (BIBFS_NO_VERTEX if none / src==dst)         */
    uint32_t depth_s, depth_t; /* radius reached by each side                               */
    uint32_t levels;         /* single-side expansions performed                            */
    uint32_t levels_bottom_up;
    uint32_t path_len;       /* vertices written to the path buffer (length+1), 0 if none   */
    int32_t  status;         /* 0 ok; 1 watchdog abort; 2 bad arguments; 3 path found but longer than the
                                path buffer (length is valid, path_len = 0: retry with a larger max_path / path_cap) */
    uint64_t edges_scanned;  /* directed adjacency entries inspected                        */
    uint64_t vertices_visited;
    double   search_ms;      /* host variants: wall; device variants: cudaEvent search kernel */
    double   path_ms;
    uint32_t levels_grid;    /* levels that needed the whole grid (grid top-down, hub phase, bottom-up) */
    uint32_t levels_paired;  /* levels run by the pair tier (both directions per round)      */
} bibfs_result;

/* ------------------------------------------------------------------------------------------
 * CPU variants
 *   serial     — v1 equivalent (v1/main-v1.cpp:50-97): smaller frontier first, discovery-time
 *                meet test, parent walk.  policy 0 = v1 schedule exactly (early break inside
 *                the level); policy 1 = finish the level (what the GPU kernels do).
 *   cpu_ranks  — v2 equivalent (v2/second_try.cpp:53-129): R worker threads ("ranks"), bitset
 *                frontiers, vertex ids strided by rank, OR-merge at a barrier.  Unlike v2 it
 *                reports the true length (v2 prints ceil(d/2)) and keeps parents.
 * ---------------------------------------------------------------------------------------- */
int bibfs_serial(uint32_t n, const uint64_t* row_ptr, const uint32_t* col,
                 uint32_t src, uint32_t dst, int policy,
                 bibfs_result* res, uint32_t* path, uint32_t path_cap);
int bibfs_cpu_ranks(uint32_t n, const uint64_t* row_ptr, const uint32_t* col,
                    uint32_t src, uint32_t dst, int ranks,
                    bibfs_result* res, uint32_t* path, uint32_t path_cap);
/* path check: endpoints, every hop an edge, hop count == length. returns 0 if valid. */
int bibfs_validate_path(uint32_t n, const uint64_t* row_ptr, const uint32_t* col,
                        uint32_t src, uint32_t dst, const uint32_t* path, uint32_t path_len);

/* ------------------------------------------------------------------------------------------
 * GPU search context (single GPU, multi GPU, and P "virtual ranks" on one GPU for testing).
 *
 * Vertices are block-partitioned: rank r owns [r*B, min(n,(r+1)*B)), B = bibfs_block_size(n,P).
 * A context hosts `vranks` consecutive ranks starting at rank_lo on the current device:
 *   - production multi-GPU: one process per GPU, vranks = 1, peers attached through CUDA IPC;
 *   - single process, several devices: one context per device, attached with attach_local;
 *   - emulation: vranks = P on one device (one cooperative launch runs all ranks' CTAs).
 * ---------------------------------------------------------------------------------------- */
typedef struct bibfs_ctx bibfs_ctx;

typedef struct {
    uint32_t n;                 /* global vertex count                                     */
    uint32_t nranks;            /* P, 1..BIBFS_MAX_RANKS                                   */
    uint32_t rank_lo;           /* first rank hosted here                                  */
    uint32_t vranks;            /* ranks hosted here                                       */
    const uint64_t* d_row_ptr[BIBFS_MAX_RANKS]; /* per hosted rank i: rows of rank rank_lo+i */
    const uint32_t* d_col[BIBFS_MAX_RANKS];     /* 16-byte aligned, and readable for 12 bytes past the last entry: rows of
                                                   >= 1024 entries are fetched with 16-byte-granular bulk copies (TMA)    */
    uint32_t max_path;          /* capacity of the device path buffer (0 -> default)       */
    uint32_t solo_threshold;    /* frontier <= this is expanded by one CTA (0 -> default)  */
    uint32_t bottom_up_div;     /* go bottom-up when frontier > n/div (0 -> default, ~0u never) */
    uint32_t max_ctas;          /* cap on CTAs per rank (0 -> all SMs)                     */
    uint32_t watchdog_ms;       /* spin-loop timeout (0 -> default 4000)                   */
    uint32_t max_degree;        /* longest row of the whole graph, 0 = unknown                */
    const void* d_vrec[BIBFS_MAX_RANKS];        /* per hosted rank: vertex records (bibfs_vrec_build_device) or NULL.
                                                   With records, small searches run on the pair tier: both directions
                                                   per round, one memory round trip per round                            */
} bibfs_ctx_desc;

uint32_t bibfs_block_size(uint32_t n, uint32_t nranks);

/* Sharded CSR construction (one rank per GPU): a rank walks only its slice [chunk_lo, chunk_hi) of the edge stream (the
 * generator's chunk space, bibfs_gen_chunks, or its M/P slice of the edge records) and routes each half-edge into the
 * OWNER's degree histogram / cursor / column arrays through peer mappings (pointer tables indexed by rank; entries of
 * other processes come from bibfs_ipc_import).  Sequence: route_degree -> barrier -> every rank scans its histogram
 * (bibfs_scan_u32_device), allocates col, copies the row offsets into its cursors -> barrier -> route_fill -> barrier.
 * Replaces the reference's MPI_Bcast of the whole edge list (v4/mpi_bas.cpp:39-42). */
uint64_t bibfs_gen_chunks(const bibfs_gen_spec* spec);
int bibfs_csr_route_degree_device(uint32_t n, uint64_t m, const uint32_t* d_edges, const bibfs_gen_spec* spec,
                                  uint64_t chunk_lo, uint64_t chunk_hi, uint32_t nranks,
                                  uint32_t* const* d_deg_of_rank, uint32_t* d_bad_flag, void* stream);
int bibfs_csr_route_fill_device(uint32_t n, uint64_t m, const uint32_t* d_edges, const bibfs_gen_spec* spec,
                                uint64_t chunk_lo, uint64_t chunk_hi, uint32_t nranks,
                                uint64_t* const* d_cursor_of_rank, uint32_t* const* d_col_of_rank, void* stream);
int bibfs_csr_sort_rows_device(const uint64_t* d_row_ptr, uint32_t* d_col, uint32_t rows, void* stream);

/* Vertex records of a CSR shard: 64 bytes per row = {degree, row start (32 bits), first six neighbours} + {neighbours 7..14}.
 * d_vrec: bibfs_vrec_bytes(rows) bytes, 32-byte aligned; nnz = stored entries of the shard (< 2^32).
 * bibfs_ctx_desc::max_degree (largest row of the WHOLE graph, 0 = unknown) lets the search skip the handling of long
 * rows on graphs that have none.
 * No counterpart in the reference. */
size_t bibfs_vrec_bytes(uint32_t rows);
int bibfs_vrec_build_device(const uint64_t* d_row_ptr, const uint32_t* d_col, uint32_t rows, uint64_t nnz,
                            void* d_vrec, uint32_t* d_max_degree /* device word, atomicMax'ed; may be NULL */, void* stream);

int bibfs_ctx_create(bibfs_ctx** out, const bibfs_ctx_desc* desc);
int bibfs_ctx_destroy(bibfs_ctx* ctx);
/* 64-byte CUDA IPC handle of hosted rank i's arena. */
int bibfs_ctx_ipc_handle(bibfs_ctx* ctx, uint32_t i, void* handle64);
int bibfs_ctx_attach_ipc(bibfs_ctx* ctx, uint32_t rank, const void* handle64);
int bibfs_ctx_attach_local(bibfs_ctx* ctx, uint32_t rank, bibfs_ctx* owner);
size_t bibfs_ctx_arena_bytes(const bibfs_ctx* ctx);
/* adjacency entries / vertex records of a rank hosted elsewhere, as mapped on this device (peer access or an IPC
 * mapping): with them for every rank, small queries are answered by rank 0 alone (front kernel, csrc/front.cuh). */
int bibfs_ctx_set_peer_graph(bibfs_ctx* ctx, uint32_t rank, const uint32_t* d_col, const void* d_vrec);
/* Share a device buffer allocated by somebody else (a framework's caching allocator) with another process: the IPC
 * handle of the allocation it lies in plus its offset; the importer maps each allocation once per process and device
 * and keeps it mapped.  Used for the col / vertex-record tensors of the one-process-per-GPU deployment. */
int bibfs_ipc_export(const void* d_ptr, void* handle64, uint64_t* offset);
/* device-to-device copy between raw pointers (e.g. from an imported peer mapping into a local tensor) */
int bibfs_copy_device(void* d_dst, const void* d_src, size_t bytes, void* stream);
int bibfs_ipc_import(const void* handle64, uint64_t offset, void** d_ptr);
/* continue a query that returned BIBFS_STATUS_HANDED_OVER: call it on every context of the group, then fetch the results */
int bibfs_ctx_resume(bibfs_ctx* ctx, void* stream);
/* re-zero all per-query state (needed only after a watchdog abort; on every rank of the group). */
int bibfs_ctx_reset(bibfs_ctx* ctx, void* stream);

/* flags for bibfs_ctx_query */
#define BIBFS_F_TRACE        1u   /* record per-level trace (side, mode, frontier, ns)      */
#define BIBFS_F_FORCE_GRID   2u   /* never use the one-CTA path                             */
#define BIBFS_F_FORCE_TD     4u   /* never go bottom-up                                     */
#define BIBFS_F_FORCE_BU     8u   /* always go bottom-up                                    */
#define BIBFS_F_NO_PATH     16u   /* skip the path kernel                                   */
#define BIBFS_F_DEVICE_CLOCK 32u  /* one rank: no CUDA events around the launch; search_ms / path_ms come from the
                                     device's own clock and completion is detected by polling the pinned result    */
#define BIBFS_F_NO_FRONT    64u   /* go straight to the general kernel (no front kernel / pair tier launch)                */
#define BIBFS_F_ALL        127u   /* every defined flag bit; bibfs_ctx_query rejects any other bit */
/* bibfs_result::status of a context whose query was handed over by the front kernel and must be continued with
 * bibfs_ctx_resume on every context of the group (only contexts of ONE process spanning several devices see it) */
#define BIBFS_STATUS_HANDED_OVER 100

/* Enqueue one query: (optional) H2D of the query from pinned host memory, the persistent
 * search kernel, and the path kernel.  h_query_pinned = {src,dst} in pinned memory (8-byte aligned), or NULL
 * to pass src/dst by value.  Asynchronous on `stream`.  Every rank must call it with the same
 * arguments. */
int bibfs_ctx_query(bibfs_ctx* ctx, uint32_t src, uint32_t dst, const uint32_t* h_query_pinned,
                    uint32_t flags, void* stream);
/* Wait for the query and fetch the result; path may be NULL.  When h_pinned_scratch is given
 * (>= 64 + 4*path_cap bytes of pinned memory) the D2H copies land there. */
int bibfs_ctx_result(bibfs_ctx* ctx, bibfs_result* res, uint32_t* path, uint32_t path_cap,
                     void* stream);
/* bibfs_ctx_query + bibfs_ctx_result in one call. */
int bibfs_ctx_query_result(bibfs_ctx* ctx, uint32_t src, uint32_t dst, const uint32_t* h_query_pinned, uint32_t flags,
                           bibfs_result* res, uint32_t* path, uint32_t path_cap, void* stream);
/* device timing of the last query as measured by cudaEvents around each kernel */
int bibfs_ctx_timing(bibfs_ctx* ctx, float* search_ms, float* path_ms);

typedef struct {
    uint32_t side;       /* 0 = source side, 1 = target side                               */
    uint32_t mode;       /* 0 solo top-down, 1 grid top-down, 2 bottom-up; +8 hub phase; +16 paired round */
    uint32_t frontier;   /* global frontier vertices expanded at this level                 */
    uint32_t discovered; /* global vertices discovered                                      */
    uint64_t t_ns;       /* %globaltimer at the end of the level (rank-local)               */
    uint64_t edges;      /* adjacency entries inspected by this rank in this level          */
    uint64_t t_expand_ns;/* %globaltimer when this rank finished expanding (sync starts)    */
} bibfs_level_trace;
int bibfs_ctx_trace(bibfs_ctx* ctx, uint32_t i, bibfs_level_trace* out, uint32_t cap, uint32_t* count);

/* ------------------------------------------------------------------------------------------
 * batched queries: bit-parallel multi-source bidirectional BFS, up to 64 queries per pass
 * (single device, whole graph).  lengths[i] = hops or -1.  d_ws: bibfs_batch_workspace_bytes.
 * No counterpart in the reference (one query per process).
 * ---------------------------------------------------------------------------------------- */
size_t bibfs_batch_workspace_bytes(uint32_t n);
int bibfs_batch_query_device(uint32_t n, const uint64_t* d_row_ptr, const uint32_t* d_col,
                             const uint32_t* h_src, const uint32_t* h_dst, uint32_t nq,
                             int32_t* h_lengths, uint32_t* h_meet, void* d_ws, size_t ws_bytes,
                             float* search_ms, void* stream);

/* Query pool: `lanes` search arenas; a batch of nq queries is ONE launch whose thread-block clusters (2 or 4 CTAs) are
 * the lanes: each draws queries from a shared counter until the list is used up and writes its 64-byte results straight
 * into pinned host memory.  A bidirectional search is bound by the latency of its level chain, so independent queries
 * scale with the number of SMs.  (Without vertex records the lanes fall back to one general-kernel launch per query on
 * their own streams, ctas_per_lane CTAs each, 0 -> 1.)  Outputs may be NULL.  ms = device time of the batch.
 * Out-of-range queries report length -1.  No counterpart in the reference. */
typedef struct bibfs_pool bibfs_pool;
int bibfs_pool_create(bibfs_pool** out, const bibfs_ctx_desc* desc, uint32_t lanes, uint32_t ctas_per_lane);
/* Pool on a PARTITIONED graph (desc->nranks > 1, created in the process that hosts rank 0): every lane is rank 0's front
 * kernel working alone through the peer mappings, so the lanes need the other ranks' col / vertex records as mapped on
 * this device.  h_len[q] = -2 marks a query the lanes could not finish (frontier beyond 65536 vertices per side): answer
 * it with bibfs_ctx_query on the group. */
int bibfs_pool_set_peer_graph(bibfs_pool* pool, uint32_t rank, const uint32_t* d_col, const void* d_vrec);
int bibfs_pool_query(bibfs_pool* pool, const uint32_t* h_src, const uint32_t* h_dst, uint32_t nq,
                     int32_t* h_len, uint32_t* h_meet, uint32_t* h_levels, uint64_t* h_edges, float* ms);
int bibfs_pool_destroy(bibfs_pool* pool);

/* Resident solver (single rank): ONE front kernel that stays on the GPU and takes its queries from a mailbox in pinned
 * host memory — no launch per query.  bibfs_resident_query writes {src, dst} into the mailbox, the kernel (polling it
 * over PCIe) runs the search, walks the path and writes [result | path] back, the host polls the completion word: a
 * query costs the search plus three PCIe trips instead of a launch and a teardown.  The kernel occupies `cluster` SMs
 * (1, 2, 4, 8 or 16; 0 -> 1) while it is resident and leaves by itself after idle_ms without a query (0 -> 50; the next
 * query launches it again), so a device-wide synchronisation elsewhere in the process is delayed by at most that long.
 * A search beyond the front kernel (frontier > 65536 per side) is answered by the general kernel after the resident
 * kernel has been sent home.  flags: 0 or BIBFS_F_NO_PATH.  Needs vertex records (desc->d_vrec[0]).  One caller
 * thread per solver.  No counterpart in the reference (one process per query). */
typedef struct bibfs_resident bibfs_resident;
int bibfs_resident_start(bibfs_resident** out, const bibfs_ctx_desc* desc, uint32_t cluster, uint32_t idle_ms);
int bibfs_resident_query(bibfs_resident* r, uint32_t src, uint32_t dst, uint32_t flags,
                         bibfs_result* res, uint32_t* path, uint32_t path_cap);
int bibfs_resident_launches(const bibfs_resident* r, uint64_t* launches);   /* how often the kernel had to be (re)launched */
int bibfs_resident_stop(bibfs_resident* r);                                  /* sends the kernel home and frees everything */

/* ------------------------------------------------------------------------------------------
 * reference-style level-synchronous GPU baseline (one thread per vertex over all N, int flag
 * arrays, host in the loop) — OUR restatement of the v3 algorithm (v3/bibfs_cuda_only.cu:173-203)
 * used only as an in-process comparison point; the real reference binaries are built by
 * baseline/build_ref.sh.
 * ---------------------------------------------------------------------------------------- */
int bibfs_naive_levelsync_device(uint32_t n, const uint64_t* d_row_ptr, const uint32_t* d_col,
                                 uint32_t src, uint32_t dst, int32_t* length, float* ms, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* BIBFS_B200_H */
