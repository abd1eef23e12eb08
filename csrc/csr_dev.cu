// csr_dev.cu — sm_100a CSR builder for one 1-D partition shard (rows [row_lo,row_hi)).
//
// Replaces the reference's host loops (v3/bibfs_cuda_only.cu:89-99, v4/mpi_bas.cpp:45-57: degree
// count -> prefix sum -> cursor fill on the CPU, whole graph replicated on every rank) with
//   k_degree   edge stream -> atomic degree histogram of the owned rows
//   scan       hand-written reduce-then-scan (scan.cuh) -> 64-bit row offsets
//   k_fill     edge stream again -> atomic cursor per row -> column indices
//   k_sort_*   optional per-row sort (warp bitonic in registers / CTA bitonic in shared or
//              global memory) so that results are deterministic
// The edge stream is either a device edge array (.bin files) or the generator itself, in which
// case every rank regenerates the stream and keeps only what it owns: no edge list is ever
// materialised or exchanged, which is what lets a scale-26 Kronecker graph be built in place.
#include "gen.h"
#include "scan.cuh"

using namespace bibfs;

namespace {

__global__ void __launch_bounds__(256) k_degree(const Gen g, u32 n, u32 lo, u32 hi, u32* __restrict__ deg, u32* __restrict__ bad) {
    for (u64 c = (u64)blockIdx.x * blockDim.x + threadIdx.x; c < g.nchunks; c += (u64)gridDim.x * blockDim.x) {
        g.for_each(c, [&](u32 u, u32 v) {
            if (u >= n || v >= n) { *bad = 1; return; }
            if (u == v) return;
            if (u >= lo && u < hi) atomicAdd(&deg[u - lo], 1u);
            if (v >= lo && v < hi) atomicAdd(&deg[v - lo], 1u);
        });
    }
}

__global__ void __launch_bounds__(256) k_fill(const Gen g, u32 n, u32 lo, u32 hi, unsigned long long* __restrict__ cursor,
                                              u32* __restrict__ col) {
    for (u64 c = (u64)blockIdx.x * blockDim.x + threadIdx.x; c < g.nchunks; c += (u64)gridDim.x * blockDim.x) {
        g.for_each(c, [&](u32 u, u32 v) {
            if (u == v || u >= n || v >= n) return;
            if (u >= lo && u < hi) col[atomicAdd(&cursor[u - lo], 1ull)] = v;
            if (v >= lo && v < hi) col[atomicAdd(&cursor[v - lo], 1ull)] = u;
        });
    }
}

// ---- sharded construction: a rank walks only ITS slice of the edge stream and routes every half-edge to the rank that
// owns the row, straight into that rank's degree histogram / cursor array / column array through the peer mappings
// (the reference broadcasts the whole edge list to every rank instead, v4/mpi_bas.cpp:39-42).
struct RoutePtrs {
    u32* deg[BIBFS_MAX_RANKS];                       // owner's degree histogram (u32 per owned row)
    unsigned long long* cursor[BIBFS_MAX_RANKS];     // owner's fill cursors (u64 per owned row, start = row offset)
    u32* col[BIBFS_MAX_RANKS];                       // owner's column array
};

__global__ void __launch_bounds__(256) k_degree_routed(const Gen g, u64 c_lo, u64 c_hi, u32 n, u32 B, u64 div_b, const RoutePtrs R, u32* __restrict__ bad) {
    for (u64 c = c_lo + (u64)blockIdx.x * blockDim.x + threadIdx.x; c < c_hi; c += (u64)gridDim.x * blockDim.x) {
        g.for_each(c, [&](u32 u, u32 v) {
            if (u >= n || v >= n) { *bad = 1; return; }
            if (u == v) return;
            const u32 ou = (u32)__umul64hi(div_b, (u64)u), ov = (u32)__umul64hi(div_b, (u64)v);
            atomicAdd_system(R.deg[ou] + (u - ou * B), 1u);
            atomicAdd_system(R.deg[ov] + (v - ov * B), 1u);
        });
    }
}

__global__ void __launch_bounds__(256) k_fill_routed(const Gen g, u64 c_lo, u64 c_hi, u32 n, u32 B, u64 div_b, const RoutePtrs R) {
    for (u64 c = c_lo + (u64)blockIdx.x * blockDim.x + threadIdx.x; c < c_hi; c += (u64)gridDim.x * blockDim.x) {
        g.for_each(c, [&](u32 u, u32 v) {
            if (u == v || u >= n || v >= n) return;
            const u32 ou = (u32)__umul64hi(div_b, (u64)u), ov = (u32)__umul64hi(div_b, (u64)v);
            // both reservations first (two NVLink round trips in flight), then the two stores
            const unsigned long long su = atomicAdd_system(R.cursor[ou] + (u - ou * B), 1ull);
            const unsigned long long sv = atomicAdd_system(R.cursor[ov] + (v - ov * B), 1ull);
            R.col[ou][su] = v;
            R.col[ov][sv] = u;
        });
    }
}

// rows with <= 32 entries: one warp per row, bitonic sort in registers
__global__ void __launch_bounds__(256) k_sort_rows_warp(const u64* __restrict__ row_ptr, u32* __restrict__ col, u32 rows) {
    const u32 lane = threadIdx.x & 31;
    const u64 wid = ((u64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const u64 nw = ((u64)gridDim.x * blockDim.x) >> 5;
    for (u64 r = wid; r < rows; r += nw) {
        const u64 b = row_ptr[r], e = row_ptr[r + 1];
        const u32 d = (u32)(e - b);
        if (d < 2 || d > 32) continue;
        u32 x = lane < d ? col[b + lane] : 0xFFFFFFFFu;
#pragma unroll
        for (u32 k = 2; k <= 32; k <<= 1) {
#pragma unroll
            for (u32 j = k >> 1; j > 0; j >>= 1) {
                u32 y = __shfl_xor_sync(0xffffffffu, x, j);
                bool up = ((lane & k) == 0);
                bool lower = ((lane & j) == 0);
                u32 mn = x < y ? x : y, mx = x < y ? y : x;
                x = (lower == up) ? mn : mx;
            }
        }
        if (lane < d) col[b + lane] = x;
    }
}

// rows with > 32 entries: one CTA per row.  Bitonic network in its "all ascending" form (first
// step of each stage pairs i with i ^ (k-1), later steps with i ^ j), so indices >= d can be
// treated as +inf padding that never moves: no compare-exchange with a missing partner is needed.
// Shared memory up to 4096 entries, in place in global memory above (slow, rare: hub rows).
constexpr u32 SORT_SMEM = 4096;
__global__ void __launch_bounds__(512) k_sort_rows_cta(const u64* __restrict__ row_ptr, u32* __restrict__ col, u32 rows) {
    __shared__ u32 s[SORT_SMEM];
    for (u32 r = blockIdx.x; r < rows; r += gridDim.x) {
        const u64 b = row_ptr[r], e = row_ptr[r + 1];
        const u64 d = e - b;
        if (d <= 32) continue;
        u64 p2 = 64;
        while (p2 < d) p2 <<= 1;
        const bool in_smem = p2 <= SORT_SMEM;
        u32* a = in_smem ? s : (col + b);
        if (in_smem) {
            for (u32 i = threadIdx.x; i < d; i += blockDim.x) s[i] = col[b + i];
            __syncthreads();
        }
        for (u64 k = 2; k <= p2; k <<= 1) {
            for (u64 j = k >> 1; j > 0; j >>= 1) {
                const bool first = (j == (k >> 1));
                for (u64 i = threadIdx.x; i < d; i += blockDim.x) {
                    const u64 l = first ? (i ^ (k - 1)) : (i ^ j);
                    if (l > i && l < d) {
                        const u32 x = a[i], y = a[l];
                        if (x > y) { a[i] = y; a[l] = x; }
                    }
                }
                __syncthreads();
            }
        }
        if (in_smem) {
            for (u32 i = threadIdx.x; i < d; i += blockDim.x) col[b + i] = s[i];
            __syncthreads();
        }
    }
}

unsigned grid_for(u64 items, int threads) {
    u64 b = (items + threads - 1) / threads;
    const u64 cap = 148ull * 32;
    if (b > cap) b = cap;
    if (b == 0) b = 1;
    return (unsigned)b;
}

int resolve_gen_dev(u32 n, u64 m, const u32* d_edges, const bibfs_gen_spec* spec, Gen& g) {
    if (spec) {
        std::string why;
        if (!make_gen(*spec, g, why)) return fail("generator spec: %s", why.c_str());
        if (g.n != n) return fail("generator n != n");
        return 0;
    }
    if (!d_edges && m) return fail("no edges and no generator spec");
    g = make_edge_gen(n, m, d_edges);
    return 0;
}

size_t align256(size_t x) { return (x + 255) / 256 * 256; }

}  // namespace

extern "C" size_t bibfs_scan_workspace_bytes(uint64_t count) { return scan_workspace_bytes(count) + 256; }

extern "C" int bibfs_scan_u32_device(const uint32_t* d_in, uint64_t* d_out, uint64_t count,
                                     void* d_ws, size_t ws_bytes, void* stream) {
    return scan_u32_to_u64(d_in, d_out, count, d_ws, ws_bytes, (cudaStream_t)stream);
}

// workspace: deg u32[rows+1] | cursor u64[rows] | scan scratch | flag
extern "C" size_t bibfs_csr_workspace_bytes(uint32_t rows) {
    return align256((size_t)(rows + 1) * 4) + align256((size_t)rows * 8 + 8) + align256(scan_workspace_bytes(rows)) + 256;
}

extern "C" int bibfs_csr_count_device(uint32_t n, uint64_t m, const uint32_t* d_edges, const bibfs_gen_spec* spec,
                                      uint32_t row_lo, uint32_t row_hi, uint64_t* d_row_ptr,
                                      void* d_ws, size_t ws_bytes, uint64_t* nnz_out, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (row_lo > row_hi || row_hi > n) return fail("bad row range [%u,%u) for n=%u", row_lo, row_hi, n);
    const u32 rows = row_hi - row_lo;
    if (ws_bytes < bibfs_csr_workspace_bytes(rows)) return fail("csr workspace too small");
    Gen g;
    if (resolve_gen_dev(n, m, d_edges, spec, g)) return 1;
    char* w = (char*)d_ws;
    u32* deg = (u32*)w;
    w += align256((size_t)(rows + 1) * 4);
    w += align256((size_t)rows * 8 + 8);
    void* sws = w;
    w += align256(scan_workspace_bytes(rows));
    u32* bad = (u32*)w;
    BIBFS_CUDA_TRY(cudaMemsetAsync(deg, 0, (size_t)(rows + 1) * 4, st));
    BIBFS_CUDA_TRY(cudaMemsetAsync(bad, 0, 4, st));
    if (g.nchunks) {
        k_degree<<<grid_for(g.nchunks, 256), 256, 0, st>>>(g, n, row_lo, row_hi, deg, bad);
        BIBFS_KERNEL_CHECK();
    }
    if (scan_u32_to_u64(deg, d_row_ptr, rows, sws, align256(scan_workspace_bytes(rows)), st)) return 1;
    u64 nnz = 0;
    u32 hbad = 0;
    BIBFS_CUDA_TRY(cudaMemcpyAsync(&nnz, d_row_ptr + rows, 8, cudaMemcpyDeviceToHost, st));
    BIBFS_CUDA_TRY(cudaMemcpyAsync(&hbad, bad, 4, cudaMemcpyDeviceToHost, st));
    BIBFS_CUDA_TRY(cudaStreamSynchronize(st));
    if (hbad) return fail("edge endpoint >= n");
    if (nnz_out) *nnz_out = nnz;
    return 0;
}

extern "C" int bibfs_csr_fill_device(uint32_t n, uint64_t m, const uint32_t* d_edges, const bibfs_gen_spec* spec,
                                     uint32_t row_lo, uint32_t row_hi, const uint64_t* d_row_ptr, uint32_t* d_col,
                                     void* d_ws, size_t ws_bytes, int sort_rows, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (row_lo > row_hi || row_hi > n) return fail("bad row range [%u,%u) for n=%u", row_lo, row_hi, n);
    const u32 rows = row_hi - row_lo;
    if (ws_bytes < bibfs_csr_workspace_bytes(rows)) return fail("csr workspace too small");
    Gen g;
    if (resolve_gen_dev(n, m, d_edges, spec, g)) return 1;
    char* w = (char*)d_ws;
    w += align256((size_t)(rows + 1) * 4);
    unsigned long long* cursor = (unsigned long long*)w;
    if (rows) BIBFS_CUDA_TRY(cudaMemcpyAsync(cursor, d_row_ptr, (size_t)rows * 8, cudaMemcpyDeviceToDevice, st));
    if (g.nchunks) {
        k_fill<<<grid_for(g.nchunks, 256), 256, 0, st>>>(g, n, row_lo, row_hi, cursor, d_col);
        BIBFS_KERNEL_CHECK();
    }
    if (sort_rows && rows) {
        k_sort_rows_warp<<<grid_for((u64)rows * 32, 256), 256, 0, st>>>(d_row_ptr, d_col, rows);
        BIBFS_KERNEL_CHECK();
        k_sort_rows_cta<<<grid_for((u64)rows * 512, 512), 512, 0, st>>>(d_row_ptr, d_col, rows);
        BIBFS_KERNEL_CHECK();
    }
    return 0;
}


// ------------------------------------------------------------------------------------------
// sharded construction (one process per GPU): see k_degree_routed / k_fill_routed.
//   1. every rank zeroes its degree histogram (u32[B]); all ranks exchange the pointers (IPC)
//   2. route_degree on the rank's chunk slice            -> barrier
//   3. every rank scans ITS histogram (bibfs_scan_u32_device) -> row offsets, allocates col, copies the offsets
//      into its cursor array; pointers exchanged           -> barrier
//   4. route_fill on the rank's chunk slice               -> barrier;  optional bibfs_csr_sort_rows_device
// chunk_lo / chunk_hi: slice of the generator's chunk space (bibfs_gen_chunks), or of the edge records.
// ------------------------------------------------------------------------------------------
extern "C" uint64_t bibfs_gen_chunks(const bibfs_gen_spec* spec) {
    Gen g;
    std::string why;
    if (!spec || !make_gen(*spec, g, why)) return 0;
    return g.nchunks;
}

static int route_args(uint32_t n, uint32_t nranks, uint64_t chunk_lo, uint64_t chunk_hi, const Gen& g) {
    if (nranks < 1 || nranks > BIBFS_MAX_RANKS) return fail("nranks must be 1..%d", BIBFS_MAX_RANKS);
    if (chunk_lo > chunk_hi || chunk_hi > g.nchunks) return fail("bad chunk slice [%llu,%llu) of %llu", (unsigned long long)chunk_lo, (unsigned long long)chunk_hi, (unsigned long long)g.nchunks);
    return 0;
}

extern "C" int bibfs_csr_route_degree_device(uint32_t n, uint64_t m, const uint32_t* d_edges, const bibfs_gen_spec* spec,
                                             uint64_t chunk_lo, uint64_t chunk_hi, uint32_t nranks,
                                             uint32_t* const* d_deg_of_rank, uint32_t* d_bad_flag, void* stream) {
    Gen g;
    if (resolve_gen_dev(n, m, d_edges, spec, g) || route_args(n, nranks, chunk_lo, chunk_hi, g)) return 1;
    RoutePtrs R = {};
    for (u32 r = 0; r < nranks; ++r) R.deg[r] = d_deg_of_rank[r];
    const u32 B = bibfs_block_size(n, nranks);
    if (chunk_hi > chunk_lo) {
        k_degree_routed<<<grid_for(chunk_hi - chunk_lo, 256), 256, 0, (cudaStream_t)stream>>>(g, chunk_lo, chunk_hi, n, B, ~0ull / B + 1ull, R, d_bad_flag);
        BIBFS_KERNEL_CHECK();
    }
    return 0;
}

extern "C" int bibfs_csr_route_fill_device(uint32_t n, uint64_t m, const uint32_t* d_edges, const bibfs_gen_spec* spec,
                                           uint64_t chunk_lo, uint64_t chunk_hi, uint32_t nranks,
                                           uint64_t* const* d_cursor_of_rank, uint32_t* const* d_col_of_rank, void* stream) {
    Gen g;
    if (resolve_gen_dev(n, m, d_edges, spec, g) || route_args(n, nranks, chunk_lo, chunk_hi, g)) return 1;
    RoutePtrs R = {};
    for (u32 r = 0; r < nranks; ++r) { R.cursor[r] = (unsigned long long*)d_cursor_of_rank[r]; R.col[r] = d_col_of_rank[r]; }
    const u32 B = bibfs_block_size(n, nranks);
    if (chunk_hi > chunk_lo) {
        k_fill_routed<<<grid_for(chunk_hi - chunk_lo, 256), 256, 0, (cudaStream_t)stream>>>(g, chunk_lo, chunk_hi, n, B, ~0ull / B + 1ull, R);
        BIBFS_KERNEL_CHECK();
    }
    return 0;
}

extern "C" int bibfs_csr_sort_rows_device(const uint64_t* d_row_ptr, uint32_t* d_col, uint32_t rows, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!rows) return 0;
    k_sort_rows_warp<<<grid_for((u64)rows * 32, 256), 256, 0, st>>>(d_row_ptr, d_col, rows);
    BIBFS_KERNEL_CHECK();
    k_sort_rows_cta<<<grid_for((u64)rows * 512, 512), 512, 0, st>>>(d_row_ptr, d_col, rows);
    BIBFS_KERNEL_CHECK();
    return 0;
}
