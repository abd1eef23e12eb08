// gen_dev.cu — sm_100a graph generator kernels: materialise an edge list on the device.
//
// Erdos-Renyi / R-MAT (Graph500 Kronecker) / 2-D grid, all through Gen::for_each (gen.h).  The
// reference generates on the CPU with networkx (graphs/generate_graph.py:31) and allocates an
// unused dense N x N matrix (:32); these kernels write straight into device memory
// so a graph can go generator -> CSR builder -> search without touching the host or the disk.
// R-MAT and grid have closed-form output offsets; ER needs count -> scan -> emit.
#include "gen.h"
#include "scan.cuh"

using namespace bibfs;

namespace {

__global__ void __launch_bounds__(256) k_gen_count(const Gen g, u32* __restrict__ counts) {
    for (u64 c = (u64)blockIdx.x * blockDim.x + threadIdx.x; c < g.nchunks; c += (u64)gridDim.x * blockDim.x) {
        u32 k = 0;
        g.for_each(c, [&](u32, u32) { ++k; });
        counts[c] = k;
    }
}

__global__ void __launch_bounds__(256) k_gen_emit_scanned(const Gen g, const u64* __restrict__ offs, u32* __restrict__ edges, u64 cap) {
    for (u64 c = (u64)blockIdx.x * blockDim.x + threadIdx.x; c < g.nchunks; c += (u64)gridDim.x * blockDim.x) {
        u64 o = offs[c];
        g.for_each(c, [&](u32 u, u32 v) {
            if (o < cap) { edges[2 * o] = u; edges[2 * o + 1] = v; }
            ++o;
        });
    }
}

__device__ __forceinline__ u64 grid_offset(const Gen& g, u64 c) {
    const u64 W = g.width, H = g.height;
    const u64 x = c % W, y = c / W;
    const u64 right = y * (W - 1) + x;
    const u64 down = (y < H - 1) ? (y * W + x) : (H - 1) * W;
    return right + down;
}

__global__ void __launch_bounds__(256) k_gen_emit_direct(const Gen g, u32* __restrict__ edges, u64 cap) {
    for (u64 c = (u64)blockIdx.x * blockDim.x + threadIdx.x; c < g.nchunks; c += (u64)gridDim.x * blockDim.x) {
        u64 o = (g.kind == GEN_GRID) ? grid_offset(g, c) : c;
        g.for_each(c, [&](u32 u, u32 v) {
            if (o < cap) {
                // one 8-byte store per record
                *reinterpret_cast<uint2*>(edges + 2 * o) = make_uint2(u, v);
            }
            ++o;
        });
    }
}

unsigned grid_for(u64 items, int threads) {
    u64 b = (items + threads - 1) / threads;
    const u64 cap = 148ull * 32;
    if (b > cap) b = cap;
    if (b == 0) b = 1;
    return (unsigned)b;
}

u64 closed_form_count(const Gen& g) {
    if (g.kind == GEN_GRID) return (u64)g.height * (g.width - 1) + (u64)(g.height - 1) * g.width;
    return g.nchunks;  // RMAT / EDGES: one record per chunk
}

}  // namespace

extern "C" size_t bibfs_gen_workspace_bytes(const bibfs_gen_spec* spec) {
    Gen g;
    std::string why;
    if (!make_gen(*spec, g, why)) return 0;
    if (g.kind != GEN_ER) return 256;
    // counts u32[nchunks] (padded to 8 B) + offsets u64[nchunks+1] + scan scratch
    size_t counts = ((size_t)g.nchunks * 4 + 255) / 256 * 256;
    size_t offs = ((size_t)(g.nchunks + 1) * 8 + 255) / 256 * 256;
    return counts + offs + scan_workspace_bytes(g.nchunks) + 256;
}

static int er_count_and_scan(const Gen& g, void* d_ws, size_t ws_bytes, u64** offs_out, u64* m_out, cudaStream_t st) {
    size_t counts_b = ((size_t)g.nchunks * 4 + 255) / 256 * 256;
    size_t offs_b = ((size_t)(g.nchunks + 1) * 8 + 255) / 256 * 256;
    if (ws_bytes < counts_b + offs_b + scan_workspace_bytes(g.nchunks)) return fail("generator workspace too small");
    u32* counts = (u32*)d_ws;
    u64* offs = (u64*)((char*)d_ws + counts_b);
    void* sws = (char*)d_ws + counts_b + offs_b;
    if (g.nchunks) {
        k_gen_count<<<grid_for(g.nchunks, 256), 256, 0, st>>>(g, counts);
        BIBFS_KERNEL_CHECK();
    }
    if (scan_u32_to_u64(counts, offs, g.nchunks, sws, ws_bytes - counts_b - offs_b, st)) return 1;
    BIBFS_CUDA_TRY(cudaMemcpyAsync(m_out, offs + g.nchunks, sizeof(u64), cudaMemcpyDeviceToHost, st));
    BIBFS_CUDA_TRY(cudaStreamSynchronize(st));
    *offs_out = offs;
    return 0;
}

extern "C" int bibfs_gen_count_device(const bibfs_gen_spec* spec, void* d_ws, size_t ws_bytes, uint64_t* m_out, void* stream) {
    Gen g;
    std::string why;
    if (!make_gen(*spec, g, why)) return fail("generator spec: %s", why.c_str());
    if (g.kind != GEN_ER) { *m_out = closed_form_count(g); return 0; }
    u64* offs;
    return er_count_and_scan(g, d_ws, ws_bytes, &offs, m_out, (cudaStream_t)stream);
}

extern "C" int bibfs_gen_edges_device(const bibfs_gen_spec* spec, uint32_t* d_edges, uint64_t cap,
                                      void* d_ws, size_t ws_bytes, uint64_t* m_out, void* stream) {
    Gen g;
    std::string why;
    cudaStream_t st = (cudaStream_t)stream;
    if (!make_gen(*spec, g, why)) return fail("generator spec: %s", why.c_str());
    if (g.kind != GEN_ER) {
        *m_out = closed_form_count(g);
        if (*m_out > cap) return fail("edge buffer too small: need %llu records", (unsigned long long)*m_out);
        if (g.nchunks) {
            k_gen_emit_direct<<<grid_for(g.nchunks, 256), 256, 0, st>>>(g, d_edges, cap);
            BIBFS_KERNEL_CHECK();
        }
        return 0;
    }
    u64* offs;
    if (er_count_and_scan(g, d_ws, ws_bytes, &offs, m_out, st)) return 1;
    if (*m_out > cap) return fail("edge buffer too small: need %llu records", (unsigned long long)*m_out);
    if (g.nchunks) {
        k_gen_emit_scanned<<<grid_for(g.nchunks, 256), 256, 0, st>>>(g, offs, d_edges, cap);
        BIBFS_KERNEL_CHECK();
    }
    return 0;
}
