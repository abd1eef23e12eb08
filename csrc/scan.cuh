// scan.cuh — hand-written device-wide exclusive scan (uint32 -> uint64), reduce-then-scan.
//
// Three launches: per-tile sums, one-CTA scan of the tile sums, per-tile scan with the tile
// offset added.  Scratch = one uint64 per tile, provided by the caller.  No CUB/Thrust.
#pragma once
#include "common.h"

namespace bibfs {

constexpr int SCAN_THREADS = 1024;
constexpr int SCAN_ITEMS = 4;                       // per thread
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

inline u64 scan_tiles(u64 count) { return (count + SCAN_TILE - 1) / SCAN_TILE; }
inline size_t scan_workspace_bytes(u64 count) { return (size_t)(scan_tiles(count) + 1) * sizeof(u64); }

#if defined(__CUDACC__)
__device__ __forceinline__ u64 warp_incl_scan(u64 v, int lane) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        u64 t = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d) v += t;
    }
    return v;
}

// CTA-wide exclusive scan of one value per thread; returns the exclusive prefix, total in *tot.
__device__ __forceinline__ u64 block_excl_scan(u64 v, u64* smem_warp /*[33]*/, u64* tot) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nwarps = (blockDim.x + 31) >> 5;
    u64 inc = warp_incl_scan(v, lane);
    if (lane == 31) smem_warp[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        u64 w = lane < nwarps ? smem_warp[lane] : 0;
        u64 winc = warp_incl_scan(w, lane);
        smem_warp[lane] = winc - w;
        if (lane == 31) smem_warp[32] = winc;
    }
    __syncthreads();
    u64 r = smem_warp[warp] + inc - v;
    *tot = smem_warp[32];
    __syncthreads();
    return r;
}

static __global__ void __launch_bounds__(SCAN_THREADS) k_scan_tile_sums(const u32* __restrict__ in, u64 count, u64* __restrict__ tile_sums) {
    __shared__ u64 sw[33];
    const u64 base = (u64)blockIdx.x * SCAN_TILE + (u64)threadIdx.x * SCAN_ITEMS;
    u64 s = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i)
        if (base + i < count) s += in[base + i];
    u64 tot;
    block_excl_scan(s, sw, &tot);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = tot;
}

static __global__ void __launch_bounds__(SCAN_THREADS) k_scan_tile_offsets(u64* __restrict__ tile_sums, u64 ntiles) {
    __shared__ u64 sw[33];
    u64 carry = 0;
    for (u64 base = 0; base < ntiles; base += SCAN_THREADS) {
        u64 i = base + threadIdx.x;
        u64 v = i < ntiles ? tile_sums[i] : 0;
        u64 tot;
        u64 ex = block_excl_scan(v, sw, &tot);
        if (i < ntiles) tile_sums[i] = carry + ex;
        carry += tot;
    }
    if (threadIdx.x == 0) tile_sums[ntiles] = carry;
}

static __global__ void __launch_bounds__(SCAN_THREADS) k_scan_apply(const u32* __restrict__ in, u64 count, const u64* __restrict__ tile_offs,
                                                            u64* __restrict__ out) {
    __shared__ u64 sw[33];
    const u64 base = (u64)blockIdx.x * SCAN_TILE + (u64)threadIdx.x * SCAN_ITEMS;
    u32 v[SCAN_ITEMS];
    u64 s = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) {
        v[i] = base + i < count ? in[base + i] : 0;
        s += v[i];
    }
    u64 tot;
    u64 ex = block_excl_scan(s, sw, &tot) + tile_offs[blockIdx.x];
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) {
        if (base + i < count) out[base + i] = ex;
        ex += v[i];
    }
    if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) out[count] = tile_offs[gridDim.x];
}

// host launcher; ws must hold scan_workspace_bytes(count)
inline int scan_u32_to_u64(const u32* d_in, u64* d_out, u64 count, void* d_ws, size_t ws_bytes, cudaStream_t st) {
    if (count == 0) {
        BIBFS_CUDA_TRY(cudaMemsetAsync(d_out, 0, sizeof(u64), st));
        return 0;
    }
    const u64 nt = scan_tiles(count);
    if (ws_bytes < scan_workspace_bytes(count)) return fail("scan workspace too small: %zu < %zu", ws_bytes, scan_workspace_bytes(count));
    if (nt > 0x7FFFFFFFull) return fail("scan: too many tiles");
    u64* ts = (u64*)d_ws;
    k_scan_tile_sums<<<(unsigned)nt, SCAN_THREADS, 0, st>>>(d_in, count, ts);
    BIBFS_KERNEL_CHECK();
    k_scan_tile_offsets<<<1, SCAN_THREADS, 0, st>>>(ts, nt);
    BIBFS_KERNEL_CHECK();
    k_scan_apply<<<(unsigned)nt, SCAN_THREADS, 0, st>>>(d_in, count, ts, d_out);
    BIBFS_KERNEL_CHECK();
    return 0;
}
#endif

}  // namespace bibfs
