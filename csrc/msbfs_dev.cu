// msbfs_dev.cu — batched queries: bit-parallel bidirectional multi-source BFS (sm_100a).
//
// Up to 64 (src,dst) queries share one sweep of the graph per level: bit i of every 64-bit word
// belongs to query i (Then et al., "The More the Merrier", VLDB'14, applied to both directions).
// The reference answers one query per process launch ("Batched / multi-source
// queries: No"); this is the engine behind bibfs_batch() and BASELINE config 5.
//
// One persistent cooperative kernel runs all levels.  Levels alternate source side / target side
// for all still-active queries; a level either PUSHES (vertices with frontier bits scatter them
// to their neighbours with atomicOr) or PULLS (vertices that some query has not reached gather
// their neighbours' frontier bits; no atomics, early exit once every active query reached the
// vertex), chosen from the number of vertices carrying frontier bits.  The meet test is fused:
// newly set bits are ANDed with the other side's visited word at discovery time; the first level
// in which query i meets gives length_i = depth_s + depth_t (levels are complete, so it is exact).
#include "common.h"
#include "ptx.cuh"
#include <cooperative_groups.h>

namespace cg = cooperative_groups;
using namespace bibfs;

namespace {

constexpr int MS_THREADS = 512;

struct MsCtrl {
    u64 active;            // queries still searching
    u64 found_new;         // queries that met in the current level
    u64 any_new;           // queries that discovered at least one vertex in the current level
    u32 nonzero;           // vertices that received frontier bits in the current level
    u32 depth[2];
    u32 front_nonzero[2];  // vertices carrying frontier bits, per side
    u32 mode;              // 0 push, 1 pull
    u32 levels;
    u32 cur[2];            // which buffer is the current frontier, per side
    u32 dirty[2];          // the "next" buffer of that side holds stale bits
    i32 length[64];
    u32 meet[64];
    u64 edges;
};

struct MsParams {
    u32 n, nq;
    const u64* row_ptr;
    const u32* col;
    MsCtrl* C;
    u64* vis[2];
    u64* fr[2][2];
    u32 pull_div;
};

__device__ __forceinline__ void note_meet(MsCtrl* C, u64 meetbits, u32 v) {
    atomicOr((unsigned long long*)&C->found_new, meetbits);
    while (meetbits) {
        const int b = __ffsll((long long)meetbits) - 1;
        meetbits &= meetbits - 1;
        atomicMin(&C->meet[b], v);
    }
}

__global__ void __launch_bounds__(MS_THREADS, 1) k_msbfs(const __grid_constant__ MsParams P) {
    cg::grid_group grid = cg::this_grid();
    MsCtrl* C = P.C;
    const u32 lane = threadIdx.x & 31;
    const u64 gtid = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    const u64 gthreads = (u64)gridDim.x * blockDim.x;
    const u32 gwarp = (u32)(gtid >> 5), nwarps = (u32)(gthreads >> 5);
    const u32 nblk32 = (P.n + 31) / 32;

    for (u32 it = 0; it < 2 * P.n + 2; ++it) {
        const u32 side = it & 1u, other = side ^ 1u;
        const u64 active = C->active;
        if (!active) break;
        const u32 mode = C->mode;
        u64* front = P.fr[side][C->cur[side]];
        u64* next = P.fr[side][C->cur[side] ^ 1u];
        u64* vis = P.vis[side];
        const u64* ovis = P.vis[other];
        if (mode == 0 && C->dirty[side]) {            // a pull level left stale bits in `next`
            for (u64 i = gtid; i < P.n; i += gthreads) next[i] = 0;
            grid.sync();
        }
        u64 any = 0, edges = 0;
        u32 nz = 0;
        if (mode == 0) {
            // ---- push: warp per 32 vertices; long rows by the whole warp, short rows per lane
            for (u32 blk = gwarp; blk < nblk32; blk += nwarps) {
                const u32 u = blk * 32 + lane;
                u64 F = 0, r0 = 0, r1 = 0;
                if (u < P.n) {
                    F = front[u] & active;
                    front[u] = 0;                       // becomes the clean `next` of the level after
                    if (F) { r0 = __ldg(P.row_ptr + u); r1 = __ldg(P.row_ptr + u + 1); }
                }
                const u64 deg = r1 - r0;
                u32 big = __ballot_sync(0xffffffffu, F != 0 && deg >= 64);
                while (big) {
                    const u32 sl = __ffs(big) - 1;
                    big &= big - 1;
                    const u64 FF = __shfl_sync(0xffffffffu, F, sl);
                    const u64 b = __shfl_sync(0xffffffffu, r0, sl), e = __shfl_sync(0xffffffffu, r1, sl);
                    for (u64 p = b + lane; p < e; p += 32) {
                        const u32 v = __ldg(P.col + p);
                        ++edges;
                        const u64 nb = FF & ~vis[v];
                        if (nb) {
                            const u64 old = atomicOr((unsigned long long*)&vis[v], nb);
                            const u64 truly = nb & ~old;
                            if (truly) {
                                const u64 prev = atomicOr((unsigned long long*)&next[v], truly);
                                if (!prev) ++nz;
                                any |= truly;
                                const u64 mb = truly & ovis[v];
                                if (mb) note_meet(C, mb, v);
                            }
                        }
                    }
                }
                if (F != 0 && deg < 64) {
                    for (u64 p = r0; p < r1; ++p) {
                        const u32 v = __ldg(P.col + p);
                        ++edges;
                        const u64 nb = F & ~vis[v];
                        if (nb) {
                            const u64 old = atomicOr((unsigned long long*)&vis[v], nb);
                            const u64 truly = nb & ~old;
                            if (truly) {
                                const u64 prev = atomicOr((unsigned long long*)&next[v], truly);
                                if (!prev) ++nz;
                                any |= truly;
                                const u64 mb = truly & ovis[v];
                                if (mb) note_meet(C, mb, v);
                            }
                        }
                    }
                }
            }
        } else {
            // ---- pull: every vertex some active query has not reached gathers frontier bits
            for (u64 v = gtid; v < P.n; v += gthreads) {
                const u64 seen = vis[v];
                u64 want = active & ~seen;
                u64 got = 0;
                if (want) {
                    const u64 r0 = __ldg(P.row_ptr + v), r1 = __ldg(P.row_ptr + v + 1);
                    for (u64 p = r0; p < r1; ++p) {
                        got |= front[__ldg(P.col + p)];
                        ++edges;
                        if (!(want & ~got)) break;
                    }
                    got &= want;
                }
                next[v] = got;
                if (got) {
                    vis[v] = seen | got;
                    ++nz;
                    any |= got;
                    const u64 mb = got & ovis[v];
                    if (mb) note_meet(C, mb, (u32)v);
                }
            }
        }
        // warp-reduce and publish the level's summaries
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            any |= __shfl_xor_sync(0xffffffffu, any, d);
            edges += __shfl_xor_sync(0xffffffffu, edges, d);
            nz += __shfl_xor_sync(0xffffffffu, nz, d);
        }
        if (lane == 0) {
            if (any) atomicOr((unsigned long long*)&C->any_new, any);
            if (edges) atomicAdd((unsigned long long*)&C->edges, (unsigned long long)edges);
            if (nz) atomicAdd(&C->nonzero, nz);
        }
        grid.sync();
        if (gtid == 0) {
            C->depth[side] += 1;
            C->levels += 1;
            u64 act = C->active;
            u64 fn = C->found_new & act;
            const u32 len = C->depth[0] + C->depth[1];
            while (fn) {
                const int b = __ffsll((long long)fn) - 1;
                fn &= fn - 1;
                C->length[b] = (i32)len;
            }
            act &= ~C->found_new;
            act &= C->any_new;                 // a side that found nothing new is exhausted: no path
            C->active = act;
            C->dirty[side] = (mode == 1) ? 1u : 0u;   // pull reads `front` everywhere: it stays dirty
            C->cur[side] ^= 1u;
            C->front_nonzero[side] = C->nonzero;
            C->found_new = 0; C->any_new = 0; C->nonzero = 0;
            const u32 nxt = other;
            C->mode = (P.pull_div && C->front_nonzero[nxt] > P.n / P.pull_div) ? 1u : 0u;
            __threadfence();
        }
        grid.sync();
    }
}

__global__ void k_ms_seed(MsParams P, const u32* __restrict__ src, const u32* __restrict__ dst) {
    const u32 i = threadIdx.x;
    MsCtrl* C = P.C;
    if (i < 64) { C->length[i] = -1; C->meet[i] = 0xFFFFFFFFu; }
    __syncthreads();
    if (i < P.nq) {
        const u32 s = src[i], d = dst[i];
        if (s < P.n && d < P.n) {
            if (s == d) { C->length[i] = 0; C->meet[i] = s; }
            else {
                atomicOr((unsigned long long*)&P.vis[0][s], 1ull << i);
                atomicOr((unsigned long long*)&P.fr[0][0][s], 1ull << i);
                atomicOr((unsigned long long*)&P.vis[1][d], 1ull << i);
                atomicOr((unsigned long long*)&P.fr[1][0][d], 1ull << i);
                atomicOr((unsigned long long*)&C->active, 1ull << i);
            }
        }
    }
    if (i == 0) { C->front_nonzero[0] = P.nq; C->front_nonzero[1] = P.nq; }
}

size_t al256(size_t x) { return (x + 255) / 256 * 256; }

}  // namespace

extern "C" size_t bibfs_batch_workspace_bytes(uint32_t n) {
    return al256(sizeof(MsCtrl)) + 6 * al256((size_t)n * 8) + al256(64 * 4) * 2 + 256;
}

extern "C" int bibfs_batch_query_device(uint32_t n, const uint64_t* d_row_ptr, const uint32_t* d_col,
                                        const uint32_t* h_src, const uint32_t* h_dst, uint32_t nq,
                                        int32_t* h_lengths, uint32_t* h_meet, void* d_ws, size_t ws_bytes,
                                        float* search_ms, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (nq == 0 || nq > 64) return fail("batch size must be 1..64");
    if (ws_bytes < bibfs_batch_workspace_bytes(n)) return fail("batch workspace too small");
    char* w = (char*)d_ws;
    MsParams P;
    P.n = n; P.nq = nq; P.row_ptr = d_row_ptr; P.col = d_col;
    P.C = (MsCtrl*)w; w += al256(sizeof(MsCtrl));
    const size_t vb = al256((size_t)n * 8);
    P.vis[0] = (u64*)w; w += vb;
    P.vis[1] = (u64*)w; w += vb;
    for (int s = 0; s < 2; ++s) for (int b = 0; b < 2; ++b) { P.fr[s][b] = (u64*)w; w += vb; }
    u32* d_src = (u32*)w; w += al256(64 * 4);
    u32* d_dst = (u32*)w; w += al256(64 * 4);
    P.pull_div = 32;
    BIBFS_CUDA_TRY(cudaMemsetAsync(d_ws, 0, al256(sizeof(MsCtrl)) + 6 * vb, st));
    BIBFS_CUDA_TRY(cudaMemcpyAsync(d_src, h_src, nq * 4, cudaMemcpyHostToDevice, st));
    BIBFS_CUDA_TRY(cudaMemcpyAsync(d_dst, h_dst, nq * 4, cudaMemcpyHostToDevice, st));
    k_ms_seed<<<1, 64, 0, st>>>(P, d_src, d_dst);
    BIBFS_KERNEL_CHECK();
    int dev = 0, sms = 0, per = 0;
    BIBFS_CUDA_TRY(cudaGetDevice(&dev));
    BIBFS_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    BIBFS_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, k_msbfs, MS_THREADS, 0));
    if (per < 1) return fail("batch kernel does not fit");
    if (per > 2) per = 2;
    cudaEvent_t e0, e1;
    BIBFS_CUDA_TRY(cudaEventCreate(&e0));
    BIBFS_CUDA_TRY(cudaEventCreate(&e1));
    BIBFS_CUDA_TRY(cudaEventRecord(e0, st));
    void* args[] = {(void*)&P};
    BIBFS_CUDA_TRY(cudaLaunchCooperativeKernel((void*)k_msbfs, dim3(sms * per), dim3(MS_THREADS), args, 0, st));
    count_launch();
    BIBFS_CUDA_TRY(cudaEventRecord(e1, st));
    MsCtrl hc;
    BIBFS_CUDA_TRY(cudaMemcpyAsync(&hc, P.C, sizeof(MsCtrl), cudaMemcpyDeviceToHost, st));
    BIBFS_CUDA_TRY(cudaStreamSynchronize(st));
    if (search_ms) cudaEventElapsedTime(search_ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    for (u32 i = 0; i < nq; ++i) {
        if (h_lengths) h_lengths[i] = hc.length[i];
        if (h_meet) h_meet[i] = hc.meet[i];
    }
    return 0;
}
