// naive_dev.cu — reference-STYLE level-synchronous GPU Bi-BFS, used only as an in-process
// comparison point for benchmarks ("what the reference's design costs on a B200").
//
// This is our restatement of the v3 scheme (v3/bibfs_cuda_only.cu:13-62,173-203), not its code:
// one int flag per vertex for frontier / next / visited, one thread per vertex over all N every
// level, a separate intersect kernel, and the host in the loop (memset + launch + sync + D2H
// flag per level).  Unlike v3 the intersect test runs after each single-side expansion so the
// hop count is exact (v3 prints none).  The real v3/v4 binaries are built and timed
// by baseline/ (bench.py --impl reference).
#include "common.h"

using namespace bibfs;

namespace {
__global__ void k_naive_expand(u32 n, const u64* __restrict__ row_ptr, const u32* __restrict__ col,
                               const int* __restrict__ fin, int* __restrict__ fout, int* __restrict__ vis, int* changed) {
    const u32 u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= n || !fin[u]) return;
    for (u64 e = row_ptr[u]; e < row_ptr[u + 1]; ++e) {
        const u32 v = col[e];
        if (atomicExch(&vis[v], 1) == 0) { fout[v] = 1; *changed = 1; }
    }
}
__global__ void k_naive_intersect(u32 n, const int* __restrict__ a, const int* __restrict__ b, int* found) {
    const u32 u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u < n && a[u] && b[u]) *found = 1;
}
}  // namespace

extern "C" int bibfs_naive_levelsync_device(uint32_t n, const uint64_t* d_row_ptr, const uint32_t* d_col,
                                            uint32_t src, uint32_t dst, int32_t* length, float* ms, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (src >= n || dst >= n) return fail("src/dst out of range");
    *length = -1;
    if (src == dst) { *length = 0; if (ms) *ms = 0; return 0; }
    int *vis[2], *fr[2], *next, *flags;
    const size_t nb = (size_t)n * sizeof(int);
    BIBFS_CUDA_TRY(cudaMalloc(&vis[0], nb)); BIBFS_CUDA_TRY(cudaMalloc(&vis[1], nb));
    BIBFS_CUDA_TRY(cudaMalloc(&fr[0], nb));  BIBFS_CUDA_TRY(cudaMalloc(&fr[1], nb));
    BIBFS_CUDA_TRY(cudaMalloc(&next, nb));   BIBFS_CUDA_TRY(cudaMalloc(&flags, 2 * sizeof(int)));
    for (int s = 0; s < 2; ++s) { cudaMemsetAsync(vis[s], 0, nb, st); cudaMemsetAsync(fr[s], 0, nb, st); }
    const int one = 1;
    cudaMemcpyAsync(vis[0] + src, &one, 4, cudaMemcpyHostToDevice, st);
    cudaMemcpyAsync(fr[0] + src, &one, 4, cudaMemcpyHostToDevice, st);
    cudaMemcpyAsync(vis[1] + dst, &one, 4, cudaMemcpyHostToDevice, st);
    cudaMemcpyAsync(fr[1] + dst, &one, 4, cudaMemcpyHostToDevice, st);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0, st);
    const unsigned blocks = (n + 255) / 256;
    int depth[2] = {0, 0};
    int rc = 0;
    for (u32 it = 0; it < 2 * n; ++it) {
        const int s = it & 1;
        int h[2] = {0, 0};
        cudaMemsetAsync(flags, 0, 8, st);
        cudaMemsetAsync(next, 0, nb, st);
        k_naive_expand<<<blocks, 256, 0, st>>>(n, d_row_ptr, d_col, fr[s], next, vis[s], flags);
        count_launch();
        k_naive_intersect<<<blocks, 256, 0, st>>>(n, vis[0], vis[1], flags + 1);
        count_launch();
        cudaMemcpyAsync(fr[s], next, nb, cudaMemcpyDeviceToDevice, st);
        cudaMemcpyAsync(h, flags, 8, cudaMemcpyDeviceToHost, st);
        if (cudaStreamSynchronize(st) != cudaSuccess) { rc = fail("naive baseline: %s", cudaGetErrorString(cudaGetLastError())); break; }
        ++depth[s];
        if (h[1]) { *length = depth[0] + depth[1]; break; }
        if (!h[0]) break;   // this side is exhausted: no path
    }
    cudaEventRecord(e1, st);
    cudaEventSynchronize(e1);
    if (ms) cudaEventElapsedTime(ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(vis[0]); cudaFree(vis[1]); cudaFree(fr[0]); cudaFree(fr[1]); cudaFree(next); cudaFree(flags);
    return rc;
}
