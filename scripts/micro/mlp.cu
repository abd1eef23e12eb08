// Do several atomics-with-return issued by one lane overlap?  conditional (branchy) vs unconditional issue.
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64; typedef unsigned u32;
__device__ u64* g_sink; __device__ u64 g_magic;
__device__ __forceinline__ u64 clk_dep(u64 dep) { if (dep == g_magic) *g_sink = dep; u64 t; asm volatile("mov.u64 %0, %%clock64;" : "=l"(t) :: "memory"); return t; }
__device__ __forceinline__ u64 atom_or(u64* p, u64 v) { u64 o; asm volatile("atom.relaxed.gpu.global.or.b64 %0, [%1], %2;" : "=l"(o) : "l"(p), "l"(v) : "memory"); return o; }
__device__ __forceinline__ u64 atom_or_sys(u64* p, u64 v) { u64 o; asm volatile("atom.relaxed.sys.global.or.b64 %0, [%1], %2;" : "=l"(o) : "l"(p), "l"(v) : "memory"); return o; }

template <int N, int MODE>   // MODE 0 unconditional, 1 conditional on a runtime mask (all true), 2 sys-scope unconditional (ptr may be remote)
__device__ u64 run(u64* bits, u32 seed, u32 mask) {
    u64 old[N]; u64 acc = 0;
    u64 t0 = clk_dep(seed);
#pragma unroll
    for (int k = 0; k < N; ++k) {
        u32 idx = ((seed + k * 7919u + threadIdx.x * 104729u) * 2654435761u) >> 15;   // 17-bit word index
        if (MODE == 1) { old[k] = 0; if ((mask >> k) & 1u) old[k] = atom_or(bits + idx, 1ull << (k & 63)); }
        else if (MODE == 2) old[k] = atom_or_sys(bits + idx, 1ull << (k & 63));
        else old[k] = atom_or(bits + idx, 1ull << (k & 63));
    }
#pragma unroll
    for (int k = 0; k < N; ++k) acc += old[k];
    u64 t1 = clk_dep(acc);
    return t1 - t0;
}
__global__ void k(u64* bits, u64* remote, u64* out, u32 mask) {
    u64 r[8];
    for (int rep = 0; rep < 3; ++rep) {
        r[0] = run<1, 0>(bits, 1 + rep, mask);
        r[1] = run<8, 0>(bits, 11 + rep, mask);
        r[2] = run<8, 1>(bits, 21 + rep, mask);
        r[3] = run<1, 2>(remote, 31 + rep, mask);
        r[4] = run<8, 2>(remote, 41 + rep, mask);
    }
    if (threadIdx.x == 0) for (int i = 0; i < 5; ++i) out[i] = r[i];
}
int main() {
    int n = 0; cudaGetDeviceCount(&n);
    u64 *bits, *remote, *out;
    cudaSetDevice(0); cudaMalloc(&bits, 1 << 20); cudaMemset(bits, 0, 1 << 20); cudaMalloc(&out, 128);
    remote = bits;
    { u64 magic = 0x123456789abcdefull; cudaMemcpyToSymbol(g_magic, &magic, 8); cudaMemcpyToSymbol(g_sink, &out, 8); }
    if (n > 1) { cudaDeviceEnablePeerAccess(1, 0); cudaSetDevice(1); cudaMalloc(&remote, 1 << 20); cudaMemset(remote, 0, 1 << 20); cudaDeviceSynchronize(); cudaSetDevice(0); }
    for (int lanes : {1, 32}) {
        k<<<1, lanes>>>(bits, remote, out, 0xFF); k<<<1, lanes>>>(bits, remote, out, 0xFF);
        cudaDeviceSynchronize();
        u64 h[5]; cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
        printf("lanes=%2d (cycles @1.965GHz): local x1=%llu  x8 unconditional=%llu  x8 conditional=%llu | %s x1=%llu x8=%llu\n", lanes, h[0], h[1], h[2], n > 1 ? "remote(sys)" : "local(sys)", h[3], h[4]);
    }
    return 0;
}
