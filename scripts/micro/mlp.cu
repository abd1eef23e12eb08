// memory-level parallelism of atomics-with-return and loads from one thread / one warp (B200)
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64; typedef unsigned u32;
__device__ __forceinline__ u64 clk() { u64 t; asm volatile("mov.u64 %0, %%clock64;" : "=l"(t) :: "memory"); return t; }
__device__ __forceinline__ u64 atom_or(u64* p, u64 v) { u64 o; asm volatile("atom.relaxed.gpu.global.or.b64 %0, [%1], %2;" : "=l"(o) : "l"(p), "l"(v) : "memory"); return o; }

template <int N, bool COND>
__device__ u64 run_atoms(u64* bits, u32 seed, u32 lanes_active) {
    u64 old[N]; u64 acc = 0;
    u32 lane = threadIdx.x & 31;
    u64 t0 = clk();
#pragma unroll
    for (int k = 0; k < N; ++k) {
        u32 idx = ((seed + k * 7919u + lane * 104729u) * 2654435761u) >> 12;   // 20-bit random word index
        if (!COND || ((lane + k) & 1) == 0 || lanes_active == 1) old[k] = atom_or(bits + idx, 1ull << (k & 63)); else old[k] = 0;
    }
#pragma unroll
    for (int k = 0; k < N; ++k) acc += old[k];
    u64 t1 = clk() + (acc & 0);
    return (t1 - t0) | ((acc & 1) << 63 >> 63 << 62 >> 62 & 0);
}
template <int N>
__device__ u64 run_loads(const u32* data, u32 seed) {
    u32 v[N]; u32 acc = 0; u32 lane = threadIdx.x & 31;
    u64 t0 = clk();
#pragma unroll
    for (int k = 0; k < N; ++k) { u32 idx = ((seed + k * 7919u + lane * 104729u) * 2654435761u) >> 12; v[k] = __ldcg(data + idx); }
#pragma unroll
    for (int k = 0; k < N; ++k) acc += v[k];
    u64 t1 = clk() + (acc & 0);
    return t1 - t0;
}
__global__ void k(u64* bits, const u32* data, u64* out, int lanes) {
    if ((int)(threadIdx.x & 31) >= lanes) return;
    u64 r[10];
    for (int rep = 0; rep < 3; ++rep) {
        r[0] = run_atoms<1, false>(bits, 1 + rep, lanes);
        r[1] = run_atoms<2, false>(bits, 11 + rep, lanes);
        r[2] = run_atoms<4, false>(bits, 21 + rep, lanes);
        r[3] = run_atoms<8, false>(bits, 31 + rep, lanes);
        r[4] = run_atoms<8, true>(bits, 41 + rep, lanes);
        r[5] = run_loads<1>(data, 51 + rep);
        r[6] = run_loads<8>(data, 61 + rep);
    }
    if (threadIdx.x == 0) for (int i = 0; i < 7; ++i) out[i] = r[i];
}
int main() {
    u64* bits; u32* data; u64* out;
    cudaMalloc(&bits, 8u << 20); cudaMalloc(&data, 4u << 20); cudaMalloc(&out, 128);
    cudaMemset(bits, 0, 8u << 20); cudaMemset(data, 0, 4u << 20);
    for (int lanes : {1, 2, 8, 32}) {
        for (int pass = 0; pass < 2; ++pass) k<<<1, 32>>>(bits, data, out, lanes);
        cudaDeviceSynchronize();
        u64 h[7]; cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
        printf("lanes=%2d: atoms x1=%llu x2=%llu x4=%llu x8=%llu x8(divergent)=%llu | loads x1=%llu x8=%llu cycles\n", lanes, h[0], h[1], h[2], h[3], h[4], h[5], h[6]);
    }
    return 0;
}
