// latency microbenchmarks for the solo path (B200): dependent loads, atomics, barriers, pollers
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned long long u64; typedef unsigned u32;
__device__ __forceinline__ u64 gt() { u64 t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t) :: "memory"); return t; }
__device__ __forceinline__ u64 clk() { u64 t; asm volatile("mov.u64 %0, %%clock64;" : "=l"(t) :: "memory"); return t; }

__global__ void k(u32* chain, u64* bits, u64* out, volatile u32* flag, int pollers, int iters) {
    if (blockIdx.x > 0) {   // pollers
        if (threadIdx.x == 0 && pollers) { while (*flag == 0) { asm volatile("nanosleep.u32 32;"); } }
        return;
    }
    __shared__ u32 sa;
    if (threadIdx.x == 0) {
        sa = 0;
        // 1. dependent global loads (pointer chase, random, L2-resident after warm pass)
        u32 p = 0;
        for (int w = 0; w < 2; ++w) {
            u64 t0 = clk();
            for (int i = 0; i < iters; ++i) p = __ldcg(chain + p);
            u64 t1 = clk();
            out[w] = (t1 - t0) / iters;
        }
        out[7] = p;
        // 2. dependent atomics with return (address depends on previous result)
        u64 a = 1;
        u64 t0 = clk();
        for (int i = 0; i < iters; ++i) { u64 old = atomicOr(bits + (a & 1023) * 16, 1ull << (i & 63)); a = a * 1664525u + (u32)old + 1013904223u; }
        u64 t1 = clk();
        out[2] = (t1 - t0) / iters; out[8] = a;
        // 3. shared atomics with return
        t0 = clk();
        u32 s = 0;
        for (int i = 0; i < iters; ++i) s += atomicAdd(&sa, 1 + (s & 1));
        t1 = clk();
        out[3] = (t1 - t0) / iters; out[9] = s;
        // 4. globaltimer granularity
        u64 g0 = gt(), g1 = g0; while (g1 == g0) g1 = gt();
        out[5] = g1 - g0;
    }
    // 5. __syncthreads cost with the whole CTA
    __syncthreads();
    u64 t0 = clk();
    for (int i = 0; i < iters; ++i) __syncthreads();
    u64 t1 = clk();
    if (threadIdx.x == 0) { out[4] = (t1 - t0) / iters; *flag = 1; __threadfence(); }
}

int main() {
    const int N = 1 << 20, iters = 2000;
    u32* h = (u32*)malloc(N * 4);
    // random cycle
    u32* perm = (u32*)malloc(N * 4);
    for (int i = 0; i < N; ++i) perm[i] = i;
    srand(1);
    for (int i = N - 1; i > 0; --i) { int j = rand() % (i + 1); u32 t = perm[i]; perm[i] = perm[j]; perm[j] = t; }
    for (int i = 0; i < N; ++i) h[perm[i]] = perm[(i + 1) % N];
    u32* d; u64 *bits, *out; u32* flag;
    cudaMalloc(&d, N * 4); cudaMalloc(&bits, 1024 * 16 * 8); cudaMalloc(&out, 16 * 8); cudaMalloc(&flag, 4);
    cudaMemcpy(d, h, N * 4, cudaMemcpyHostToDevice); cudaMemset(bits, 0, 1024 * 16 * 8);
    for (int pollers = 0; pollers < 2; ++pollers)
        for (int threads : {32, 1024}) {
            cudaMemset(flag, 0, 4);
            k<<<148, threads>>>(d, bits, out, flag, pollers, iters);
            cudaError_t e = cudaDeviceSynchronize();
            u64 ho[16]; cudaMemcpy(ho, out, sizeof(ho), cudaMemcpyDeviceToHost);
            printf("pollers=%d threads=%d: ldcg chase cold=%llu warm=%llu cyc | atom.or.b64 ret=%llu cyc | smem atomicAdd ret=%llu cyc | syncthreads=%llu cyc | globaltimer tick=%llu ns (%s)\n",
                   pollers, threads, ho[0], ho[1], ho[2], ho[3], ho[4], ho[5], cudaGetErrorString(e));
        }
    return 0;
}
