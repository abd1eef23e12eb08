// is the SM clock what nvidia-smi says while a single CTA runs?  ratio of %clock64 to %globaltimer
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64; typedef unsigned u32;
__device__ __forceinline__ u64 gt() { u64 t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t) :: "memory"); return t; }
__device__ __forceinline__ u64 clk() { u64 t; asm volatile("mov.u64 %0, %%clock64;" : "=l"(t) :: "memory"); return t; }
__global__ void k(u64* out, int spin_others) {
    if (blockIdx.x > 0) { if (spin_others) { u64 t0 = gt(); while (gt() - t0 < 300000) {} } return; }
    if (threadIdx.x == 0) {
        u64 g0 = gt(), c0 = clk();
        // dependent ALU chain: 20000 iterations x (1 IMAD + 1 LOP) 
        u32 x = (u32)g0;
        for (int i = 0; i < 20000; ++i) { x = x * 1664525u + 1013904223u; x ^= x >> 7; }
        u64 c1 = clk(), g1 = gt();
        out[0] = c1 - c0; out[1] = g1 - g0; out[2] = x;
    }
}
int main() {
    u64* out; cudaMalloc(&out, 64);
    for (int spin : {0, 1}) for (int grid : {1, 148}) {
        for (int r = 0; r < 3; ++r) k<<<grid, 1024>>>(out, spin);
        cudaDeviceSynchronize();
        u64 h[3]; cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
        printf("grid=%3d others_spin=%d: %llu cycles in %llu ns -> %.0f MHz; %.2f cycles per dependent-iteration (3 ops)\n", grid, spin, h[0], h[1], h[0] * 1e3 / h[1], h[0] / 20000.0);
    }
    return 0;
}
