// Does a warp that polled a flag for T us run slower right afterwards?  CTA 0 polls (one lane, ld.relaxed.sys on local
// memory, like the cross-rank exchange), CTA 1 sets the flag T us after the start; then CTA 0 times a fixed chain of
// dependent instructions (shared-memory + ALU, no global memory).
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64; typedef unsigned u32;
__device__ __forceinline__ u64 gt() { u64 t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t) :: "memory"); return t; }
__device__ __forceinline__ u64 ld_rlx_sys(const u64* p) { u64 v; asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ void st_rlx_sys(u64* p, u64 v) { asm volatile("st.relaxed.sys.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory"); }
__global__ void k(u64* flag, u64* out, u64 T_ns, u64 seq, int others_at_barrier) {
    __shared__ u32 s[256];
    if (blockIdx.x == 1) {
        if (threadIdx.x == 0) { const u64 t0 = gt(); while (gt() - t0 < T_ns) {} st_rlx_sys(flag, seq); }
        return;
    }
    s[threadIdx.x & 255] = threadIdx.x;
    __syncthreads();
    if (threadIdx.x < 32) {
        u64 t0 = gt(), spins = 0;
        if (threadIdx.x == 0) { while (ld_rlx_sys(flag) != seq) ++spins; }
        __syncwarp();
        u64 t1 = gt();
        u32 x = threadIdx.x;
        if (threadIdx.x == 0) {
#pragma unroll 1
            for (int i = 0; i < 200; ++i) { x = s[x & 255] * 1664525u + 1013904223u; s[(x >> 8) & 255] = x; }
        }
        u64 t2 = gt();
        if (threadIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t1; out[2] = spins; out[3] = x; }
    }
    if (others_at_barrier) __syncthreads();
}
int main() {
    u64 *flag, *out; cudaMalloc(&flag, 256); cudaMemset(flag, 0, 256); cudaMalloc(&out, 64);
    u64 seq = 0;
    for (int oab = 0; oab < 2; ++oab)
        for (u64 T : {500, 1000, 2000, 3000, 3500, 4000, 4500, 5000, 6000, 8000, 12000, 20000}) {
            u64 best[3] = {0, ~0ull, 0}, worst = 0;
            for (int rep = 0; rep < 20; ++rep) {
                k<<<2, 512>>>(flag, out, T, ++seq, oab); cudaDeviceSynchronize();
                u64 h[4]; cudaMemcpy(h, out, 32, cudaMemcpyDeviceToHost);
                if (h[1] < best[1]) { best[0] = h[0]; best[1] = h[1]; best[2] = h[2]; }
                if (h[1] > worst) worst = h[1];
            }
            printf("others_at_barrier=%d T=%5llu ns: polled %5llu ns (%3llu spins), fixed chain min %5llu ns max %5llu ns\n", oab, T, best[0], best[2], best[1], worst);
        }
    return 0;
}
