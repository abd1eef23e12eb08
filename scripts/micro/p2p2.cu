// What makes the one-way delivery of a flag over NVLink slow?  Ping-pong variants (see main()).
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64; typedef unsigned u32;
__device__ __forceinline__ u64 gt() { u64 t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t) :: "memory"); return t; }
__device__ __forceinline__ u64 ld_rlx_sys(const u64* p) { u64 v; asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ u64 ld_acq_gpu(const u64* p) { u64 v; asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ void st_rlx_sys(u64* p, u64 v) { asm volatile("st.relaxed.sys.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory"); }
__device__ __forceinline__ void st_weak(u32* p, u32 v) { asm volatile("st.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ u64 atom_or_sys(u64* p, u64 v) { u64 o; asm volatile("atom.relaxed.sys.global.or.b64 %0, [%1], %2;" : "=l"(o) : "l"(p), "l"(v) : "memory"); return o; }

// mine/peer: 1 MB regions; flag words at offset 0 (4 words), scratch beyond 4096 B
__global__ void k_pp(u64* mine, u64* peer, u64* out, int iters, int first, int mode, int bg) {
    if (blockIdx.x != 0) {           // background CTAs: parked like the non-leader CTAs of the search kernel
        if (threadIdx.x == 0 && bg) { while (ld_acq_gpu(mine + 64) == 0) { asm volatile("nanosleep.u32 32;"); } }
        return;
    }
    const u32 lane = threadIdx.x;
    u64 t0 = gt(), acc = 0;
    for (int i = 1; i <= iters; ++i) {
        for (int half = 0; half < 2; ++half) {
            const bool send = (half == 0) == (first != 0);
            if (send) {
                if (lane == 0) {
                    if (mode & 1) acc += atom_or_sys(peer + 1024 + (i & 63) * 16, 1ull << (i & 63));
                    if (mode & 2) {
                        for (int k = 0; k < 8; ++k) st_rlx_sys(peer + 4096 + ((i * 8 + k) & 1023) * 16, i);
                        for (int k = 0; k < 8; ++k) st_weak(reinterpret_cast<u32*>(peer + 32768) + ((i * 8 + k) & 1023) * 32, i + (u32)acc);
                    }
                }
                if (mode & 8) __syncthreads();
                if (lane == 0) {
                    st_rlx_sys(peer + 0, i);
                    if (mode & 4) { st_rlx_sys(peer + 1, i); st_rlx_sys(peer + 2, i); st_rlx_sys(peer + 3, i); }
                }
            } else {
                if (lane == 0) {
                    if (mode & 4) { while (ld_rlx_sys(mine) < (u64)i || ld_rlx_sys(mine + 1) < (u64)i || ld_rlx_sys(mine + 2) < (u64)i || ld_rlx_sys(mine + 3) < (u64)i) {} }
                    else { while (ld_rlx_sys(mine) < (u64)i) {} }
                }
                if (mode & 8) __syncthreads();
            }
        }
    }
    u64 t1 = gt();
    if (lane == 0) { if (first) out[0] = (t1 - t0) / iters; out[1] = acc; st_rlx_sys(mine + 64, 1); }
}
int main() {
    int n = 0; cudaGetDeviceCount(&n);
    if (n < 2) { printf("need 2 GPUs\n"); return 0; }
    u64 *m0, *m1, *out0, *out1;
    cudaSetDevice(0); cudaDeviceEnablePeerAccess(1, 0); cudaMalloc(&m0, 1 << 20); cudaMalloc(&out0, 128);
    cudaSetDevice(1); cudaDeviceEnablePeerAccess(0, 0); cudaMalloc(&m1, 1 << 20); cudaMalloc(&out1, 128);
    for (int bg = 0; bg < 2; ++bg)
        for (int mode = 0; mode < 16; ++mode) {
            cudaSetDevice(0); cudaMemset(m0, 0, 1 << 20); cudaSetDevice(1); cudaMemset(m1, 0, 1 << 20); cudaDeviceSynchronize(); cudaSetDevice(0); cudaDeviceSynchronize();
            const int grid = bg ? 148 : 1, threads = (mode & 8) ? 512 : 32;
            cudaSetDevice(1); k_pp<<<grid, threads>>>(m1, m0, out1, 2000, 0, mode, bg);
            cudaSetDevice(0); k_pp<<<grid, threads>>>(m0, m1, out0, 2000, 1, mode, bg);
            cudaSetDevice(0); cudaError_t e0 = cudaDeviceSynchronize(); cudaSetDevice(1); cudaError_t e1 = cudaDeviceSynchronize();
            u64 h = 0; cudaMemcpy(&h, out0, 8, cudaMemcpyDeviceToHost);
            printf("bg=%d atom=%d stores=%d words4=%d cta512=%d : round trip %llu ns (%s %s)\n", bg, mode & 1, (mode >> 1) & 1, (mode >> 2) & 1, (mode >> 3) & 1, h,
                   cudaGetErrorString(e0), cudaGetErrorString(e1));
        }
    return 0;
}
