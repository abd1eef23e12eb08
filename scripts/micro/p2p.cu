// NVLink latency microbenchmarks between GPU0 and GPU1 (one process, peer access)
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64; typedef unsigned u32;
__device__ __forceinline__ u64 gt() { u64 t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t) :: "memory"); return t; }
__device__ __forceinline__ u64 ld_acq_sys(const u64* p) { u64 v; asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ u64 ld_rlx_sys(const u64* p) { u64 v; asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ void st_rel_sys(u64* p, u64 v) { asm volatile("st.release.sys.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory"); }
__device__ __forceinline__ void st_rlx_sys(u64* p, u64 v) { asm volatile("st.relaxed.sys.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory"); }
__device__ __forceinline__ u64 atom_or_sys(u64* p, u64 v) { u64 o; asm volatile("atom.relaxed.sys.global.or.b64 %0, [%1], %2;" : "=l"(o) : "l"(p), "l"(v) : "memory"); return o; }

// GPU0: remote atomic round trip, remote store + fence cost
__global__ void k_single(u64* remote, u64* local, u64* out, int iters) {
    u64 a = 0, t0, t1;
    t0 = gt(); for (int i = 0; i < iters; ++i) a += atom_or_sys(remote + ((a + i) & 255) * 16, 1ull << (i & 63)); t1 = gt(); out[0] = (t1 - t0) / iters;
    t0 = gt(); for (int i = 0; i < iters; ++i) a += atom_or_sys(local + ((a + i) & 255) * 16, 1ull << (i & 63)); t1 = gt(); out[1] = (t1 - t0) / iters;
    t0 = gt(); for (int i = 0; i < iters; ++i) { st_rlx_sys(remote + 8192 + (i & 255), i); asm volatile("fence.acq_rel.sys;" ::: "memory"); } t1 = gt(); out[2] = (t1 - t0) / iters;
    t0 = gt(); for (int i = 0; i < iters; ++i) { asm volatile("fence.acq_rel.sys;" ::: "memory"); } t1 = gt(); out[3] = (t1 - t0) / iters;
    t0 = gt(); for (int i = 0; i < iters; ++i) { st_rel_sys(remote + 8192 + (i & 255), i); } t1 = gt(); out[4] = (t1 - t0) / iters;
    t0 = gt(); for (int i = 0; i < iters; ++i) { asm volatile("fence.acq_rel.gpu;" ::: "memory"); } t1 = gt(); out[5] = (t1 - t0) / iters;
    t0 = gt(); for (int i = 0; i < iters; ++i) a += ld_rlx_sys(remote + ((a + i) & 255) * 16); t1 = gt(); out[6] = (t1 - t0) / iters;
    out[7] = a;
}
// ping-pong: each side writes its flag into the PEER's memory and polls its own memory
__global__ void k_pingpong(u64* my_flag, u64* peer_flag, u64* out, int iters, int first, int mode) {
    u64 t0 = gt();
    for (int i = 1; i <= iters; ++i) {
        if (first) {
            if (mode == 0) st_rel_sys(peer_flag, i); else st_rlx_sys(peer_flag, i);
            if (mode == 0) { while (ld_acq_sys(my_flag) < (u64)i) {} } else { while (ld_rlx_sys(my_flag) < (u64)i) {} }
        } else {
            if (mode == 0) { while (ld_acq_sys(my_flag) < (u64)i) {} } else { while (ld_rlx_sys(my_flag) < (u64)i) {} }
            if (mode == 0) st_rel_sys(peer_flag, i); else st_rlx_sys(peer_flag, i);
        }
    }
    u64 t1 = gt();
    if (first) out[0] = (t1 - t0) / iters;   // round trip
}
int main() {
    int n = 0; cudaGetDeviceCount(&n);
    if (n < 2) { printf("need 2 GPUs\n"); return 0; }
    u64 *m0, *m1, *out0, *out1;
    cudaSetDevice(0); cudaDeviceEnablePeerAccess(1, 0); cudaMalloc(&m0, 1 << 20); cudaMemset(m0, 0, 1 << 20); cudaMalloc(&out0, 128);
    cudaSetDevice(1); cudaDeviceEnablePeerAccess(0, 0); cudaMalloc(&m1, 1 << 20); cudaMemset(m1, 0, 1 << 20); cudaMalloc(&out1, 128);
    cudaSetDevice(0);
    k_single<<<1, 1>>>(m1, m0, out0, 2000); cudaDeviceSynchronize();
    k_single<<<1, 1>>>(m1, m0, out0, 2000); cudaDeviceSynchronize();
    u64 h[8]; cudaMemcpy(h, out0, sizeof(h), cudaMemcpyDeviceToHost);
    printf("remote atom.or.sys round trip %llu ns | local atom.sys %llu ns | remote st + fence.sys %llu ns | bare fence.sys %llu ns | st.release.sys (remote) %llu ns | fence.gpu %llu ns | remote ld %llu ns\n",
           h[0], h[1], h[2], h[3], h[4], h[5], h[6]);
    for (int mode = 0; mode < 2; ++mode) {
        cudaSetDevice(0); cudaMemset(m0, 0, 64); cudaSetDevice(1); cudaMemset(m1, 0, 64); cudaDeviceSynchronize(); cudaSetDevice(0); cudaDeviceSynchronize();
        cudaSetDevice(1); k_pingpong<<<1, 1>>>(m1, m0, out1, 5000, 0, mode);
        cudaSetDevice(0); k_pingpong<<<1, 1>>>(m0, m1, out0, 5000, 1, mode);
        cudaSetDevice(0); cudaError_t e0 = cudaDeviceSynchronize(); cudaSetDevice(1); cudaError_t e1 = cudaDeviceSynchronize();
        cudaMemcpy(h, out0, 8, cudaMemcpyDeviceToHost);
        printf("ping-pong (%s): round trip %llu ns  (%s %s)\n", mode == 0 ? "st.release.sys / ld.acquire.sys" : "st.relaxed.sys / ld.relaxed.sys", h[0], cudaGetErrorString(e0), cudaGetErrorString(e1));
    }
    return 0;
}
