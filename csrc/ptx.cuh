// ptx.cuh — inline-PTX wrappers for sm_100a: scoped acquire/release accesses and system-scope
// atomics used by the device-side barriers and the peer-memory frontier exchange, bulk async
// copies (TMA 1-D, cp.async.bulk) with mbarrier completion, cluster / DSMEM helpers.
#pragma once
#include "common.h"

namespace bibfs {
namespace ptx {

__device__ __forceinline__ u64 globaltimer_ns() {
    u64 t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// ---- gpu scope (CTAs of one kernel on one device) ---------------------------------------
__device__ __forceinline__ u64 ld_acquire_gpu(const u64* p) {
    u64 v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ u32 ld_acquire_gpu(const u32* p) {
    u32 v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_gpu(u64* p, u64 v) {
    asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_gpu(u32* p, u32 v) {
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void red_release_gpu_add(u32* p, u32 v) {
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ u32 ld_relaxed_gpu(const u32* p) {
    u32 v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ u64 ld_relaxed_gpu(const u64* p) {
    u64 v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

// ---- system scope (peer GPUs over NVLink) -----------------------------------------------
__device__ __forceinline__ u64 ld_relaxed_sys(const u64* p) {
    u64 v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_sys(u64* p, u64 v) {
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void fence_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }
__device__ __forceinline__ void fence_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }

// visited test-and-set on the owner's bitmap word: returns the previous word so the caller learns
// "newly discovered?" and "already reached by the other side?" in a single NVLink round trip.
__device__ __forceinline__ u64 atom_or_sys(u64* p, u64 v) {
    u64 old;
    asm volatile("atom.relaxed.sys.global.or.b64 %0, [%1], %2;" : "=l"(old) : "l"(p), "l"(v) : "memory");
    return old;
}
__device__ __forceinline__ u64 atom_or_gpu(u64* p, u64 v) {
    u64 old;
    asm volatile("atom.relaxed.gpu.global.or.b64 %0, [%1], %2;" : "=l"(old) : "l"(p), "l"(v) : "memory");
    return old;
}
// Predicated forms for branch-free batches: the instruction is issued under a predicate (no branch, so the batch stays
// one straight line of independent memory operations) and idle slots cost no memory transaction at all.
// atom_or_gpu_if returns all ones when not executed ("already marked").
__device__ __forceinline__ u64 atom_or_gpu_if(bool on, u64* p, u64 v) {
    u64 old;
    asm volatile(
        "{\n\t.reg .pred q;\n\t"
        "setp.ne.u32 q, %3, 0;\n\t"
        "mov.u64 %0, 0xffffffffffffffff;\n\t"
        "@q atom.relaxed.gpu.global.or.b64 %0, [%1], %2;\n\t}"
        : "=&l"(old)
        : "l"(p), "l"(v), "r"((u32)on)
        : "memory");
    return old;
}
__device__ __forceinline__ uint4 ld_v4_gpu_if(bool on, const uint4* p) {
    uint4 r;
    asm volatile(
        "{\n\t.reg .pred q;\n\t"
        "setp.ne.u32 q, %5, 0;\n\t"
        "mov.u32 %0, 0;\n\tmov.u32 %1, 0;\n\tmov.u32 %2, 0;\n\tmov.u32 %3, 0;\n\t"
        "@q ld.relaxed.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];\n\t}"
        : "=&r"(r.x), "=&r"(r.y), "=&r"(r.z), "=&r"(r.w)
        : "l"(p), "r"((u32)on)
        : "memory");
    return r;
}
__device__ __forceinline__ uint4 ld_v4_sys(const uint4* p) {
    uint4 r;
    asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
    return r;
}
__device__ __forceinline__ u32 ld_u32_gpu_if(bool on, const u32* p) {
    u32 r;
    asm volatile(
        "{\n\t.reg .pred q;\n\t"
        "setp.ne.u32 q, %2, 0;\n\t"
        "mov.u32 %0, 0;\n\t"
        "@q ld.relaxed.gpu.global.u32 %0, [%1];\n\t}"
        : "=&r"(r)
        : "l"(p), "r"((u32)on)
        : "memory");
    return r;
}
__device__ __forceinline__ u32 ld_relaxed_sys(const u32* p) {
    u32 v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_or_sys(u32* p, u32 v) {
    asm volatile("red.relaxed.sys.global.or.b32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// fire-and-forget stores into peer memory (weak; ordered by the fence before the barrier flag)
__device__ __forceinline__ void st_peer(u32* p, u32 v) {
    asm volatile("st.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_peer_v2(u32* p, u32 a, u32 b) {
    asm volatile("st.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(a), "r"(b) : "memory");
}

// Scheduling fence for batches of independent atomics.  ptxas likes to consume the result of the first
// atomic of a batch before it has issued the last one (seen in SASS: ATOMG x2, SHF on the first result,
// ATOMG x6), which turns N overlapping 285 ns round trips into two back-to-back ones.  This returns a
// zero that ptxas cannot prove to be zero and that depends on `last`; adding it to the operands of the
// consumers pins them behind the whole batch.  (popc of a 64-bit word is at most 64.)
__device__ __forceinline__ u32 zero_after(u64 last) {
    u32 z;
    asm volatile("{\n\t.reg .u32 t;\n\tpopc.b64 t, %1;\n\tshr.u32 %0, t, 7;\n\t}" : "=r"(z) : "l"(last));
    return z;
}
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ void nanosleep(u32 ns) { asm volatile("nanosleep.u32 %0;" ::"r"(ns)); }

// ---- mbarrier + 1-D bulk async copy (TMA engine, SASS UBLKCP) -----------------------------
__device__ __forceinline__ u32 smem_addr(const void* p) { return (u32)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(u64* bar, u32 count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(u64* bar, u32 bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(u64* bar, u32 parity) {
    u32 ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_addr(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(u64* bar, u32 parity) {
    while (!mbar_try_wait(bar, parity)) {}
}
// global -> shared, bytes % 16 == 0, both addresses 16-byte aligned
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, u32 bytes, u64* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_addr(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_addr(bar))
                 : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// shared-memory atomics through the shared window (a generic-address atomic on a shared location
// takes the slow generic path)
__device__ __forceinline__ u32 atoms_add(u32* p, u32 v) {
    u32 old;
    asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(old) : "r"(smem_addr(p)), "r"(v) : "memory");
    return old;
}
__device__ __forceinline__ void reds_add(u32* p, u32 v) {
    asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(smem_addr(p)), "r"(v) : "memory");
}
__device__ __forceinline__ void reds_add64(unsigned long long* p, u64 v) {
    asm volatile("red.shared.add.u64 [%0], %1;" ::"r"(smem_addr(p)), "l"(v) : "memory");
}
__device__ __forceinline__ void reds_min(u32* p, u32 v) {
    asm volatile("red.shared.min.u32 [%0], %1;" ::"r"(smem_addr(p)), "r"(v) : "memory");
}

// ---- clusters / DSMEM ---------------------------------------------------------------------
__device__ __forceinline__ u32 cluster_ctarank() {
    u32 r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ u32 cluster_nctarank() {
    u32 r;
    asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
    return r;
}
// split form: the leader CTA of a query arrives at kernel start and collects the barrier only when it first needs
// its cluster (the other CTAs of a cluster start up to ~2 us after the first one)
__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ u32 dsmem_map(u32 smem_a, u32 cta) {
    u32 r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_a), "r"(cta));
    return r;
}
__device__ __forceinline__ u32 dsmem_ld(u32 a) {
    u32 v;
    asm volatile("ld.shared::cluster.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
    return v;
}
// remote shared memory of another CTA of the cluster (addresses from dsmem_map)
__device__ __forceinline__ u32 dsmem_atom_add(u32 a, u32 v) {
    u32 old;
    asm volatile("atom.relaxed.cluster.shared::cluster.add.u32 %0, [%1], %2;" : "=r"(old) : "r"(a), "r"(v) : "memory");
    return old;
}
__device__ __forceinline__ void dsmem_red_min(u32 a, u32 v) {
    asm volatile("red.relaxed.cluster.shared::cluster.min.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory");
}
__device__ __forceinline__ void dsmem_red_add64(u32 a, u64 v) {
    asm volatile("red.relaxed.cluster.shared::cluster.add.u64 [%0], %1;" ::"r"(a), "l"(v) : "memory");
}
__device__ __forceinline__ u32 dsmem_ld_acquire(u32 a) {
    u32 v;
    asm volatile("ld.acquire.cluster.shared::cluster.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ void dsmem_st_release(u32 a, u32 v) {
    asm volatile("st.release.cluster.shared::cluster.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory");
}
__device__ __forceinline__ void dsmem_st(u32 a, u32 v) { asm volatile("st.shared::cluster.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }

}  // namespace ptx
}  // namespace bibfs
