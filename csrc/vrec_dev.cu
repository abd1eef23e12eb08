// vrec_dev.cu — vertex records for the latency tier of the search kernel (pair_tier.cuh).
//
// record(v) = 64 bytes: sector 0 = { degree, row start, first six neighbours }, sector 1 = neighbours 7..14.
// One 32-byte sector answers
// "how long is v's row, where does it start, and who are its first neighbours", which a CSR
// answers with two dependent round trips (row_ptr, then col).  Rows longer than six entries keep
// the rest in col[].  No counterpart in the reference (its kernels read row_ptr / col_ind per
// level, v3/bibfs_cuda_only.cu:27-33); the structure exists to halve the dependent-load chain of
// a level.  Row starts are stored as 32 bits: shards with >= 2^32 stored entries get no records
// (the search then runs on the general tiers only).
#include "common.h"

namespace bibfs {

__global__ void __launch_bounds__(256) k_vrec_build(const u64* __restrict__ row_ptr, const u32* __restrict__ col, u32 rows, uint4* __restrict__ out,
                                                     u32* __restrict__ max_deg) {
    u32 mx = 0;
    for (u32 i = blockIdx.x * blockDim.x + threadIdx.x; i < rows; i += gridDim.x * blockDim.x) {
        const u64 r0 = row_ptr[i], r1 = row_ptr[i + 1];
        const u64 d64 = r1 - r0;
        const u32 deg = d64 > 0xFFFFFFFFull ? 0xFFFFFFFFu : (u32)d64;
        u32 nb[14];
#pragma unroll
        for (int k = 0; k < 14; ++k) nb[k] = (u32)k < deg ? col[r0 + k] : 0u;
        out[4 * (size_t)i] = make_uint4(deg, (u32)r0, nb[0], nb[1]);
        out[4 * (size_t)i + 1] = make_uint4(nb[2], nb[3], nb[4], nb[5]);
        // second sector: neighbours 7..14, fetched only for rows longer than six entries — so that such rows need no trip to
        // col[] (which, for rank 0 working alone on a partitioned graph, lives on another GPU)
        out[4 * (size_t)i + 2] = make_uint4(nb[6], nb[7], nb[8], nb[9]);
        out[4 * (size_t)i + 3] = make_uint4(nb[10], nb[11], nb[12], nb[13]);
        mx = deg > mx ? deg : mx;
    }
    if (max_deg) {
        mx = __reduce_max_sync(0xffffffffu, mx);
        if ((threadIdx.x & 31) == 0 && mx) atomicMax(max_deg, mx);
    }
}

}  // namespace bibfs

using namespace bibfs;

extern "C" size_t bibfs_vrec_bytes(uint32_t rows) { return (size_t)rows * 64; }

extern "C" int bibfs_vrec_build_device(const uint64_t* d_row_ptr, const uint32_t* d_col, uint32_t rows, uint64_t nnz,
                                       void* d_vrec, uint32_t* d_max_degree, void* stream) {
    if (!d_row_ptr || !d_vrec) return fail("NULL argument");
    if (nnz >= (1ull << 32)) return fail("vertex records need < 2^32 stored entries per shard (have %llu)", (unsigned long long)nnz);
    if (reinterpret_cast<uintptr_t>(d_vrec) & 31u) return fail("d_vrec must be 32-byte aligned");
    if (rows == 0) return 0;
    const u32 blocks = (rows + 255u) / 256u;
    k_vrec_build<<<blocks < 4096u ? blocks : 4096u, 256, 0, (cudaStream_t)stream>>>(d_row_ptr, d_col, rows, (uint4*)d_vrec, d_max_degree);
    BIBFS_KERNEL_CHECK();
    return 0;
}
