// scan_noalloc.cu — the exclusive scan again, for callers that are not allowed to allocate.
//
// bibfs_scan_u32_device() (csr_dev.cu) needs one uint64 of scratch per 4096-element tile and asks
// the caller to bring it.  Some callers have nowhere to get it from: a custom operator running
// inside an inference server or another framework's stream pipeline is typically handed its
// inputs and outputs and nothing else — it must stay inside a memory budget fixed at start-up,
// and it must not call cudaMalloc/cudaFree because those can synchronise the whole device and
// stall every other stream of the host application.  For them the scratch lives in a fixed
// array in this library's device image, so the call itself allocates nothing.
//
// The price: the array exists on every GPU that loads this library, from load until process
// exit, whether or not the entry point is ever called (128 KB + 8 bytes per GPU); the input size
// is capped at BIBFS_SCAN_NOALLOC_MAX elements; and all calls on a device share the one array.
#include "scan.cuh"
#include <mutex>

using namespace bibfs;

namespace {

constexpr u64 kMaxCount = 1ull << 26;                                   // 67,108,864 elements
constexpr u64 kScratchWords = kMaxCount / SCAN_TILE + 1;                // 16,385 tile sums (+ the total)

// Scratch of the tile-sum pass.  In HOST code this name is only a handle for the symbol — it is
// not an address the GPU can use; the device address is looked up with cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned long long g_scan_scratch[kScratchWords];

constexpr int kMaxDevices = 64;
std::mutex g_setup_lock;
void* g_scratch_addr[kMaxDevices] = {};     // per device ordinal, resolved on first use

// Look up (once per device, with that device current) where the array lives on the current GPU.
int scratch_for_current_device(void** out) {
    int dev = -1;
    BIBFS_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return fail("device ordinal %d not supported by the no-alloc scan", dev);
    std::lock_guard<std::mutex> hold(g_setup_lock);
    if (!g_scratch_addr[dev]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_scan_scratch);
        if (e != cudaSuccess || !p)
            return fail("cudaGetSymbolAddress(scan scratch) on device %d: %s", dev, cudaGetErrorString(e));
        g_scratch_addr[dev] = p;
    }
    *out = g_scratch_addr[dev];
    return 0;
}

}  // namespace

extern "C" uint64_t bibfs_scan_noalloc_max_count(void) { return kMaxCount; }

// See include/bibfs_b200.h for the contract.
extern "C" int bibfs_scan_u32_device_noalloc(const uint32_t* d_in, uint64_t* d_out, uint64_t count, void* stream) {
    if (count > kMaxCount)
        return fail("bibfs_scan_u32_device_noalloc: count %llu exceeds the built-in scratch (max %llu elements); "
                    "use bibfs_scan_u32_device with caller-provided scratch",
                    (unsigned long long)count, (unsigned long long)kMaxCount);
    if (!d_out || (!d_in && count)) return fail("bibfs_scan_u32_device_noalloc: NULL input/output");
    void* scratch = nullptr;
    if (scratch_for_current_device(&scratch)) return 1;
    // Nothing in the scratch has to start at zero: every launch writes tile sums [0, tiles] before
    // it reads them, so whatever the previous call left behind is overwritten, never consumed.
    return scan_u32_to_u64(d_in, d_out, count, scratch, sizeof(unsigned long long) * kScratchWords, (cudaStream_t)stream);
}
