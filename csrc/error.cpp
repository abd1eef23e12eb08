// error.cpp — thread-local error string, version string, launch counter.
#include "common.h"
#include <atomic>

namespace bibfs {

static thread_local char g_err[1024] = "";
static std::atomic<u64> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int fail(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return 1;
}

void count_launch(u64 k) { g_launches.fetch_add(k, std::memory_order_relaxed); }

}  // namespace bibfs

extern "C" const char* bibfs_last_error(void) { return bibfs::g_err; }
extern "C" const char* bibfs_version(void) { return "bibfs_b200 0.1 (sm_100a)"; }
extern "C" uint64_t bibfs_launch_count(void) { return bibfs::g_launches.load(); }
