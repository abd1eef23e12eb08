// common.h — shared types, error plumbing and launch helpers for libbibfs_b200.
#pragma once
#include <cstdint>
#include <cstddef>
#include <cstdio>
#include <cstdarg>
#include <string>

#include "../include/bibfs_b200.h"

#if defined(__CUDACC__)
#define BIBFS_HD __host__ __device__ __forceinline__
#else
#define BIBFS_HD inline
#endif

typedef uint32_t u32;
typedef uint64_t u64;
typedef int32_t i32;
typedef int64_t i64;

namespace bibfs {

// thread-local error string behind bibfs_last_error()
void set_error(const char* fmt, ...);
int fail(const char* fmt, ...);   // sets the error and returns 1

// process-wide kernel launch counter behind bibfs_launch_count()
void count_launch(u64 k = 1);

}  // namespace bibfs

#if defined(__CUDACC__) || defined(BIBFS_WITH_CUDA_RUNTIME)
#include <cuda_runtime.h>
#define BIBFS_CUDA_TRY(expr)                                                                   \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess)                                                                 \
            return bibfs::fail("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
    } while (0)
#define BIBFS_KERNEL_CHECK()                                                                   \
    do {                                                                                       \
        bibfs::count_launch();                                                                 \
        cudaError_t _e = cudaGetLastError();                                                   \
        if (_e != cudaSuccess)                                                                 \
            return bibfs::fail("%s:%d: kernel launch -> %s", __FILE__, __LINE__, cudaGetErrorString(_e)); \
    } while (0)
#endif
