/*
 * mpi.h — single-node stand-in for the handful of MPI calls the reference uses, so that its v2
 * (v2/second_try.cpp, v2/main-v2.cpp) and v4 (v4/mpi_bas.cpp) build and run UNMODIFIED on a box
 * that has no MPI installation.  Ranks are processes started by the `mpirun` script
 * next to this file; collectives go through one POSIX shared-memory segment:
 *     header { arrive counter, generation }  +  one staging slot per rank.
 * Implemented exactly: MPI_Init, MPI_Finalize, MPI_Comm_rank, MPI_Comm_size, MPI_Bcast,
 * MPI_Allreduce (BOR, LOR, SUM, MAX, MIN, BAND, LAND; MPI_IN_PLACE), MPI_Allgather,
 * MPI_Allgatherv, MPI_Barrier, MPI_Wtime, MPI_Abort.
 * A shared-memory allreduce is FASTER than a real MPI over a network, so timings obtained with
 * this shim flatter the reference.  Not a general MPI: one communicator, blocking collectives,
 * every rank must call the same collective in the same order.
 */
#ifndef BIBFS_MPI_SHIM_H
#define BIBFS_MPI_SHIM_H

#include <fcntl.h>
#include <sched.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef struct { int dummy; } MPI_Status;

#define MPI_COMM_WORLD 0
#define MPI_SUCCESS 0
#define MPI_IN_PLACE ((void*)(intptr_t)-1)

enum { MPI_CHAR = 1, MPI_UNSIGNED_CHAR, MPI_BYTE, MPI_INT, MPI_UNSIGNED, MPI_LONG, MPI_UNSIGNED_LONG,
       MPI_LONG_LONG, MPI_UNSIGNED_LONG_LONG, MPI_FLOAT, MPI_DOUBLE, MPI_C_BOOL, MPI_UINT32_T, MPI_UINT64_T,
       MPI_INT32_T, MPI_INT64_T };
#define MPI_LONG_LONG_INT MPI_LONG_LONG
enum { MPI_BOR = 1, MPI_LOR, MPI_SUM, MPI_MAX, MPI_MIN, MPI_BAND, MPI_LAND };

typedef struct {
    volatile uint32_t count;
    volatile uint32_t gen;
    uint64_t slot_bytes;
} mpishim_hdr;

static int mpishim_rank = 0, mpishim_size = 1;
static mpishim_hdr* mpishim_h = 0;
static char* mpishim_slots = 0;
static size_t mpishim_slot_bytes = 0;

static inline size_t mpishim_tsize(MPI_Datatype t) {
    switch (t) {
        case MPI_CHAR: case MPI_UNSIGNED_CHAR: case MPI_BYTE: case MPI_C_BOOL: return 1;
        case MPI_INT: case MPI_UNSIGNED: case MPI_FLOAT: case MPI_UINT32_T: case MPI_INT32_T: return 4;
        default: return 8;
    }
}

static inline void mpishim_barrier(void) {
    if (mpishim_size == 1) return;
    uint32_t g = __atomic_load_n(&mpishim_h->gen, __ATOMIC_ACQUIRE);
    if (__atomic_add_fetch(&mpishim_h->count, 1, __ATOMIC_ACQ_REL) == (uint32_t)mpishim_size) {
        __atomic_store_n(&mpishim_h->count, 0, __ATOMIC_RELAXED);
        __atomic_store_n(&mpishim_h->gen, g + 1, __ATOMIC_RELEASE);
    } else {
        unsigned spins = 0;
        while (__atomic_load_n(&mpishim_h->gen, __ATOMIC_ACQUIRE) == g)
            if (++spins > 2000) sched_yield();
    }
}

static inline int MPI_Init(int* argc, char*** argv) {
    (void)argc; (void)argv;
    const char* r = getenv("MPISHIM_RANK");
    const char* s = getenv("MPISHIM_SIZE");
    const char* n = getenv("MPISHIM_NAME");
    if (!r || !s || !n || atoi(s) <= 1) { mpishim_rank = 0; mpishim_size = 1; return MPI_SUCCESS; }
    mpishim_rank = atoi(r);
    mpishim_size = atoi(s);
    int fd = shm_open(n, O_RDWR, 0600);
    if (fd < 0) { perror("mpi shim: shm_open"); exit(2); }
    struct stat st;
    fstat(fd, &st);
    void* p = mmap(0, (size_t)st.st_size, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    if (p == MAP_FAILED) { perror("mpi shim: mmap"); exit(2); }
    close(fd);
    mpishim_h = (mpishim_hdr*)p;
    mpishim_slot_bytes = (size_t)mpishim_h->slot_bytes;
    mpishim_slots = (char*)p + 4096;
    mpishim_barrier();
    return MPI_SUCCESS;
}
static inline int MPI_Finalize(void) { mpishim_barrier(); return MPI_SUCCESS; }
static inline int MPI_Comm_rank(MPI_Comm c, int* r) { (void)c; *r = mpishim_rank; return MPI_SUCCESS; }
static inline int MPI_Comm_size(MPI_Comm c, int* s) { (void)c; *s = mpishim_size; return MPI_SUCCESS; }
static inline int MPI_Barrier(MPI_Comm c) { (void)c; mpishim_barrier(); return MPI_SUCCESS; }
static inline double MPI_Wtime(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}
static inline int MPI_Abort(MPI_Comm c, int code) { (void)c; fprintf(stderr, "MPI_Abort(%d) on rank %d\n", code, mpishim_rank); _exit(code ? code : 1); return 0; }

static inline char* mpishim_slot(int r) { return mpishim_slots + (size_t)r * mpishim_slot_bytes; }

static inline int MPI_Bcast(void* buf, int count, MPI_Datatype t, int root, MPI_Comm c) {
    (void)c;
    if (mpishim_size == 1) return MPI_SUCCESS;
    size_t bytes = (size_t)count * mpishim_tsize(t), off = 0;
    while (off < bytes || bytes == 0) {
        size_t k = bytes - off < mpishim_slot_bytes ? bytes - off : mpishim_slot_bytes;
        if (mpishim_rank == root) memcpy(mpishim_slot(root), (char*)buf + off, k);
        mpishim_barrier();
        if (mpishim_rank != root) memcpy((char*)buf + off, mpishim_slot(root), k);
        mpishim_barrier();
        off += k;
        if (bytes == 0) break;
    }
    return MPI_SUCCESS;
}

#define MPISHIM_REDUCE(T)                                                                      \
    do {                                                                                       \
        T* d = (T*)dst; const T* s = (const T*)src;                                            \
        switch (op) {                                                                          \
            case MPI_BOR:  for (size_t i = 0; i < n; ++i) d[i] = (T)(d[i] | s[i]); break;      \
            case MPI_BAND: for (size_t i = 0; i < n; ++i) d[i] = (T)(d[i] & s[i]); break;      \
            case MPI_LOR:  for (size_t i = 0; i < n; ++i) d[i] = (T)((d[i] || s[i]) ? 1 : 0); break; \
            case MPI_LAND: for (size_t i = 0; i < n; ++i) d[i] = (T)((d[i] && s[i]) ? 1 : 0); break; \
            case MPI_SUM:  for (size_t i = 0; i < n; ++i) d[i] = (T)(d[i] + s[i]); break;      \
            case MPI_MAX:  for (size_t i = 0; i < n; ++i) d[i] = d[i] > s[i] ? d[i] : s[i]; break; \
            case MPI_MIN:  for (size_t i = 0; i < n; ++i) d[i] = d[i] < s[i] ? d[i] : s[i]; break; \
        }                                                                                      \
    } while (0)

static inline void mpishim_combine(void* dst, const void* src, size_t n, MPI_Datatype t, MPI_Op op) {
    switch (t) {
        case MPI_CHAR: MPISHIM_REDUCE(signed char); break;
        case MPI_UNSIGNED_CHAR: case MPI_BYTE: case MPI_C_BOOL: MPISHIM_REDUCE(unsigned char); break;
        case MPI_INT: case MPI_INT32_T: MPISHIM_REDUCE(int32_t); break;
        case MPI_UNSIGNED: case MPI_UINT32_T: MPISHIM_REDUCE(uint32_t); break;
        case MPI_LONG: case MPI_LONG_LONG: case MPI_INT64_T: MPISHIM_REDUCE(int64_t); break;
        case MPI_UNSIGNED_LONG: case MPI_UNSIGNED_LONG_LONG: case MPI_UINT64_T: MPISHIM_REDUCE(uint64_t); break;
        case MPI_FLOAT: if (op == MPI_SUM || op == MPI_MAX || op == MPI_MIN) { float* d = (float*)dst; const float* s = (const float*)src;
                for (size_t i = 0; i < n; ++i) d[i] = op == MPI_SUM ? d[i] + s[i] : op == MPI_MAX ? (d[i] > s[i] ? d[i] : s[i]) : (d[i] < s[i] ? d[i] : s[i]); } break;
        case MPI_DOUBLE: if (op == MPI_SUM || op == MPI_MAX || op == MPI_MIN) { double* d = (double*)dst; const double* s = (const double*)src;
                for (size_t i = 0; i < n; ++i) d[i] = op == MPI_SUM ? d[i] + s[i] : op == MPI_MAX ? (d[i] > s[i] ? d[i] : s[i]) : (d[i] < s[i] ? d[i] : s[i]); } break;
    }
}

static inline int MPI_Allreduce(const void* send, void* recv, int count, MPI_Datatype t, MPI_Op op, MPI_Comm c) {
    (void)c;
    size_t ts = mpishim_tsize(t), total = (size_t)count;
    if (mpishim_size == 1) { if (send != MPI_IN_PLACE && send != recv) memcpy(recv, send, total * ts); return MPI_SUCCESS; }
    const char* in = (send == MPI_IN_PLACE) ? (const char*)recv : (const char*)send;
    size_t per = mpishim_slot_bytes / ts, off = 0;
    while (off < total) {
        size_t k = total - off < per ? total - off : per;
        memcpy(mpishim_slot(mpishim_rank), in + off * ts, k * ts);
        mpishim_barrier();
        memcpy((char*)recv + off * ts, mpishim_slot(0), k * ts);
        for (int r = 1; r < mpishim_size; ++r) mpishim_combine((char*)recv + off * ts, mpishim_slot(r), k, t, op);
        mpishim_barrier();
        off += k;
    }
    return MPI_SUCCESS;
}

static inline int MPI_Allgather(const void* send, int scount, MPI_Datatype st, void* recv, int rcount, MPI_Datatype rt, MPI_Comm c) {
    (void)c; (void)rcount; (void)rt;
    size_t b = (size_t)scount * mpishim_tsize(st);
    if (mpishim_size == 1) { if (send != recv) memcpy(recv, send, b); return MPI_SUCCESS; }
    if (b > mpishim_slot_bytes) { fprintf(stderr, "mpi shim: Allgather block too large\n"); exit(2); }
    memcpy(mpishim_slot(mpishim_rank), send, b);
    mpishim_barrier();
    for (int r = 0; r < mpishim_size; ++r) memcpy((char*)recv + (size_t)r * b, mpishim_slot(r), b);
    mpishim_barrier();
    return MPI_SUCCESS;
}

static inline int MPI_Allgatherv(const void* send, int scount, MPI_Datatype st, void* recv, const int* rcounts, const int* displs,
                                 MPI_Datatype rt, MPI_Comm c) {
    (void)c;
    size_t ts = mpishim_tsize(rt), b = (size_t)scount * mpishim_tsize(st);
    if (mpishim_size == 1) { memcpy((char*)recv + (size_t)displs[0] * ts, send, b); return MPI_SUCCESS; }
    if (b > mpishim_slot_bytes) { fprintf(stderr, "mpi shim: Allgatherv block too large\n"); exit(2); }
    memcpy(mpishim_slot(mpishim_rank), send, b);
    mpishim_barrier();
    for (int r = 0; r < mpishim_size; ++r) memcpy((char*)recv + (size_t)displs[r] * ts, mpishim_slot(r), (size_t)rcounts[r] * ts);
    mpishim_barrier();
    return MPI_SUCCESS;
}

#endif /* BIBFS_MPI_SHIM_H */
