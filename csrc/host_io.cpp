// host_io.cpp — .bin edge-list and legacy dense-matrix files (mmap based).
//
// Format written by the reference generator (graphs/generate_graph.py:35-39) and read by every
// reference variant (v1/main-v1.cpp:26-34, v3/bibfs_cuda_only.cu:72-87, v4/mpi_bas.cpp:29-42):
//   <u32 N> <u32 M> then M x <u32 u, u32 v>, little endian.  The reference readers validate
// nothing; here sizes and vertex ids are checked.
#include "common.h"
#include <cerrno>
#include <cstring>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

using namespace bibfs;

namespace {
struct Mapping {
    void* p = nullptr;
    size_t len = 0;
    int fd = -1;
    ~Mapping() {
        if (p && p != MAP_FAILED) munmap(p, len);
        if (fd >= 0) close(fd);
    }
    int open_ro(const char* path) {
        fd = ::open(path, O_RDONLY);
        if (fd < 0) return fail("cannot open %s: %s", path, strerror(errno));
        struct stat st;
        if (fstat(fd, &st) != 0) return fail("cannot stat %s: %s", path, strerror(errno));
        len = (size_t)st.st_size;
        if (len == 0) return fail("%s is empty", path);
        p = mmap(nullptr, len, PROT_READ, MAP_PRIVATE, fd, 0);
        if (p == MAP_FAILED) return fail("mmap %s: %s", path, strerror(errno));
        madvise(p, len, MADV_SEQUENTIAL);
        return 0;
    }
};
}  // namespace

extern "C" int bibfs_bin_header(const char* path, uint32_t* n, uint64_t* m) {
    Mapping mp;
    if (mp.open_ro(path)) return 1;
    if (mp.len < 8) return fail("%s: shorter than the 8-byte header", path);
    const u32* w = (const u32*)mp.p;
    u64 mm = w[1];
    if (mp.len < 8 + mm * 8) return fail("%s: header says M=%llu but file has room for %llu edges", path,
                                         (unsigned long long)mm, (unsigned long long)((mp.len - 8) / 8));
    *n = w[0];
    *m = mm;
    return 0;
}

extern "C" int bibfs_bin_read(const char* path, uint32_t* edges, uint64_t m) {
    Mapping mp;
    if (mp.open_ro(path)) return 1;
    if (mp.len < 8 + m * 8) return fail("%s: unexpected EOF reading edges", path);
    const u32* w = (const u32*)mp.p;
    const u32 n = w[0];
    const u32* e = w + 2;
    u32 bad = 0;
    for (u64 i = 0; i < 2 * m; ++i) {
        u32 x = e[i];
        bad |= (x >= n);
        edges[i] = x;
    }
    if (bad) return fail("%s: vertex id >= N=%u in the edge list", path, n);
    return 0;
}

extern "C" int bibfs_bin_write(const char* path, uint32_t n, uint64_t m, const uint32_t* edges) {
    if (m > 0xFFFFFFFFull) return fail("the .bin format stores M as uint32; M=%llu does not fit", (unsigned long long)m);
    FILE* f = fopen(path, "wb");
    if (!f) return fail("cannot create %s: %s", path, strerror(errno));
    u32 hdr[2] = {n, (u32)m};
    bool ok = fwrite(hdr, 4, 2, f) == 2;
    const size_t CH = 1u << 22;
    for (u64 off = 0; ok && off < 2 * m; off += CH) {
        size_t k = (size_t)((2 * m - off) < CH ? (2 * m - off) : CH);
        ok = fwrite(edges + off, 4, k, f) == k;
    }
    ok = (fclose(f) == 0) && ok;
    if (!ok) return fail("short write to %s", path);
    return 0;
}

extern "C" int bibfs_dense_header(const char* path, uint32_t* n) {
    Mapping mp;
    if (mp.open_ro(path)) return 1;
    if (mp.len < 4) return fail("%s: too short", path);
    u32 nn = *(const u32*)mp.p;
    if (mp.len != 4 + (size_t)nn * nn) return fail("%s: size %zu != 4 + N*N for N=%u (not a dense-matrix file)", path, mp.len, nn);
    *n = nn;
    return 0;
}

extern "C" int bibfs_dense_read(const char* path, uint8_t* out, uint32_t n) {
    Mapping mp;
    if (mp.open_ro(path)) return 1;
    if (mp.len != 4 + (size_t)n * n) return fail("%s: size mismatch for N=%u", path, n);
    memcpy(out, (const char*)mp.p + 4, (size_t)n * n);
    return 0;
}
