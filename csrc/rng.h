// rng.h — Philox-4x32-10 counter-based RNG and a bit-reproducible natural log.
//
// Everything here is __host__ __device__ and uses only IEEE-exact operations (integer ops,
// +, *, /, sqrt and explicit fma), so the CPU generator (tests, no GPU) and the sm_100a
// generator kernels produce the same graph bit for bit.  The reference generator
// (graphs/generate_graph.py:31) is networkx's fast_gnp_random_graph, unseeded.
#pragma once
#include "common.h"
#include <cmath>

namespace bibfs {

struct Philox {
    u32 c[4];
};

BIBFS_HD void mulhilo32(u32 a, u32 b, u32& hi, u32& lo) {
    u64 p = (u64)a * (u64)b;
    hi = (u32)(p >> 32);
    lo = (u32)p;
}

// Philox4x32-10 (Salmon et al., SC'11).  counter = (c0..c3), key = (k0,k1).
BIBFS_HD Philox philox4x32(u32 c0, u32 c1, u32 c2, u32 c3, u32 k0, u32 k1) {
    const u32 M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        u32 hi0, lo0, hi1, lo1;
        mulhilo32(M0, c0, hi0, lo0);
        mulhilo32(M1, c2, hi1, lo1);
        u32 n0 = hi1 ^ c1 ^ k0;
        u32 n1 = lo1;
        u32 n2 = hi0 ^ c3 ^ k1;
        u32 n3 = lo0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += W0; k1 += W1;
    }
    Philox out;
    out.c[0] = c0; out.c[1] = c1; out.c[2] = c2; out.c[3] = c3;
    return out;
}

BIBFS_HD u64 double_bits(double x) {
#if defined(__CUDA_ARCH__)
    return (u64)__double_as_longlong(x);
#else
    union { double d; u64 u; } v; v.d = x; return v.u;
#endif
}
BIBFS_HD double bits_double(u64 b) {
#if defined(__CUDA_ARCH__)
    return __longlong_as_double((long long)b);
#else
    union { double d; u64 u; } v; v.u = b; return v.d;
#endif
}

// ln(x) for normal positive x.  atanh series on the mantissa with explicit fma: every step is
// a correctly-rounded IEEE operation, so host and device agree exactly (|rel err| < 1e-15).
BIBFS_HD double det_log(double x) {
    u64 b = double_bits(x);
    int e = (int)((b >> 52) & 0x7ff) - 1023;
    double m = bits_double((b & 0x000fffffffffffffull) | 0x3ff0000000000000ull);  // [1,2)
    if (m > 1.4142135623730951) { m = m * 0.5; e += 1; }
    double s = (m - 1.0) / (m + 1.0);
    double z = s * s;
    double p = 1.0 / 23.0;
    p = fma(p, z, 1.0 / 21.0);
    p = fma(p, z, 1.0 / 19.0);
    p = fma(p, z, 1.0 / 17.0);
    p = fma(p, z, 1.0 / 15.0);
    p = fma(p, z, 1.0 / 13.0);
    p = fma(p, z, 1.0 / 11.0);
    p = fma(p, z, 1.0 / 9.0);
    p = fma(p, z, 1.0 / 7.0);
    p = fma(p, z, 1.0 / 5.0);
    p = fma(p, z, 1.0 / 3.0);
    p = fma(p, z, 1.0);
    double lm = 2.0 * s * p;
    return fma((double)e, 0.6931471805599453, lm);
}

// uniform in (0,1] with 53 bits from two 32-bit words
BIBFS_HD double u01_open_closed(u32 hi, u32 lo) {
    u64 r = (((u64)hi << 32) | lo) >> 11;      // 53 bits
    return (double)(r + 1) * (1.0 / 9007199254740992.0);
}

}  // namespace bibfs
