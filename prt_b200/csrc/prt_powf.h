// powf exactly as glibc 2.39 evaluates it (sysdeps/ieee754/flt-32/e_powf.c — the ARM optimized-routines algorithm):
//   log2(x) = k + logc[i] + poly4(z*invc[i] - 1)   in fp64, 16-entry table indexed by the top mantissa bits,
//   x^y     = 2^(y*log2 x) = 2^(k/32 + r) = tab[k % 32] * 2^(k / 32) * poly3(r)   in fp64, rounded ONCE to fp32.
// The reference calls powf for the texture de-gamma (material.cpp:24-28, y = 2.2f) and the PPM gamma (image.cpp:16-19,
// y = 1/2.2f).  Rounding an accurate fp64 pow() to fp32 is not the same function: it differs from glibc in the last
// bit on a fraction of inputs, which is what kept the textured config and the PPM payload from being bit-identical.
// x86-64 glibc dispatches to its FMA build (e_powf-fma) on every CPU with FMA3, where gcc contracts each a*b+c of the
// source into one fma; the fma() calls below are those contractions.  tests/test_powf.py checks this restatement
// against the host libm's powf on 2e8 arguments (every fp32 in [0,1] for both exponents, plus random pairs).
// Tables: prt_powf_tables.h (recovered from libm by scripts/extract_powf_tables.py).
//
// Domain restated: any x; y finite, non-zero and NOT an integer (all the reference uses).  Negative x therefore gives
// NaN; zero / inf / NaN / subnormal x follow e_powf.c.
#pragma once
#include <stdint.h>
#include <string.h>
#include <math.h>
#include "prt_powf_tables.h"

#if defined(__CUDACC__)
#define PRT_HD __host__ __device__ __forceinline__
#else
#define PRT_HD static inline
#endif

namespace prt_powf {

static const double kLog2TabH[16][2] = PRT_POWF_LOG2_TAB;
static const uint64_t kExp2TabH[32] = PRT_EXP2F_TAB;
#if defined(__CUDACC__)
__device__ __constant__ const double kLog2TabD[16][2] = PRT_POWF_LOG2_TAB;
__device__ __constant__ const uint64_t kExp2TabD[32] = PRT_EXP2F_TAB;
#endif
#if defined(__CUDA_ARCH__)
#define PRT_POWF_LOG2(i, j) kLog2TabD[i][j]
#define PRT_POWF_EXP2(i) kExp2TabD[i]
#else
#define PRT_POWF_LOG2(i, j) kLog2TabH[i][j]
#define PRT_POWF_EXP2(i) kExp2TabH[i]
#endif

PRT_HD uint32_t asU32(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
PRT_HD float asF32(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
PRT_HD uint64_t asU64(double d) { uint64_t u; memcpy(&u, &d, 8); return u; }
PRT_HD double asF64(uint64_t u) { double d; memcpy(&d, &u, 8); return d; }

PRT_HD double log2Inline(uint32_t ix) {
    const double A[5] = PRT_POWF_LOG2_POLY;
    const uint32_t tmp = ix - 0x3f330000u;
    const int i = (int)((tmp >> (23 - 4)) % 16u);
    const uint32_t top = tmp & 0xff800000u;
    const uint32_t iz = ix - top;
    const int k = (int32_t)top >> 23;
    const double invc = PRT_POWF_LOG2(i, 0), logc = PRT_POWF_LOG2(i, 1);
    const double z = (double)asF32(iz);
    const double r = fma(z, invc, -1.0);
    const double y0 = logc + (double)k;
    const double r2 = r * r;
    double y = fma(A[0], r, A[1]);
    const double p = fma(A[2], r, A[3]);
    const double r4 = r2 * r2;
    double q = fma(A[4], r, y0);
    q = fma(p, r2, q);
    y = fma(y, r4, q);
    return y;
}

PRT_HD float exp2Inline(double xd) {
    const double C[3] = PRT_EXP2F_POLY;
    const double shift = 0x1.8p+52 / 32.0;
    double kd = xd + shift;
    const uint64_t ki = asU64(kd);
    kd -= shift;
    const double r = xd - kd;
    uint64_t t = PRT_POWF_EXP2(ki % 32u);
    t += ki << (52 - 5);
    const double s = asF64(t);
    const double z = fma(C[0], r, C[1]);
    const double r2 = r * r;
    double y = fma(C[2], r, 1.0);
    y = fma(z, r2, y);
    y = y * s;
    return (float)y;
}

PRT_HD float powfRef(float x, float y) {
    uint32_t ix = asU32(x);
    if (ix - 0x00800000u >= 0x7f800000u - 0x00800000u) {          // zero, subnormal, inf, NaN or negative
        if (2u * ix == 0u || 2u * ix >= 2u * 0x7f800000u) {       // zeroinfnan(ix)
            if (2u * ix > 2u * 0x7f800000u) return x + y;         // NaN
            const float x2 = x * x;                               // y is not an odd integer: the sign is dropped
            return (asU32(y) & 0x80000000u) ? 1.0f / x2 : x2;
        }
        if (ix & 0x80000000u) return asF32(0x7fc00000u);          // x < 0 with a non-integer y: invalid
        ix = asU32(x * 0x1p23f);                                  // subnormal: normalise
        ix &= 0x7fffffffu;
        ix -= 23u << 23;
    }
    const double logx = log2Inline(ix);
    const double ylogx = (double)y * logx;
    if (((asU64(ylogx) >> 47) & 0xffffu) >= (asU64(126.0) >> 47)) {          // |y*log2(x)| >= 126
        if (ylogx > 0x1.fffffffd1d571p+6) return asF32(0x7f800000u);         // overflow
        if (ylogx <= -150.0) return 0.0f;                                    // underflow
    }
    return exp2Inline(ylogx);
}

}  // namespace prt_powf
