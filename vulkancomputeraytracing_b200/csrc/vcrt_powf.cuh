/* @file vcrt_powf.cuh

    pow(x, 5.) of functions.glsl:61 with the bits of the reference build.
    SPDX-License-Identifier: WTFPL

    GLSL leaves the precision of pow() to the driver.  The parity anchor of this repository is the
    reference's shader code compiled as C++ (the reference build of the test tree), where pow() is C `powf` of the C library
    in the image: GNU libc 2.39, x86-64, the FMA variant its ifunc picks on every CPU with AVX2+FMA
    (sysdeps/ieee754/flt-32/e_powf.c, built with -mfma -mavx2 for sysdeps/x86_64/fpu/multiarch).
    That routine is NOT correctly rounded (it differs from the correctly rounded x^5 by one ulp on
    0.071 % of the inputs schlick() sees, measured on 2e7 of them), and three binary32 multiplies
    differ from it on 51 %.  To be bit-identical, powf5() below restates glibc's published algorithm
    for the one exponent the path uses:

        log2(x)  = k + log2(c_i) + log1p(z/c_i - 1)/ln 2      16-entry table of (1/c, log2 c), degree-5 polynomial
        x^5      = 2^(5 log2 x) = 2^(m/32) 2^r                32-entry table of 2^(i/32), degree-3 polynomial

    in binary64, every multiply-add pair fused exactly where gcc -mfma fuses it, and one final
    rounding to binary32.  The tables are glibc's published constants (__powf_log2_data,
    __exp2f_data; LGPL-2.1+, values are not copyrightable facts).  tests/test_powf.py compiles this
    header for the host and compares it with the libm of the machine at hand (0 mismatches on 8e7
    inputs here); the GPU test compares the device routine with `ref_schlick` of that reference build.

    Outside the main path of powf (x zero, subnormal, infinite or NaN, or |5 log2|x|| >= 126, i.e.
    |x| beyond 2^25.2 or below 2^-25.2) the result is x^5 from binary64 products rounded once: equal
    to powf wherever powf is exact about it (0, inf, NaN, sign), and in the remaining corner (results
    below 1e-38 or overflowing) the value has no influence on `rand < reflect_prob`.
*/
#pragma once
#include <stdint.h>

// nvcc: a device routine with its tables in the constant bank; any other compiler: the same code for the host (tests)
#if defined(__CUDACC__)
#define VCRT_POWF_FN __device__ __forceinline__
#define VCRT_POWF_TABLE __constant__
#else
#define VCRT_POWF_FN static inline
#define VCRT_POWF_TABLE static const
#endif

namespace vcrt {

#if defined(__CUDACC__)
VCRT_POWF_FN double pw_fma(double a, double b, double c) { return __fma_rn(a, b, c); }
VCRT_POWF_FN double pw_mul(double a, double b) { return __dmul_rn(a, b); }
VCRT_POWF_FN double pw_add(double a, double b) { return __dadd_rn(a, b); }
VCRT_POWF_FN uint32_t pw_bits(float f) { return __float_as_uint(f); }
VCRT_POWF_FN float pw_float(uint32_t u) { return __uint_as_float(u); }
VCRT_POWF_FN unsigned long long pw_bits64(double d) { return (unsigned long long)__double_as_longlong(d); }
VCRT_POWF_FN double pw_double(unsigned long long u) { return __longlong_as_double((long long)u); }
VCRT_POWF_FN float pw_round(double d) { return __double2float_rn(d); }
#else
}  // namespace vcrt
#include <cmath>
#include <cstring>
namespace vcrt {
VCRT_POWF_FN double pw_fma(double a, double b, double c) { return std::fma(a, b, c); }
VCRT_POWF_FN double pw_mul(double a, double b) { volatile double r = a * b; return r; }
VCRT_POWF_FN double pw_add(double a, double b) { volatile double r = a + b; return r; }
VCRT_POWF_FN uint32_t pw_bits(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
VCRT_POWF_FN float pw_float(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }
VCRT_POWF_FN unsigned long long pw_bits64(double d) { unsigned long long u; std::memcpy(&u, &d, 8); return u; }
VCRT_POWF_FN double pw_double(unsigned long long u) { double d; std::memcpy(&d, &u, 8); return d; }
VCRT_POWF_FN float pw_round(double d) { return (float)d; }
#endif

// __powf_log2_data.tab: (1/c, log2 c) for the 16 subintervals of [0x1.66p-1, 0x1.66p+0)
VCRT_POWF_TABLE double kPowfLog2Table[16][2] = {
    {0x1.661ec79f8f3bep+0, -0x1.efec65b963019p-2}, {0x1.571ed4aaf883dp+0, -0x1.b0b6832d4fca4p-2},
    {0x1.49539f0f010b0p+0, -0x1.7418b0a1fb77bp-2}, {0x1.3c995b0b80385p+0, -0x1.39de91a6dcf7bp-2},
    {0x1.30d190c8864a5p+0, -0x1.01d9bf3f2b631p-2}, {0x1.25e227b0b8ea0p+0, -0x1.97c1d1b3b7af0p-3},
    {0x1.1bb4a4a1a343fp+0, -0x1.2f9e393af3c9fp-3}, {0x1.12358f08ae5bap+0, -0x1.960cbbf788d5cp-4},
    {0x1.0953f419900a7p+0, -0x1.a6f9db6475fcep-5}, {0x1.0000000000000p+0, 0x0.0p+0},
    {0x1.e608cfd9a47acp-1, 0x1.338ca9f24f53dp-4},  {0x1.ca4b31f026aa0p-1, 0x1.476a9543891bap-3},
    {0x1.b2036576afce6p-1, 0x1.e840b4ac4e4d2p-3},  {0x1.9c2d163a1aa2dp-1, 0x1.40645f0c6651cp-2},
    {0x1.886e6037841edp-1, 0x1.88e9c2c1b9ff8p-2},  {0x1.767dcf5534862p-1, 0x1.ce0a44eb17bccp-2}};
// __exp2f_data.tab: bits of 2^(i/32) with i << 47 subtracted
VCRT_POWF_TABLE unsigned long long kPowfExp2Table[32] = {
    0x3ff0000000000000ull, 0x3fefd9b0d3158574ull, 0x3fefb5586cf9890full, 0x3fef9301d0125b51ull,
    0x3fef72b83c7d517bull, 0x3fef54873168b9aaull, 0x3fef387a6e756238ull, 0x3fef1e9df51fdee1ull,
    0x3fef06fe0a31b715ull, 0x3feef1a7373aa9cbull, 0x3feedea64c123422ull, 0x3feece086061892dull,
    0x3feebfdad5362a27ull, 0x3feeb42b569d4f82ull, 0x3feeab07dd485429ull, 0x3feea47eb03a5585ull,
    0x3feea09e667f3bcdull, 0x3fee9f75e8ec5f74ull, 0x3feea11473eb0187ull, 0x3feea589994cce13ull,
    0x3feeace5422aa0dbull, 0x3feeb737b0cdc5e5ull, 0x3feec49182a3f090ull, 0x3feed503b23e255dull,
    0x3feee89f995ad3adull, 0x3feeff76f2fb5e47ull, 0x3fef199bdd85529cull, 0x3fef3720dcef9069ull,
    0x3fef5818dcfba487ull, 0x3fef7c97337b9b5full, 0x3fefa4afa2a490daull, 0x3fefd0765b6e4540ull};

// powf(x, 5.0f) as glibc 2.39 (x86-64, FMA variant) returns it; see the file header for the domain.
VCRT_POWF_FN float powf5(float x) {
    const uint32_t bits = pw_bits(x), mag = bits & 0x7fffffffu;
    const double xd = (double)x;
    const double chain = pw_mul(pw_mul(pw_mul(xd, xd), pw_mul(xd, xd)), xd);   // exact special values, the fallback elsewhere
    if (mag - 0x00800000u >= 0x7f800000u - 0x00800000u) return pw_round(chain);   // zero, subnormal, inf, NaN
    // log2_inline(|x|): |x| = 2^k z, z in [0x1.66p-1, 0x1.66p+0)
    const uint32_t tmp = mag - 0x3f330000u;
    const int i = (int)((tmp >> 19) & 15u);
    const uint32_t top = tmp & 0xff800000u;
    const int k = (int32_t)top >> 23;
    const double z = (double)pw_float(mag - top);
    const double r = pw_fma(z, kPowfLog2Table[i][0], -1.0);
    const double y0 = pw_add(kPowfLog2Table[i][1], (double)k);
    const double r2 = pw_mul(r, r);
    double y = pw_fma(0x1.27616c9496e0bp-2, r, -0x1.71969a075c67ap-2);
    const double p = pw_fma(0x1.ec70a6ca7baddp-2, r, -0x1.7154748bef6c8p-1);
    const double r4 = pw_mul(r2, r2);
    double q = pw_fma(0x1.71547652ab82bp+0, r, y0);
    q = pw_fma(p, r2, q);
    y = pw_fma(y, r4, q);
    const double ylogx = pw_mul(5.0, y);
    if (!(ylogx < 126.0 && ylogx > -126.0)) return pw_round(chain);                // powf's overflow / underflow paths
    // exp2_inline(ylogx, sign): ylogx = m/32 + s, |s| <= 1/64
    const double shift = 0x1.8p+47;
    double kd = pw_add(ylogx, shift);
    const unsigned long long ki = pw_bits64(kd);
    kd = pw_add(kd, -shift);
    const double s = pw_add(ylogx, -kd);
    const unsigned long long signBias = (bits >> 31) ? 0x10000ull : 0ull;          // odd exponent: the sign of x
    const unsigned long long t = kPowfExp2Table[ki & 31ull] + ((ki + signBias) << 47);
    const double zz = pw_fma(0x1.c6af84b912394p-5, s, 0x1.ebfce50fac4f3p-3);
    const double s2 = pw_mul(s, s);
    double e = pw_fma(0x1.62e42ff0c52d6p-1, s, 1.0);
    e = pw_fma(zz, s2, e);
    return pw_round(pw_mul(e, pw_double(t)));
}

}  // namespace vcrt
