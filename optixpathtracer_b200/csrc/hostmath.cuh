// hostmath.cuh — sinf / cosf / expf that return, bit for bit, what the host C
// library returns for the reference's device code when it is compiled as host
// C++ (the oracle, oracle/ref_harness.cpp): PBRT/SamplingFunctions.h:16-34
// (cos/sin of the disk angle), PBRT/Microfacet.h (Sample_wm) and
// PBRT/LayeredBxDF.h:97-106 (Tr = exp(-|dz / w.z|)).
//
// CUDA's own sinf/cosf/expf are 1-2 ulp functions and disagree with the host
// library on a fifth of all arguments; each disagreement moves a sampled
// direction by an ulp, and on mirror-like lobes that is enough to send the path
// to another triangle.  The host library's algorithms are short (published as
// the "ARM optimized routines" single-precision sin/cos/exp, which glibc >= 2.28
// ships): a double-precision reduction by pi/2 and two minimax polynomials for
// sin/cos, a 32-entry 2^(i/32) table and a cubic for exp.  They are restated here
// from the published description and pinned empirically: 5e7 random arguments
// each against glibc 2.39, zero mismatches (tests/test_bsdf_cpu.py repeats a
// smaller sweep on whatever libm the test host has).
//
// Valid for the arguments the path produces: |x| < 120 for sin/cos (the path
// passes angles in [0, 2 pi]); any x for exp (overflow/underflow/NaN handled like
// the library, minus errno).  The shading translation units are built with
// -fmad=false, so the doubles below are rounded after every operation as in the
// library's non-FMA build.
//
// Provenance of the constants: bit parity with the library REQUIRES the library's own numbers, so the polynomial
// coefficients, the pi/2 reduction constants and the 32-entry exp table below are numerically identical to the data
// of the Arm Optimized Routines single-precision sinf / cosf / expf (math/sincosf_data.c `__sincosf_table`,
// math/exp2f_data.c `__exp2f_data`; Copyright (c) Arm Limited, SPDX-License-Identifier: MIT OR Apache-2.0 WITH
// LLVM-exception), the same data glibc >= 2.28 ships in sysdeps/ieee754/flt-32/ under the LGPL-2.1-or-later.  They
// are mathematical constants (minimax coefficients; T[i] = bits(2^(i/32)) - (i << 47)); the code around them is
// written here against the algorithm's description, not transcribed.
#pragma once

#include "vecmath.cuh"

namespace b2pt {
namespace hostmath {

B2PT_HD uint32_t f32_bits(float x) {
#if defined(__CUDA_ARCH__)
    return __float_as_uint(x);
#else
    union { float f; uint32_t u; } c;
    c.f = x;
    return c.u;
#endif
}
B2PT_HD uint64_t f64_bits(double x) {
#if defined(__CUDA_ARCH__)
    return (uint64_t)__double_as_longlong(x);
#else
    union { double f; uint64_t u; } c;
    c.f = x;
    return c.u;
#endif
}
B2PT_HD double bits_f64(uint64_t u) {
#if defined(__CUDA_ARCH__)
    return __longlong_as_double((long long)u);
#else
    union { double f; uint64_t u; } c;
    c.u = u;
    return c.f;
#endif
}
B2PT_HD uint32_t abstop12(float x) { return (f32_bits(x) >> 20) & 0x7ffu; }

// sin on [-pi/4, pi/4] (quadrant even) or +-cos (quadrant odd), x in double
B2PT_HD float sincos_poly(double x, double x2, bool negate_cos, int n) {
    if ((n & 1) == 0) {
        const double s1c = -0x1.555545995a603p-3, s2c = 0x1.1107605230bc4p-7, s3c = -0x1.994eb3774cf24p-13;
        const double x3 = x * x2;
        const double s1 = s2c + x2 * s3c;
        const double x5 = x3 * x2;
        const double s = x + x3 * s1c;
        return (float)(s + x5 * s1);
    }
    const double sg = negate_cos ? -1.0 : 1.0;
    const double c0 = sg * 0x1p0, c1c = sg * -0x1.ffffffd0c621cp-2, c2c = sg * 0x1.55553e1068f19p-5,
                 c3 = sg * -0x1.6c087e89a359dp-10, c4 = sg * 0x1.99343027bf8c3p-16;
    const double x4 = x2 * x2;
    const double c2 = c3 + x2 * c4;
    const double c1 = c0 + x2 * c1c;
    const double x6 = x4 * x2;
    const double c = c1 + x4 * c2c;
    return (float)(c + x6 * c2);
}

// quadrant n and remainder of x / (pi/2): 2/pi is pre-scaled by 2^24 so that the quadrant is
// bits 24..31 of the truncated product (rounds to nearest for either sign)
B2PT_HD double reduce_pio2(double x, int& n) {
    const double r = x * 0x1.45F306DC9C883p+23;
    n = ((int32_t)r + 0x800000) >> 24;
    return x - n * 0x1.921FB54442D18p0;
}

template <bool COS>
B2PT_HD float sincosf_host(float y) {
    double x = (double)y;
    if (abstop12(y) < abstop12(0x1.921FB6p-1f)) {   // |y| < pi/4
        if (abstop12(y) < abstop12(0x1p-12f)) return COS ? 1.0f : y;
        return sincos_poly(x, x * x, false, COS ? 1 : 0);
    }
    if (abstop12(y) < abstop12(120.0f)) {
        int n;
        x = reduce_pio2(x, n);
        const double s = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;
        return sincos_poly(x * s, x * x, (n & 2) != 0, COS ? (n ^ 1) : n);
    }
    return COS ? ::cosf(y) : ::sinf(y);   // outside the path's range: the platform's function
}

#if defined(B2PT_RELAXED_MATH) && defined(__CUDA_ARCH__)
// wavefront_relaxed.cu (the A/B build of the Philox kernels, B2PT_RELAXED=1): the hardware approximations, evaluated
// in [-pi, pi) where sin.approx / cos.approx are accurate (the path passes angles in [0, 2 pi])
B2PT_HD float sinf_host(float y) { return -__sinf(y - 3.14159265358979323846f); }
B2PT_HD float cosf_host(float y) { return -__cosf(y - 3.14159265358979323846f); }
B2PT_HD float expf_host(float x) { return __expf(x); }
B2PT_HD float expf_host_exact(float x) {
#else
B2PT_HD float sinf_host(float y) { return sincosf_host<false>(y); }
B2PT_HD float cosf_host(float y) { return sincosf_host<true>(y); }

B2PT_HD float expf_host(float x) {
#endif
    // 2^(i/32) with i subtracted from the exponent field (i << 47), so that adding k << 47 for
    // k = 32 e + i lands on 2^e * 2^(i/32)
    const uint64_t tab[32] = {
        0x3ff0000000000000ull, 0x3fefd9b0d3158574ull, 0x3fefb5586cf9890full, 0x3fef9301d0125b51ull,
        0x3fef72b83c7d517bull, 0x3fef54873168b9aaull, 0x3fef387a6e756238ull, 0x3fef1e9df51fdee1ull,
        0x3fef06fe0a31b715ull, 0x3feef1a7373aa9cbull, 0x3feedea64c123422ull, 0x3feece086061892dull,
        0x3feebfdad5362a27ull, 0x3feeb42b569d4f82ull, 0x3feeab07dd485429ull, 0x3feea47eb03a5585ull,
        0x3feea09e667f3bcdull, 0x3fee9f75e8ec5f74ull, 0x3feea11473eb0187ull, 0x3feea589994cce13ull,
        0x3feeace5422aa0dbull, 0x3feeb737b0cdc5e5ull, 0x3feec49182a3f090ull, 0x3feed503b23e255dull,
        0x3feee89f995ad3adull, 0x3feeff76f2fb5e47ull, 0x3fef199bdd85529cull, 0x3fef3720dcef9069ull,
        0x3fef5818dcfba487ull, 0x3fef7c97337b9b5full, 0x3fefa4afa2a490daull, 0x3fefd0765b6e4540ull};
    if (abstop12(x) >= abstop12(88.0f)) {
        if (f32_bits(x) == 0xff800000u) return 0.0f;
        if (abstop12(x) >= 0x7f8u) return x + x;           // +inf, NaN
        if (x > 0x1.62e42ep6f) return INFINITY;            // > log(2^128)
        if (x < -0x1.9fe368p6f) return 0.0f;               // < log(2^-150)
    }
    const double inv_ln2_n = 0x1.71547652b82fep+0 * 32.0, shift = 0x1.8p52;
    const double c0 = 0x1.c6af84b912394p-5 / 32.0 / 32.0 / 32.0, c1 = 0x1.ebfce50fac4f3p-3 / 32.0 / 32.0,
                 c2 = 0x1.62e42ff0c52d6p-1 / 32.0;
    double z = inv_ln2_n * (double)x;
    double kd = z + shift;             // round to nearest integer in the low mantissa bits
    const uint64_t ki = f64_bits(kd);
    kd -= shift;
    const double r = z - kd;
    const double s = bits_f64(tab[ki & 31u] + (ki << 47));
    z = c0 * r + c1;
    const double r2 = r * r;
    double y = c2 * r + 1.0;
    y = z * r2 + y;
    return (float)(y * s);
}

}  // namespace hostmath
}  // namespace b2pt
