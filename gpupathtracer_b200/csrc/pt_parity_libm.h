// pt_parity_libm.h -- sin / cos / log / exp / pow with ONE definition for every compiler ("parity libm").
//
// GLSL pins neither the precision nor the algorithm of its transcendental functions (the reference calls sin, cos in
// InitRay / the importance samplers, log, exp in the Beckmann model, pow in Present.frag), and CUDA's sincosf / logf / expf /
// powf differ from glibc's by 1-2 ulp on some arguments. Everywhere else the parity contract is "IEEE binary32, every
// operation rounded separately", which every compiler evaluates to the same bits. The functions below extend that to the
// transcendentals: the argument is widened to binary64, the function is evaluated with IEEE +, -, *, / and bit moves only
// (range reduction by exact products, Taylor / atanh series long enough for ~1e-16), and the binary64 result is rounded
// to binary32 once. Compiled without FMA contraction (nvcc -fmad=false, gcc -ffp-contract=off) the CUDA kernels, the CPU
// oracle and the reference's shader source compiled as C++ all get the SAME BITS from them, so frames can be compared for
// equality instead of "99.5 % of pixels". Results are faithfully rounded (<= 1 ulp of binary32, the correctly rounded
// value on all but ~1e-8 of arguments; tests/test_parity_libm.py measures it against glibc).
//
// Selected at run time by pt_config.parity_flags & PT_PARITY_LIBM (the default build keeps CUDA's fast libm); the oracle and
// the GLSL harness switch with or_set_parity_libm / model code | 4.
#pragma once
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define PTPL_HD __host__ __device__ __forceinline__
#else
#define PTPL_HD inline
#endif

namespace ptpl {

PTPL_HD double from_bits(uint64_t u) {
#if defined(__CUDA_ARCH__)
    return __longlong_as_double((long long)u);
#else
    double d; memcpy(&d, &u, 8); return d;
#endif
}
PTPL_HD uint64_t to_bits(double d) {
#if defined(__CUDA_ARCH__)
    return (uint64_t)__double_as_longlong(d);
#else
    uint64_t u; memcpy(&u, &d, 8); return u;
#endif
}
PTPL_HD double nan_d() { return from_bits(0x7ff8000000000000ull); }
PTPL_HD double inf_d() { return from_bits(0x7ff0000000000000ull); }
// nearest integer (ties to even) of |x| < 2^51 in the default rounding mode: two IEEE additions
PTPL_HD double rint_small(double x) { const double M = 6755399441055744.0; double t = x + M; return t - M; }

// sin and cos of |x| < 2^17 * pi/2 (NaN beyond: GLSL gives no precision guarantee there either, the path never gets near).
// k = rint(x * 2/pi); r = x - k * pi/2 with pi/2 cut into 33 + 33 + 53 bits, so the first two products are exact;
// Taylor series in r (|r| <= pi/4 + 1e-11): sin to r^17 (remainder < 9e-20), cos to r^16 (< 2e-18).
PTPL_HD void sincos_d(double x, double* s, double* c) {
    if (!(x > -205887.0 && x < 205887.0)) { *s = nan_d(); *c = nan_d(); return; }
    const double kd = rint_small(x * 0x1.45f306dc9c883p-1);
    const double r = ((x - kd * 0x1.921fb54400000p+0) - kd * 0x1.0b4611a600000p-34) - kd * 0x1.3198a2e037073p-69;
    const double z = r * r;
    double ps = 2.8114572543455206e-15;                       // 1/17!
    ps = ps * z - 7.647163731819816e-13;                      // 1/15!
    ps = ps * z + 1.6059043836821613e-10;                     // 1/13!
    ps = ps * z - 2.505210838544172e-08;                      // 1/11!
    ps = ps * z + 2.7557319223985893e-06;                     // 1/9!
    ps = ps * z - 0.0001984126984126984;                      // 1/7!
    ps = ps * z + 0.008333333333333333;                       // 1/5!
    ps = ps * z - 0.16666666666666666;                        // 1/3!
    const double sr = r + r * (z * ps);
    double pc = 4.779477332387385e-14;                        // 1/16!
    pc = pc * z - 1.1470745597729725e-11;                     // 1/14!
    pc = pc * z + 2.08767569878681e-09;                       // 1/12!
    pc = pc * z - 2.755731922398589e-07;                      // 1/10!
    pc = pc * z + 2.48015873015873e-05;                       // 1/8!
    pc = pc * z - 0.001388888888888889;                       // 1/6!
    pc = pc * z + 0.041666666666666664;                       // 1/4!
    pc = pc * z - 0.5;
    const double cr = 1.0 + z * pc;
    const int q = (int)kd & 3;
    *s = (q == 0) ? sr : (q == 1) ? cr : (q == 2) ? -sr : -cr;
    *c = (q == 0) ? cr : (q == 1) ? -sr : (q == 2) ? -cr : sr;
}

// log of a finite x > 0: x = 2^e * m with m in [sqrt(1/2), sqrt(2)), f = m - 1 (exact), s = f / (2 + f),
// log m = 2 s (1 + s^2/3 + s^4/5 + ... + s^20/21) (|s| < 0.1716: remainder < 2e-19 relative), ln 2 cut into 32 + 53 bits.
PTPL_HD double log_pos_d(double x) {
    uint64_t b = to_bits(x);
    int e = 0;
    if ((b >> 52) == 0) { x = x * 0x1p+54; b = to_bits(x); e = -54; }          // binary64 subnormal
    e += (int)(b >> 52) - 1023;
    double m = from_bits((b & 0x000fffffffffffffull) | 0x3ff0000000000000ull);
    if (m > 1.4142135623730951) { m = m * 0.5; e += 1; }
    const double f = m - 1.0;
    const double s = f / (2.0 + f);
    const double z = s * s;
    double p = 0.047619047619047616;                          // 1/21
    p = p * z + 0.05263157894736842;                          // 1/19
    p = p * z + 0.058823529411764705;                         // 1/17
    p = p * z + 0.06666666666666667;                          // 1/15
    p = p * z + 0.07692307692307693;                          // 1/13
    p = p * z + 0.09090909090909091;                          // 1/11
    p = p * z + 0.1111111111111111;                           // 1/9
    p = p * z + 0.14285714285714285;                          // 1/7
    p = p * z + 0.2;                                          // 1/5
    p = p * z + 0.3333333333333333;                           // 1/3
    const double lm = 2.0 * s + 2.0 * s * (z * p);
    const double ed = (double)e;
    return (ed * 0x1.a39ef35793c76p-33 + lm) + ed * 0x1.62e42fee00000p-1;
}

// exp of -745 < x < 709.7: k = rint(x / ln 2), r = x - k ln 2 (|r| <= 0.3466), Taylor to r^14 (remainder < 4e-18), times 2^k
PTPL_HD double exp_d(double x) {
    if (x != x) return x;
    if (x > 709.7) return inf_d();
    if (x < -745.0) return 0.0;
    const double kd = rint_small(x * 1.4426950408889634);
    const double r = (x - kd * 0x1.62e42fee00000p-1) - kd * 0x1.a39ef35793c76p-33;
    double p = 1.1470745597729725e-11;                        // 1/14!
    p = p * r + 1.6059043836821613e-10;
    p = p * r + 2.08767569878681e-09;
    p = p * r + 2.505210838544172e-08;
    p = p * r + 2.755731922398589e-07;
    p = p * r + 2.7557319223985893e-06;
    p = p * r + 2.48015873015873e-05;
    p = p * r + 0.0001984126984126984;
    p = p * r + 0.001388888888888889;
    p = p * r + 0.008333333333333333;
    p = p * r + 0.041666666666666664;
    p = p * r + 0.16666666666666666;
    p = p * r + 0.5;
    p = p * r + 1.0;
    p = p * r + 1.0;
    int k = (int)kd;
    if (k < -1000) { p = p * 0x1p-100; k += 100; }            // results below the binary64 normal range: scale in two steps
    return p * from_bits((uint64_t)(k + 1023) << 52);
}

PTPL_HD void sincosf_(float x, float* s, float* c) {
    double sd, cd;
    sincos_d((double)x, &sd, &cd);
    *s = (float)sd; *c = (float)cd;
}
PTPL_HD float sinf_(float x) { float s, c; sincosf_(x, &s, &c); return s; }
PTPL_HD float cosf_(float x) { float s, c; sincosf_(x, &s, &c); return c; }
PTPL_HD float logf_(float x) {
    if (x != x) return x;
    if (x < 0.0f) return (float)nan_d();
    if (x == 0.0f) return (float)-inf_d();
    if (x > 3.4028234663852886e+38f) return x;                 // +inf
    return (float)log_pos_d((double)x);
}
PTPL_HD float expf_(float x) { return (float)exp_d((double)x); }
// pow(x, y) for x >= 0 (GLSL leaves x < 0 undefined: NaN here); exp(y log x) in binary64 (error ~1e-14 relative at most)
PTPL_HD float powf_(float x, float y) {
    if (x != x || y != y) return (float)nan_d();
    if (y == 0.0f) return 1.0f;
    if (x < 0.0f) return (float)nan_d();
    if (x == 0.0f) return y > 0.0f ? 0.0f : (float)inf_d();
    if (x > 3.4028234663852886e+38f) return y > 0.0f ? x : 0.0f;
    return (float)exp_d((double)y * log_pos_d((double)x));
}

}  // namespace ptpl
