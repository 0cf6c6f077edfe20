// Transcendentals of the path (sin/cos in cosineHemisphereDir raytrace.comp:411-413, pow in
// schlick :180 and getEnvironmentLight :358-360, tan in main :547) evaluated in double
// precision with explicit polynomials and rounded once to float. In the reference these are
// SPIR-V GLSL.std.450 built-ins whose last-ulp behaviour is driver-defined and pinned by
// nothing; evaluating them this way makes the backend's results
// reproducible operation-for-operation on any IEEE-754 machine (no libm/libdevice call, no
// FMA: the kernels are compiled with -fmad=false), and within 1 ulp of the correctly rounded
// value. tests/test_portable_math.py sweeps them against libm.
#pragma once
#include "pt_math.h"

namespace pt {

PT_HD double pm_as_double(uint64_t u) {
#if defined(__CUDA_ARCH__)
    return __longlong_as_double((long long)u);
#else
    union { uint64_t u; double d; } c; c.u = u; return c.d;
#endif
}
PT_HD uint64_t pm_as_u64(double d) {
#if defined(__CUDA_ARCH__)
    return (uint64_t)__double_as_longlong(d);
#else
    union { uint64_t u; double d; } c; c.d = d; return c.u;
#endif
}

// sin and cos of r, |r| <= pi/4 (+ slack), Taylor polynomials in Horner form (error < 1e-16)
PT_HD double pm_sin_kernel(double r) {
    double z = r * r;
    double p = -7.6471637318198164759e-13;            // -1/15!
    p = p * z + 1.6059043836821614599e-10;            //  1/13!
    p = p * z - 2.5052108385441718775e-08;            // -1/11!
    p = p * z + 2.7557319223985890653e-06;            //  1/9!
    p = p * z - 1.9841269841269841270e-04;            // -1/7!
    p = p * z + 8.3333333333333333333e-03;            //  1/5!
    p = p * z - 1.6666666666666666667e-01;            // -1/3!
    return r + r * (z * p);
}
PT_HD double pm_cos_kernel(double r) {
    double z = r * r;
    double p = 4.7794773323873852974e-14;             //  1/16!
    p = p * z - 1.1470745597729724714e-11;            // -1/14!
    p = p * z + 2.0876756987868098979e-09;            //  1/12!
    p = p * z - 2.7557319223985890653e-07;            // -1/10!
    p = p * z + 2.4801587301587301587e-05;            //  1/8!
    p = p * z - 1.3888888888888888889e-03;            // -1/6!
    p = p * z + 4.1666666666666666667e-02;            //  1/4!
    return 1.0 - z * (0.5 - z * p);
}

// Quadrant reduction for |x| up to ~1e5 (the path only needs [0, 2*pi] and small angles).
PT_HD void pm_sincos_d(double x, double* s, double* c) {
    double kf = floor(x * 0.63661977236758134308 + 0.5);          // x * 2/pi
    double r = (x - kf * 1.5707963267341256142) - kf * 6.0771005065061922e-11;   // pi/2 = hi + lo (hi has 33 bits)
    long long k = (long long)kf;
    double sr = pm_sin_kernel(r), cr = pm_cos_kernel(r);
    switch ((int)(k & 3)) {
        case 0: *s = sr; *c = cr; break;
        case 1: *s = cr; *c = -sr; break;
        case 2: *s = -sr; *c = -cr; break;
        default: *s = -cr; *c = sr; break;
    }
}
PT_HD float pt_sinf(float x) { double s, c; pm_sincos_d((double)x, &s, &c); return (float)s; }
PT_HD float pt_cosf(float x) { double s, c; pm_sincos_d((double)x, &s, &c); return (float)c; }
PT_HD void pt_sincosf(float x, float* sn, float* cs) {
    double s, c; pm_sincos_d((double)x, &s, &c); *sn = (float)s; *cs = (float)c;
}
PT_HD float pt_tanf(float x) { double s, c; pm_sincos_d((double)x, &s, &c); return (float)(s / c); }

// ln(x) for finite x > 0 (double): x = m * 2^e, m in [sqrt(1/2), sqrt(2)); ln m = 2 atanh((m-1)/(m+1))
PT_HD double pm_log_d(double x) {
    uint64_t b = pm_as_u64(x);
    int e = (int)((b >> 52) & 0x7ff);
    if (e == 0) {                                   // subnormal double (cannot come from a float, kept for safety)
        x = x * 18014398509481984.0;                // 2^54
        b = pm_as_u64(x);
        e = (int)((b >> 52) & 0x7ff) - 54;
    }
    e -= 1023;
    double m = pm_as_double((b & 0x000fffffffffffffULL) | 0x3ff0000000000000ULL);   // [1, 2)
    if (m > 1.4142135623730951) { m = m * 0.5; e += 1; }
    double s = (m - 1.0) / (m + 1.0);
    double z = s * s;
    double p = 1.0 / 23.0;
    p = p * z + 1.0 / 21.0;
    p = p * z + 1.0 / 19.0;
    p = p * z + 1.0 / 17.0;
    p = p * z + 1.0 / 15.0;
    p = p * z + 1.0 / 13.0;
    p = p * z + 1.0 / 11.0;
    p = p * z + 1.0 / 9.0;
    p = p * z + 1.0 / 7.0;
    p = p * z + 1.0 / 5.0;
    p = p * z + 1.0 / 3.0;
    double lnm = 2.0 * (s + s * (z * p));
    return (double)e * 0.69314718055994530942 + lnm;
}

// exp(t) for |t| < 745 (double)
PT_HD double pm_exp_d(double t) {
    double nf = floor(t * 1.4426950408889634074 + 0.5);
    double r = (t - nf * 0.693147180369123816490) - nf * 1.90821492927058770002e-10;   // ln2 = hi + lo
    double p = 1.0 / 6227020800.0;                    // 1/13!
    p = p * r + 1.0 / 479001600.0;
    p = p * r + 1.0 / 39916800.0;
    p = p * r + 1.0 / 3628800.0;
    p = p * r + 1.0 / 362880.0;
    p = p * r + 1.0 / 40320.0;
    p = p * r + 1.0 / 5040.0;
    p = p * r + 1.0 / 720.0;
    p = p * r + 1.0 / 120.0;
    p = p * r + 1.0 / 24.0;
    p = p * r + 1.0 / 6.0;
    p = p * r + 0.5;
    double er = 1.0 + (r + r * r * p);
    int n = (int)nf;
    // scale by 2^n in two steps so that results in the float range (incl. float subnormals) stay exact
    int n1 = n / 2, n2 = n - n1;
    double s1 = pm_as_double((uint64_t)(n1 + 1023) << 52);
    double s2 = pm_as_double((uint64_t)(n2 + 1023) << 52);
    return er * s1 * s2;
}

// pow(x, y) with IEEE/C99 special cases (what GLSL leaves undefined for x < 0 follows C99 here).
PT_HD float pt_powf(float x, float y) {
    if (y == 0.f || x == 1.f) return 1.f;
    if (x != x || y != y) return x + y;
    float ax = fabsf(x);
    bool y_is_int = (floorf(y) == y) && !isinf1(y);
    bool y_is_odd = y_is_int && fabsf(y) < 16777216.f && (((long long)fabsf(y)) & 1);
    float sgn = (x < 0.f || (x == 0.f && as_uint(x) != 0u)) && y_is_odd ? -1.f : 1.f;
    if (ax == 0.f) return y > 0.f ? sgn * 0.f : sgn * INFINITY;
    if (isinf1(y)) {
        if (ax == 1.f) return 1.f;
        return ((ax > 1.f) == (y > 0.f)) ? INFINITY : 0.f;
    }
    if (isinf1(ax)) return y > 0.f ? sgn * INFINITY : sgn * 0.f;
    if (x < 0.f && !y_is_int) return as_float(0x7fc00000u);
    double t = (double)y * pm_log_d((double)ax);
    if (t > 100.0) return sgn * INFINITY;              // > ln(FLT_MAX) = 88.7
    if (t < -110.0) return sgn * 0.f;                  // < ln(min float subnormal) = -103.3
    return sgn * (float)pm_exp_d(t);
}

}  // namespace pt
