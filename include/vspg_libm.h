/* vspg_libm.h -- the elementary functions the render pass computes with, as portable IEEE-754 code.
 *
 * Why this is part of the boundary.  The reference integrator calls the host C library for log / sin / cos / asin /
 * acos / atan2 / pow (util/sampling.h:222-225 SampleExponential, util/sampling.cpp:163-345 spherical-rectangle
 * sampling, shapes.h:1053-1173 spherical-triangle sampling, openpgl ParallaxAwareVonMisesFisherMixture.h:1086-1129
 * sincosf, vspg.cpp:1026 pow, media_sampleTMaj.h:341-468 log).  Their last bit is not specified: glibc picks an FMA or
 * an SSE2 variant per CPU (IFUNC), other platforms ship other algorithms, CUDA's device functions are different again.
 * A sampled position that differs in its last bit re-seeds the shadow-ray tracker (RNG(Hash(ray.o), Hash(ray.d)),
 * vspg.cpp:1193), so "the same result as the reference" is only defined together with a libm.  This header is that
 * libm: every function is written with IEEE double / float +, -, *, /, sqrt, fma and integer bit operations only, so
 * the device build (nvcc, -fmad=false) and a host build (gcc, -ffp-contract=off) return identical bits.  The CUDA
 * kernels call these functions; oracle/portm_shim.cpp exports the same code under the C library's names so that the
 * reference itself can be run with it (LD_PRELOAD) -- that pairing is what the golden vectors in tests/golden/pbrt/
 * were produced with, and it is how a maintainer reproduces the product's images with the reference binary.
 *
 * Accuracy.  logf and sincosf (|x| <= 16) -- the two the kernels call per tentative collision / per sampled direction --
 * are float-only algorithms (short dependency chains on the GPU): logf is within 1 ulp; sin / cos are within 1 ulp
 * except next to their zeros, where the absolute error stays below 2e-13.  Every other float function evaluates in
 * double to < 1e-13 relative error and rounds once (the correctly rounded float except when the exact value lies
 * within ~1e-13 of a rounding boundary).  sin / cos / tan beyond 16 reduce in double with a three-part pi/2 and are
 * accurate for |x| < 2^20; beyond that they stay deterministic but lose accuracy.
 */
#ifndef VSPG_LIBM_H
#define VSPG_LIBM_H

#include <stdint.h>
#include <string.h>
#if defined(__CUDACC__)
#define VSPG_HD __host__ __device__ static inline
#else
#include <math.h>
#define VSPG_HD static inline
#endif

VSPG_HD double vspg_m_fma(double a, double b, double c) { return fma(a, b, c); }
VSPG_HD double vspg_m_sqrt(double a) { return sqrt(a); }
VSPG_HD uint64_t vspg_m_dbits(double x) { uint64_t u; memcpy(&u, &x, 8); return u; }
VSPG_HD double vspg_m_bitsd(uint64_t u) { double x; memcpy(&x, &u, 8); return x; }
VSPG_HD uint32_t vspg_m_fbits(float x) { uint32_t u; memcpy(&u, &x, 4); return u; }
VSPG_HD float vspg_m_bitsf(uint32_t u) { float x; memcpy(&x, &u, 4); return x; }
VSPG_HD int vspg_m_isnan(double x) { return x != x; }
VSPG_HD double vspg_m_inf(void) { return vspg_m_bitsd(0x7ff0000000000000ull); }
VSPG_HD double vspg_m_nan(void) { return vspg_m_bitsd(0x7ff8000000000000ull); }
/* 2^k for -1022 <= k <= 1023 */
VSPG_HD double vspg_m_pow2(int k) { return vspg_m_bitsd((uint64_t)(k + 1023) << 52); }
/* nearest integer (ties to even) of |x| < 2^51 without relying on the rounding-mode intrinsics */
VSPG_HD double vspg_m_rint(double x) {
    const double big = 6755399441055744.0; /* 1.5 * 2^52 */
    return (x + big) - big;
}

/* ------------------------------------------------------------------------------------------------ log */
/* log(x) for finite x > 0 (double, normal or subnormal).  x = 2^k m, m in [sqrt(1/2), sqrt(2)); with f = (m-1)/(m+1),
 * log m = 2 f (1 + f^2/3 + f^4/5 + ...); |f| <= 0.1716, the series is cut after f^22 (term < 3e-18). */
VSPG_HD double vspg_k_log(double x) {
    uint64_t ix = vspg_m_dbits(x);
    int k = 0;
    if (ix < 0x0010000000000000ull) { x *= 18014398509481984.0; ix = vspg_m_dbits(x); k = -54; } /* subnormal: * 2^54 */
    k += (int)(ix >> 52) - 1023;
    double m = vspg_m_bitsd((ix & 0x000fffffffffffffull) | 0x3ff0000000000000ull);
    if (m > 1.4142135623730951) { m *= 0.5; k += 1; }
    double f = (m - 1.0) / (m + 1.0), f2 = f * f;
    double p = 1.0 / 23.0;
    p = vspg_m_fma(p, f2, 1.0 / 21.0);
    p = vspg_m_fma(p, f2, 1.0 / 19.0);
    p = vspg_m_fma(p, f2, 1.0 / 17.0);
    p = vspg_m_fma(p, f2, 1.0 / 15.0);
    p = vspg_m_fma(p, f2, 1.0 / 13.0);
    p = vspg_m_fma(p, f2, 1.0 / 11.0);
    p = vspg_m_fma(p, f2, 1.0 / 9.0);
    p = vspg_m_fma(p, f2, 1.0 / 7.0);
    p = vspg_m_fma(p, f2, 1.0 / 5.0);
    p = vspg_m_fma(p, f2, 1.0 / 3.0);
    p = p * f2; /* log m = 2f + 2f p */
    const double ln2_hi = 6.93147180369123816490e-01, ln2_lo = 1.90821492927058770002e-10; /* hi has 32 significant bits */
    double dk = (double)k;
    double lo = vspg_m_fma(2.0 * f, p, dk * ln2_lo);
    return dk * ln2_hi + (2.0 * f + lo);
}
VSPG_HD double vspg_log(double x) {
    if (vspg_m_isnan(x)) return x;
    if (x < 0) return vspg_m_nan();
    if (x == 0) return -vspg_m_inf();
    if (x == vspg_m_inf()) return x;
    return vspg_k_log(x);
}
/* logf in float arithmetic only (the walk through a medium takes one per tentative collision): x = 2^e m with
 * m in [2/3, 4/3), f = m - 1, log m = f - f^2/2 + f^3 P(f) with a degree-7 fit of (log(1+f) - f + f^2/2) / f^3 on
 * [-1/3, 1/3] (relative error 1.6e-8 before rounding); every step is an IEEE float operation or an fmaf, so host and
 * device agree bit for bit.  Zero, subnormal, negative, inf and NaN arguments take the double path above. */
VSPG_HD float vspg_logf(float x) {
    uint32_t ix = vspg_m_fbits(x);
    if (ix - 0x00800000u >= 0x7f000000u) return (float)vspg_log((double)x);
    int32_t e = (int32_t)(ix - 0x3f2aaaabu) >> 23;
    float m = vspg_m_bitsf(ix - ((uint32_t)e << 23));
    float f = m - 1.0f, f2 = f * f;
    float p = -0.124309964f;
    p = fmaf(p, f, 0.134348407f);
    p = fmaf(p, f, -0.122870483f);
    p = fmaf(p, f, 0.141168296f);
    p = fmaf(p, f, -0.16673407f);
    p = fmaf(p, f, 0.200038403f);
    p = fmaf(p, f, -0.249999434f);
    p = fmaf(p, f, 0.333333194f);
    float r = fmaf(f2 * f, p, -0.5f * f2);
    float fe = (float)e;
    r = fmaf(fe, 1.42860677e-06f, r);          /* e * ln2_lo */
    r = r + f;
    return fmaf(fe, 0.693145752f, r);          /* e * ln2_hi (15 significant bits: the product is exact) */
}

/* ------------------------------------------------------------------------------------------------ exp */
/* exp(x), |x| <= 745.2: x = k ln2 + r, |r| <= 0.3466, Taylor through r^13 (remainder < 5e-18), scaled in two steps */
VSPG_HD double vspg_k_exp(double x) {
    const double inv_ln2 = 1.44269504088896338700e+00;
    const double ln2_hi = 6.93147180369123816490e-01, ln2_lo = 1.90821492927058770002e-10;
    double dk = vspg_m_rint(x * inv_ln2);
    double r = vspg_m_fma(-dk, ln2_hi, x);
    r = vspg_m_fma(-dk, ln2_lo, r);
    double p = 1.0 / 6227020800.0;
    p = vspg_m_fma(p, r, 1.0 / 479001600.0);
    p = vspg_m_fma(p, r, 1.0 / 39916800.0);
    p = vspg_m_fma(p, r, 1.0 / 3628800.0);
    p = vspg_m_fma(p, r, 1.0 / 362880.0);
    p = vspg_m_fma(p, r, 1.0 / 40320.0);
    p = vspg_m_fma(p, r, 1.0 / 5040.0);
    p = vspg_m_fma(p, r, 1.0 / 720.0);
    p = vspg_m_fma(p, r, 1.0 / 120.0);
    p = vspg_m_fma(p, r, 1.0 / 24.0);
    p = vspg_m_fma(p, r, 1.0 / 6.0);
    p = vspg_m_fma(p, r, 0.5);
    p = vspg_m_fma(p, r, 1.0);
    p = vspg_m_fma(p, r, 1.0);
    int k = (int)dk;
    int k1 = k / 2, k2 = k - k1; /* both within the normal exponent range for |k| <= 1080 */
    return p * vspg_m_pow2(k1) * vspg_m_pow2(k2);
}
VSPG_HD double vspg_exp_d(double x) {
    if (vspg_m_isnan(x)) return x;
    if (x > 709.8) return vspg_m_inf();
    if (x < -745.2) return 0.0;
    return vspg_k_exp(x);
}
VSPG_HD float vspg_expf(float x) { return (float)vspg_exp_d((double)x); }

/* ------------------------------------------------------------------------------------------------ pow */
VSPG_HD int vspg_m_is_int(double y) { return y == vspg_m_rint(y) || (y > 4503599627370496.0 || y < -4503599627370496.0); }
VSPG_HD int vspg_m_is_odd_int(double y) {
    if (y > 9007199254740992.0 || y < -9007199254740992.0) return 0;
    if (y != vspg_m_rint(y)) return 0;
    double h = y * 0.5;
    return h != vspg_m_rint(h);
}
/* C99 pow semantics for the special cases; the general case is exp(y log x) in double, rounded once to float */
VSPG_HD float vspg_powf(float xf, float yf) {
    double x = xf, y = yf;
    const double inf = vspg_m_inf();
    if (y == 0 || x == 1) return 1.f;
    if (vspg_m_isnan(x) || vspg_m_isnan(y)) return (float)vspg_m_nan();
    if (x == 0) {
        int odd = vspg_m_is_odd_int(y);
        int neg = (vspg_m_dbits(x) >> 63) != 0;
        if (y < 0) return (float)((odd && neg) ? -inf : inf);
        return (float)((odd && neg) ? -0.0 : 0.0);
    }
    if (y == inf || y == -inf) {
        double ax = x < 0 ? -x : x;
        if (ax == 1) return 1.f;
        return (float)(((ax > 1) == (y > 0)) ? inf : 0.0);
    }
    if (x == inf) return (float)(y > 0 ? inf : 0.0);
    if (x == -inf) {
        int odd = vspg_m_is_odd_int(y);
        if (y > 0) return (float)(odd ? -inf : inf);
        return (float)(odd ? -0.0 : 0.0);
    }
    double sign = 1.0;
    if (x < 0) {
        if (!vspg_m_is_int(y)) return (float)vspg_m_nan();
        if (vspg_m_is_odd_int(y)) sign = -1.0;
        x = -x;
    }
    double z = y * vspg_k_log(x);
    return (float)(sign * vspg_exp_d(z));
}

/* ------------------------------------------------------------------------------------------------ sin / cos / tan */
/* x = k pi/2 + r, |r| <= pi/4 (+ reduction error); pi/2 in three parts of 33 + 33 + 53 bits (exact products for |k| < 2^20) */
VSPG_HD double vspg_k_rem_pio2(double x, int *quadrant) {
    const double two_over_pi = 6.36619772367581382433e-01;
    const double p1 = 1.57079632673412561417e+00, p2 = 6.07710050630396597660e-11, p3 = 2.02226624871116645580e-21;
    double dk = vspg_m_rint(x * two_over_pi);
    double r = vspg_m_fma(-dk, p1, x);
    r = vspg_m_fma(-dk, p2, r);
    r = vspg_m_fma(-dk, p3, r);
    /* quadrant = k mod 4 taken on the double so that huge k stays deterministic */
    double q = dk - 4.0 * vspg_m_rint(dk * 0.25);
    int iq = (int)q;
    *quadrant = iq & 3;
    return r;
}
VSPG_HD double vspg_k_sin(double r) { /* |r| <= 0.79, Taylor through r^17 */
    double z = r * r;
    double p = 1.0 / 355687428096000.0;
    p = vspg_m_fma(p, z, -1.0 / 1307674368000.0);
    p = vspg_m_fma(p, z, 1.0 / 6227020800.0);
    p = vspg_m_fma(p, z, -1.0 / 39916800.0);
    p = vspg_m_fma(p, z, 1.0 / 362880.0);
    p = vspg_m_fma(p, z, -1.0 / 5040.0);
    p = vspg_m_fma(p, z, 1.0 / 120.0);
    p = vspg_m_fma(p, z, -1.0 / 6.0);
    return vspg_m_fma(r * z, p, r);
}
VSPG_HD double vspg_k_cos(double r) { /* |r| <= 0.79, Taylor through r^18 */
    double z = r * r;
    double p = -1.0 / 6402373705728000.0;
    p = vspg_m_fma(p, z, 1.0 / 20922789888000.0);
    p = vspg_m_fma(p, z, -1.0 / 87178291200.0);
    p = vspg_m_fma(p, z, 1.0 / 479001600.0);
    p = vspg_m_fma(p, z, -1.0 / 3628800.0);
    p = vspg_m_fma(p, z, 1.0 / 40320.0);
    p = vspg_m_fma(p, z, -1.0 / 720.0);
    p = vspg_m_fma(p, z, 1.0 / 24.0);
    p = vspg_m_fma(p, z, -0.5);
    return vspg_m_fma(p, z, 1.0);
}
VSPG_HD void vspg_sincos_d(double x, double *s, double *c) {
    if (vspg_m_isnan(x) || x == vspg_m_inf() || x == -vspg_m_inf()) { *s = *c = vspg_m_nan(); return; }
    int q;
    double r = vspg_k_rem_pio2(x, &q);
    double sr = vspg_k_sin(r), cr = vspg_k_cos(r);
    switch (q) {
    case 0: *s = sr; *c = cr; break;
    case 1: *s = cr; *c = -sr; break;
    case 2: *s = -sr; *c = -cr; break;
    default: *s = -cr; *c = sr; break;
    }
}
/* sincosf in float arithmetic only for |x| <= 16 (every phase / polar angle of the path is below 7): three-part
 * Cody-Waite reduction by pi/2 (the first two parts have 11 significant bits, so j * part is exact for |j| < 2^12),
 * then r + r^3 S(r^2) and 1 - r^2/2 + r^4 C(r^2) with degree-3 fits (relative error 3.5e-9 / 7.8e-10 before rounding).
 * Larger arguments take the double path. */
VSPG_HD void vspg_sincosf(float x, float *s, float *c) {
    if (x == 0) { *s = x; *c = 1.f; return; } /* keeps the sign of zero */
    float ax = x < 0 ? -x : x;
    if (!(ax <= 16.f)) {
        double ds, dc;
        vspg_sincos_d((double)x, &ds, &dc);
        *s = (float)ds; *c = (float)dc;
        return;
    }
    float j = (x * 0.636619747f + 12582912.f) - 12582912.f;   /* nearest integer of x * 2/pi (ties to even) */
    float r = fmaf(-j, 1.5703125f, x);
    r = fmaf(-j, 0.0004837512969970703f, r);
    r = fmaf(-j, 7.54979013e-08f, r);
    float z = r * r;
    float ps = 2.72342027e-06f;
    ps = fmaf(ps, z, -0.000198399619f);
    ps = fmaf(ps, z, 0.00833333191f);
    ps = fmaf(ps, z, -0.166666672f);
    float sr = fmaf(r * z, ps, r);
    float pc = -2.7287831e-07f;
    pc = fmaf(pc, z, 2.48004962e-05f);
    pc = fmaf(pc, z, -0.00138888869f);
    pc = fmaf(pc, z, 0.0416666679f);
    float cr = fmaf(z * z, pc, fmaf(-0.5f, z, 1.0f));
    int q = (int)j & 3;
    switch (q) {
    case 0: *s = sr; *c = cr; break;
    case 1: *s = cr; *c = -sr; break;
    case 2: *s = -sr; *c = -cr; break;
    default: *s = -cr; *c = sr; break;
    }
}
VSPG_HD float vspg_sinf(float x) { float s, c; vspg_sincosf(x, &s, &c); return s; }
VSPG_HD float vspg_cosf(float x) { float s, c; vspg_sincosf(x, &s, &c); return c; }
VSPG_HD float vspg_tanf(float x) {
    if (x == 0) return x;
    double ds, dc;
    vspg_sincos_d((double)x, &ds, &dc);
    return (float)(ds / dc);
}

/* ------------------------------------------------------------------------------------------------ atan / asin / acos */
/* atan(t), 0 <= t <= 1: t > tan(pi/8) is moved to (t-1)/(t+1) (+ pi/4), one half-angle step a = t/(1+sqrt(1+t^2)) brings
 * |a| <= tan(pi/16) = 0.1989, Taylor through a^25 (remainder < 1e-19).  Returns hi + lo split as value (lo folded in). */
VSPG_HD double vspg_k_atan01(double t) {
    double base = 0.0;
    if (t > 0.41421356237309503) { t = (t - 1.0) / (t + 1.0); base = 7.85398163397448278999e-01; }
    double a = t / (1.0 + vspg_m_sqrt(vspg_m_fma(t, t, 1.0)));
    double z = a * a;
    double p = 1.0 / 25.0;
    p = vspg_m_fma(p, z, -1.0 / 23.0);
    p = vspg_m_fma(p, z, 1.0 / 21.0);
    p = vspg_m_fma(p, z, -1.0 / 19.0);
    p = vspg_m_fma(p, z, 1.0 / 17.0);
    p = vspg_m_fma(p, z, -1.0 / 15.0);
    p = vspg_m_fma(p, z, 1.0 / 13.0);
    p = vspg_m_fma(p, z, -1.0 / 11.0);
    p = vspg_m_fma(p, z, 1.0 / 9.0);
    p = vspg_m_fma(p, z, -1.0 / 7.0);
    p = vspg_m_fma(p, z, 1.0 / 5.0);
    p = vspg_m_fma(p, z, -1.0 / 3.0);
    double at = vspg_m_fma(a * z, p, a);
    return base + 2.0 * at;
}
/* atan2 for finite y >= 0, x real, not both zero: result in [0, pi] */
VSPG_HD double vspg_k_atan2_pos(double y, double x) {
    const double pi = 3.14159265358979311600e+00, pio2 = 1.57079632679489655800e+00;
    double ax = x < 0 ? -x : x;
    double r;
    if (ax >= y) r = vspg_k_atan01(y / ax);          /* angle from the x axis, <= pi/4 */
    else r = pio2 - vspg_k_atan01(ax / y);           /* above the diagonal */
    return x < 0 ? pi - r : r;
}
VSPG_HD float vspg_atanf(float xf) {
    double x = xf;
    if (vspg_m_isnan(x)) return xf;
    if (x == 0) return xf;
    double ax = x < 0 ? -x : x;
    double r = (ax == vspg_m_inf()) ? 1.57079632679489655800e+00 : vspg_k_atan2_pos(ax, 1.0);
    return (float)(x < 0 ? -r : r);
}
VSPG_HD float vspg_atan2f(float yf, float xf) {
    double y = yf, x = xf;
    const double pi = 3.14159265358979311600e+00, pio2 = 1.57079632679489655800e+00, inf = vspg_m_inf();
    if (vspg_m_isnan(x) || vspg_m_isnan(y)) return (float)vspg_m_nan();
    int yneg = (vspg_m_dbits(y) >> 63) != 0, xneg = (vspg_m_dbits(x) >> 63) != 0;
    double r;
    if (y == 0) r = xneg ? pi : 0.0;
    else if (x == 0) r = pio2;
    else {
        double ay = yneg ? -y : y;
        if (ay == inf) r = (x == inf) ? pi / 4 : (x == -inf) ? 3 * pi / 4 : pio2;
        else if (x == inf) r = 0.0;
        else if (x == -inf) r = pi;
        else r = vspg_k_atan2_pos(ay, x);
    }
    return (float)(yneg ? -r : r);
}
VSPG_HD float vspg_asinf(float xf) {
    double x = xf;
    if (vspg_m_isnan(x)) return xf;
    double a = x < 0 ? -x : x;
    if (a > 1) return (float)vspg_m_nan();
    if (a == 0) return xf;
    double c = vspg_m_sqrt((1.0 - a) * (1.0 + a));   /* exact factors: a is a float */
    double r = vspg_k_atan2_pos(a, c);
    return (float)(x < 0 ? -r : r);
}
VSPG_HD float vspg_acosf(float xf) {
    double x = xf;
    if (vspg_m_isnan(x)) return xf;
    double a = x < 0 ? -x : x;
    if (a > 1) return (float)vspg_m_nan();
    double s = vspg_m_sqrt((1.0 - a) * (1.0 + a));
    return (float)vspg_k_atan2_pos(s, x);
}

/* ------------------------------------------------------------------------------------------------ hyperbolic */
VSPG_HD float vspg_atanhf(float xf) {
    double x = xf;
    if (vspg_m_isnan(x)) return xf;
    double a = x < 0 ? -x : x;
    if (a > 1) return (float)vspg_m_nan();
    if (a == 1) return (float)(x < 0 ? -vspg_m_inf() : vspg_m_inf());
    if (a < 1e-4) return (float)vspg_m_fma(x * x * x, 1.0 / 3.0, x);   /* x + x^3/3, next term < 1e-21 */
    double r = 0.5 * vspg_k_log((1.0 + a) / (1.0 - a));
    return (float)(x < 0 ? -r : r);
}
VSPG_HD float vspg_coshf(float xf) {
    double x = xf;
    if (vspg_m_isnan(x)) return xf;
    double a = x < 0 ? -x : x;
    if (a > 90) return (float)vspg_m_inf();
    double e = vspg_k_exp(a);
    return (float)(0.5 * (e + 1.0 / e));
}
VSPG_HD float vspg_sinhf(float xf) {
    double x = xf;
    if (vspg_m_isnan(x) || x == 0) return xf;
    double a = x < 0 ? -x : x, r;
    if (a > 90) r = vspg_m_inf();
    else if (a < 0.25) { /* x + x^3/3! + ... + x^13/13! (remainder < 2e-21) */
        double z = a * a;
        double p = 1.0 / 6227020800.0;
        p = vspg_m_fma(p, z, 1.0 / 39916800.0);
        p = vspg_m_fma(p, z, 1.0 / 362880.0);
        p = vspg_m_fma(p, z, 1.0 / 5040.0);
        p = vspg_m_fma(p, z, 1.0 / 120.0);
        p = vspg_m_fma(p, z, 1.0 / 6.0);
        r = vspg_m_fma(a * z, p, a);
    } else { double e = vspg_k_exp(a); r = 0.5 * (e - 1.0 / e); }
    return (float)(x < 0 ? -r : r);
}
VSPG_HD float vspg_tanhf(float xf) {
    double x = xf;
    if (vspg_m_isnan(x) || x == 0) return xf;
    double a = x < 0 ? -x : x, r;
    if (a > 20) r = 1.0;
    else if (a < 0.25) { /* sinh / cosh through the two series above, both accurate here */
        double z = a * a;
        double ps = 1.0 / 6227020800.0;
        ps = vspg_m_fma(ps, z, 1.0 / 39916800.0);
        ps = vspg_m_fma(ps, z, 1.0 / 362880.0);
        ps = vspg_m_fma(ps, z, 1.0 / 5040.0);
        ps = vspg_m_fma(ps, z, 1.0 / 120.0);
        ps = vspg_m_fma(ps, z, 1.0 / 6.0);
        double sh = vspg_m_fma(a * z, ps, a);
        double pc = 1.0 / 87178291200.0;
        pc = vspg_m_fma(pc, z, 1.0 / 479001600.0);
        pc = vspg_m_fma(pc, z, 1.0 / 3628800.0);
        pc = vspg_m_fma(pc, z, 1.0 / 40320.0);
        pc = vspg_m_fma(pc, z, 1.0 / 720.0);
        pc = vspg_m_fma(pc, z, 1.0 / 24.0);
        pc = vspg_m_fma(pc, z, 0.5);
        double ch = vspg_m_fma(pc, z, 1.0);
        r = sh / ch;
    } else { double e = vspg_k_exp(2.0 * a); r = (e - 1.0) / (e + 1.0); }
    return (float)(x < 0 ? -r : r);
}

#endif /* VSPG_LIBM_H */
