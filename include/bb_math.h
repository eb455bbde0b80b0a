/*
 * bb_math.h — deterministic float math for the basketball step path.
 *
 * The reference (src/game.cpp, src/helper.cpp, src/gen.cpp) calls libm's
 * sinf/cosf/atan2f/acosf/atan/erf/exp (call-site table: SURVEY.md App. E).
 * glibc and CUDA libdevice disagree in the last ulp on a fraction of inputs,
 * and those results feed discrete state (shot-going-in, spawn position ->
 * collisions -> resets).  This header is the arithmetic SPEC both sides use:
 * every function is evaluated in IEEE binary64 with only +,-,*,/,sqrt,fma and
 * rint (all correctly rounded on x86-64 and on sm_100a), in one fixed
 * operation order, then rounded once to binary32.  The same inputs therefore
 * give the same bits on the host (oracle/) and on the device (csrc/).
 * Accuracy versus the true function is < 1e-12 relative before the final
 * rounding, i.e. the float result is the correctly rounded one except on
 * ~1e-5 of inputs; tests/test_math.py measures the deviation from glibc.
 *
 * Build rules that keep it bit-stable:
 *   host  : -ffp-contract=off  (fma only where written), no -ffast-math
 *   device: -fmad=false, no --use_fast_math; sqrt/div via the _rn intrinsics
 *
 * Polynomial provenance: sin/cos/atan kernels use the classic fdlibm
 * (Sun Microsystems, freely redistributable) coefficients; exp is a Taylor
 * polynomial; erf is a Taylor series below 0.5 and, above, the form
 * erfc(x) = t*exp(-x^2 + P(s)) with P fitted by tools/gen_erfc_coeffs.py.
 */
#ifndef BB_MATH_H
#define BB_MATH_H

#include <stdint.h>
#if !defined(__CUDACC__)
#include <string.h>
#endif

#if defined(__CUDACC__)
#define BB_HD __host__ __device__ __forceinline__
/* the big polynomial kernels are shared subroutines on the device: the step kernel
 * is instruction-fetch bound, so one copy per kernel beats one copy per call site */
#define BB_HD_SUB __host__ __device__ __noinline__
#else
#define BB_HD static inline
#define BB_HD_SUB static inline
#endif

/* ---------------------------------------------------------------- primitives */

BB_HD double bb_fma(double a, double b, double c)
{
#if defined(__CUDA_ARCH__)
    return __fma_rn(a, b, c);
#else
    return __builtin_fma(a, b, c);
#endif
}

BB_HD double bb_sqrt(double x)
{
#if defined(__CUDA_ARCH__)
    return __dsqrt_rn(x);
#else
    return __builtin_sqrt(x);
#endif
}

BB_HD double bb_div(double a, double b)
{
#if defined(__CUDA_ARCH__)
    return __ddiv_rn(a, b);
#else
    return a / b;
#endif
}

BB_HD double bb_rint(double x)
{
#if defined(__CUDA_ARCH__)
    return rint(x);
#else
    return __builtin_rint(x);
#endif
}

BB_HD float bb_sqrtf(float x)
{
#if defined(__CUDA_ARCH__)
    return __fsqrt_rn(x);
#else
    return __builtin_sqrtf(x);
#endif
}

BB_HD float bb_divf(float a, float b)
{
#if defined(__CUDA_ARCH__)
    return __fdiv_rn(a, b);
#else
    return a / b;
#endif
}

BB_HD double bb_bits_to_double(int64_t bits)
{
#if defined(__CUDA_ARCH__)
    return __longlong_as_double(bits);
#else
    double d;
    memcpy(&d, &bits, sizeof d);
    return d;
#endif
}

/* ------------------------------------------------------------------ sin / cos */

#define BB_TWO_OVER_PI 0x1.45f306dc9c883p-1 /* 0.63661977236758138 */
#define BB_PIO2_HI 0x1.921fb54442d18p+0     /* 1.57079632679489656 */
#define BB_PIO2_LO 0x1.1a62633145c07p-54    /* 6.12323399573676604e-17 */
#define BB_PI_D 0x1.921fb54442d18p+1        /* 3.14159265358979312 */

/* sin on |r| <= ~pi/4 (fdlibm __kernel_sin coefficients). */
BB_HD double bb_ksin(double r)
{
    const double S1 = -1.66666666666666324348e-01, S2 = 8.33333333332248946124e-03,
                 S3 = -1.98412698298579493134e-04, S4 = 2.75573137070700676789e-06,
                 S5 = -2.50507602534068634195e-08, S6 = 1.58969099521155010221e-10;
    double z = r * r;
    double p = bb_fma(z, S6, S5);
    p = bb_fma(z, p, S4);
    p = bb_fma(z, p, S3);
    p = bb_fma(z, p, S2);
    p = bb_fma(z, p, S1);
    return bb_fma(z * r, p, r);
}

/* cos on |r| <= ~pi/4 (fdlibm __kernel_cos coefficients). */
BB_HD double bb_kcos(double r)
{
    const double C1 = 4.16666666666666019037e-02, C2 = -1.38888888888741095749e-03,
                 C3 = 2.48015872894767294178e-05, C4 = -2.75573143513906633035e-07,
                 C5 = 2.08757232129817482790e-09, C6 = -1.13596475577881948265e-11;
    double z = r * r;
    double p = bb_fma(z, C6, C5);
    p = bb_fma(z, p, C4);
    p = bb_fma(z, p, C3);
    p = bb_fma(z, p, C2);
    p = bb_fma(z, p, C1);
    p = bb_fma(z, p, -0.5);
    return bb_fma(z, p, 1.0);
}

/* Cody-Waite reduction, valid for |x| up to ~1e5 at float-output accuracy.
 * Call sites only pass |x| < 8. */
BB_HD double bb_reduce_pio2(double x, int *quadrant)
{
    double k = bb_rint(x * BB_TWO_OVER_PI);
    double r = bb_fma(-k, BB_PIO2_HI, x);
    r = bb_fma(-k, BB_PIO2_LO, r);
    *quadrant = (int)k;
    return r;
}

/* sin(x + qoff * pi/2) for qoff = 0 / 1: bb_sinf and bb_cosf share one reduction and
 * one pair of polynomials (cos x takes the sine rule with the quadrant advanced by 1). */
BB_HD_SUB double bb_sincos_quadrant(double x, int qoff)
{
    int q;
    double r = bb_reduce_pio2(x, &q);
    q += qoff;
    double s = bb_ksin(r), c = bb_kcos(r);
    double v = (q & 1) ? c : s;
    if (q & 2) v = -v;
    return v;
}

BB_HD float bb_sinf(float x)
{
    if (x == 0.0f) return x; /* keep sinf(-0) == -0 like libm */
    return (float)bb_sincos_quadrant((double)x, 0);
}

BB_HD float bb_cosf(float x) { return (float)bb_sincos_quadrant((double)x, 1); }

/* ----------------------------------------------------------------------- atan */

/* atan for x >= 0 (x may be +inf); fdlibm s_atan.c reduction and coefficients. */
BB_HD_SUB double bb_atan_pos(double x)
{
    const double aT0 = 3.33333333333329318027e-01, aT1 = -1.99999999998764832476e-01,
                 aT2 = 1.42857142725034663711e-01, aT3 = -1.11111104054623557880e-01,
                 aT4 = 9.09088713343650656196e-02, aT5 = -7.69187620504482999495e-02,
                 aT6 = 6.66107313738753120669e-02, aT7 = -5.83357013379057348645e-02,
                 aT8 = 4.97687799461593236017e-02, aT9 = -3.65315727442169155270e-02,
                 aT10 = 1.62858201153657823623e-02;
    double hi = 0.0, lo = 0.0;
    int reduced = 1;
    if (x < 0.4375) {
        reduced = 0;
    } else if (x < 1.1875) {
        if (x < 0.6875) {
            hi = 4.63647609000806093515e-01; lo = 2.26987774529616870924e-17;
            x = bb_div(bb_fma(2.0, x, -1.0), 2.0 + x);
        } else {
            hi = 7.85398163397448278999e-01; lo = 3.06161699786838301793e-17;
            x = bb_div(x - 1.0, x + 1.0);
        }
    } else {
        if (x < 2.4375) {
            hi = 9.82793723247329054082e-01; lo = 1.39033110312309984516e-17;
            x = bb_div(x - 1.5, bb_fma(1.5, x, 1.0));
        } else {
            hi = 1.57079632679489655800e+00; lo = 6.12323399573676603587e-17;
            x = bb_div(-1.0, x);
        }
    }
    double z = x * x;
    double w = z * z;
    double s1 = bb_fma(w, aT10, aT8);
    s1 = bb_fma(w, s1, aT6);
    s1 = bb_fma(w, s1, aT4);
    s1 = bb_fma(w, s1, aT2);
    s1 = bb_fma(w, s1, aT0);
    s1 = z * s1;
    double s2 = bb_fma(w, aT9, aT7);
    s2 = bb_fma(w, s2, aT5);
    s2 = bb_fma(w, s2, aT3);
    s2 = bb_fma(w, s2, aT1);
    s2 = w * s2;
    double xs = x * (s1 + s2);
    if (!reduced) return x - xs;
    return hi - ((xs - lo) - x);
}

/* std::atan(float) for x >= 0  (game.cpp:806 passes 0.1/dist >= 0). Negative
 * inputs are handled by odd symmetry. */
BB_HD float bb_atanf(float x)
{
    double xd = (double)x;
    int neg = xd < 0.0;
    double a = bb_atan_pos(neg ? -xd : xd);
    return (float)(neg ? -a : a);
}

/* atan2f(y, x), finite inputs.  Quadrant rules follow C99 Annex F. */
BB_HD float bb_atan2f(float y, float x)
{
    double xd = (double)x, yd = (double)y;
    int xneg = (x < 0.0f) || (x == 0.0f && (1.0f / x) < 0.0f);
    int yneg = (y < 0.0f) || (y == 0.0f && (1.0f / y) < 0.0f);
    double ax = xneg ? -xd : xd, ay = yneg ? -yd : yd;
    double a;
    if (ay == 0.0) a = 0.0;
    else a = bb_atan_pos(bb_div(ay, ax)); /* ax == 0 -> +inf -> pi/2 */
    if (xneg) a = (BB_PI_D - a) + 1.2246467991473532e-16;
    return (float)(yneg ? -a : a);
}

/* acosf(x); x is clamped to [-1,1] (every reference call site guarantees it:
 * helper.cpp:24-34 rejects |dot|>0.999999, game.cpp:746 clamps explicitly). */
BB_HD float bb_acosf(float x)
{
    double xd = (double)x;
    if (xd > 1.0) xd = 1.0;
    if (xd < -1.0) xd = -1.0;
    double s = bb_sqrt(bb_fma(-xd, xd, 1.0));
    int neg = xd < 0.0;
    double a;
    if (s == 0.0) a = 0.0;
    else a = bb_atan_pos(bb_div(s, neg ? -xd : xd));
    if (neg) a = (BB_PI_D - a) + 1.2246467991473532e-16;
    return (float)a;
}

/* ------------------------------------------------------------------------ exp */

BB_HD_SUB double bb_exp(double x)
{
    const double LOG2E = 1.44269504088896338700e+00;
    const double LN2_HI = 6.93147180369123816490e-01, LN2_LO = 1.90821492927058770002e-10;
    if (x < -700.0) return 0.0;
    if (x > 700.0) x = 700.0;
    double k = bb_rint(x * LOG2E);
    double r = bb_fma(-k, LN2_HI, x);
    r = bb_fma(-k, LN2_LO, r);
    /* Taylor to r^13; |r| <= 0.3466 -> truncation < 5e-18 */
    double p = 1.0 / 6227020800.0;
    p = bb_fma(r, p, 1.0 / 479001600.0);
    p = bb_fma(r, p, 1.0 / 39916800.0);
    p = bb_fma(r, p, 1.0 / 3628800.0);
    p = bb_fma(r, p, 1.0 / 362880.0);
    p = bb_fma(r, p, 1.0 / 40320.0);
    p = bb_fma(r, p, 1.0 / 5040.0);
    p = bb_fma(r, p, 1.0 / 720.0);
    p = bb_fma(r, p, 1.0 / 120.0);
    p = bb_fma(r, p, 1.0 / 24.0);
    p = bb_fma(r, p, 1.0 / 6.0);
    p = bb_fma(r, p, 0.5);
    p = bb_fma(r, p, 1.0);
    p = bb_fma(r, p, 1.0);
    int64_t e = (int64_t)((int)k + 1023);
    return p * bb_bits_to_double(e << 52);
}

BB_HD float bb_expf(float x) { return (float)bb_exp((double)x); }

/* ------------------------------------------------------------------------ erf */

BB_HD float bb_erff(float x)
{
    double xd = (double)x;
    int neg = xd < 0.0;
    double ax = neg ? -xd : xd;
    double r;
    if (!(ax == ax)) return x; /* NaN passthrough */
    if (ax < 0.5) {
        /* erf(x) = 2/sqrt(pi) * x * sum (-1)^n t^n / (n! (2n+1)),  t = x^2 */
        double t = ax * ax;
        double p = 1.0 / 76204800.0;
        p = bb_fma(t, p, -1.0 / 6894720.0);
        p = bb_fma(t, p, 1.0 / 685440.0);
        p = bb_fma(t, p, -1.0 / 75600.0);
        p = bb_fma(t, p, 1.0 / 9360.0);
        p = bb_fma(t, p, -1.0 / 1320.0);
        p = bb_fma(t, p, 1.0 / 216.0);
        p = bb_fma(t, p, -1.0 / 42.0);
        p = bb_fma(t, p, 1.0 / 10.0);
        p = bb_fma(t, p, -1.0 / 3.0);
        p = bb_fma(t, p, 1.0);
        r = (1.12837916709551255856e+00 * ax) * p;
    } else if (ax < 6.5) {
        /* erfc(x) = t exp(-x^2 + P(s)), t = 2/(2+x), s affine in t on [-1,1];
         * coefficients from tools/gen_erfc_coeffs.py (degree 12). */
        const double T_LO = 2.0 / 8.5, T_HI = 0.8;
        const double A = 2.0 / (T_HI - T_LO), B = -(T_HI + T_LO) / (T_HI - T_LO);
        double t = bb_div(2.0, 2.0 + ax);
        double s = bb_fma(t, A, B);
        double p = 0x1.00558f477655cp-24;
        p = bb_fma(s, p, -0x1.e770e46628e74p-23);
        p = bb_fma(s, p, -0x1.01755d548aff3p-21);
        p = bb_fma(s, p, 0x1.e70c43d5867c7p-19);
        p = bb_fma(s, p, 0x1.4a21578e1877ap-21);
        p = bb_fma(s, p, -0x1.69eabca5ed14ep-15);
        p = bb_fma(s, p, 0x1.4028324192e38p-15);
        p = bb_fma(s, p, 0x1.130e8966e8bedp-11);
        p = bb_fma(s, p, -0x1.b9c8a84076e8dp-11);
        p = bb_fma(s, p, -0x1.1c4afd84bfaa5p-7);
        p = bb_fma(s, p, 0x1.ba18bf64c271cp-7);
        p = bb_fma(s, p, 0x1.86cac39badd50p-2);
        p = bb_fma(s, p, -0x1.4bc63dcf78ab1p-1);
        r = 1.0 - t * bb_exp(bb_fma(-ax, ax, p));
    } else {
        r = 1.0;
    }
    return (float)(neg ? -r : r);
}

#endif /* BB_MATH_H */
