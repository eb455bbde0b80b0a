/* Accuracy probe for include/bb_math.h against glibc (double functions rounded
 * to float = the correctly rounded float result up to double-rounding).
 * gcc -O2 -ffp-contract=off -mfma -Iinclude tools/check_math.c -lm -o /tmp/check_math */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include "bb_math.h"

static unsigned long long s = 88172645463325252ull;
static double urand(void) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return (s >> 11) * (1.0 / 9007199254740992.0); }
static int ulpdiff(float a, float b) { int ia, ib; memcpy(&ia,&a,4); memcpy(&ib,&b,4); if (ia<0) ia=0x80000000-ia; if (ib<0) ib=0x80000000-ib; return abs(ia-ib); }

#define PROBE(name, mine, ref_d, ref_f, lo, hi) do { \
    long n = 20000000, mism_d = 0, mism_f = 0; int maxu_d = 0, maxu_f = 0; \
    for (long i = 0; i < n; i++) { float x = (float)((lo) + ((hi)-(lo)) * urand()); \
        float m = mine(x); float rd = (float)ref_d((double)x); float rf = ref_f(x); \
        int ud = ulpdiff(m, rd), uf = ulpdiff(m, rf); \
        if (ud) mism_d++; if (uf) mism_f++; if (ud > maxu_d) maxu_d = ud; if (uf > maxu_f) maxu_f = uf; } \
    printf("%-8s [%g,%g]: vs (float)f(double): mismatch %.2e max %d ulp | vs glibc float fn: mismatch %.2e max %d ulp\n", \
        name, (double)(lo), (double)(hi), (double)mism_d / n, maxu_d, (double)mism_f / n, maxu_f); } while (0)

static float my_atan2_fixed(float y) { return bb_atan2f(y, 0.37f); }
static double ref_atan2_d(double y) { return atan2(y, (double)0.37f); }
static float ref_atan2_f(float y) { return atan2f(y, 0.37f); }
static float my_atan2_neg(float y) { return bb_atan2f(y, -1.7f); }
static double ref_atan2n_d(double y) { return atan2(y, (double)-1.7f); }
static float ref_atan2n_f(float y) { return atan2f(y, -1.7f); }

int main(void)
{
    PROBE("sinf", bb_sinf, sin, sinf, -8.0, 8.0);
    PROBE("cosf", bb_cosf, cos, cosf, -8.0, 8.0);
    PROBE("atanf", bb_atanf, atan, atanf, 0.0, 64.0);
    PROBE("atanf", bb_atanf, atan, atanf, -1.0, 1.0);
    PROBE("acosf", bb_acosf, acos, acosf, -1.0, 1.0);
    PROBE("expf", bb_expf, exp, expf, -20.0, 2.0);
    PROBE("erff", bb_erff, erf, erff, 0.0, 0.6);
    PROBE("erff", bb_erff, erf, erff, -5.0, 5.0);
    PROBE("atan2f+", my_atan2_fixed, ref_atan2_d, ref_atan2_f, -30.0, 30.0);
    PROBE("atan2f-", my_atan2_neg, ref_atan2n_d, ref_atan2n_f, -30.0, 30.0);
    printf("edge: sin(0)=%g sin(-0)=%g cos(0)=%g atan2(0,1)=%g atan2(0,-1)=%.9g atan2(1,0)=%.9g acos(1)=%g acos(-1)=%.9g acos(0)=%.9g erf(0)=%g erf(7)=%g erf(-7)=%g exp(0)=%g atan(inf)=%.9g\n",
        bb_sinf(0.f), bb_sinf(-0.f), bb_cosf(0.f), bb_atan2f(0.f,1.f), bb_atan2f(0.f,-1.f), bb_atan2f(1.f,0.f), bb_acosf(1.f), bb_acosf(-1.f), bb_acosf(0.f), bb_erff(0.f), bb_erff(7.f), bb_erff(-7.f), bb_expf(0.f), bb_atanf(INFINITY));
    return 0;
}
