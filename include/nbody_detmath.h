/* nbody_detmath.h -- deterministic atan2 / sincos for the exact-mode N-body step.
 *
 * Why this exists: the reference's per-pair formula calls libm atan2, cos and
 * sin (reference nbody/nbody-seq.c:91,97,98 and :123-125; nbody/nbody-par.c:341,
 * 347,348,382-384).  glibc's results are < 1 ULP but not correctly rounded and
 * change with glibc version and CPU ifunc variant, and CUDA's libdevice differs
 * again, so "bit-exact with nbody-seq" is only well defined against a fixed
 * evaluation scheme.  This header IS that scheme: every operation is an IEEE-754
 * binary64 (or, for one table-index estimate, binary32) add, subtract, multiply,
 * divide or comparison in a fixed order, with no fused multiply-add.  It is
 * compiled by gcc (-ffp-contract=off) into the interposer that the unmodified
 * reference source is linked against (oracle/), and by nvcc into the exact-mode
 * kernels (where the operations are spelled __dadd_rn/__dmul_rn/... so that no
 * compiler flag can contract them).  Same inputs -> same 64 bits on both.
 *
 * Accuracy (measured in tests/test_detmath.py against mpmath and libm): about
 * 1 ULP for atan2 on finite inputs, about 1 ULP for sin/cos on |angle| <= 4.
 * Domain: finite inputs; sincos is meant for atan2 results (|angle| <= pi) and
 * stays accurate to |angle| < 2^19.
 *
 * Written from scratch; constants come from tools/gen_detmath_tables.py (mpmath).
 */
#ifndef NBODY_DETMATH_H
#define NBODY_DETMATH_H

#include <stdint.h>
#include <string.h>
#include "nbody_detmath_tables.h"

#if defined(__CUDACC__)
#define NBODY_DM_FN static __host__ __device__ __forceinline__
#else
#define NBODY_DM_FN static inline
#endif

#if defined(__CUDA_ARCH__)
#define NBODY_DM_ADD(a, b) __dadd_rn((a), (b))
#define NBODY_DM_SUB(a, b) __dsub_rn((a), (b))
#define NBODY_DM_MUL(a, b) __dmul_rn((a), (b))
#define NBODY_DM_DIV(a, b) __ddiv_rn((a), (b))
#define NBODY_DM_FADD(a, b) __fadd_rn((a), (b))
#define NBODY_DM_FMUL(a, b) __fmul_rn((a), (b))
#define NBODY_DM_FDIV(a, b) __fdiv_rn((a), (b))
#else
#define NBODY_DM_ADD(a, b) ((a) + (b))
#define NBODY_DM_SUB(a, b) ((a) - (b))
#define NBODY_DM_MUL(a, b) ((a) * (b))
#define NBODY_DM_DIV(a, b) ((a) / (b))
#define NBODY_DM_FADD(a, b) ((a) + (b))
#define NBODY_DM_FMUL(a, b) ((a) * (b))
#define NBODY_DM_FDIV(a, b) ((a) / (b))
#endif

NBODY_DM_FN uint64_t nbody_dm_bits(double x)
{
#if defined(__CUDA_ARCH__)
    return (uint64_t)__double_as_longlong(x);
#else
    uint64_t u;
    memcpy(&u, &x, sizeof u);
    return u;
#endif
}

NBODY_DM_FN double nbody_dm_from_bits(uint64_t u)
{
#if defined(__CUDA_ARCH__)
    return __longlong_as_double((long long)u);
#else
    double x;
    memcpy(&x, &u, sizeof x);
    return x;
#endif
}

NBODY_DM_FN float nbody_dm_float_from_bits(uint32_t u)
{
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    float x;
    memcpy(&x, &u, sizeof x);
    return x;
#endif
}

/* Table index for atan(a/b), 0 <= a <= b, b > 0: k = round(8*a/b) from a 20-bit
 * estimate of the ratio.  Integer and binary32 work only (keeps the FP64 pipe
 * free on the GPU); any k in 0..8 gives a valid reduction, k only sets |u|. */
NBODY_DM_FN int nbody_dm_atan_index(double a, double b)
{
    uint32_t ha = (uint32_t)(nbody_dm_bits(a) >> 32);
    uint32_t hb = (uint32_t)(nbody_dm_bits(b) >> 32);
    int de = (int)(ha >> 20) - (int)(hb >> 20);            /* exponent difference, <= 0 */
    uint32_t fb = (127u << 23) | ((hb & 0xFFFFFu) << 3);   /* b scaled into [1,2) */
    uint32_t fa = ((uint32_t)(127 + de) << 23) | ((ha & 0xFFFFFu) << 3);
    if (de < -100 || ha < 0x00100000u) fa = 0u;            /* ratio ~ 0 (or a zero/subnormal) */
    float rf = NBODY_DM_FDIV(nbody_dm_float_from_bits(fa), nbody_dm_float_from_bits(fb));
    int k = (int)NBODY_DM_FADD(NBODY_DM_FMUL(rf, 8.0f), 0.5f);
    if (k > 8) k = 8;
    if (k < 0) k = 0;
    return k;
}

/* atan2(y, x), C99 semantics for finite arguments including signed zeros.
 * tab = the 72-entry combination table (NBODY_DM_ATAN_TAB_INIT); the caller
 * chooses where it lives (static array on the host, shared memory on the GPU). */
NBODY_DM_FN double nbody_dm_atan2(double y, double x, const double *tab)
{
    const uint64_t SIGN = 0x8000000000000000ull;
    uint64_t by = nbody_dm_bits(y), bx = nbody_dm_bits(x);
    double ay = nbody_dm_from_bits(by & ~SIGN);
    double ax = nbody_dm_from_bits(bx & ~SIGN);
    int swap = ay > ax;
    double a = swap ? ax : ay;                 /* min */
    double b = swap ? ay : ax;                 /* max */
    int q = swap + (int)((bx >> 63) << 1);
    int k = 0;
    double p = 0.0;
    if (b > 0.0) {
        if (b < 0x1p-900) {   /* subnormal neighbourhood: scale up (exact) so the ratio estimate sees normal numbers */
            a = NBODY_DM_MUL(a, 0x1p200);
            b = NBODY_DM_MUL(b, 0x1p200);
        }
        k = nbody_dm_atan_index(a, b);
        double c = NBODY_DM_MUL((double)k, 0.125);
        double num = NBODY_DM_SUB(a, NBODY_DM_MUL(c, b));
        double den = NBODY_DM_ADD(b, NBODY_DM_MUL(c, a));
        double u = NBODY_DM_DIV(num, den);
        double z = NBODY_DM_MUL(u, u);
        double s = NBODY_DM_A7;
        s = NBODY_DM_ADD(NBODY_DM_MUL(s, z), NBODY_DM_A6);
        s = NBODY_DM_ADD(NBODY_DM_MUL(s, z), NBODY_DM_A5);
        s = NBODY_DM_ADD(NBODY_DM_MUL(s, z), NBODY_DM_A4);
        s = NBODY_DM_ADD(NBODY_DM_MUL(s, z), NBODY_DM_A3);
        s = NBODY_DM_ADD(NBODY_DM_MUL(s, z), NBODY_DM_A2);
        s = NBODY_DM_ADD(NBODY_DM_MUL(s, z), NBODY_DM_A1);
        p = NBODY_DM_ADD(u, NBODY_DM_MUL(u, NBODY_DM_MUL(z, s)));
    }
    const double *e = tab + ((q * 9 + k) << 1);
    double hi = e[0], lo = e[1];
    /* q = 0,3 add the reduced arctangent, q = 1,2 subtract it */
    double sp = (q == 1 || q == 2) ? -p : p;
    double r = NBODY_DM_ADD(hi, NBODY_DM_ADD(lo, sp));
    return nbody_dm_from_bits(nbody_dm_bits(r) | (by & SIGN));
}

/* sin and cos of the same angle (gcc merges the reference's cos(angle), sin(angle)
 * into one sincos call; the interposer exports all three entry points). */
NBODY_DM_FN void nbody_dm_sincos(double x, double *sn, double *cs)
{
    double qf = NBODY_DM_MUL(x, NBODY_DM_2OPI);
    int k = (int)NBODY_DM_ADD(qf, (qf < 0.0) ? -0.5 : 0.5);   /* round half away; |k| small */
    double kd = (double)k;
    /* three-part Cody-Waite: k*PIO2_1 and k*PIO2_2 are exact (33-bit constants), so
     * the first two subtractions cancel exactly when x is close to k*pi/2 */
    double r = NBODY_DM_SUB(NBODY_DM_SUB(NBODY_DM_SUB(x, NBODY_DM_MUL(kd, NBODY_DM_PIO2_1)),
                                         NBODY_DM_MUL(kd, NBODY_DM_PIO2_2)),
                            NBODY_DM_MUL(kd, NBODY_DM_PIO2_2T));
    double z = NBODY_DM_MUL(r, r);
    double s = NBODY_DM_S7;
    s = NBODY_DM_ADD(NBODY_DM_MUL(s, z), NBODY_DM_S6);
    s = NBODY_DM_ADD(NBODY_DM_MUL(s, z), NBODY_DM_S5);
    s = NBODY_DM_ADD(NBODY_DM_MUL(s, z), NBODY_DM_S4);
    s = NBODY_DM_ADD(NBODY_DM_MUL(s, z), NBODY_DM_S3);
    s = NBODY_DM_ADD(NBODY_DM_MUL(s, z), NBODY_DM_S2);
    s = NBODY_DM_ADD(NBODY_DM_MUL(s, z), NBODY_DM_S1);
    double sr = NBODY_DM_ADD(r, NBODY_DM_MUL(r, NBODY_DM_MUL(z, s)));
    double c = NBODY_DM_C6;
    c = NBODY_DM_ADD(NBODY_DM_MUL(c, z), NBODY_DM_C5);
    c = NBODY_DM_ADD(NBODY_DM_MUL(c, z), NBODY_DM_C4);
    c = NBODY_DM_ADD(NBODY_DM_MUL(c, z), NBODY_DM_C3);
    c = NBODY_DM_ADD(NBODY_DM_MUL(c, z), NBODY_DM_C2);
    c = NBODY_DM_ADD(NBODY_DM_MUL(c, z), NBODY_DM_C1);
    /* cos r = (1 - z/2) + z*z*c, with the rounding error of (1 - z/2) recovered:
     * h = z/2 is exact, w = 1 - h, (1 - w) - h is the exact residual. */
    double h = NBODY_DM_MUL(z, 0.5);
    double w = NBODY_DM_SUB(1.0, h);
    double res = NBODY_DM_SUB(NBODY_DM_SUB(1.0, w), h);
    double cr = NBODY_DM_ADD(w, NBODY_DM_ADD(res, NBODY_DM_MUL(NBODY_DM_MUL(z, z), c)));
    switch (k & 3) {
    case 0:  *sn = sr;  *cs = cr;  break;
    case 1:  *sn = cr;  *cs = -sr; break;
    case 2:  *sn = -sr; *cs = -cr; break;
    default: *sn = -cr; *cs = sr;  break;
    }
}

#endif /* NBODY_DETMATH_H */
