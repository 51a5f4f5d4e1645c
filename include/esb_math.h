/* esb_math.h — deterministic elementary functions and counter-based noise.
 *
 * The reference (engine-sim, absent submodule) calls libm's sin/cos/pow and
 * libc rand().  libm and CUDA's libdevice differ in the last ulp and rand() is
 * platform-specific, so a GPU render could never be compared bit-for-bit with
 * a CPU render.  These replacements are built from IEEE-754 +,-,*,/,sqrt,fma
 * and integer bit operations only, so the same inputs give the same bits on
 * x86 (g++ -ffp-contract=off) and on sm_100a (nvcc -fmad=false).  Accuracy is
 * < 2 ulp for sin/cos/log/exp and ~|y ln x| ulp for pow — far inside the
 * reference's own sensitivity.  tests/test_math.py checks them against libm.
 *
 * Both the CUDA kernels and the CPU oracle include this header; nothing else
 * is shared between them.
 */
#ifndef ESB_MATH_H
#define ESB_MATH_H

#include <stdint.h>
#include <math.h>
#include <string.h>

#if defined(__CUDACC__)
#define ESB_HD __host__ __device__ __forceinline__
#else
#define ESB_HD static inline
#endif

ESB_HD int64_t esb_d2bits(double x) {
#if defined(__CUDA_ARCH__)
    return __double_as_longlong(x);
#else
    int64_t b; memcpy(&b, &x, 8); return b;
#endif
}
ESB_HD double esb_bits2d(int64_t b) {
#if defined(__CUDA_ARCH__)
    return __longlong_as_double(b);
#else
    double x; memcpy(&x, &b, 8); return x;
#endif
}

ESB_HD double esb_fma(double a, double b, double c) {
#if defined(__CUDA_ARCH__)
    return __fma_rn(a, b, c);
#else
    return __builtin_fma(a, b, c);
#endif
}

/* ---- sin / cos --------------------------------------------------------- */
/* Cody-Waite reduction by pi/2 in three fma steps, then the classic minimax
 * kernels on [-pi/4, pi/4].  Valid for |x| < ~1e8. */
ESB_HD void esb_sincos(double x, double *s_out, double *c_out) {
    const double INV_PIO2 = 6.36619772367581382433e-01;
    const double PIO2_1 = 1.57079632673412561417e+00;
    const double PIO2_2 = 6.07710050630396597660e-11;
    const double PIO2_3 = 2.02226624879595063154e-21;
    double kf = floor(x * INV_PIO2 + 0.5);
    double r = esb_fma(-kf, PIO2_1, x);
    r = esb_fma(-kf, PIO2_2, r);
    r = esb_fma(-kf, PIO2_3, r);
    int64_t k = (int64_t)kf;
    double z = r * r;
    /* sin kernel */
    double sp = 1.58969099521155010221e-10;
    sp = esb_fma(sp, z, -2.50507602534068634195e-08);
    sp = esb_fma(sp, z, 2.75573137070700676789e-06);
    sp = esb_fma(sp, z, -1.98412698298579493134e-04);
    sp = esb_fma(sp, z, 8.33333333332248946124e-03);
    sp = esb_fma(sp, z, -1.66666666666666324348e-01);
    double sn = esb_fma(r * z, sp, r);
    /* cos kernel */
    double cp = -1.13596475577881948265e-11;
    cp = esb_fma(cp, z, 2.08757232129817482790e-09);
    cp = esb_fma(cp, z, -2.75573143513906633035e-07);
    cp = esb_fma(cp, z, 2.48015872894767294178e-05);
    cp = esb_fma(cp, z, -1.38888888888741095749e-03);
    cp = esb_fma(cp, z, 4.16666666666666019037e-02);
    double cs = esb_fma(z * z, cp, esb_fma(-0.5, z, 1.0));
    switch ((int)(k & 3)) {
    case 0: *s_out = sn; *c_out = cs; break;
    case 1: *s_out = cs; *c_out = -sn; break;
    case 2: *s_out = -sn; *c_out = -cs; break;
    default: *s_out = -cs; *c_out = sn; break;
    }
}
ESB_HD double esb_sin(double x) { double s, c; esb_sincos(x, &s, &c); return s; }
ESB_HD double esb_cos(double x) { double s, c; esb_sincos(x, &s, &c); return c; }

/* Polynomial coefficients of log / exp: literals on the host; on the device a __constant__ table, so that the FP64 instructions take
 * them as constant-bank operands instead of holding every coefficient in a register pair for the whole of a caller (same values). */
#if defined(__CUDA_ARCH__) && defined(ESB_MATH_CONSTANT_TABLE)
#define ESB_K(i, literal) esbMathK[i]
#else
#define ESB_K(i, literal) (literal)
#endif
#if defined(__CUDACC__) && defined(ESB_MATH_CONSTANT_TABLE)
static __constant__ double esbMathK[16] = {
    6.93147180369123816490e-01, 1.90821492927058770002e-10,
    1.531383769920937332e-01, 2.222219843214978396e-01, 3.999999999940941908e-01,
    1.479819860511658591e-01, 1.818357216161805012e-01, 2.857142874366239149e-01, 6.666666666666735130e-01,
    1.44269504088896338700e+00,
    4.13813679705723846039e-08, -1.65339022054652515390e-06, 6.61375632143793436117e-05, -2.77777777770155933842e-03, 1.66666666666666019037e-01,
    0.0};
#endif

/* ---- log / exp / pow ------------------------------------------------------ */
/* Each contains one division.  The bodies are written once, over a divider `dv`, so the CUDA
 * physics kernel can supply its straight-line correctly-rounded division (physics_core.cuh
 * FastAr) while the oracle and everything else use the plain IEEE `/`: same value either way. */
#ifdef __cplusplus
struct EsbIeeeDiv {
    ESB_HD double div(double a, double b) { return a / b; }
};
#define ESB_DIV_TEMPLATE template <class DV>
#define ESB_DIV_ARG , DV &dv
#define ESB_DIV(a, b) dv.div((a), (b))
#else
#define ESB_DIV_TEMPLATE
#define ESB_DIV_ARG
#define ESB_DIV(a, b) ((a) / (b))
#endif

/* x > 0, finite, normal. */
ESB_DIV_TEMPLATE ESB_HD double esb_log_t(double x ESB_DIV_ARG) {
    const double LN2_HI = ESB_K(0, 6.93147180369123816490e-01);
    const double LN2_LO = ESB_K(1, 1.90821492927058770002e-10);
    int64_t b = esb_d2bits(x);
    int64_t e = ((b >> 52) & 0x7ff) - 1023;
    int64_t mant = b & 0x000fffffffffffffLL;
    /* m in [sqrt(2)/2, sqrt(2)) */
    if (mant > 0x6a09e667f3bcdLL) { e += 1; mant |= 0x3fe0000000000000LL; }
    else { mant |= 0x3ff0000000000000LL; }
    double m = esb_bits2d(mant);
    double f = m - 1.0;
    double s = ESB_DIV(f, 2.0 + f);
    double z = s * s;
    double w = z * z;
    double t1 = w * esb_fma(w, esb_fma(w, ESB_K(2, 1.531383769920937332e-01), ESB_K(3, 2.222219843214978396e-01)), ESB_K(4, 3.999999999940941908e-01));
    double t2 = z * esb_fma(w, esb_fma(w, esb_fma(w, ESB_K(5, 1.479819860511658591e-01), ESB_K(6, 1.818357216161805012e-01)), ESB_K(7, 2.857142874366239149e-01)), ESB_K(8, 6.666666666666735130e-01));
    double R = t2 + t1;
    double hfsq = 0.5 * f * f;
    double dk = (double)e;
    return dk * LN2_HI - ((hfsq - (s * (hfsq + R) + dk * LN2_LO)) - f);
}

/* Flushes to 0 below -700 and saturates above 700 (no subnormal results). */
ESB_DIV_TEMPLATE ESB_HD double esb_exp_t(double x ESB_DIV_ARG) {
    const bool under = x < -700.0;
    if (x > 700.0) x = 700.0;
    if (under) x = -700.0;
    const double LN2_HI = ESB_K(0, 6.93147180369123816490e-01);
    const double LN2_LO = ESB_K(1, 1.90821492927058770002e-10);
    const double INV_LN2 = ESB_K(9, 1.44269504088896338700e+00);
    double kf = floor(x * INV_LN2 + 0.5);
    double hi = esb_fma(-kf, LN2_HI, x);
    double lo = kf * LN2_LO;
    double r = hi - lo;
    double t = r * r;
    double c = r - t * esb_fma(t, esb_fma(t, esb_fma(t, esb_fma(t, ESB_K(10, 4.13813679705723846039e-08), ESB_K(11, -1.65339022054652515390e-06)), ESB_K(12, 6.61375632143793436117e-05)), ESB_K(13, -2.77777777770155933842e-03)), ESB_K(14, 1.66666666666666019037e-01));
    double y = 1.0 - ((lo - ESB_DIV(r * c, 2.0 - c)) - hi);
    int64_t k = (int64_t)kf;
    /* scale by 2^k (k within the normal range for our arguments) */
    y = y * esb_bits2d((k + 1023) << 52);
    return under ? 0.0 : y;
}

/* pow for x >= 0 (the only case the gas model needs). */
ESB_DIV_TEMPLATE ESB_HD double esb_pow_t(double x, double y ESB_DIV_ARG) {
    const bool nonpos = x <= 0.0;
    const double xs = nonpos ? 1.0 : x;
#ifdef __cplusplus
    const double r = esb_exp_t(y * esb_log_t(xs, dv), dv);
#else
    const double r = esb_exp_t(y * esb_log_t(xs));
#endif
    return nonpos ? ((y == 0.0) ? 1.0 : 0.0) : r;
}

#ifdef __cplusplus
ESB_HD double esb_log(double x) { EsbIeeeDiv d; return esb_log_t(x, d); }
ESB_HD double esb_exp(double x) { EsbIeeeDiv d; return esb_exp_t(x, d); }
ESB_HD double esb_pow(double x, double y) { EsbIeeeDiv d; return esb_pow_t(x, y, d); }
#else
ESB_HD double esb_log(double x) { return esb_log_t(x); }
ESB_HD double esb_exp(double x) { return esb_exp_t(x); }
ESB_HD double esb_pow(double x, double y) { return esb_pow_t(x, y); }
#endif

/* ---- counter-based noise ---------------------------------------------- */
/* Replaces libc rand() (platform-specific stream).  A splitmix64-style mix of
 * (seed, stream, counter): stateless, so oracle and kernels draw identical
 * numbers without sharing a generator state. */
ESB_HD uint64_t esb_hash64(uint64_t seed, uint64_t stream, uint64_t counter) {
    uint64_t z = seed * 0x9E3779B97F4A7C15ULL + stream * 0xD1B54A32D192ED03ULL + counter * 0xBF58476D1CE4E5B9ULL + 0x94D049BB133111EBULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    z = z ^ (z >> 31);
    z = (z ^ (z >> 33)) * 0xFF51AFD7ED558CCDULL;
    return z ^ (z >> 29);
}
/* uniform in [0,1), 53 bits */
ESB_HD double esb_rand_u01(uint64_t seed, uint64_t stream, uint64_t counter) {
    return (double)(esb_hash64(seed, stream, counter) >> 11) * (1.0 / 9007199254740992.0);
}
/* uniform in [0,1), 24 bits, as float */
ESB_HD float esb_rand_u01f(uint64_t seed, uint64_t stream, uint64_t counter) {
    return (float)(esb_hash64(seed, stream, counter) >> 40) * (1.0f / 16777216.0f);
}

/* noise stream ids */
#define ESB_STREAM_JITTER 1   /* + channel */
#define ESB_STREAM_AIR 17     /* + channel */
#define ESB_STREAM_BURN 33    /* + cylinder */

#endif
