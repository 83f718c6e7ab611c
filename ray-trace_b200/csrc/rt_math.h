/* rt_math.h — deterministic FP64 elementary functions, one source for host (gcc, C99/C++) and
 * device (nvcc, sm_100a).
 *
 * Why this exists: the reference differentiates the refractive index with central differences of
 * width 1e-11 (RayTrace_lat.jl:210-222, RayTrace_3D.jl:235-247), which amplifies a 1-ulp change of
 * mu by ~1e5.  The GPU path and the CPU oracle therefore have to evaluate sin/cos/tan/atan/acos/
 * exp/10^x/log to the SAME BITS.  Vendor libms (glibc, CUDA libdevice, Julia's openlibm) do not
 * agree with each other in the last bit, so both sides compile this header instead.
 *
 * Rules that make it bit-reproducible:
 *   - only IEEE-754 correctly rounded primitives: + - * / sqrt and fma (RT_FMA);
 *   - every fused multiply-add is written explicitly; the translation units that include this
 *     header are compiled with contraction off (gcc -ffp-contract=off, nvcc -fmad=false);
 *   - no table lookups that depend on memory layout, no fmin/fmax, no float->int conversions of
 *     unbounded values (x86 and PTX disagree on out-of-range conversions).
 *
 * Accuracy (measured in tests/test_rt_math.py against mpmath): <= 1 ulp for exp, exp10, log, sin,
 * cos, atan, acos; <= 2 ulp for tan (sin/cos quotient) and log10; <= 4 ulp for x^-4.5.
 * Domain note: trigonometric arguments with |x| >= 2^45 return NaN (documented in DESIGN.md).
 */
#ifndef RT_MATH_H
#define RT_MATH_H

#include <stdint.h>

#if defined(__CUDACC__)
#define RT_HD __host__ __device__ __forceinline__
#else
#define RT_HD static inline
#endif

#if defined(__CUDA_ARCH__)
#define RT_FMA(a, b, c) __fma_rn((a), (b), (c))
#define RT_SQRT(a) __dsqrt_rn(a)
#else
#define RT_FMA(a, b, c) __builtin_fma((a), (b), (c))
#define RT_SQRT(a) __builtin_sqrt(a)
#endif

/* Polynomial coefficients and reduction constants.  On the device they are read from ONE __constant__ table
 * (pairs of them arrive with a single uniform load; as literals every FP64 constant costs two UMOV instructions,
 * 11 % of the first profile's instruction stream); on the host they are the literals.  One list, so the two
 * cannot drift apart. */
#define RT_CONSTS(X) \
  X(EXP0, 1.6059043836821613e-10) \
  X(EXP1, 2.08767569878681e-09) \
  X(EXP2, 2.505210838544172e-08) \
  X(EXP3, 2.755731922398589e-07) \
  X(EXP4, 2.7557319223985893e-06) \
  X(EXP5, 2.48015873015873e-05) \
  X(EXP6, 0.0001984126984126984) \
  X(EXP7, 0.001388888888888889) \
  X(EXP8, 0.008333333333333333) \
  X(EXP9, 0.041666666666666664) \
  X(EXP10, 0.16666666666666666) \
  X(SIN0, 1.58969099521155010221e-10) \
  X(SIN1, -2.50507602534068634195e-08) \
  X(SIN2, 2.75573137070700676789e-06) \
  X(SIN3, -1.98412698298579493134e-04) \
  X(SIN4, 8.33333333332248946124e-03) \
  X(SIN5, 1.66666666666666324348e-01) \
  X(COS0, -1.13596475577881948265e-11) \
  X(COS1, 2.08757232129817482790e-09) \
  X(COS2, -2.75573143513906633035e-07) \
  X(COS3, 2.48015872894767294178e-05) \
  X(COS4, -1.38888888888741095749e-03) \
  X(COS5, 4.16666666666666019037e-02) \
  X(LOG2E, 1.4426950408889634) \
  X(LN2HI, 0.6931471805599453) \
  X(LN2LO, 2.3190468138462996e-17) \
  X(LOG2_10, 3.321928094887362) \
  X(LG2HI, 0.3010299956639812) \
  X(LG2LO, -2.8037281277851704e-18) \
  X(LN10HI, 2.302585092994046) \
  X(LN10LO, -2.1707562233822494e-16) \
  X(TWO_OVER_PI, 0.6366197723675814) \
  X(PIO2_1, 1.5707963267948966) \
  X(PIO2_2, 6.123233995736766e-17) \
  X(PIO2_3, -1.4973849048591698e-33) \
  X(AS_P0, 3.47933107596021167570e-05) \
  X(AS_P1, 7.91534994289814532176e-04) \
  X(AS_P2, -4.00555345006794114027e-02) \
  X(AS_P3, 2.01212532134862925881e-01) \
  X(AS_P4, -3.25565818622400915405e-01) \
  X(AS_P5, 1.66666666666666657415e-01) \
  X(AS_Q0, 7.70381505559019352791e-02) \
  X(AS_Q1, -6.88283971605453293030e-01) \
  X(AS_Q2, 2.02094576023350569471e+00) \
  X(AS_Q3, -2.40339491173441421878e+00) \
  X(AT_A0, 1.62858201153657823623e-02) \
  X(AT_A1, 4.97687799461593236017e-02) \
  X(AT_A2, 6.66107313738753120669e-02) \
  X(AT_A3, 9.09088713343650656196e-02) \
  X(AT_A4, 1.42857142725034663711e-01) \
  X(AT_A5, 3.33333333333329318027e-01) \
  X(AT_B0, -3.65315727442169155270e-02) \
  X(AT_B1, -5.83357013379057348645e-02) \
  X(AT_B2, -7.69187620504482999495e-02) \
  X(AT_B3, -1.11111104054623557880e-01) \
  X(AT_B4, -1.99999999998764832476e-01)
#define RT_X_ENUM(n, v) RT_CI_##n,
enum { RT_CONSTS(RT_X_ENUM) RT_CI_COUNT };
#define RT_X_HOST(n, v) static const double RT_CV_##n __attribute__((unused)) = v;
RT_CONSTS(RT_X_HOST)
#if defined(__CUDACC__)
#define RT_X_VAL(n, v) v,
static __constant__ double rt_ctab[RT_CI_COUNT] = {RT_CONSTS(RT_X_VAL)};
#endif
#if defined(__CUDA_ARCH__)
#define RT_CK(n) rt_ctab[RT_CI_##n]
#else
#define RT_CK(n) RT_CV_##n
#endif

RT_HD uint64_t rt_d2u(double x) {
#if defined(__CUDA_ARCH__)
  return (uint64_t)__double_as_longlong(x);
#else
  uint64_t u;
  __builtin_memcpy(&u, &x, 8);
  return u;
#endif
}
RT_HD double rt_u2d(uint64_t u) {
#if defined(__CUDA_ARCH__)
  return __longlong_as_double((long long)u);
#else
  double x;
  __builtin_memcpy(&x, &u, 8);
  return x;
#endif
}

RT_HD double rt_nan(void) { return rt_u2d(0x7ff8000000000000ULL); }
RT_HD double rt_inf(void) { return rt_u2d(0x7ff0000000000000ULL); }
RT_HD double rt_abs(double x) { return rt_u2d(rt_d2u(x) & 0x7fffffffffffffffULL); }
RT_HD int rt_isnan(double x) { return x != x; }
RT_HD int rt_isfinite(double x) { return (rt_d2u(x) & 0x7ff0000000000000ULL) != 0x7ff0000000000000ULL; }

/* round to nearest integer (ties to even), valid for |x| < 2^51 */
RT_HD double rt_rint_small(double x) {
  const double magic = 6755399441055744.0; /* 1.5 * 2^52 */
  return (x + magic) - magic;
}

/* y * 2^k for k in [-2200, 2200]; two exact-ish scalings so that the factors stay normal */
RT_HD double rt_scale2(double y, int k) {
  int k1, k2;
  if (k > 1023) {
    k1 = 1023;
    k2 = k - 1023;
    if (k2 > 1023) k2 = 1023;
  } else if (k < -1022) {
    k1 = -1022;
    k2 = k + 1022;
    if (k2 < -1022) k2 = -1022;
  } else {
    k1 = k;
    k2 = 0;
  }
  double s1 = rt_u2d((uint64_t)(k1 + 1023) << 52);
  double s2 = rt_u2d((uint64_t)(k2 + 1023) << 52);
  return (y * s1) * s2;
}

/* ---------------------------------------------------------------- exp ---- */
/* exp(r) on |r| <= 0.35 as 1 + r + r^2 q(r), q = Taylor to r^11 (truncation < 2^-57) */
RT_HD double rt_exp_kernel(double r, double tail) {
  /* q(r) = EXP0 r^11 + ... + EXP10 r + 1/2 as two interleaved Horner chains in z = r^2 (even and odd powers): half the
   * dependent depth of one chain in r */
  double z = r * r;
  double qo = RT_CK(EXP0);
  double qe = RT_CK(EXP1);
  qo = RT_FMA(qo, z, RT_CK(EXP2));
  qe = RT_FMA(qe, z, RT_CK(EXP3));
  qo = RT_FMA(qo, z, RT_CK(EXP4));
  qe = RT_FMA(qe, z, RT_CK(EXP5));
  qo = RT_FMA(qo, z, RT_CK(EXP6));
  qe = RT_FMA(qe, z, RT_CK(EXP7));
  qo = RT_FMA(qo, z, RT_CK(EXP8));
  qe = RT_FMA(qe, z, RT_CK(EXP9));
  qo = RT_FMA(qo, z, RT_CK(EXP10));
  qe = RT_FMA(qe, z, 0.5);
  double q = RT_FMA(qo, r, qe);
  double lo = RT_FMA(z, q, tail);
  return 1.0 + (r + lo);
}

RT_HD double rt_exp(double x) {
  if (x != x) return x;
  if (x > 709.782712893384) return rt_inf();
  if (x < -745.2) return 0.0;
  double kd = rt_rint_small(x * 1.4426950408889634);
  double r = RT_FMA(-kd, 0.6931471805599453, x);
  r = RT_FMA(-kd, 2.3190468138462996e-17, r);
  return rt_scale2(rt_exp_kernel(r, 0.0), (int)kd);
}

/* 10^x (Julia `10^x`, RayTrace_lat.jl:74,76) */
RT_HD double rt_exp10(double x) {
  if (x != x) return x;
  if (x > 308.25471555991675) return rt_inf();
  if (x < -323.7) return 0.0;
  double kd = rt_rint_small(x * 3.321928094887362);
  double r = RT_FMA(-kd, 0.3010299956639812, x);
  r = RT_FMA(-kd, -2.8037281277851704e-18, r);
  /* t = r * ln(10) as th + tl */
  double th = r * 2.302585092994046;
  double tl = RT_FMA(r, 2.302585092994046, -th);
  tl = RT_FMA(r, -2.1707562233822494e-16, tl);
  return rt_scale2(rt_exp_kernel(th, tl), (int)kd);
}

/* ---------------------------------------------------------------- log ---- */
/* natural log; x = 2^k (1+f), sqrt(1/2) < 1+f <= sqrt(2); log(1+f) = 2s + s*R(s^2) with
 * s = f/(2+f).  Used by the step-size controller (EEst^beta = exp(beta log EEst)) and the
 * initial-step heuristic only — never inside the refractive index. */
RT_HD double rt_log(double x) {
  if (x != x) return x;
  if (x < 0.0) return rt_nan();
  if (x == 0.0) return -rt_inf();
  if (!rt_isfinite(x)) return x;
  int k = 0;
  uint64_t u = rt_d2u(x);
  if ((u >> 52) == 0) { /* subnormal: scale up by 2^54 */
    x = x * 18014398509481984.0;
    u = rt_d2u(x);
    k = -54;
  }
  k += (int)(u >> 52) - 1023;
  uint64_t m = u & 0x000fffffffffffffULL;
  /* mantissa in [1,2); move to [sqrt(1/2), sqrt(2)) */
  if (m >= 0x6a09e667f3bcdULL) { /* >= sqrt(2) */
    k += 1;
    x = rt_u2d(m | 0x3fe0000000000000ULL);
  } else {
    x = rt_u2d(m | 0x3ff0000000000000ULL);
  }
  double f = x - 1.0;
  double s = f / (2.0 + f);
  double z = s * s;
  double p = 1.479819860511658591e-01;
  p = RT_FMA(p, z, 1.531383769920937332e-01);
  p = RT_FMA(p, z, 1.818357216161805012e-01);
  p = RT_FMA(p, z, 2.222219843214978396e-01);
  p = RT_FMA(p, z, 2.857142874366239149e-01);
  p = RT_FMA(p, z, 3.999999999940941908e-01);
  p = RT_FMA(p, z, 6.666666666666735130e-01);
  double R = p * z;
  double hfsq = 0.5 * f * f;
  double dk = (double)k;
  /* log(1+f) = f - (hfsq - s*(hfsq+R)) */
  double lo = RT_FMA(dk, 2.3190468138462996e-17, s * (hfsq + R));
  return RT_FMA(dk, 0.6931471805599453, f - (hfsq - lo));
}

RT_HD double rt_log10(double x) {
  double l = rt_log(x);
  double h = l * 0.4342944819032518;
  double e = RT_FMA(l, 0.4342944819032518, -h);
  e = RT_FMA(l, 1.098319650216765e-17, e);
  return h + e;
}

/* x^p for x > 0 via exp(p log x): controller use only (relative error ~ |p log x| ulp) */
RT_HD double rt_pow_pos(double x, double p) { return rt_exp(p * rt_log(x)); }

/* --------------------------------------------------------------- trig ---- */
/* Cody–Waite with exact FMA products: x - n*(pi/2) = r + tail, |r| <= pi/4 (+eps); q = n mod 4.
 * The first product-difference is exact (both terms are multiples of 2^-52 and the result is < 1),
 * the rounding error of the second is recovered into `tail`.  Valid for |x| < 2^45; caller guards. */
RT_HD double rt_rem_pio2(double x, double *tail, int *q) {
  double n = rt_rint_small(x * 0.6366197723675814);
  double r1 = RT_FMA(-n, 1.5707963267948966, x);
  double r2 = RT_FMA(-n, 6.123233995736766e-17, r1);
  double t = RT_FMA(-n, 6.123233995736766e-17, r1 - r2);
  t = RT_FMA(-n, -1.4973849048591698e-33, t);
  double r = r2 + t;
  *tail = (r2 - r) + t;
  *q = (int)(((int64_t)n) & 3);
  return r;
}

/* sin(r + y), |r| <= pi/4, y the reduction tail */
RT_HD double rt_sin_kernel(double r, double y) {
  double z = r * r;
  double p = RT_CK(SIN0);
  p = RT_FMA(p, z, RT_CK(SIN1));
  p = RT_FMA(p, z, RT_CK(SIN2));
  p = RT_FMA(p, z, RT_CK(SIN3));
  p = RT_FMA(p, z, RT_CK(SIN4));
  double v = z * r;
  /* r + v*(S1 + z*p) + y*(1 - z/2) */
  double corr = RT_FMA(z, RT_FMA(-v, p, 0.5 * y), -y); /* z*(y/2 - v*p) - y */
  return r - RT_FMA(v, RT_CK(SIN5), corr);
}

/* cos(r + y) */
RT_HD double rt_cos_kernel(double r, double y) {
  double z = r * r;
  double p = RT_CK(COS0);
  p = RT_FMA(p, z, RT_CK(COS1));
  p = RT_FMA(p, z, RT_CK(COS2));
  p = RT_FMA(p, z, RT_CK(COS3));
  p = RT_FMA(p, z, RT_CK(COS4));
  p = RT_FMA(p, z, RT_CK(COS5));
  double hz = 0.5 * z;
  double w = 1.0 - hz;
  double tailc = RT_FMA(z * z, p, -(r * y)); /* z^2 p - r y */
  return w + (((1.0 - w) - hz) + tailc);
}

RT_HD void rt_sincos(double x, double *s, double *c) {
  if (!(rt_abs(x) < 35184372088832.0)) { /* 2^45, also catches NaN/Inf */
    *s = rt_nan();
    *c = rt_nan();
    return;
  }
  int q;
  double y;
  double r = rt_rem_pio2(x, &y, &q);
  double sk = rt_sin_kernel(r, y);
  double ck = rt_cos_kernel(r, y);
  double ss = (q & 1) ? ck : sk;
  double cc = (q & 1) ? sk : ck;
  if (q & 2) ss = -ss;
  if ((q + 1) & 2) cc = -cc;
  *s = ss;
  *c = cc;
}

RT_HD double rt_sin(double x) {
  double s, c;
  rt_sincos(x, &s, &c);
  return s;
}
RT_HD double rt_cos(double x) {
  double s, c;
  rt_sincos(x, &s, &c);
  return c;
}
RT_HD double rt_tan(double x) {
  double s, c;
  rt_sincos(x, &s, &c);
  return s / c;
}

/* --------------------------------------------------------------- atan ---- */
/* odd polynomial on |t| <= 7/16 after the classic 4-breakpoint reduction
 * atan(x) = atan(c) + atan((x-c)/(1+cx)), c in {0.5, 1, 1.5, inf} */
RT_HD double rt_atan(double x) {
  if (x != x) return x;
  double ax = rt_abs(x);
  double hi, lo, t;
  int id;
  if (ax < 0.4375) {
    id = -1;
    t = ax;
    hi = 0.0;
    lo = 0.0;
  } else if (ax < 0.6875) {
    id = 0;
    t = (2.0 * ax - 1.0) / (2.0 + ax);
    hi = 0.4636476090008061;
    lo = 2.2698777452961687e-17;
  } else if (ax < 1.1875) {
    id = 1;
    t = (ax - 1.0) / (ax + 1.0);
    hi = 0.7853981633974483;
    lo = 3.061616997868383e-17;
  } else if (ax < 2.4375) {
    id = 2;
    t = (ax - 1.5) / (1.0 + 1.5 * ax);
    hi = 0.982793723247329;
    lo = 1.3903311031230998e-17;
  } else {
    id = 3;
    t = -1.0 / ax; /* ax = inf gives -0 */
    hi = 1.5707963267948966;
    lo = 6.123233995736766e-17;
  }
  double z = t * t;
  double w = z * z;
  double s1 = 1.62858201153657823623e-02;
  s1 = RT_FMA(s1, w, 4.97687799461593236017e-02);
  s1 = RT_FMA(s1, w, 6.66107313738753120669e-02);
  s1 = RT_FMA(s1, w, 9.09088713343650656196e-02);
  s1 = RT_FMA(s1, w, 1.42857142725034663711e-01);
  s1 = RT_FMA(s1, w, 3.33333333333329318027e-01);
  s1 = z * s1;
  double s2 = -3.65315727442169155270e-02;
  s2 = RT_FMA(s2, w, -5.83357013379057348645e-02);
  s2 = RT_FMA(s2, w, -7.69187620504482999495e-02);
  s2 = RT_FMA(s2, w, -1.11111104054623557880e-01);
  s2 = RT_FMA(s2, w, -1.99999999998764832476e-01);
  s2 = w * s2;
  double r;
  if (id < 0) {
    r = t - t * (s1 + s2);
  } else {
    r = hi - ((t * (s1 + s2) - lo) - t);
  }
  return (x < 0.0) ? -r : r;
}

/* --------------------------------------------------------------- acos ---- */
/* R(z) ~ (asin(x)-x)/x^3 as a rational P/Q on z = x^2 <= 1/4 */
RT_HD double rt_asin_R(double z) {
  double p = 3.47933107596021167570e-05;
  p = RT_FMA(p, z, 7.91534994289814532176e-04);
  p = RT_FMA(p, z, -4.00555345006794114027e-02);
  p = RT_FMA(p, z, 2.01212532134862925881e-01);
  p = RT_FMA(p, z, -3.25565818622400915405e-01);
  p = RT_FMA(p, z, 1.66666666666666657415e-01);
  p = p * z;
  double q = 7.70381505559019352791e-02;
  q = RT_FMA(q, z, -6.88283971605453293030e-01);
  q = RT_FMA(q, z, 2.02094576023350569471e+00);
  q = RT_FMA(q, z, -2.40339491173441421878e+00);
  q = RT_FMA(q, z, 1.0);
  return p / q;
}

RT_HD double rt_acos(double x) {
  if (x != x) return x;
  double ax = rt_abs(x);
  if (ax > 1.0) return rt_nan();
  if (ax == 1.0) return (x > 0.0) ? 0.0 : 3.141592653589793;
  if (ax < 0.5) {
    /* acos(x) = pi/2 - (x + x*R(x^2)) */
    double z = x * x;
    double r = rt_asin_R(z);
    return 1.5707963267948966 - (x - (6.123233995736766e-17 - x * r));
  }
  if (x < 0.0) {
    /* acos(x) = pi - 2 asin(sqrt((1+x)/2)) */
    double z = (1.0 + x) * 0.5;
    double s = RT_SQRT(z);
    double r = rt_asin_R(z);
    double w = r * s - 6.123233995736766e-17;
    return 3.141592653589793 - 2.0 * (s + w);
  }
  /* x >= 0.5: acos(x) = 2 asin(sqrt((1-x)/2)); split s to recover the sqrt rounding error */
  double z = (1.0 - x) * 0.5;
  double s = RT_SQRT(z);
  double df = rt_u2d(rt_d2u(s) & 0xffffffff00000000ULL);
  double c = (z - df * df) / (s + df);
  double r = rt_asin_R(z);
  double w = r * s + c;
  return 2.0 * (df + w);
}

/* ------------------------------------------------------------- powers ---- */
/* L^(-4.5), the only non-integer power on the path (RayTrace_lat.jl:78, RayTrace_3D.jl:114) */
RT_HD double rt_pow_m4p5(double x) { return 1.0 / (((x * x) * (x * x)) * RT_SQRT(x)); }

#endif /* RT_MATH_H */
