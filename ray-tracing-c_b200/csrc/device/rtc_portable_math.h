/*
 * rtc_portable_math.h -- transcendental functions written only with IEEE-754 binary32 + - * / and integer ops,
 * so that gcc (-O2 -ffp-contract=off) and nvcc (-fmad=false) produce THE SAME BITS for the same argument.
 *
 * Why: the reference calls glibc sinf/cosf/logf/powf/atan2f/acosf on the hot path (material.c:26-27,73;
 * hittable.c:146-147,171-173,413; texture.c:50).  glibc and CUDA's libdevice round differently in the last
 * place, which makes every path fork after its first diffuse bounce.  The "exact" kernel (RT_KERNEL_EXACT) and
 * the oracle's PORTABLE mode both use these functions instead, which makes whole images bit-comparable between
 * CPU and GPU; the oracle's LIBM mode (glibc, bit-identical to the compiled reference) shows the substitution is
 * statistically invisible.  Accuracy: <= 2 ulp on the ranges the path uses (tests/test_portable_math.py).
 *
 * Polynomials are the classic single-precision Cephes minimax sets (public domain, S. Moshier).
 * Usable from C11 and CUDA C++.
 */
#ifndef RTC_PORTABLE_MATH_H
#define RTC_PORTABLE_MATH_H

#include <stdint.h>

#if defined(__CUDACC__)
#define RTC_PM_FN __host__ __device__ __forceinline__
#else
#define RTC_PM_FN static inline
#endif

RTC_PM_FN float pm_from_bits(uint32_t u) {
  union {
    uint32_t u;
    float f;
  } c;
  c.u = u;
  return c.f;
}
RTC_PM_FN uint32_t pm_to_bits(float f) {
  union {
    uint32_t u;
    float f;
  } c;
  c.f = f;
  return c.u;
}
RTC_PM_FN float pm_abs(float x) { return pm_from_bits(pm_to_bits(x) & 0x7fffffffu); }

#define PM_PI_F 3.14159274101257324f   /* (float)M_PI */
#define PM_PIO2_F 1.57079637050628662f /* (float)(M_PI/2) */
#define PM_PIO4_F 0.785398185253143311f

/* Shared argument reduction: |x| = j*(pi/4) + r with j even, r in [-pi/4, pi/4]; returns j & 7. */
RTC_PM_FN int pm_reduce_pio4(float ax, float *r) {
  const float FOPI = 1.27323954473516f;
  const float DP1 = 0.78515625f, DP2 = 2.4187564849853515625e-4f, DP3 = 3.77489497744594108e-8f;
  int j = (int)(FOPI * ax);
  j += (j & 1);
  float y = (float)j;
  *r = ((ax - y * DP1) - y * DP2) - y * DP3;
  return j & 7;
}
RTC_PM_FN float pm_sin_poly(float r) {
  float z = r * r;
  return r + r * z * ((-1.9515295891e-4f * z + 8.3321608736e-3f) * z - 1.6666654611e-1f);
}
RTC_PM_FN float pm_cos_poly(float r) {
  float z = r * r;
  return (1.0f - 0.5f * z) + z * z * ((2.443315711809948e-5f * z - 1.388731625493765e-3f) * z + 4.166664568298827e-2f);
}

RTC_PM_FN float pm_sinf(float x) {
  uint32_t sign = pm_to_bits(x) & 0x80000000u;
  float ax = pm_abs(x);
  if (!(ax < 1.0e9f))
    return ax - ax; /* inf, nan, or hopeless: 0 or NaN, same on both sides */
  float r;
  int j = pm_reduce_pio4(ax, &r);
  if (j > 3) {
    sign ^= 0x80000000u;
    j -= 4;
  }
  float y = (j == 2) ? pm_cos_poly(r) : pm_sin_poly(r);
  return pm_from_bits(pm_to_bits(y) ^ sign);
}

RTC_PM_FN float pm_cosf(float x) {
  float ax = pm_abs(x);
  if (!(ax < 1.0e9f))
    return ax - ax;
  float r;
  int j = pm_reduce_pio4(ax, &r);
  uint32_t sign = 0;
  if (j > 3) {
    sign = 0x80000000u;
    j -= 4;
  }
  if (j > 1)
    sign ^= 0x80000000u;
  float y = (j == 2) ? pm_sin_poly(r) : pm_cos_poly(r);
  return pm_from_bits(pm_to_bits(y) ^ sign);
}

/* natural log; log(0) = -inf, log(<0) = NaN, log(inf) = inf */
RTC_PM_FN float pm_logf(float x) {
  uint32_t bits = pm_to_bits(x);
  if ((bits << 1) == 0u)
    return pm_from_bits(0xff800000u);
  if (bits >= 0x7f800000u) /* negative, inf or nan */
    return (bits == 0x7f800000u) ? x : pm_from_bits(0x7fc00000u);
  int e = 0;
  if (bits < 0x00800000u) { /* subnormal: scale into the normal range first */
    x = x * 33554432.0f;    /* 2^25 */
    bits = pm_to_bits(x);
    e = -25;
  }
  e += (int)(bits >> 23) - 126;
  float m = pm_from_bits((bits & 0x007fffffu) | 0x3f000000u); /* mantissa in [0.5, 1) */
  if (m < 0.707106781186547524f) {
    e -= 1;
    m = (m + m) - 1.0f;
  } else {
    m = m - 1.0f;
  }
  float z = m * m;
  float y = ((((((((7.0376836292e-2f * m - 1.1514610310e-1f) * m + 1.1676998740e-1f) * m - 1.2420140846e-1f) * m +
                1.4249322787e-1f) * m - 1.6668057665e-1f) * m + 2.0000714765e-1f) * m - 2.4999993993e-1f) * m +
             3.3333331174e-1f) * m * z;
  float fe = (float)e;
  y = y + -2.12194440e-4f * fe;
  y = y - 0.5f * z;
  return (m + y) + 0.693359375f * fe;
}

/* x^5 by repeated multiplication: stands in for powf(x, 5.0f) in the Schlick term (material.c:73) */
RTC_PM_FN float pm_pow5f(float x) {
  float x2 = x * x;
  return (x2 * x2) * x;
}

/* atan on the whole line */
RTC_PM_FN float pm_atanf(float x) {
  uint32_t sign = pm_to_bits(x) & 0x80000000u;
  float ax = pm_abs(x);
  float base;
  if (ax > 2.414213562373095f) { /* tan(3pi/8) */
    base = PM_PIO2_F;
    ax = -(1.0f / ax);
  } else if (ax > 0.4142135623730950f) { /* tan(pi/8) */
    base = PM_PIO4_F;
    ax = (ax - 1.0f) / (ax + 1.0f);
  } else {
    base = 0.0f;
  }
  float z = ax * ax;
  float y = base + ((((8.05374449538e-2f * z - 1.38776856032e-1f) * z + 1.99777106478e-1f) * z - 3.33329491539e-1f) *
                        z * ax + ax);
  return pm_from_bits(pm_to_bits(y) ^ sign);
}

RTC_PM_FN float pm_atan2f(float y, float x) {
  if (x != x || y != y)
    return x + y;
  if (y == 0.0f) {
    int x_neg = (pm_to_bits(x) >> 31) != 0u;
    float r = x_neg ? PM_PI_F : 0.0f;
    return pm_from_bits(pm_to_bits(r) | (pm_to_bits(y) & 0x80000000u));
  }
  if (x == 0.0f)
    return y > 0.0f ? PM_PIO2_F : -PM_PIO2_F;
  float a = pm_atanf(y / x);
  if (x > 0.0f)
    return a;
  return y > 0.0f ? a + PM_PI_F : a - PM_PI_F;
}

RTC_PM_FN float pm_sqrtf(float x); /* provided per compiler below */

/* asin on [-1, 1]; arguments a hair outside (|x| <= 1 + 2^-20, from rounding of a unit normal) are clamped,
 * anything further out gives NaN like libm */
RTC_PM_FN float pm_asinf(float x) {
  uint32_t sign = pm_to_bits(x) & 0x80000000u;
  float ax = pm_abs(x);
  if (ax > 1.0f) {
    if (!(ax <= 1.00000095367431640625f))
      return pm_from_bits(0x7fc00000u);
    ax = 1.0f;
  }
  float z, w;
  int big = ax > 0.5f;
  if (big) {
    z = 0.5f * (1.0f - ax);
    w = pm_sqrtf(z);
  } else {
    w = ax;
    z = w * w;
  }
  float y = ((((4.2163199048e-2f * z + 2.4181311049e-2f) * z + 4.5470025998e-2f) * z + 7.4953002686e-2f) * z +
             1.6666752422e-1f) * z * w + w;
  if (big)
    y = PM_PIO2_F - (y + y);
  return pm_from_bits(pm_to_bits(y) ^ sign);
}

RTC_PM_FN float pm_acosf(float x) {
  if (x != x)
    return x;
  if (x < -0.5f)
    return PM_PI_F - 2.0f * pm_asinf(pm_sqrtf(0.5f * (1.0f + (x < -1.0f && x >= -1.00000095367431640625f ? -1.0f : x))));
  if (x > 0.5f)
    return 2.0f * pm_asinf(pm_sqrtf(0.5f * (1.0f - (x > 1.0f && x <= 1.00000095367431640625f ? 1.0f : x))));
  return PM_PIO2_F - pm_asinf(x);
}

/* IEEE-correct square root on both compilers */
#if defined(__CUDA_ARCH__)
RTC_PM_FN float pm_sqrtf(float x) { return __fsqrt_rn(x); }
#else
#include <math.h>
RTC_PM_FN float pm_sqrtf(float x) { return sqrtf(x); }
#endif

#endif /* RTC_PORTABLE_MATH_H */
