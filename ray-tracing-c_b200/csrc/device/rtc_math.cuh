/*
 * rtc_math.cuh -- device-side Vec3 arithmetic and the PCG32 stream.
 *
 * Two flavours of every operation:
 *   * the plain operators/functions (`v_add`, `v_dot`, ...) follow the reference's operation ORDER
 *     (/root/reference/src/vec3.c:6-30) and are subject to whatever contraction the translation unit allows:
 *     none in rtc_exact.cu (compiled -fmad=false), FMA in rtc_mega.cu / rtc_wave.cu;
 *   * the `s_` ("strict") functions are built from __fmul_rn/__fadd_rn/__fdiv_rn/__fsqrt_rn, which nvcc never
 *     contracts, so ray/primitive intersection produces the reference's bits in EVERY translation unit.
 */
#ifndef RTC_MATH_CUH
#define RTC_MATH_CUH

#include <cuda_runtime.h>
#include <stdint.h>

#define RTC_DEV __device__ __forceinline__

#define RTC_PI 3.14159274101257324f     /* (float)M_PI   */
#define RTC_1_PI 0.318309873342514038f  /* (float)M_1_PI */
#define RTC_INF __int_as_float(0x7f800000)

struct V3 {
  float x, y, z;
};

RTC_DEV V3 v3(float x, float y, float z) {
  V3 r;
  r.x = x, r.y = y, r.z = z;
  return r;
}
RTC_DEV V3 v_load(const float *p) { return v3(p[0], p[1], p[2]); }
RTC_DEV V3 v_neg(V3 a) { return v3(-a.x, -a.y, -a.z); }
RTC_DEV V3 v_add(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
RTC_DEV V3 v_sub(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); } /* a + (-b) == a - b in IEEE */
RTC_DEV V3 v_mul(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
RTC_DEV V3 v_scale(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
RTC_DEV float v_dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
RTC_DEV float v_len2(V3 a) { return v_dot(a, a); }
RTC_DEV V3 v_cross(V3 a, V3 b) { return v3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y); }
/* Division and square root of the SHADING code.  IEEE-rounded by default (the exact kernel must match gcc's bits);
 * a translation unit that defines RTC_FAST_SHADING before including this header (rtc_mega.cu, rtc_wave.cu) gets the
 * two-instruction approximations instead -- its images are judged by RMSE.  The primitive tests never use these:
 * they go through the strict s_div / __fsqrt_rn below, so hit distances keep the reference's bits everywhere. */
#ifdef RTC_FAST_SHADING
RTC_DEV float f_div(float a, float b) { return __fdividef(a, b); }
RTC_DEV float f_sqrt(float x) {
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
#else
RTC_DEV float f_div(float a, float b) { return __fdiv_rn(a, b); }
RTC_DEV float f_sqrt(float x) { return __fsqrt_rn(x); }
#endif
/* vec3.c:19: division is reciprocal-then-multiply; vec3.c:29 normalises through it */
RTC_DEV V3 v_div(V3 a, float s) { return v_scale(a, f_div(1.0f, s)); }
RTC_DEV float v_len(V3 a) { return f_sqrt(v_len2(a)); }
RTC_DEV V3 v_unit(V3 a) { return v_div(a, v_len(a)); }

/* ---- strict (never contracted) ---- */
RTC_DEV float s_mul(float a, float b) { return __fmul_rn(a, b); }
RTC_DEV float s_add(float a, float b) { return __fadd_rn(a, b); }
RTC_DEV float s_sub(float a, float b) { return __fsub_rn(a, b); }
RTC_DEV float s_div(float a, float b) { return __fdiv_rn(a, b); }
RTC_DEV V3 sv_add(V3 a, V3 b) { return v3(s_add(a.x, b.x), s_add(a.y, b.y), s_add(a.z, b.z)); }
RTC_DEV V3 sv_sub(V3 a, V3 b) { return v3(s_sub(a.x, b.x), s_sub(a.y, b.y), s_sub(a.z, b.z)); }
RTC_DEV V3 sv_scale(V3 a, float s) { return v3(s_mul(a.x, s), s_mul(a.y, s), s_mul(a.z, s)); }
RTC_DEV float sv_dot(V3 a, V3 b) { return s_add(s_add(s_mul(a.x, b.x), s_mul(a.y, b.y)), s_mul(a.z, b.z)); }
RTC_DEV float sv_len2(V3 a) { return sv_dot(a, a); }
RTC_DEV V3 sv_cross(V3 a, V3 b) {
  return v3(s_sub(s_mul(a.y, b.z), s_mul(b.y, a.z)), s_sub(s_mul(a.z, b.x), s_mul(b.z, a.x)),
            s_sub(s_mul(a.x, b.y), s_mul(b.x, a.y)));
}
RTC_DEV V3 sv_div(V3 a, float s) { return sv_scale(a, s_div(1.0f, s)); }
RTC_DEV V3 sv_at(V3 o, V3 d, float t) { return sv_add(o, sv_scale(d, t)); } /* hittable.c:7 */
/* hittable.c:347-353 */
RTC_DEV V3 sv_rot_y(V3 u, float cs, float sn) {
  return v3(s_sub(s_mul(cs, u.x), s_mul(sn, u.z)), u.y, s_add(s_mul(sn, u.x), s_mul(cs, u.z)));
}
RTC_DEV V3 sv_rot_y_inv(V3 u, float cs, float sn) {
  return v3(s_add(s_mul(cs, u.x), s_mul(sn, u.z)), u.y, s_add(s_mul(-sn, u.x), s_mul(cs, u.z)));
}

/* ------------------------------------------------------------------ pcg32.c:3-22 ------------------------- */

struct Rng {
  uint64_t state, inc;
};

RTC_DEV uint32_t rng_u32(Rng &g) {
  uint64_t s = g.state;
  g.state = s * 6364136223846793005ULL + g.inc;
  uint32_t xs = (uint32_t)(((s >> 18u) ^ s) >> 27u);
  uint32_t rot = (uint32_t)(s >> 59u);
  return __funnelshift_r(xs, xs, rot); /* rotate right */
}
RTC_DEV void rng_seed(Rng &g, uint64_t initstate, uint64_t initseq) {
  g.state = 0;
  g.inc = (initseq << 1u) | 1u;
  rng_u32(g);
  g.state += initstate;
  rng_u32(g);
}
RTC_DEV float rng_f32(Rng &g) { return (float)(rng_u32(g) >> 8) * (1.0f / 16777216.0f); } /* exact: power of two */
RTC_DEV float rng_between(Rng &g, float lo, float hi) { return lo + rng_f32(g) * (hi - lo); }

/* The pixel / sample seeding contract (raytracing.c:93-94; RT_RNG_SAMPLE extends it, rtc_b200.h) */
RTC_DEV void rng_seed_pixel(Rng &g, int i, int j) { rng_seed(g, 17u + (uint64_t)j, 23u + (uint64_t)i); }
RTC_DEV void rng_seed_sample(Rng &g, int i, int j, int s) {
  rng_seed(g, ((uint64_t)(uint32_t)s << 32) | (uint64_t)(17 + j), 23u + (uint64_t)i);
}

#endif /* RTC_MATH_CUH */
