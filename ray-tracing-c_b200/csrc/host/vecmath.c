/*
 * vecmath.c -- host-side Vec3 / PCG32 / allocation helpers behind rtc_api.h.
 *
 * Only used while a scene is being BUILT (constructors, BVH build, Perlin tables, scene builders); the render
 * loop itself runs on the GPU.  The arithmetic keeps the operation order of the reference so that a scene built
 * here is bit-identical to one built by the reference with gcc -O2:
 *   PCG32          /root/reference/src/pcg32.c:3-22   (pcg-c-basic XSH-RR 64/32)
 *   Vec3 algebra   /root/reference/src/vec3.c:6-30    (sub = add(neg), div = mul by reciprocal)
 *   random vectors /root/reference/src/vec3.c:32-47   (gcc evaluates call arguments right-to-left, so the first
 *                                                      draw lands in z; written out explicitly here)
 */
#include "rtc_api.h"
#include <math.h>

void *my_malloc(size_t size) {
  void *p = malloc(size);
  assert(p != NULL && "Failed to allocate memory");
  return p;
}

/* ---- PCG32 ---- */

#define PCG32_MULT 6364136223846793005ULL

uint32_t pcg32_u32(PCG32 *rng) {
  uint64_t s = rng->state;
  rng->state = s * PCG32_MULT + rng->inc;
  uint32_t xs = (uint32_t)(((s >> 18u) ^ s) >> 27u);
  uint32_t rot = (uint32_t)(s >> 59u);
  return (xs >> rot) | (xs << ((0u - rot) & 31u));
}

void pcg32_seed(PCG32 *rng, uint64_t initstate, uint64_t initseq) {
  rng->state = 0u;
  rng->inc = (initseq << 1u) | 1u;
  (void)pcg32_u32(rng);
  rng->state += initstate;
  (void)pcg32_u32(rng);
}

/* Jump `delta` draws ahead in O(log delta) (not in the reference, whose pcg32.c:11-17 only steps by one):
 * the step state' = A * state + inc composed with itself delta times is state' = A^delta * state + inc * (A^delta - 1) /
 * (A - 1), accumulated by repeated squaring of the affine map (Brown, "Random number generation with arbitrary
 * strides").  After it the stream continues exactly as if pcg32_u32 had been called delta times. */
void pcg32_advance(PCG32 *rng, uint64_t delta) {
  uint64_t cur_mult = 6364136223846793005ULL, cur_plus = rng->inc, acc_mult = 1u, acc_plus = 0u;
  while (delta > 0) {
    if (delta & 1) {
      acc_mult *= cur_mult;
      acc_plus = acc_plus * cur_mult + cur_plus;
    }
    cur_plus = (cur_mult + 1) * cur_plus;
    cur_mult *= cur_mult;
    delta >>= 1;
  }
  rng->state = acc_mult * rng->state + acc_plus;
}
uint32_t pcg32_u32_between(PCG32 *rng, uint32_t lo, uint32_t hi) { return lo + pcg32_u32(rng) % (hi - lo); }
float pcg32_f32(PCG32 *rng) { return (float)(pcg32_u32(rng) >> 8) / 16777216.0f; }
float pcg32_f32_between(PCG32 *rng, float lo, float hi) { return lo + pcg32_f32(rng) * (hi - lo); }

/* ---- Vec3 ---- */

const Vec3 VEC3_ZERO = {{0.0f, 0.0f, 0.0f}};

Vec3 vec3(float x, float y, float z) {
  Vec3 r;
  r.x = x;
  r.y = y;
  r.z = z;
  return r;
}
Vec3 *Vec3_new(float x, float y, float z) {
  Vec3 *r = my_malloc(sizeof(Vec3));
  *r = vec3(x, y, z);
  return r;
}
Vec3 vec3_neg(Vec3 a) { return vec3(-a.x, -a.y, -a.z); }
Vec3 vec3_inv(Vec3 a) { return vec3(1.0f / a.x, 1.0f / a.y, 1.0f / a.z); }

Vec3 vec3_add_vec3(Vec3 a, Vec3 b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
Vec3 vec3_mul_vec3(Vec3 a, Vec3 b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
Vec3 vec3_sub_vec3(Vec3 a, Vec3 b) { return vec3_add_vec3(a, vec3_neg(b)); }
Vec3 vec3_div_vec3(Vec3 a, Vec3 b) { return vec3_mul_vec3(a, vec3_inv(b)); }

Vec3 vec3_add_float(Vec3 a, float s) { return vec3(a.x + s, a.y + s, a.z + s); }
Vec3 vec3_mul_float(Vec3 a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
Vec3 vec3_sub_float(Vec3 a, float s) { return vec3_add_float(a, -s); }
Vec3 vec3_div_float(Vec3 a, float s) { return vec3_mul_float(a, 1.0f / s); }

Vec3 vec3_lerp(Vec3 a, Vec3 b, float w) {
  return vec3_add_vec3(vec3_mul_float(a, 1.0f - w), vec3_mul_float(b, w));
}
Vec3 vec3_min(Vec3 a, Vec3 b) { return vec3(fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z)); }
Vec3 vec3_max(Vec3 a, Vec3 b) { return vec3(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z)); }

float vec3_dot(Vec3 a, Vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
float vec3_length2(Vec3 a) { return vec3_dot(a, a); }
float vec3_length(Vec3 a) { return sqrtf(vec3_length2(a)); }
Vec3 vec3_cross(Vec3 a, Vec3 b) {
  return vec3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
Vec3 vec3_normalize(Vec3 a) { return vec3_div_float(a, vec3_length(a)); }
bool vec3_near_zero(Vec3 a) { return fabsf(a.x) < 1e-8f && fabsf(a.y) < 1e-8f && fabsf(a.z) < 1e-8f; }

Vec3 vec3_rand(PCG32 *rng) {
  float z = pcg32_f32(rng);
  float y = pcg32_f32(rng);
  float x = pcg32_f32(rng);
  return vec3(x, y, z);
}
Vec3 vec3_rand_between(PCG32 *rng, float lo, float hi) {
  float z = pcg32_f32_between(rng, lo, hi);
  float y = pcg32_f32_between(rng, lo, hi);
  float x = pcg32_f32_between(rng, lo, hi);
  return vec3(x, y, z);
}
Vec3 vec3_rand_unit_vector(PCG32 *rng) {
  for (;;) {
    Vec3 c = vec3_rand_between(rng, -1.0f, 1.0f);
    float l2 = vec3_length2(c);
    if (l2 < 1.0f)
      return vec3_div_float(c, sqrtf(l2));
  }
}
Vec3 vec3_rand_hemisphere(Vec3 normal, PCG32 *rng) {
  Vec3 d = vec3_rand_unit_vector(rng);
  return vec3_dot(d, normal) > 0.0f ? d : vec3_neg(d);
}

Vec3 ray_at(const Ray *ray, float t) { return vec3_add_vec3(ray->origin, vec3_mul_float(ray->direction, t)); }
