/*
 * rtc_oracle.c -- TEST INFRASTRUCTURE: a plain-C, CPU restatement of the reference's per-pixel render loop,
 * evaluated over the flattened scene (rt_scene_desc) that the GPU library consumes.
 *
 * Nothing in the product links, loads or calls this file.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs do, and only as the checker.
 *
 * Pinning (see DESIGN.md "Oracle"):
 *   * math = ORACLE_MATH_LIBM  uses glibc exactly where the reference does; built -O2 -ffp-contract=off it gives
 *     BYTE-IDENTICAL images and bit-identical HitRecords to the unmodified reference compiled with the same
 *     flags (oracle/_ref/ref_tool; tests/test_oracle_vs_reference.py; committed vectors in tests/golden/).
 *   * math = ORACLE_MATH_PORTABLE swaps the seven libm calls for rtc_portable_math.h; this is what the GPU's
 *     exact kernel must match bit for bit.
 *
 * Every function cites the reference lines it restates (paths relative to /root/reference/src).  The structure
 * is recursive like the reference's (the GPU's exact kernel is an explicit-stack machine instead, so the two
 * implementations are independent); operation order inside each formula is the reference's, because
 * float addition/multiplication are not associative and parity is judged on bits.
 */
#include "rtc_b200.h"
#include "../ray-tracing-c_b200/csrc/device/rtc_portable_math.h"
#include <math.h>
#include <stdbool.h>
#include <stdlib.h>
#include <string.h>

enum { ORACLE_MATH_LIBM = 0, ORACLE_MATH_PORTABLE = 1 };

/* ------------------------------------------------------------------ vec3.c:6-30 -------------------------- */

typedef struct V3 {
  float x, y, z;
} V3;

static inline V3 v3(float x, float y, float z) { return (V3){x, y, z}; }
static inline V3 v3_load(const float *p) { return (V3){p[0], p[1], p[2]}; }
static inline V3 v3_neg(V3 a) { return v3(-a.x, -a.y, -a.z); }
static inline V3 v3_add(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline V3 v3_sub(V3 a, V3 b) { return v3_add(a, v3_neg(b)); } /* vec3.c:13: add(u, neg(v)) */
static inline V3 v3_mul(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline V3 v3_scale(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
static inline V3 v3_div(V3 a, float s) { return v3_scale(a, 1.0f / s); } /* vec3.c:19: reciprocal, then multiply */
static inline float v3_dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline float v3_len2(V3 a) { return v3_dot(a, a); }
static inline float v3_len(V3 a) { return sqrtf(v3_len2(a)); }
static inline V3 v3_cross(V3 a, V3 b) {
  return v3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
static inline V3 v3_unit(V3 a) { return v3_div(a, v3_len(a)); }

/* ------------------------------------------------------------------ pcg32.c:3-22 ------------------------- */

typedef struct Rng {
  uint64_t state, inc;
} Rng;

static inline uint32_t rng_u32(Rng *g) {
  uint64_t s = g->state;
  g->state = s * 6364136223846793005ULL + g->inc;
  uint32_t xs = (uint32_t)(((s >> 18u) ^ s) >> 27u);
  uint32_t rot = (uint32_t)(s >> 59u);
  return (xs >> rot) | (xs << ((0u - rot) & 31u));
}
static inline void rng_seed(Rng *g, uint64_t initstate, uint64_t initseq) {
  g->state = 0;
  g->inc = (initseq << 1u) | 1u;
  rng_u32(g);
  g->state += initstate;
  rng_u32(g);
}
static inline float rng_f32(Rng *g) { return (float)(rng_u32(g) >> 8) / 16777216.0f; }
static inline float rng_between(Rng *g, float lo, float hi) { return lo + rng_f32(g) * (hi - lo); }

/* vec3.c:33-43 under gcc: the three coordinates are drawn z, y, x */
static V3 rng_unit_vector(Rng *g) {
  for (;;) {
    float z = rng_between(g, -1.0f, 1.0f);
    float y = rng_between(g, -1.0f, 1.0f);
    float x = rng_between(g, -1.0f, 1.0f);
    V3 c = v3(x, y, z);
    float l2 = v3_len2(c);
    if (l2 < 1.0f)
      return v3_div(c, sqrtf(l2));
  }
}

/* ------------------------------------------------------------------ context ------------------------------ */

typedef struct Ctx {
  const rt_scene_desc *d;
  int math;
  uint64_t segments, light_pdf_evals;
  uint64_t work[5]; /* visits of the reference's own traversal: AABB, sphere, quad tests, medium evals, transforms */
} Ctx;
enum { W_AABB = 0, W_SPHERE = 1, W_QUAD = 2, W_MEDIUM = 3, W_XFORM = 4 };

static inline float m_sin(const Ctx *c, float x) { return c->math == ORACLE_MATH_LIBM ? sinf(x) : pm_sinf(x); }
static inline float m_cos(const Ctx *c, float x) { return c->math == ORACLE_MATH_LIBM ? cosf(x) : pm_cosf(x); }
static inline float m_log(const Ctx *c, float x) { return c->math == ORACLE_MATH_LIBM ? logf(x) : pm_logf(x); }
static inline float m_pow5(const Ctx *c, float x) { return c->math == ORACLE_MATH_LIBM ? powf(x, 5.0f) : pm_pow5f(x); }
static inline float m_atan2(const Ctx *c, float y, float x) {
  return c->math == ORACLE_MATH_LIBM ? atan2f(y, x) : pm_atan2f(y, x);
}
static inline float m_acos(const Ctx *c, float x) { return c->math == ORACLE_MATH_LIBM ? acosf(x) : pm_acosf(x); }

typedef struct Ray {
  V3 o, d;
} Ray;

/* material.h:10-18 plus the id of the node that produced the hit */
typedef struct Hit {
  V3 p, n;
  uint32_t material;
  float t, u, v;
  int front;
  uint32_t node;
} Hit;

static inline V3 ray_at(const Ray *r, float t) { return v3_add(r->o, v3_scale(r->d, t)); } /* hittable.c:7 */

/* ------------------------------------------------------------------ hittable.c:38-55 --------------------- */

static bool aabb_hit(Ctx *c, const float bbox[6], const Ray *r, float t_min, float t_max) {
  c->work[W_AABB]++;
  const float o[3] = {r->o.x, r->o.y, r->o.z}, d[3] = {r->d.x, r->d.y, r->d.z};
  for (int a = 0; a < 3; a++) {
    float inv = 1.0f / d[a];
    float t0 = (bbox[2 * a] - o[a]) * inv;
    float t1 = (bbox[2 * a + 1] - o[a]) * inv;
    if (inv < 0) {
      float s = t0;
      t0 = t1;
      t1 = s;
    }
    t_min = fmaxf(t_min, t0);
    t_max = fminf(t_max, t1);
    if (t_max <= t_min)
      return false;
  }
  return true;
}

/* ------------------------------------------------------------------ hittable.c:120-151 ------------------- */

static bool sphere_hit(Ctx *c, const rt_sphere *s, const Ray *r, float t_min, float t_max, Hit *rec) {
  c->work[W_SPHERE]++;
  V3 center = v3_load(s->center);
  V3 oc = v3_sub(r->o, center);
  float a = v3_len2(r->d);
  float b = v3_dot(oc, r->d);
  float cc = v3_len2(oc) - s->radius * s->radius;
  float disc = b * b - a * cc;
  if (disc < 0)
    return false;
  float sq = sqrtf(disc);
  float root = (-b - sq) / a;
  if (root <= t_min || root >= t_max) {
    root = (-b + sq) / a;
    if (root <= t_min || root >= t_max)
      return false;
  }
  rec->t = root;
  rec->p = ray_at(r, root);
  V3 outward = v3_div(v3_sub(rec->p, center), s->radius);
  rec->front = v3_dot(r->d, outward) < 0.0f;
  rec->n = rec->front ? outward : v3_neg(outward);
  rec->u = (m_atan2(c, -outward.z, outward.x) + (float)M_PI) * (float)M_1_PI * 0.5f;
  rec->v = m_acos(c, -outward.y) * (float)M_1_PI;
  rec->material = s->material;
  rec->node = s->node;
  return true;
}

/* ------------------------------------------------------------------ hittable.c:186-214 ------------------- */

static bool quad_hit(Ctx *c, const rt_quad *q, const Ray *r, float t_min, float t_max, Hit *rec) {
  c->work[W_QUAD]++;
  V3 n = v3_load(q->normal);
  float denom = v3_dot(n, r->d);
  if (fabsf(denom) < 1e-8f)
    return false;
  float t = (q->D - v3_dot(n, r->o)) / denom;
  if (t < t_min || t > t_max) /* closed interval: an equal-t later quad replaces the earlier hit */
    return false;
  V3 p = ray_at(r, t);
  V3 hp = v3_sub(p, v3_load(q->Q));
  V3 w = v3_load(q->w);
  float alpha = v3_dot(w, v3_cross(hp, v3_load(q->v)));
  float beta = v3_dot(w, v3_cross(v3_load(q->u), hp));
  if (alpha < 0 || alpha > 1 || beta < 0 || beta > 1)
    return false;
  rec->u = alpha;
  rec->v = beta;
  rec->t = t;
  rec->p = p;
  rec->material = q->material;
  rec->front = v3_dot(r->d, n) < 0.0f;
  rec->n = rec->front ? n : v3_neg(n);
  rec->node = q->node;
  return true;
}

/* ------------------------------------------------------------------ the Hittable graph ------------------- */

static bool node_hit(Ctx *c, uint32_t id, const Ray *r, float t_min, float t_max, Hit *rec, Rng *g);

/* hittable.c:74-88: items in order, t_max shrinks to the latest hit */
static bool list_hit(Ctx *c, const rt_node *n, const Ray *r, float t_min, float t_max, Hit *rec, Rng *g) {
  bool any = false;
  for (uint32_t i = 0; i < n->b; i++)
    if (node_hit(c, c->d->children[n->a + i], r, t_min, t_max, rec, g)) {
      t_max = rec->t;
      any = true;
    }
  return any;
}

/* hittable.c:266-277: own box, then left ALWAYS before right, right sees the shrunk t_max */
static bool bvh_hit(Ctx *c, const rt_node *n, const Ray *r, float t_min, float t_max, Hit *rec, Rng *g) {
  if (!aabb_hit(c, n->bbox, r, t_min, t_max))
    return false;
  bool left = node_hit(c, n->a, r, t_min, t_max, rec, g);
  if (left)
    t_max = rec->t;
  bool right = node_hit(c, n->b, r, t_min, t_max, rec, g);
  return left || right;
}

/* hittable.c:325-334: no box test of its own; only p is moved back */
static bool translate_hit(Ctx *c, const rt_node *n, const Ray *r, float t_min, float t_max, Hit *rec, Rng *g) {
  c->work[W_XFORM]++;
  V3 offset = v3_load(n->f);
  Ray local = {v3_sub(r->o, offset), r->d};
  if (!node_hit(c, n->a, &local, t_min, t_max, rec, g))
    return false;
  rec->p = v3_add(rec->p, offset);
  return true;
}

/* hittable.c:347-367 */
static inline V3 rot_y(V3 u, float cs, float sn) { return v3(cs * u.x - sn * u.z, u.y, sn * u.x + cs * u.z); }
static inline V3 rot_y_inv(V3 u, float cs, float sn) { return v3(cs * u.x + sn * u.z, u.y, -sn * u.x + cs * u.z); }
static bool rotate_y_hit(Ctx *c, const rt_node *n, const Ray *r, float t_min, float t_max, Hit *rec, Rng *g) {
  c->work[W_XFORM]++;
  float sn = n->f[0], cs = n->f[1];
  Ray local = {rot_y(r->o, cs, sn), rot_y(r->d, cs, sn)};
  if (!node_hit(c, n->a, &local, t_min, t_max, rec, g))
    return false;
  rec->p = rot_y_inv(rec->p, cs, sn);
  rec->n = rot_y_inv(rec->n, cs, sn);
  return true;
}

/* hittable.c:392-423: two boundary queries, one random draw only if the clipped interval is non-empty;
 * normal / front / u / v of `rec` are left as earlier candidates wrote them */
static bool medium_hit(Ctx *c, uint32_t id, const rt_node *n, const Ray *r, float t_min, float t_max, Hit *rec,
                       Rng *g) {
  c->work[W_MEDIUM]++;
  Hit in, out;
  memset(&in, 0, sizeof(in));
  memset(&out, 0, sizeof(out));
  if (!node_hit(c, n->a, r, -INFINITY, INFINITY, &in, g))
    return false;
  if (!node_hit(c, n->a, r, in.t + 0.0001f, INFINITY, &out, g))
    return false;
  in.t = fmaxf(in.t, t_min);
  out.t = fminf(out.t, t_max);
  if (in.t >= out.t)
    return false;
  in.t = in.t > 0.0f ? in.t : 0.0f; /* the `max` macro of utils.h:15 */
  float ray_len = v3_len(r->d);
  float inside = (out.t - in.t) * ray_len;
  float travel = n->f[0] * m_log(c, rng_f32(g));
  if (travel > inside)
    return false;
  rec->t = in.t + travel / ray_len;
  rec->p = ray_at(r, rec->t);
  rec->material = n->b;
  rec->node = id;
  return true;
}

static bool node_hit(Ctx *c, uint32_t id, const Ray *r, float t_min, float t_max, Hit *rec, Rng *g) {
  const rt_node *n = &c->d->nodes[id];
  switch (n->kind) {
  case RT_NODE_LIST: return list_hit(c, n, r, t_min, t_max, rec, g);
  case RT_NODE_SPHERE: return sphere_hit(c, &c->d->spheres[n->a], r, t_min, t_max, rec);
  case RT_NODE_QUAD: return quad_hit(c, &c->d->quads[n->a], r, t_min, t_max, rec);
  case RT_NODE_BVH: return bvh_hit(c, n, r, t_min, t_max, rec, g);
  case RT_NODE_TRANSLATE: return translate_hit(c, n, r, t_min, t_max, rec, g);
  case RT_NODE_ROTATE_Y: return rotate_y_hit(c, n, r, t_min, t_max, rec, g);
  case RT_NODE_MEDIUM: return medium_hit(c, id, n, r, t_min, t_max, rec, g);
  default: return false;
  }
}

/* ------------------------------------------------------------------ material.c:144-152 ------------------- */

typedef struct Basis {
  V3 u, v, w;
} Basis;
static Basis basis_from_w(V3 w) {
  Basis b;
  b.w = v3_unit(w);
  V3 a = fabsf(b.w.x) > 0.9f ? v3(0, 1, 0) : v3(1, 0, 0);
  b.v = v3_unit(v3_cross(b.w, a));
  b.u = v3_cross(b.w, b.v);
  return b;
}
static V3 basis_local(const Basis *b, V3 a) {
  return v3_add(v3_add(v3_scale(b->u, a.x), v3_scale(b->v, a.y)), v3_scale(b->w, a.z));
}

/* ------------------------------------------------------------------ light sampling: hittable.c:89-107,152-178,215-228 */

static float node_pdf(Ctx *c, uint32_t id, const Ray *r, Rng *g);
static V3 node_rand(Ctx *c, uint32_t id, V3 origin, Rng *g);

static float sphere_pdf(Ctx *c, const rt_sphere *s, const Ray *r) {
  Hit rec;
  if (!sphere_hit(c, s, r, 0.001f, INFINITY, &rec))
    return 0.0f;
  V3 oc = v3_sub(v3_load(s->center), r->o);
  float cos_max = sqrtf(1.0f - s->radius * s->radius / v3_len2(oc));
  float solid_angle = 2.0f * (float)M_PI * (1.0f - cos_max);
  return 1.0f / solid_angle;
}
static V3 sphere_rand(Ctx *c, const rt_sphere *s, V3 origin, Rng *g) {
  V3 oc = v3_sub(v3_load(s->center), origin);
  float r1 = rng_f32(g);
  float r2 = rng_f32(g);
  float z = 1.0f + r2 * (sqrtf(1.0f - s->radius * s->radius / v3_len2(oc)) - 1);
  float phi = 2.0f * (float)M_PI * r1;
  float x = m_cos(c, phi) * sqrtf(1.0f - z * z);
  float y = m_sin(c, phi) * sqrtf(1.0f - z * z);
  Basis b = basis_from_w(oc);
  return basis_local(&b, v3(x, y, z));
}
static float quad_pdf(Ctx *c, const rt_quad *q, const Ray *r) {
  Hit rec;
  if (!quad_hit(c, q, r, 0.001f, INFINITY, &rec))
    return 0.0f;
  float d2 = rec.t * rec.t * v3_len2(r->d);
  float cos_theta = fabsf(v3_dot(rec.n, v3_unit(r->d)));
  return d2 / (cos_theta * q->area);
}
static V3 quad_rand(const rt_quad *q, V3 origin, Rng *g) {
  float along_v = rng_f32(g); /* gcc evaluates the v term's argument first (hittable.c:227) */
  float along_u = rng_f32(g);
  V3 p = v3_add(v3_add(v3_load(q->Q), v3_scale(v3_load(q->u), along_u)), v3_scale(v3_load(q->v), along_v));
  return v3_add(p, v3_neg(origin));
}
static float node_pdf(Ctx *c, uint32_t id, const Ray *r, Rng *g) {
  const rt_node *n = &c->d->nodes[id];
  switch (n->kind) {
  case RT_NODE_SPHERE: return sphere_pdf(c, &c->d->spheres[n->a], r);
  case RT_NODE_QUAD: return quad_pdf(c, &c->d->quads[n->a], r);
  case RT_NODE_LIST: {
    float sum = 0.0f, count = 0.0f;
    for (uint32_t i = 0; i < n->b; i++) {
      uint32_t item = c->d->children[n->a + i];
      if (c->d->nodes[item].flags & RT_NODE_HAS_PDF) {
        sum += node_pdf(c, item, r, g);
        count += 1.0f;
      }
    }
    return sum / fmaxf(count, 1.0f);
  }
  default: return 0.0f;
  }
}
static V3 node_rand(Ctx *c, uint32_t id, V3 origin, Rng *g) {
  const rt_node *n = &c->d->nodes[id];
  switch (n->kind) {
  case RT_NODE_SPHERE: return sphere_rand(c, &c->d->spheres[n->a], origin, g);
  case RT_NODE_QUAD: return quad_rand(&c->d->quads[n->a], origin, g);
  case RT_NODE_LIST:
    for (;;) { /* rejection over items without `rand` */
      uint32_t item = c->d->children[n->a + rng_u32(g) % n->b];
      if (c->d->nodes[item].flags & RT_NODE_HAS_PDF)
        return node_rand(c, item, origin, g);
    }
  default: return v3(0, 0, 0);
  }
}

/* ------------------------------------------------------------------ textures: texture.c:8-114 ------------ */

static float perlin_noise(const rt_perlin *pl, V3 p) { /* texture.c:78-103 */
  int i = (int)floorf(p.x), j = (int)floorf(p.y), k = (int)floorf(p.z);
  float t1 = p.x - (float)i, t2 = p.y - (float)j, t3 = p.z - (float)k;
  float tt1 = t1 * t1 * (3.0f - 2.0f * t1);
  float tt2 = t2 * t2 * (3.0f - 2.0f * t2);
  float tt3 = t3 * t3 * (3.0f - 2.0f * t3);
  float value = 0;
  for (int di = 0; di < 2; di++)
    for (int dj = 0; dj < 2; dj++)
      for (int dk = 0; dk < 2; dk++) {
        int cell = pl->perm_x[(i + di) & 255] ^ pl->perm_y[(j + dj) & 255] ^ pl->perm_z[(k + dk) & 255];
        V3 grad = v3_load(&pl->grad[3 * cell]);
        V3 weight = v3(t1 - di, t2 - dj, t3 - dk);
        value += v3_dot(grad, weight) * (di * tt1 + (1 - di) * (1.0f - tt1)) * (dj * tt2 + (1 - dj) * (1.0f - tt2)) *
                 (dk * tt3 + (1 - dk) * (1.0f - tt3));
      }
  return value;
}
static float perlin_turbulence(const rt_perlin *pl, int depth, V3 p) { /* texture.c:105-114 */
  float value = 0.0f, weight = 1.0f;
  for (int i = 0; i < depth; i++) {
    value += weight * perlin_noise(pl, p);
    weight *= 0.5f;
    p = v3_scale(p, 2.0f);
  }
  return fabsf(value);
}

static V3 texture_value(const Ctx *c, uint32_t id, float u, float v, V3 p) {
  for (;;) {
    const rt_texture *t = &c->d->textures[id];
    switch (t->kind) {
    case RT_TEX_SOLID: return v3_load(t->color);
    case RT_TEX_CHECKER: { /* texture.c:12-22: on (u,v); C's % keeps the sign, so -1 selects `odd` too */
      int iu = (int)floorf(u / t->scale), iv = (int)floorf(v / t->scale);
      id = ((iu + iv) % 2) ? t->b : t->a;
      continue;
    }
    case RT_TEX_IMAGE: { /* texture.c:28-37, nearest neighbour; indices clamped where the reference has UB */
      const rt_image *im = &c->d->images[t->a];
      float fi = roundf(u * (float)(im->width - 1));
      float fj = roundf((1.0f - v) * (float)(im->height - 1));
      int i = (fi >= 0.0f) ? (fi <= (float)(im->width - 1) ? (int)fi : im->width - 1) : 0;   /* NaN -> 0 */
      int j = (fj >= 0.0f) ? (fj <= (float)(im->height - 1) ? (int)fj : im->height - 1) : 0;
      const uint8_t *px = c->d->texels + im->offset + ((size_t)j * im->width + i) * 3;
      return v3((float)px[0] / 255.0f, (float)px[1] / 255.0f, (float)px[2] / 255.0f);
    }
    case RT_TEX_PERLIN: { /* texture.c:47-52 */
      V3 q = v3_scale(p, t->scale);
      float s = 0.5f * (1.0f + m_sin(c, q.z + 10.0f * perlin_turbulence(&c->d->perlins[t->a], (int)t->b, q)));
      return v3(s, s, s);
    }
    default: return v3(0, 0, 0);
    }
  }
}

/* ------------------------------------------------------------------ materials: material.c:23-142 --------- */

static V3 reflect(V3 incident, V3 normal) { /* material.c:45-47 */
  return v3_sub(incident, v3_scale(normal, 2.0f * v3_dot(incident, normal)));
}

/* material.c:103-120.  Returns false when the material does not scatter. */
static bool material_scatter(const Ctx *c, const Hit *rec, V3 r_in, V3 *r_out, V3 *albedo, bool *skip_pdf, Rng *g) {
  const rt_material *m = &c->d->materials[rec->material];
  switch (m->tag) {
  case RT_MAT_LAMBERTIAN: { /* material.c:23-37 */
    Basis b = basis_from_w(rec->n);
    float r1 = rng_f32(g);
    float r2 = rng_f32(g);
    float phi = 2.0f * (float)M_PI * r1;
    V3 local = v3(m_cos(c, phi) * sqrtf(r2), m_sin(c, phi) * sqrtf(r2), sqrtf(1.0f - r2));
    *r_out = basis_local(&b, local);
    *albedo = texture_value(c, m->texture, rec->u, rec->v, rec->p);
    *skip_pdf = false;
    return true;
  }
  case RT_MAT_METAL: { /* material.c:48-58: three draws per try even when fuzz == 0 */
    V3 mirrored = reflect(v3_unit(r_in), rec->n);
    *r_out = v3_add(mirrored, v3_scale(rng_unit_vector(g), m->param));
    *albedo = texture_value(c, m->texture, rec->u, rec->v, rec->p);
    *skip_pdf = true;
    if (v3_dot(*r_out, rec->n) < 0.0f)
      *r_out = mirrored;
    return true;
  }
  case RT_MAT_DIELECTRIC: { /* material.c:62-86 */
    float eta = m->param;
    if (rec->front)
      eta = 1.0f / eta;
    V3 in = v3_unit(r_in);
    float cos_theta = fminf(-v3_dot(in, rec->n), 1.0f);
    float sin_theta = sqrtf(1.0f - cos_theta * cos_theta);
    float schlick = (1.0f - eta) / (1.0f + eta);
    schlick *= schlick;
    schlick += (1 - schlick) * m_pow5(c, 1.0f - cos_theta);
    if (eta * sin_theta > 1.0f || schlick > rng_f32(g)) { /* the draw is skipped on total internal reflection */
      *r_out = reflect(in, rec->n);
    } else {
      V3 perp = v3_scale(v3_add(in, v3_scale(rec->n, cos_theta)), eta);
      V3 para = v3_scale(rec->n, -sqrtf(fabsf(1.0f - v3_len2(perp))));
      *r_out = v3_add(perp, para);
    }
    *albedo = v3(1, 1, 1);
    *skip_pdf = true;
    return true;
  }
  case RT_MAT_ISOTROPIC: /* material.c:93-98 */
    *r_out = rng_unit_vector(g);
    *albedo = texture_value(c, m->texture, rec->u, rec->v, rec->p);
    *skip_pdf = false;
    return true;
  default: /* SURFACE_NORMAL, DIFFUSE_LIGHT */
    *skip_pdf = true;
    return false;
  }
}

static float material_scatter_pdf(const Ctx *c, uint32_t material, V3 normal, V3 r_out) { /* material.c:122-131 */
  switch (c->d->materials[material].tag) {
  case RT_MAT_LAMBERTIAN: {
    float cos_theta = v3_dot(normal, v3_unit(r_out));
    return cos_theta < 0.0f ? 0.0f : cos_theta / (float)M_PI;
  }
  case RT_MAT_ISOTROPIC: return 1.0f / (4.0f * (float)M_PI);
  default: return 0.0f;
  }
}

static V3 material_emit(const Ctx *c, const Hit *rec) { /* material.c:133-142 */
  const rt_material *m = &c->d->materials[rec->material];
  switch (m->tag) {
  case RT_MAT_SURFACE_NORMAL: return v3_scale(v3(rec->n.x + 1.0f, rec->n.y + 1.0f, rec->n.z + 1.0f), 0.5f);
  case RT_MAT_DIFFUSE_LIGHT:
    return rec->front ? texture_value(c, m->texture, rec->u, rec->v, rec->p) : v3(0, 0, 0);
  default: return v3(0, 0, 0);
  }
}

/* ------------------------------------------------------------------ integrator: raytracing.c:39-84 ------- */

static V3 ray_color(Ctx *c, const rt_camera *cam, const Ray *ray, int depth, Rng *g) {
  if (depth <= 0)
    return v3(0, 0, 0);
  c->segments++;
  Hit rec;
  memset(&rec, 0, sizeof(rec));
  if (!node_hit(c, c->d->objects_root, ray, 1e-3f, INFINITY, &rec, g))
    return v3_load(cam->background);

  Ray out = {rec.p, v3(0, 0, 0)};
  V3 albedo;
  bool skip_pdf;
  V3 emitted = material_emit(c, &rec);
  if (!material_scatter(c, &rec, ray->d, &out.d, &albedo, &skip_pdf, g))
    return emitted;

  float prob = cam->lights_sampling_prob;
  if (skip_pdf || prob == 0.0f || c->d->lights_root == RT_NONE)
    return v3_add(emitted, v3_mul(albedo, ray_color(c, cam, &out, depth - 1, g)));

  /* Book-3 mixture: half of the time aim at a light, weight by scatter_pdf / mixture_pdf */
  if (rng_f32(g) < prob)
    out.d = node_rand(c, c->d->lights_root, rec.p, g);
  float scatter_pdf = material_scatter_pdf(c, rec.material, rec.n, out.d);
  c->light_pdf_evals++;
  float sampling_pdf = (1.0f - prob) * scatter_pdf + prob * node_pdf(c, c->d->lights_root, &out, g);
  V3 incoming = ray_color(c, cam, &out, depth - 1, g);
  return v3_add(emitted, v3_scale(v3_mul(albedo, incoming), scatter_pdf / sampling_pdf));
}

/* ------------------------------------------------------------------ render: raytracing.c:86-135 ---------- */

static inline int has_nan(V3 a) { return a.x != a.x || a.y != a.y || a.z != a.z; }

static inline uint8_t tonemap(float sum, int spp) { /* raytracing.c:127-131 with utils.h's min/max/clamp macros */
  float v = sqrtf(sum / spp);
  v = v > 0.0f ? v : 0.0f; /* NaN -> 0 */
  v = v < 0.999f ? v : 0.999f;
  return (uint8_t)(int)(256.0f * v);
}

/* Visits of the last oracle_render: {AABB tests, sphere tests, quad tests, medium evaluations, instance transforms,
 * segments}.  Light-pdf self-hits are included in the sphere/quad counts, as in SURVEY.md's probe.  Basis of the
 * algorithmic-FLOP figure bench.py quotes the roofline fraction against. */
static uint64_t g_last_work[6];
void oracle_last_work(uint64_t out[6]) { memcpy(out, g_last_work, sizeof(g_last_work)); }

/* Exported.  radiance_sum: float[h][w][3] (may be NULL), rgb8: uint8[h][w][3] (may be NULL).
 * opts->kernel is ignored; opts->reserved[0] carries the math mode (ORACLE_MATH_*). */
int oracle_render(const rt_scene_desc *desc, const rt_camera *cam, const rt_opts *opts, int math, float *radiance_sum,
                  uint8_t *rgb8, rt_stats *stats) {
  const int w = cam->img_width, h = cam->img_height;
  const int s_begin = opts->sample_begin;
  const int s_count = opts->sample_count > 0 ? opts->sample_count : cam->samples_per_pixel - s_begin;
  if (opts->rng_mode == RT_RNG_PIXEL && s_begin != 0)
    return RT_ERR_INVALID; /* one stream per pixel cannot start in the middle */

  uint64_t segments = 0, pdf_evals = 0, nan_samples = 0;
  uint64_t w_aabb = 0, w_sphere = 0, w_quad = 0, w_medium = 0, w_xform = 0;
  const V3 px00 = v3_load(cam->pixel00), du = v3_load(cam->pixel_delta_u), dv = v3_load(cam->pixel_delta_v);
  const V3 eye = v3_load(cam->origin), lens_u = v3_load(cam->dof_disc_u), lens_v = v3_load(cam->dof_disc_v);

#pragma omp parallel for schedule(dynamic, 1) reduction(+ : segments, pdf_evals, nan_samples, w_aabb, w_sphere, w_quad, w_medium, w_xform)
  for (int j = 0; j < h; j++) {
    Ctx ctx = {desc, math, 0, 0, {0, 0, 0, 0, 0}};
    for (int i = 0; i < w; i++) {
      Rng rng;
      rng_seed(&rng, 17 + (uint64_t)j, 23 + (uint64_t)i);
      V3 pixel = v3_add(v3_add(px00, v3_scale(du, (float)i)), v3_scale(dv, (float)j));
      V3 sum = v3(0, 0, 0);
      for (int s = s_begin; s < s_begin + s_count; s++) {
        if (opts->rng_mode == RT_RNG_SAMPLE)
          rng_seed(&rng, ((uint64_t)s << 32) | (uint64_t)(17 + j), 23 + (uint64_t)i);
        float jx = rng_between(&rng, -0.5f, 0.5f);
        float jy = rng_between(&rng, -0.5f, 0.5f);
        Ray ray;
        if (cam->dof_angle > 0.0f) { /* thin lens: rejection-sample the unit disc */
          float a, b;
          for (;;) {
            a = rng_between(&rng, -1.0f, 1.0f);
            b = rng_between(&rng, -1.0f, 1.0f);
            if (a * a + b * b < 1.0f)
              break;
          }
          ray.o = v3_add(v3_add(eye, v3_scale(lens_u, a)), v3_scale(lens_v, b));
        } else {
          ray.o = eye;
        }
        ray.d = v3_add(v3_add(v3_add(pixel, v3_scale(du, jx)), v3_scale(dv, jy)), v3_neg(ray.o));
        V3 L = ray_color(&ctx, cam, &ray, cam->max_depth, &rng);
        if (has_nan(L)) {
          nan_samples++;
          if (opts->flags & RT_FLAG_DROP_NAN)
            continue;
        }
        sum = v3_add(sum, L);
      }
      size_t at = ((size_t)j * w + i) * 3;
      if (radiance_sum) {
        radiance_sum[at] = sum.x;
        radiance_sum[at + 1] = sum.y;
        radiance_sum[at + 2] = sum.z;
      }
      if (rgb8) {
        rgb8[at] = tonemap(sum.x, s_count);
        rgb8[at + 1] = tonemap(sum.y, s_count);
        rgb8[at + 2] = tonemap(sum.z, s_count);
      }
    }
    segments += ctx.segments;
    pdf_evals += ctx.light_pdf_evals;
    w_aabb += ctx.work[W_AABB], w_sphere += ctx.work[W_SPHERE], w_quad += ctx.work[W_QUAD];
    w_medium += ctx.work[W_MEDIUM], w_xform += ctx.work[W_XFORM];
  }
  g_last_work[0] = w_aabb, g_last_work[1] = w_sphere, g_last_work[2] = w_quad, g_last_work[3] = w_medium;
  g_last_work[4] = w_xform, g_last_work[5] = segments;
  if (stats) {
    memset(stats, 0, sizeof(*stats));
    stats->segments = segments;
    stats->light_pdf_evals = pdf_evals;
    stats->samples = (uint64_t)w * h * s_count;
    stats->nan_samples = nan_samples;
  }
  return RT_OK;
}

/* Exported: the tonemap alone (for the multi-rank host-logic tests). */
void oracle_tonemap(const float *radiance_sum, size_t n_values, int spp, uint8_t *rgb8) {
  for (size_t i = 0; i < n_values; i++)
    rgb8[i] = tonemap(radiance_sum[i], spp);
}

/* Exported: same contract as rt_hit_batch (include/rtc_b200.h). */
int oracle_hit_batch(const rt_scene_desc *desc, int math, size_t n, const float *origins, const float *dirs,
                     float t_min, float t_max, uint64_t *rng_io, float *out_t, uint32_t *out_node, float *out_p,
                     float *out_normal, float *out_uv, uint8_t *out_front, uint32_t *out_material) {
#pragma omp parallel for schedule(static)
  for (size_t i = 0; i < n; i++) {
    Ctx ctx = {desc, math, 0, 0, {0, 0, 0, 0, 0}};
    Ray r = {v3_load(origins + 3 * i), v3_load(dirs + 3 * i)};
    Rng g = {0, 1};
    if (rng_io) {
      g.state = rng_io[2 * i];
      g.inc = rng_io[2 * i + 1];
    }
    Hit rec;
    memset(&rec, 0, sizeof(rec));
    rec.material = RT_NONE;
    rec.node = RT_NONE;
    bool hit = node_hit(&ctx, desc->objects_root, &r, t_min, t_max, &rec, &g);
    if (rng_io)
      rng_io[2 * i] = g.state;
    out_t[i] = hit ? rec.t : INFINITY;
    out_node[i] = hit ? rec.node : RT_NONE;
    if (out_p)
      out_p[3 * i] = rec.p.x, out_p[3 * i + 1] = rec.p.y, out_p[3 * i + 2] = rec.p.z;
    if (out_normal)
      out_normal[3 * i] = rec.n.x, out_normal[3 * i + 1] = rec.n.y, out_normal[3 * i + 2] = rec.n.z;
    if (out_uv)
      out_uv[2 * i] = rec.u, out_uv[2 * i + 1] = rec.v;
    if (out_front)
      out_front[i] = (uint8_t)rec.front;
    if (out_material)
      out_material[i] = hit ? rec.material : RT_NONE;
  }
  return RT_OK;
}

/* Exported: PCG32 known-answer helper (pcg32.c): seeds (initstate, initseq) and writes n outputs. */
void oracle_pcg32(uint64_t initstate, uint64_t initseq, size_t n, uint32_t *out) {
  Rng g;
  rng_seed(&g, initstate, initseq);
  for (size_t i = 0; i < n; i++)
    out[i] = rng_u32(&g);
}

/* Exported: the portable math, for the accuracy test against double precision. */
float oracle_pm(int which, float x, float y) {
  switch (which) {
  case 0: return pm_sinf(x);
  case 1: return pm_cosf(x);
  case 2: return pm_logf(x);
  case 3: return pm_pow5f(x);
  case 4: return pm_atan2f(y, x);
  case 5: return pm_acosf(x);
  default: return 0.0f;
  }
}
