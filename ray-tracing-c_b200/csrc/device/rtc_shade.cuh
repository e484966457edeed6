/*
 * rtc_shade.cuh -- everything that happens at a hit point: full hit records, textures, materials, and the
 * Book-3 light sampling.  Shared by the exact, mega and wavefront kernels.
 *
 * template<bool EXACT>:
 *   true  -> the seven transcendentals come from rtc_portable_math.h (bit-identical to the oracle's PORTABLE mode
 *            when the translation unit is compiled -fmad=false);
 *   false -> CUDA's fast intrinsics (statistically equivalent; parity is judged by RMSE).
 *
 * Reference lines restated (paths relative to /root/reference/src):
 *   hit records  hittable.c:139-150, 204-213        ONB        material.c:144-152
 *   textures     texture.c:8-114                    materials  material.c:23-142
 *   light pdf / rand   hittable.c:89-107, 152-178, 215-228
 */
#ifndef RTC_SHADE_CUH
#define RTC_SHADE_CUH

#include "rtc_portable_math.h"
#include "rtc_scene.cuh"

template <bool EXACT> RTC_DEV void m_sincos(float x, float *s, float *c) {
  if (EXACT) {
    *s = pm_sinf(x);
    *c = pm_cosf(x);
  } else {
    __sincosf(x, s, c);
  }
}
template <bool EXACT> RTC_DEV float m_sin(float x) { return EXACT ? pm_sinf(x) : __sinf(x); }
template <bool EXACT> RTC_DEV float m_log(float x) { return EXACT ? pm_logf(x) : __logf(x); }
template <bool EXACT> RTC_DEV float m_atan2(float y, float x) { return EXACT ? pm_atan2f(y, x) : atan2f(y, x); }
template <bool EXACT> RTC_DEV float m_acos(float x) { return EXACT ? pm_acosf(x) : acosf(x); }

/* ------------------------------------------------------------------ full hit records --------------------- */

/* hittable.c:139-150 given the accepted root */
template <bool EXACT, bool WANT_UV> RTC_DEV void sphere_record(V3 center, float radius, V3 o, V3 d, float t, Hit &rec) {
  rec.t = t;
  rec.p = sv_at(o, d, t);
  V3 outward = sv_div(sv_sub(rec.p, center), radius);
  rec.front = sv_dot(d, outward) < 0.0f;
  rec.n = rec.front ? outward : v_neg(outward);
  if (WANT_UV) {
    rec.u = (m_atan2<EXACT>(-outward.z, outward.x) + RTC_PI) * RTC_1_PI * 0.5f;
    rec.v = m_acos<EXACT>(-outward.y) * RTC_1_PI;
  }
}
/* the (u, v) part alone, from the outward normal (hittable.c:118-123) */
template <bool EXACT> RTC_DEV void sphere_uv(V3 outward, Hit &rec) {
  rec.u = (m_atan2<EXACT>(-outward.z, outward.x) + RTC_PI) * RTC_1_PI * 0.5f;
  rec.v = m_acos<EXACT>(-outward.y) * RTC_1_PI;
}

template <bool EXACT> RTC_DEV bool sphere_hit(const rt_sphere &s, V3 o, V3 d, float t_min, float t_max, Hit &rec) {
  V3 center = v_load(s.center);
  float t;
  if (!sphere_isect(center, s.radius, o, d, t_min, t_max, &t))
    return false;
  sphere_record<EXACT, true>(center, s.radius, o, d, t, rec);
  rec.material = s.material;
  rec.node = s.node;
  return true;
}

RTC_DEV bool quad_hit(const rt_quad &q, V3 o, V3 d, float t_min, float t_max, Hit &rec) {
  V3 n = v_load(q.normal);
  float t, al, be;
  if (!quad_isect(n, q.D, v_load(q.Q), v_load(q.w), v_load(q.u), v_load(q.v), o, d, t_min, t_max, &t, &al, &be))
    return false;
  rec.u = al;
  rec.v = be;
  rec.t = t;
  rec.p = sv_at(o, d, t);
  rec.material = q.material;
  rec.front = sv_dot(d, n) < 0.0f;
  rec.n = rec.front ? n : v_neg(n);
  rec.node = q.node;
  return true;
}

/* ------------------------------------------------------------------ material.c:144-152 ------------------- */

struct Basis {
  V3 u, v, w;
};
RTC_DEV Basis basis_from_w(V3 w) {
  Basis b;
  b.w = v_unit(w);
  V3 a = fabsf(b.w.x) > 0.9f ? v3(0, 1, 0) : v3(1, 0, 0);
  b.v = v_unit(v_cross(b.w, a));
  b.u = v_cross(b.w, b.v);
  return b;
}
RTC_DEV V3 basis_local(const Basis &b, V3 a) {
  return v_add(v_add(v_scale(b.u, a.x), v_scale(b.v, a.y)), v_scale(b.w, a.z));
}

/* vec3.c:33-43 under gcc: z, y, x */
RTC_DEV V3 rng_unit_vector(Rng &g) {
  for (;;) {
    float z = rng_between(g, -1.0f, 1.0f);
    float y = rng_between(g, -1.0f, 1.0f);
    float x = rng_between(g, -1.0f, 1.0f);
    V3 c = v3(x, y, z);
    float l2 = v_len2(c);
    if (l2 < 1.0f)
      return v_div(c, f_sqrt(l2));
  }
}

/* ------------------------------------------------------------------ textures ----------------------------- */

RTC_DEV float perlin_noise(const rt_perlin *pl, V3 p) { /* texture.c:78-103 */
  float fx = floorf(p.x), fy = floorf(p.y), fz = floorf(p.z);
  int i = (int)fx, j = (int)fy, k = (int)fz;
  float t1 = p.x - (float)i, t2 = p.y - (float)j, t3 = p.z - (float)k;
  float tt1 = t1 * t1 * (3.0f - 2.0f * t1);
  float tt2 = t2 * t2 * (3.0f - 2.0f * t2);
  float tt3 = t3 * t3 * (3.0f - 2.0f * t3);
  float value = 0.0f;
#pragma unroll
  for (int di = 0; di < 2; di++)
#pragma unroll
    for (int dj = 0; dj < 2; dj++)
#pragma unroll
      for (int dk = 0; dk < 2; dk++) {
        int cell = __ldg(&pl->perm_x[(i + di) & 255]) ^ __ldg(&pl->perm_y[(j + dj) & 255]) ^
                   __ldg(&pl->perm_z[(k + dk) & 255]);
        V3 grad = v3(__ldg(&pl->grad[3 * cell]), __ldg(&pl->grad[3 * cell + 1]), __ldg(&pl->grad[3 * cell + 2]));
        V3 weight = v3(t1 - (float)di, t2 - (float)dj, t3 - (float)dk);
        value += v_dot(grad, weight) * ((float)di * tt1 + (float)(1 - di) * (1.0f - tt1)) *
                 ((float)dj * tt2 + (float)(1 - dj) * (1.0f - tt2)) *
                 ((float)dk * tt3 + (float)(1 - dk) * (1.0f - tt3));
      }
  return value;
}
RTC_DEV float perlin_turbulence(const rt_perlin *pl, int depth, V3 p) { /* texture.c:105-114 */
  float value = 0.0f, weight = 1.0f;
  for (int i = 0; i < depth; i++) {
    value += weight * perlin_noise(pl, p);
    weight *= 0.5f;
    p = v_scale(p, 2.0f);
  }
  return fabsf(value);
}

/* everything but a solid colour: out of line, so that the hot shading code stays small in the instruction cache */
template <bool EXACT> __device__ __noinline__ V3 texture_value_slow(const DScene &sc, uint32_t id, float u, float v, V3 p) {
  for (;;) {
    const rt_texture *t = &sc.textures[id];
    uint32_t kind = __ldg(&t->kind);
    if (kind == RT_TEX_SOLID)
      return v3(__ldg(&t->color[0]), __ldg(&t->color[1]), __ldg(&t->color[2]));
    if (kind == RT_TEX_CHECKER) { /* texture.c:12-22: on (u,v); % keeps the sign, so -1 selects `odd` too */
      float scale = __ldg(&t->scale);
      int iu = (int)floorf(f_div(u, scale)), iv = (int)floorf(f_div(v, scale));
      id = ((iu + iv) % 2) ? __ldg(&t->b) : __ldg(&t->a);
      continue;
    }
    if (kind == RT_TEX_IMAGE) { /* texture.c:28-37, nearest neighbour; indices clamped where the reference has UB */
      const rt_image *im = &sc.images[__ldg(&t->a)];
      int w = __ldg(&im->width), h = __ldg(&im->height);
      float fi = roundf(u * (float)(w - 1));
      float fj = roundf((1.0f - v) * (float)(h - 1));
      int i = (fi >= 0.0f) ? (fi <= (float)(w - 1) ? (int)fi : w - 1) : 0; /* NaN -> 0 */
      int j = (fj >= 0.0f) ? (fj <= (float)(h - 1) ? (int)fj : h - 1) : 0;
      const uint8_t *px = sc.texels + im->offset + ((size_t)j * w + i) * 3;
      return v3(f_div((float)__ldg(px), 255.0f), f_div((float)__ldg(px + 1), 255.0f),
                f_div((float)__ldg(px + 2), 255.0f));
    }
    if (kind == RT_TEX_PERLIN) { /* texture.c:47-52 */
      V3 q = v_scale(p, __ldg(&t->scale));
      float turb = perlin_turbulence(&sc.perlins[__ldg(&t->a)], (int)__ldg(&t->b), q);
      float s = 0.5f * (1.0f + m_sin<EXACT>(q.z + 10.0f * turb));
      return v3(s, s, s);
    }
    return v3(0, 0, 0);
  }
}

template <bool EXACT> RTC_DEV V3 texture_value(const DScene &sc, uint32_t id, float u, float v, V3 p) {
  const rt_texture *t = &sc.textures[id];
  if (__ldg(&t->kind) == RT_TEX_SOLID) /* texture.c:8 */
    return v3(__ldg(&t->color[0]), __ldg(&t->color[1]), __ldg(&t->color[2]));
  return texture_value_slow<EXACT>(sc, id, u, v, p);
}

/* A Perlin lookup a lane owes: the megakernel evaluates it later, with the whole warp (perlin_value_warp). */
struct PendingTex {
  uint32_t tex; /* RT_NONE: nothing pending.  The point is not kept: it is the hit point, i.e. the origin of the
                   scattered ray, which the caller holds anyway */
};
/* texture_value for an ALBEDO (a factor of the path throughput): a Perlin texture is not evaluated here -- 7 octaves x
 * 8 lattice corners on the one or two lanes of a warp that hit the marble sphere (ncu, Book 2 final: 8.4 % of all
 * warp instructions at 1.9 active lanes) -- but handed back as pending, with white as the factor for now. */
template <bool EXACT, bool DEFER, int FEATS = FEAT_ALL>
RTC_DEV V3 albedo_value(const DScene &sc, uint32_t id, float u, float v, V3 p, PendingTex &pend) {
  const rt_texture *t = &sc.textures[id];
  if (!(FEATS & FEAT_TEXTURES)) /* every texture of the scene is a solid colour */
    return v3(__ldg(&t->color[0]), __ldg(&t->color[1]), __ldg(&t->color[2]));
  const uint32_t kind = __ldg(&t->kind);
  if (kind == RT_TEX_SOLID)
    return v3(__ldg(&t->color[0]), __ldg(&t->color[1]), __ldg(&t->color[2]));
  if (DEFER && kind == RT_TEX_PERLIN) {
    pend.tex = id;
    return v3(1, 1, 1);
  }
  return texture_value_slow<EXACT>(sc, id, u, v, p);
}

/* Perlin_value (texture.c:47-52) for the lanes of a warp that owe one (`pending`, at the point `at`), computed by the
 * WHOLE warp: four lookups per pass, eight lanes each, lane = octave.  Every lane evaluates one octave's noise with the
 * reference's own corner order (perlin_noise above); the octaves of a lookup are then added by a butterfly over its
 * eight lanes instead of one after the other -- a fixed order, whatever the warp's population, so renders stay
 * bit-reproducible; against the sequential sum the difference is in the last bits (the megakernel's shading is
 * judged by RMSE; the exact kernel keeps the sequential form).  All 32 lanes must call it.  Multiplies the
 * throughput `beta` of each owing lane by its texture value. */
static __device__ __noinline__ void perlin_settle_warp(const DScene &sc, uint32_t tex_owed, V3 at, V3 &beta) {
  const unsigned lane = threadIdx.x & 31u, group = lane >> 3, oct0 = lane & 7u;
  unsigned m = __ballot_sync(0xffffffffu, tex_owed != RT_NONE);
  while (m) {
    /* this pass serves the four lowest owing lanes; group g takes the g-th of them */
    unsigned rest = m;
    int src = -1;
#pragma unroll
    for (unsigned g = 0; g < 4; g++) {
      const int b = rest ? __ffs(rest) - 1 : -1;
      rest &= rest - 1u;
      if (g == group)
        src = b;
    }
    const int from = src < 0 ? 0 : src;
    const uint32_t tex = __shfl_sync(0xffffffffu, tex_owed, from);
    V3 p;
    p.x = __shfl_sync(0xffffffffu, at.x, from);
    p.y = __shfl_sync(0xffffffffu, at.y, from);
    p.z = __shfl_sync(0xffffffffu, at.z, from);
    float partial = 0.0f, qz = 0.0f;
    if (src >= 0) {
      const rt_texture *t = &sc.textures[tex];
      const rt_perlin *pl = &sc.perlins[__ldg(&t->a)];
      const int depth = (int)__ldg(&t->b);
      const V3 q = v_scale(p, __ldg(&t->scale));
      qz = q.z;
      for (int oct = (int)oct0; oct < depth; oct += 8) {
        const float up = __int_as_float((127 + oct) << 23), down = __int_as_float((127 - oct) << 23); /* 2^oct, 2^-oct */
        partial += down * perlin_noise(pl, v_scale(q, up));
      }
    }
    partial += __shfl_xor_sync(0xffffffffu, partial, 1);
    partial += __shfl_xor_sync(0xffffffffu, partial, 2);
    partial += __shfl_xor_sync(0xffffffffu, partial, 4);
    const float value = 0.5f * (1.0f + __sinf(qz + 10.0f * fabsf(partial)));
    /* an owing lane of rank r < 4 among the owing lanes was served by group r */
    const unsigned rank = __popc(m & ((1u << lane) - 1u));
    const float mine = __shfl_sync(0xffffffffu, value, (rank & 3u) << 3);
    if (tex_owed != RT_NONE && rank < 4u) {
      beta = v_scale(beta, mine);
      tex_owed = RT_NONE;
    }
    m = rest;
  }
}
/* does evaluating this texture read (u, v)?  (lets the fast kernels skip atan2f/acosf on spheres) */
RTC_DEV bool texture_needs_uv(const DScene &sc, uint32_t id) {
  if (id == RT_NONE)
    return false;
  uint32_t kind = __ldg(&sc.textures[id].kind);
  return kind == RT_TEX_CHECKER || kind == RT_TEX_IMAGE;
}

/* ------------------------------------------------------------------ materials ---------------------------- */

RTC_DEV V3 reflect(V3 incident, V3 normal) { /* material.c:45-47 */
  return v_sub(incident, v_scale(normal, 2.0f * v_dot(incident, normal)));
}

/* material.c:103-120.  Returns false when the material does not scatter. */
template <bool EXACT, bool DEFER, int FEATS = FEAT_ALL>
RTC_DEV bool material_scatter(const DScene &sc, const Hit &rec, V3 r_in, V3 &r_out, V3 &albedo, bool &skip_pdf,
                              Rng &g, PendingTex &pend) {
  const rt_material *m = &sc.materials[rec.material];
  uint32_t tag = __ldg(&m->tag);
  switch (tag) {
  case RT_MAT_LAMBERTIAN: { /* material.c:23-37 */
    Basis b = basis_from_w(rec.n);
    float r1 = rng_f32(g);
    float r2 = rng_f32(g);
    float phi = 2.0f * RTC_PI * r1;
    float sn, cs;
    m_sincos<EXACT>(phi, &sn, &cs);
    float sr2 = f_sqrt(r2);
    V3 local = v3(cs * sr2, sn * sr2, f_sqrt(1.0f - r2));
    r_out = basis_local(b, local);
    albedo = albedo_value<EXACT, DEFER, FEATS>(sc, __ldg(&m->texture), rec.u, rec.v, rec.p, pend);
    skip_pdf = false;
    return true;
  }
  case RT_MAT_METAL: { /* material.c:48-58: three draws per try even when fuzz == 0 */
    V3 mirrored = reflect(v_unit(r_in), rec.n);
    r_out = v_add(mirrored, v_scale(rng_unit_vector(g), __ldg(&m->param)));
    albedo = albedo_value<EXACT, DEFER, FEATS>(sc, __ldg(&m->texture), rec.u, rec.v, rec.p, pend);
    skip_pdf = true;
    if (v_dot(r_out, rec.n) < 0.0f)
      r_out = mirrored;
    return true;
  }
  case RT_MAT_DIELECTRIC: { /* material.c:62-86 */
    float eta = __ldg(&m->param);
    if (rec.front)
      eta = f_div(1.0f, eta);
    V3 in = v_unit(r_in);
    float cos_theta = fminf(-v_dot(in, rec.n), 1.0f);
    float sin_theta = f_sqrt(1.0f - cos_theta * cos_theta);
    float schlick = f_div(1.0f - eta, 1.0f + eta);
    schlick *= schlick;
    schlick += (1.0f - schlick) * pm_pow5f(1.0f - cos_theta);
    if (eta * sin_theta > 1.0f || schlick > rng_f32(g)) { /* the draw is skipped on total internal reflection */
      r_out = reflect(in, rec.n);
    } else {
      V3 perp = v_scale(v_add(in, v_scale(rec.n, cos_theta)), eta);
      V3 para = v_scale(rec.n, -f_sqrt(fabsf(1.0f - v_len2(perp))));
      r_out = v_add(perp, para);
    }
    albedo = v3(1, 1, 1);
    skip_pdf = true;
    return true;
  }
  case RT_MAT_ISOTROPIC: /* material.c:93-98 */
    if (!(FEATS & FEAT_MEDIA)) { /* only ConstantMedium_init makes one (hittable.c:425-432) */
      skip_pdf = true;
      return false;
    }
    r_out = rng_unit_vector(g);
    albedo = albedo_value<EXACT, DEFER>(sc, __ldg(&m->texture), rec.u, rec.v, rec.p, pend);
    skip_pdf = false;
    return true;
  default: /* SURFACE_NORMAL, DIFFUSE_LIGHT */
    skip_pdf = true;
    return false;
  }
}

template <bool EXACT>
RTC_DEV bool material_scatter(const DScene &sc, const Hit &rec, V3 r_in, V3 &r_out, V3 &albedo, bool &skip_pdf,
                              Rng &g) {
  PendingTex none;
  none.tex = RT_NONE;
  return material_scatter<EXACT, false>(sc, rec, r_in, r_out, albedo, skip_pdf, g, none);
}

RTC_DEV float material_scatter_pdf(const DScene &sc, uint32_t material, V3 normal, V3 r_out) { /* material.c:122-131 */
  uint32_t tag = __ldg(&sc.materials[material].tag);
  if (tag == RT_MAT_LAMBERTIAN) {
    float cos_theta = v_dot(normal, v_unit(r_out));
    return cos_theta < 0.0f ? 0.0f : f_div(cos_theta, RTC_PI);
  }
  if (tag == RT_MAT_ISOTROPIC)
    return f_div(1.0f, 4.0f * RTC_PI);
  return 0.0f;
}

template <bool EXACT> RTC_DEV V3 material_emit(const DScene &sc, const Hit &rec) { /* material.c:133-142 */
  const rt_material *m = &sc.materials[rec.material];
  uint32_t tag = __ldg(&m->tag);
  if (tag == RT_MAT_SURFACE_NORMAL)
    return v_scale(v3(rec.n.x + 1.0f, rec.n.y + 1.0f, rec.n.z + 1.0f), 0.5f);
  if (tag == RT_MAT_DIFFUSE_LIGHT)
    return rec.front ? texture_value<EXACT>(sc, __ldg(&m->texture), rec.u, rec.v, rec.p) : v3(0, 0, 0);
  return v3(0, 0, 0);
}

/* ------------------------------------------------------------------ light sampling ----------------------- */
/* The lights list is walked in its reference form (nodes/children/spheres/quads): it is tiny, and "mean over
 * the items that have a pdf" / "uniform pick with rejection" are defined on that nesting. */

template <bool EXACT> RTC_DEV float sphere_pdf(const rt_sphere &s, V3 o, V3 d) { /* hittable.c:152-162 */
  Hit rec;
  if (!sphere_hit<EXACT>(s, o, d, 0.001f, RTC_INF, rec))
    return 0.0f;
  V3 oc = v_sub(v_load(s.center), o);
  float cos_max = f_sqrt(1.0f - f_div(s.radius * s.radius, v_len2(oc)));
  float solid_angle = 2.0f * RTC_PI * (1.0f - cos_max);
  return f_div(1.0f, solid_angle);
}
template <bool EXACT> RTC_DEV V3 sphere_rand(const rt_sphere &s, V3 origin, Rng &g) { /* hittable.c:163-178 */
  V3 oc = v_sub(v_load(s.center), origin);
  float r1 = rng_f32(g);
  float r2 = rng_f32(g);
  float z = 1.0f + r2 * (f_sqrt(1.0f - f_div(s.radius * s.radius, v_len2(oc))) - 1.0f);
  float phi = 2.0f * RTC_PI * r1;
  float sn, cs;
  m_sincos<EXACT>(phi, &sn, &cs);
  float x = cs * f_sqrt(1.0f - z * z);
  float y = sn * f_sqrt(1.0f - z * z);
  Basis b = basis_from_w(oc);
  return basis_local(b, v3(x, y, z));
}
RTC_DEV float quad_pdf(const rt_quad &q, V3 o, V3 d) { /* hittable.c:215-224 */
  Hit rec;
  if (!quad_hit(q, o, d, 0.001f, RTC_INF, rec))
    return 0.0f;
  float d2 = rec.t * rec.t * v_len2(d);
  float cos_theta = fabsf(v_dot(rec.n, v_unit(d)));
  return f_div(d2, cos_theta * q.area);
}
RTC_DEV V3 quad_rand(const rt_quad &q, V3 origin, Rng &g) { /* hittable.c:225-228; gcc draws v's coefficient first */
  float along_v = rng_f32(g);
  float along_u = rng_f32(g);
  V3 p = v_add(v_add(v_load(q.Q), v_scale(v_load(q.u), along_u)), v_scale(v_load(q.v), along_v));
  return v_add(p, v_neg(origin));
}

#define RTC_LIGHT_NESTING 4 /* lists inside the lights list, checked by rt_scene_create */

/* hittable.c:89-100: mean over the items that have a pdf; nested lists recurse */
template <bool EXACT> __device__ __noinline__ float lights_pdf(const DScene &sc, uint32_t root, V3 o, V3 d) {
  uint32_t st_node[RTC_LIGHT_NESTING], st_idx[RTC_LIGHT_NESTING];
  float st_sum[RTC_LIGHT_NESTING], st_cnt[RTC_LIGHT_NESTING];
  int sp = 0;
  const rt_node *n = &sc.nodes[root];
  if (n->kind == RT_NODE_SPHERE)
    return sphere_pdf<EXACT>(sc.spheres[n->a], o, d);
  if (n->kind == RT_NODE_QUAD)
    return quad_pdf(sc.quads[n->a], o, d);
  if (n->kind != RT_NODE_LIST)
    return 0.0f;
  st_node[0] = root, st_idx[0] = 0, st_sum[0] = 0.0f, st_cnt[0] = 0.0f;
  float ret = 0.0f;
  for (;;) {
    const rt_node *l = &sc.nodes[st_node[sp]];
    if (st_idx[sp] < l->b) {
      uint32_t item = sc.children[l->a + st_idx[sp]++];
      const rt_node *c = &sc.nodes[item];
      if (!(c->flags & RT_NODE_HAS_PDF))
        continue;
      if (c->kind == RT_NODE_LIST && sp + 1 < RTC_LIGHT_NESTING) {
        sp++;
        st_node[sp] = item, st_idx[sp] = 0, st_sum[sp] = 0.0f, st_cnt[sp] = 0.0f;
        continue;
      }
      float pdf = c->kind == RT_NODE_SPHERE ? sphere_pdf<EXACT>(sc.spheres[c->a], o, d)
                  : c->kind == RT_NODE_QUAD ? quad_pdf(sc.quads[c->a], o, d)
                                            : 0.0f;
      st_sum[sp] += pdf;
      st_cnt[sp] += 1.0f;
    } else {
      ret = f_div(st_sum[sp], fmaxf(st_cnt[sp], 1.0f));
      if (sp == 0)
        return ret;
      sp--;
      st_sum[sp] += ret;
      st_cnt[sp] += 1.0f;
    }
  }
}

/* hittable.c:101-107: uniform item, retry while the item has no `rand`; nested lists descend */
template <bool EXACT> __device__ __noinline__ V3 lights_rand(const DScene &sc, uint32_t root, V3 origin, Rng &g) {
  uint32_t id = root;
  for (;;) {
    const rt_node *n = &sc.nodes[id];
    if (n->kind == RT_NODE_SPHERE)
      return sphere_rand<EXACT>(sc.spheres[n->a], origin, g);
    if (n->kind == RT_NODE_QUAD)
      return quad_rand(sc.quads[n->a], origin, g);
    if (n->kind != RT_NODE_LIST)
      return v3(0, 0, 0);
    for (;;) {
      uint32_t item = sc.children[n->a + rng_u32(g) % n->b];
      if (sc.nodes[item].flags & RT_NODE_HAS_PDF) {
        id = item;
        break;
      }
    }
  }
}

#endif /* RTC_SHADE_CUH */
