/*
 * rtc_scene.cuh -- the device view of one flattened scene, the hit record, and the two primitive tests.
 *
 * Primitive tests are written with the strict (never contracted) operations of rtc_math.cuh in the reference's
 * operation order, so that hit distances carry the reference's bits in every kernel:
 *   sphere  /root/reference/src/hittable.c:120-151      quad  /root/reference/src/hittable.c:186-214
 */
#ifndef RTC_SCENE_CUH
#define RTC_SCENE_CUH

#include "rtc_b200.h"
#include "rtc_math.cuh"

/* ---- rebuilt acceleration structure (rtc_bvh.cpp builds it, rtc_mega.cu / rtc_wave.cu walk it) ---- */

/* child / root reference: kind in bits 30-31, instance tag in bits 22-29, index in bits 0-21.
 * A leaf is ONE primitive (measured faster than 2 or 4 per leaf) -- or one Box: the six quads of a Box_new list share
 * one tight bound, so a tree over them only adds visits (Book 2 final: 400 boxes = 2 400 of its 2 401 quads); the leaf
 * tests them one after the other in list order, exactly like HittableList_hit does (hittable.c:74-88).  tag = 0: the primitive lives in the group's own
 * frame; tag = k > 0: it lives under insts[k-1]'s transform chain -- the tree bounds it in the GROUP's frame and the
 * leaf test moves the ray into the primitive's frame with the reference's own operation sequence, so instances cost
 * no second traversal.  An inner reference is its node index (kind 0, tag 0). */
#define REF_INNER 0u
#define REF_SPHERES 1u
#define REF_QUADS 2u
#define REF_BOXES 3u /* SIX consecutive quads of mquads[] in the reference's list order: a Box (Box_new, hittable.c:246-264) */
#define REF_EMPTY 0xFFFFFFFFu
#define REF_INDEX_BITS 22
#define REF_MAX_INDEX ((1u << REF_INDEX_BITS) - 1u)
#define REF_MAX_TAG 255u
#define REF_KIND(r) ((r) >> 30)
#define REF_TAG(r) (((r) >> REF_INDEX_BITS) & 255u)
#define REF_INDEX(r) ((r)&REF_MAX_INDEX)
#define MAKE_REF(kind, tag, index) (((uint32_t)(kind) << 30) | ((uint32_t)(tag) << REF_INDEX_BITS) | (uint32_t)(index))

/* RTC_NODE_CH = 1: a child's box is stored as (centre, half extent) per axis instead of (lo, hi).  The slab test then
 * needs no pairwise min/max: near = (c - o)/d - h/|d|, far = (c - o)/d + h/|d| -- three FMAs per axis on the FMA pipe
 * instead of two FMAs plus two FMNMX on the (half as wide, and in this kernel busiest) ALU pipe.  The half extent is
 * rounded up and inflated by the rounding the extra subtraction can add (rtc_bvh.cu, make_node). */
#ifndef RTC_NODE_CH
#define RTC_NODE_CH 1
#endif

/* binary node holding BOTH children's boxes: two 32-byte loads per visit */
struct __align__(32) BNode {
  float4 xy0; /* child0: x.lo x.hi y.lo y.hi   (RTC_NODE_CH: x.c x.h y.c y.h) */
  float4 xy1; /* child1: x.lo x.hi y.lo y.hi */
  float4 zz;  /* child0 z.lo z.hi, child1 z.lo z.hi */
  uint32_t child0, child1, pad0, pad1;
};

/* 4-wide node (RT_BVH_WIDE=1, the render kernels' NW = 4 instantiations): the binary tree collapsed two levels at a
 * time (rtc_bvh.cu: wide_collapse), up to four children with their boxes as (centre, half extent), child k in lane k of
 * every float4 -- four 256-bit loads per visit for half the visits, half the dependent node fetches and half the votes
 * of the binary walk.  An absent child has child[k] = REF_EMPTY (its box is never looked at). */
struct __align__(32) WNode {
  float4 cx, hx; /* x: centres, half extents of children 0..3 */
  float4 cy, hy;
  float4 cz, hz;
  uint32_t child[4];
  uint32_t pad[4];
};

/* sphere as the traversal reads it: one 16-byte load; material/node ids live in a side array */
struct __align__(16) MSphere {
  float cx, cy, cz, r;
};
struct __align__(8) MPrimMeta {
  uint32_t material, node;
};
/* quad for the traversal: five 16-byte rows (the same numbers as rt_quad) */
struct __align__(16) MQuad {
  float4 n_D;   /* normal.xyz, D */
  float4 q_mat; /* Q.xyz, material (bits) */
  float4 w_nod; /* w.xyz, node (bits) */
  float4 u_ar;  /* u.xyz, area */
  float4 v_;    /* v.xyz, - */
};

/* one step of an instance's transform chain, outermost first (Translate: f = offset; RotateY: f0 = sin, f1 = cos) */
struct __align__(16) ChainOp {
  float f0, f1, f2;
  uint32_t is_rotate;
};
struct __align__(16) Instance {
  uint32_t op_begin, op_count, pad0, pad1;
};
struct __align__(16) MMedium {
  float bbox[6];           /* world-space bound of the boundary (conservative) */
  uint32_t root;           /* boundary group */
  uint32_t material, node; /* phase function, node id */
  float neg_inv_density;
  uint32_t op_begin, op_count; /* transforms ABOVE the medium itself */
  uint32_t in_list;            /* 1: reached through lists / transforms only (no BVHNode above it), so the reference
                                  evaluates it at its list position with the running t_max -- reproduced draw for draw
                                  (trace_media); 0: under a BVHNode, whose box test decides whether it is asked at all:
                                  evaluated after the surfaces (same distribution, other draw count) */
  uint32_t boundary_listed;    /* 1: the boundary is a plain sphere that is ALSO an item of world.objects before this
                                  medium (Book 2 final lists the glass ball, then the fog inside it): a ray that enters
                                  from outside meets that surface at exactly the entry distance, so an interval that
                                  starts behind the closest hit is empty for the reference too -- no query needed */
};

struct DScene {
  /* the reference-order graph exactly as flattened (rt_scene_desc arrays on the device) */
  const rt_node *nodes;
  const uint32_t *children;
  const rt_sphere *spheres;
  const rt_quad *quads;
  const rt_material *materials;
  const rt_texture *textures;
  const rt_perlin *perlins;
  const rt_image *images;
  const uint8_t *texels;
  uint32_t objects_root, lights_root;
  /* the rebuilt structure */
  const BNode *bnodes;
  const WNode *wnodes;   /* the world group's tree again, 4-wide (empty unless the scene was built with RT_BVH_WIDE=1) */
  uint32_t wide_root;    /* its root reference (an inner reference indexes wnodes), REF_EMPTY: no wide tree */
  const MSphere *msph;
  const MPrimMeta *msph_meta;
  const MQuad *mquads;
  const Instance *insts;
  const ChainOp *ops;
  const MMedium *media;
  uint32_t world_root, n_media;
  /* OVERSIZED primitives of the world group (box surface above half of the group's: the r = 1000 ground spheres), kept
   * out of the tree and tested before every walk (walk_pre): the walk then starts with the t_max they leave, and the
   * tree has no near-root node whose box is the whole scene */
  uint32_t n_pre, pre_ref[2];
  uint32_t n_perlins; /* 0: no Perlin texture in the scene (the megakernel skips the warp-wide lookups) */
  /* world.lights is a list of exactly ONE quad (Cornell box, Book 2 final): the quad itself, here in the kernel's
   * parameter space (constant bank, warp-uniform reads), so HittableList_pdf / _rand need no walk over the graph */
  uint32_t feats;          /* FEAT_* mask of what the scene contains */
  /* the transform chains of instances 1 and 2, when each is at most two steps (Translate(RotateY(x)): the Cornell
   * box's two blocks, Book 2 final's sphere cluster), in the kernel's parameter space: a tagged leaf then moves the ray
   * without the two dependent loads (instance record, then its steps) it otherwise starts with */
  uint32_t n_const_chains; /* 0, 1 or 2 */
  uint32_t const_count[2];
  float4 const_op[2][2];   /* ChainOp bits: f0 f1 f2 is_rotate */
  uint32_t one_quad_light; /* 1: `light` is valid */
  rt_quad light;
};

/* What a scene CONTAINS, as a compile-time mask of the render kernels (template <int FEATS>): a kernel instantiated
 * without a feature carries none of its code -- leaf tests, hit records, light sampling, media, texture lookups --
 * which is registers, spills and instruction-cache footprint the sphere-only scenes (Book 1 final) then do not pay.
 * rt_scene_create derives the scene's mask; mega_variant() only picks a lean kernel for a scene it covers. */
enum {
  FEAT_QUADS = 1,      /* quads / Box leaves in the rebuilt structure */
  FEAT_INSTANCES = 2,  /* primitives under Translate / RotateY chains */
  FEAT_LIGHTS = 4,     /* world.lights is not empty: the Book-3 mixture pdf may be evaluated */
  FEAT_MEDIA = 8,      /* ConstantMedium items (and their isotropic phase functions) */
  FEAT_TEXTURES = 16,  /* any texture that is not a solid colour (checker, image, Perlin) */
  FEAT_ALL = 31,
  FEAT_PRE = 32,       /* oversized primitives tested before the walk (DScene.pre_ref): sphere-only worlds; NOT part
                          of FEAT_ALL -- the kernels of the other scenes carry none of its code */
};

/* material.h:10-18 plus the id of the node that produced the hit */
struct Hit {
  V3 p, n;
  uint32_t material;
  float t, u, v;
  int front;
  uint32_t node;
};

/* ------------------------------------------------------------------ hittable.c:120-138 ------------------- */
/* Distance only.  Returns true and writes *t when a root lies in the OPEN interval (t_min, t_max). */
RTC_DEV bool sphere_isect(V3 center, float radius, V3 o, V3 d, float t_min, float t_max, float *t) {
  V3 oc = sv_sub(o, center);
  float a = sv_len2(d);
  float b = sv_dot(oc, d);
  float c = s_sub(sv_len2(oc), s_mul(radius, radius));
  float disc = s_sub(s_mul(b, b), s_mul(a, c));
  if (disc < 0.0f)
    return false;
  float sq = __fsqrt_rn(disc);
  float root = s_div(s_sub(-b, sq), a);
  if (root <= t_min || root >= t_max) {
    root = s_div(s_add(-b, sq), a);
    if (root <= t_min || root >= t_max)
      return false;
  }
  *t = root;
  return true;
}

/* ------------------------------------------------------------------ hittable.c:186-203 ------------------- */
/* CLOSED interval [t_min, t_max]; alpha/beta are the quad's (u, v). */
RTC_DEV bool quad_isect(V3 normal, float D, V3 Q, V3 w, V3 qu, V3 qv, V3 o, V3 d, float t_min, float t_max, float *t,
                    float *alpha, float *beta) {
  float denom = sv_dot(normal, d);
  if (fabsf(denom) < 1e-8f)
    return false;
  float tt = s_div(s_sub(D, sv_dot(normal, o)), denom);
  if (tt < t_min || tt > t_max)
    return false;
  V3 p = sv_at(o, d, tt);
  V3 hp = sv_sub(p, Q);
  float al = sv_dot(w, sv_cross(hp, qv));
  float be = sv_dot(w, sv_cross(qu, hp));
  if (al < 0.0f || al > 1.0f || be < 0.0f || be > 1.0f)
    return false;
  *t = tt;
  *alpha = al;
  *beta = be;
  return true;
}

#endif /* RTC_SCENE_CUH */
