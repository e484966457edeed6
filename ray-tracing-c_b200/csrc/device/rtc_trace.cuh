/*
 * rtc_trace.cuh -- closest hit over the REBUILT structure (rtc_bvh.h): ordered traversal of a binary BVH whose
 * nodes carry both children's boxes as (centre, half extent) (two 256-bit loads and 18 FMAs per visit), one primitive
 * per leaf read with 16-byte loads, primitives under Translate / RotateY chains bounded in the group's frame and
 * tested in their own (the ray is moved there with the reference's own operation sequence), media evaluated after
 * the surfaces.  Replaces HittableList_hit / BVHNode_hit / AABB_hit / Translate_hit / RotateY_hit /
 * ConstantMedium_hit (/root/reference/src/hittable.c:38-55, 74-88, 266-277, 325-423) on the throughput path.
 *
 * Hit distances come from strict arithmetic in the reference's operation order (sphere_isect_closed, leaf_intersect),
 * so they carry the reference's bits; what differs from the reference is only the ORDER in which candidates are met.
 * Ties are resolved as the reference's visiting order would (quads: the later node id wins, closed interval;
 * spheres: the earlier node id wins, open interval), see leaf_intersect().
 */
#ifndef RTC_TRACE_CUH
#define RTC_TRACE_CUH

#include "rtc_shade.cuh"

/* tuning switches of the node loop (defaults = what measured best on B200) */
#ifndef RTC_NODE_WIDE
#define RTC_NODE_WIDE 1 /* 256-bit node loads in the render kernels' inlined loops */
#endif
#ifndef RTC_VISITS_PER_VOTE
#define RTC_VISITS_PER_VOTE 1
#endif
#ifndef RTC_CHAIN_INLINE
#define RTC_CHAIN_INLINE 1 /* instance transforms inlined into the walk (1) or out of line (0); measured */
#endif
#if RTC_CHAIN_INLINE
#define RTC_CHAIN_ATTR RTC_DEV
#else
#define RTC_CHAIN_ATTR static __device__ __noinline__
#endif
#ifndef RTC_SLAB_RCP_APPROX
#define RTC_SLAB_RCP_APPROX 1
#endif
#ifndef RTC_MEDIA_FAST
#define RTC_MEDIA_FAST 1
#endif
#ifndef RTC_MEDIA_INLIST
#define RTC_MEDIA_INLIST 1 /* 0: media clipped against the closest surface whatever its list position (round 1) */
#endif
#ifndef RTC_NODE_LD_HINT
#define RTC_NODE_LD_HINT "" /* L1 eviction priority of the node loads, e.g. ".L1::evict_last" (A/B: no gain) */
#endif
#ifndef RTC_CONST_CHAINS
#define RTC_CONST_CHAINS 1
#endif
#ifndef RTC_POSTPONE
#define RTC_POSTPONE 1 /* 1: kernels of <= 256 threads park the first leaf of a round and keep descending; 2: all; 0: none */
#endif
#ifndef RTC_SLAB_ONE_FMA
#define RTC_SLAB_ONE_FMA 1
#endif

#define TRACE_STACK 48          /* rt_scene_create refuses deeper structures for the mega/wave kernels */
/* RTC_SMEM_STACK = K > 0: the render kernels keep the first K entries of every lane's traversal stack in SHARED memory,
 * laid out [entry][thread] (bank = lane whatever the lane's depth: one wavefront per push / pop, where the local-memory
 * stack costs one L1 wavefront per distinct depth in the warp and its lines are written back through L2 to DRAM);
 * deeper entries overflow to the lane's local array.  0: the whole stack in local memory. */
#ifndef RTC_SMEM_STACK
#define RTC_SMEM_STACK 0
#endif
#define PRIM_NONE 0xFFFFFFFFu
#define PRIM_MEDIUM_KIND 3u     /* best_prim = MAKE_REF(3, 0, medium index) for a medium hit */

struct Ray {
  V3 o, d;
};

/* world (or enclosing-frame) ray -> object frame, op by op like Translate_hit / RotateY_hit (hittable.c:325-367) */
/* (out of line: entering an instance is rare next to a node visit, and the kernels are instruction-cache bound on
 * the scenes that have instances -- see DESIGN.md) */
RTC_DEV Ray chain_step(float4 op, Ray r) {
  if (__float_as_uint(op.w) == 0u) {
    r.o = sv_sub(r.o, v3(op.x, op.y, op.z));
  } else {
    r.o = sv_rot_y(r.o, op.y, op.x);
    r.d = sv_rot_y(r.d, op.y, op.x);
  }
  return r;
}
RTC_CHAIN_ATTR Ray chain_apply(const DScene &sc, uint32_t op_begin, uint32_t op_count, Ray r) {
  for (uint32_t k = 0; k < op_count; k++)
    r = chain_step(__ldg(reinterpret_cast<const float4 *>(&sc.ops[op_begin + k])), r);
  return r;
}
/* object frame -> world for a hit point and its normal, innermost op first */
RTC_CHAIN_ATTR void chain_unapply(const DScene &sc, uint32_t op_begin, uint32_t op_count, V3 &p, V3 &n) {
  for (uint32_t k = op_count; k-- > 0;) {
    const float4 op = __ldg(reinterpret_cast<const float4 *>(&sc.ops[op_begin + k]));
    if (__float_as_uint(op.w) == 0u) {
      p = sv_add(p, v3(op.x, op.y, op.z));
    } else {
      p = sv_rot_y_inv(p, op.y, op.x);
      n = sv_rot_y_inv(n, op.y, op.x);
    }
  }
}

/* out-of-line copies for the once-per-segment users (hit record, media): the render kernels are instruction-cache
 * bound on scenes with instances and media (ncu: "no instruction" is their first stall reason), and a call per
 * segment is nothing next to the ~20 node visits of that segment */
static __device__ __noinline__ Ray chain_apply_call(const DScene &sc, uint32_t op_begin, uint32_t op_count, Ray r) {
  return chain_apply(sc, op_begin, op_count, r);
}
static __device__ __noinline__ void chain_unapply_call(const DScene &sc, uint32_t op_begin, uint32_t op_count, V3 &p,
                                                       V3 &n) {
  chain_unapply(sc, op_begin, op_count, p, n);
}

/* One 64-byte BNode.  WIDE: two 256-bit read-only loads (LDG.E.256, sm_100+) instead of four 128/64-bit ones --
 * half the load instructions and half the L1 wavefronts of the node loop.  Only the kernels' own inlined node loops
 * use it: ptxas 12.9 crashes on a 256-bit load inside a device FUNCTION that two kernels of a module call
 * (trace_group_call), so the out-of-line walks keep the 128-bit loads. */
template <bool WIDE> RTC_DEV void ldg_node(const BNode *n, float4 &xy0, float4 &xy1, float4 &zz, uint2 &ch) {
  if (WIDE) {
    uint32_t a0, a1, a2, a3, a4, a5, a6, a7, b0, b1, b2, b3, pad0, pad1;
    asm("ld.global.nc" RTC_NODE_LD_HINT ".v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=r"(a0), "=r"(a1), "=r"(a2), "=r"(a3), "=r"(a4), "=r"(a5), "=r"(a6), "=r"(a7)
        : "l"(n));
    asm("ld.global.nc" RTC_NODE_LD_HINT ".v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=r"(b0), "=r"(b1), "=r"(b2), "=r"(b3), "=r"(ch.x), "=r"(ch.y), "=r"(pad0), "=r"(pad1)
        : "l"(reinterpret_cast<const char *>(n) + 32));
    xy0 = make_float4(__uint_as_float(a0), __uint_as_float(a1), __uint_as_float(a2), __uint_as_float(a3));
    xy1 = make_float4(__uint_as_float(a4), __uint_as_float(a5), __uint_as_float(a6), __uint_as_float(a7));
    zz = make_float4(__uint_as_float(b0), __uint_as_float(b1), __uint_as_float(b2), __uint_as_float(b3));
  } else {
    xy0 = __ldg(&n->xy0), xy1 = __ldg(&n->xy1), zz = __ldg(&n->zz);
    ch = __ldg(reinterpret_cast<const uint2 *>(&n->child0));
  }
}

struct TraceHit {
  float t;
  uint32_t prim; /* PRIM_NONE, or kind/index reference into msph / mquads / media */
  uint32_t inst; /* instance the primitive was met in, RT_NONE for the group's own frame */
};

struct SlabRay { /* per-frame constants of the box test: t = plane * inv - o * inv (one FMA per plane) */
  V3 inv, oi;
#if RTC_NODE_CH
  V3 ainv; /* |inv|: the half extents' distances */
#endif
  float slack; /* bound on the rounding error of an interval comparison made with those t's */
};
/* 1/d for the box test only (hit distances never use it): the 1-ulp hardware reciprocal, one MUFU instead of the
 * ~10 instructions plus slow path of an IEEE division -- the relative slack of slab_test covers the extra half ulp. */
RTC_DEV float slab_inv(float d) {
  const float qnan = __int_as_float(0x7fc00000);
  float inv;
#if RTC_SLAB_RCP_APPROX
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv) : "f"(d)); /* denormal d -> inf, capped below; |d| > 2^126 -> 0 */
#else
  inv = 1.0f / d;
#endif
  return d == 0.0f ? qnan : copysignf(fminf(fabsf(inv), 1e30f), d);
}
RTC_DEV SlabRay slab_setup(const Ray &r) {
  SlabRay s;
  /* |1/d| is capped at 1e30 so that plane*inv and o*inv never overflow to inf - inf (scene coordinates are far
   * below 3e8).  A component of EXACTLY zero gets NaN: both planes of that axis then drop out of fminf/fmaxf, i.e.
   * the axis is ignored, which errs on the side of visiting.  (The reference's (plane - o) * inf also yields NaN
   * when the ray lies in a face plane -- hittable.c:40-42 -- and a finite stand-in would turn that into a miss.) */
  s.inv = v3(slab_inv(r.d.x), slab_inv(r.d.y), slab_inv(r.d.z));
  s.oi = v3(r.o.x * s.inv.x, r.o.y * s.inv.y, r.o.z * s.inv.z);
#if RTC_NODE_CH
  s.ainv = v3(fabsf(s.inv.x), fabsf(s.inv.y), fabsf(s.inv.z));
#endif
  /* plane*inv - o*inv cancels: each t carries an ABSOLUTE error of about eps * |o*inv| (o*inv and inv are already
   * rounded), which for coordinates in the hundreds is as large as the reference's 1e-4 padding of flat boxes.  Two
   * such errors meet in `enter <= leave`, so the exit distance is widened by 5e-7 * max|o*inv| (NaN axes drop out):
   * a box the exact ray touches is never culled.  (Budget per t: inv 1 ulp = 1.2e-7, o*inv another 6e-8.) */
  s.slack = 5e-7f * fmaxf(fmaxf(fabsf(s.oi.x), fabsf(s.oi.y)), fabsf(s.oi.z));
  return s;
}

/* Does the ray meet the box within [t_min, t_max]?  `enter` = its entry distance (for ordering the children).  NaNs
 * (0 * inf on an axis the ray is parallel to) are dropped by fminf/fmaxf, which errs on the side of visiting; the
 * exit distance carries a relative + an absolute slack so that a true hit is never culled. */
template <bool NONNEG>
RTC_DEV bool slab_test(const SlabRay &s, float xlo, float xhi, float ylo, float yhi, float zlo, float zhi, float t_min,
                       float t_max, float &enter) {
#if RTC_NODE_CH
  /* (xlo, xhi) etc. are (centre, half extent): distance of the centre plane, minus / plus the half extent's */
  const float xc = fmaf(xlo, s.inv.x, -s.oi.x), yc = fmaf(ylo, s.inv.y, -s.oi.y), zc = fmaf(zlo, s.inv.z, -s.oi.z);
  enter = fmaxf(fmaxf(fmaf(-xhi, s.ainv.x, xc), fmaf(-yhi, s.ainv.y, yc)), fmaxf(fmaf(-zhi, s.ainv.z, zc), t_min));
  float leave = fminf(fminf(fmaf(xhi, s.ainv.x, xc), fmaf(yhi, s.ainv.y, yc)), fminf(fmaf(zhi, s.ainv.z, zc), t_max));
#else
  float x0 = fmaf(xlo, s.inv.x, -s.oi.x), x1 = fmaf(xhi, s.inv.x, -s.oi.x);
  float y0 = fmaf(ylo, s.inv.y, -s.oi.y), y1 = fmaf(yhi, s.inv.y, -s.oi.y);
  float z0 = fmaf(zlo, s.inv.z, -s.oi.z), z1 = fmaf(zhi, s.inv.z, -s.oi.z);
  enter = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), t_min));
  float leave = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), t_max));
#endif
  /* NONNEG: the caller's t_min is >= 0, so a negative exit distance is a miss whatever the slack does to it and
   * one FMA serves (leave * (1 + 6e-7) only widens a non-negative exit).  6e-7 relative: the two axes that meet in
   * the comparison each carry inv (1.2e-7) and the FMA's rounding (6e-8), and so does the entry side. */
  leave = (NONNEG && RTC_SLAB_ONE_FMA) ? fmaf(leave, 1.0000006f, s.slack) : fmaf(fabsf(leave), 6e-7f, leave) + s.slack;
  return enter <= leave;
}

/* x / y, IEEE like s_div, with the one case the hardware division sends down its ~110-instruction slow path taken out:
 * a numerator of exactly zero (quotient +-0).  That case is COMMON here: a scattered ray starts on the primitive it
 * left, so the plane distance of that quad is exactly 0, and the near root of that sphere cancels to exactly 0
 * whenever a*c is below half an ulp of b*b (ncu: the slow path was 1.2 % of all warp instructions on Book 1, 2.4 % on
 * the Cornell box, at 2-4 active lanes). */
RTC_DEV float s_div_z(float x, float y) {
  if (x == 0.0f && y != 0.0f && y == y)
    return __int_as_float((__float_as_int(x) ^ __float_as_int(y)) & (int)0x80000000);
  return s_div(x, y);
}

/* Is the IEEE quotient num / den CERTAINLY outside [lo, hi]?  Decided from the one-MUFU reciprocal: its product with
 * num is within 2.4e-7 (relative) of the correctly rounded quotient, the margin asked for is 1e-6 -- so a candidate
 * that is rejected here would have been rejected by the exact comparison too, and the ~10-instruction IEEE division
 * (plus its slow path) is only paid by candidates near or inside the interval.  Most leaf candidates are not: a plane
 * hit behind the origin or beyond the best hit so far.  den must be a normal, non-zero number. */
#ifndef RTC_DIV_PRETEST
#define RTC_DIV_PRETEST 0
#endif
RTC_DEV bool quotient_outside(float num, float den, float lo, float hi) {
#if RTC_DIV_PRETEST
  float inv;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv) : "f"(den));
  const float q = num * inv;
  /* (lo = -inf or hi = +inf: never outside on that side; NaN compares false: the exact path decides) */
  return q < lo - 1e-6f * fabsf(lo) - 1e-30f || q > hi + 1e-6f * fabsf(hi) + 1e-30f;
#else
  return false;
#endif
}

/* sphere_isect with t_max INCLUDED, so that the traversal can see exact ties and resolve them like the reference.
 * `a` = |d|^2 in the reference's operation order (sv_len2), computed once per ray frame by the caller.  The far root
 * is only computed when the near one is behind t_min: both are monotone in the same rounded operands
 * (fl((-b+sq)/a) >= fl((-b-sq)/a) for a > 0), so a near root beyond t_max already decides the miss -- the accept /
 * reject outcome and the accepted bits are the reference's (hittable.c:125-137). */
RTC_DEV bool sphere_isect_closed(V3 center, float radius, V3 o, V3 d, float a, float t_min, float t_max, float *t) {
  V3 oc = sv_sub(o, center);
  float b = sv_dot(oc, d);
  float c = s_sub(sv_len2(oc), s_mul(radius, radius));
  float disc = s_sub(s_mul(b, b), s_mul(a, c));
  if (disc < 0.0f)
    return false;
  float sq = __fsqrt_rn(disc);
  const float near_num = s_sub(-b, sq);
  /* the near root certainly beyond t_max: a miss whatever the far root is (it is not smaller) */
  if (quotient_outside(near_num, a, -RTC_INF, t_max))
    return false;
  float root = s_div_z(near_num, a);
  if (root > t_max)
    return false;
  if (root <= t_min) {
    root = s_div_z(s_add(-b, sq), a);
    if (root <= t_min || root > t_max)
      return false;
  }
  *t = root;
  return true;
}

/* Intersects the ONE primitive of a leaf reference (sphere or quad) and keeps it if closer; ties as the reference.
 * `r` is the ray in the primitive's own frame, `inst` the instance that frame belongs to (RT_NONE: the group's). */
template <int FEATS>
RTC_DEV void leaf_intersect(const DScene &sc, uint32_t ref, const Ray &r, float d_len2, float t_min, uint32_t inst,
                            TraceHit &best, uint32_t &best_node) {
  const uint32_t at = REF_INDEX(ref);
  if (!(FEATS & FEAT_QUADS) || REF_KIND(ref) == REF_SPHERES) {
    const float4 s = __ldg(reinterpret_cast<const float4 *>(&sc.msph[at]));
    float t;
    if (sphere_isect_closed(v3(s.x, s.y, s.z), s.w, r.o, r.d, d_len2, t_min, best.t, &t)) {
      if (t == best.t) {
        /* exact tie.  The reference's open interval (hittable.c:134-137) keeps the sphere it visited FIRST,
         * lets any quad win, and never accepts t == t_max itself; node ids are its visiting order. */
        if (best.prim == PRIM_NONE || ((FEATS & FEAT_QUADS) && REF_KIND(best.prim) != REF_SPHERES))
          return;
        if (__ldg(&sc.msph_meta[at].node) > __ldg(&sc.msph_meta[REF_INDEX(best.prim)].node))
          return;
      }
      best.t = t;
      best.prim = MAKE_REF(REF_SPHERES, 0, at);
      best.inst = inst;
    }
  } else {
    const MQuad *q = &sc.mquads[at];
    const float4 n_D = __ldg(&q->n_D);
    /* plane distance first: most candidates die here, before the other four rows are read */
    V3 nrm = v3(n_D.x, n_D.y, n_D.z);
    float denom = sv_dot(nrm, r.d);
    if (fabsf(denom) < 1e-8f)
      return;
    const float num = s_sub(n_D.w, sv_dot(nrm, r.o));
    if (quotient_outside(num, denom, t_min, best.t))
      return;
    float tt = s_div_z(num, denom);
    if (tt < t_min || tt > best.t)
      return;
    const float4 q_mat = __ldg(&q->q_mat), w_nod = __ldg(&q->w_nod), u_ar = __ldg(&q->u_ar), v_ = __ldg(&q->v_);
    V3 hp = sv_sub(sv_at(r.o, r.d, tt), v3(q_mat.x, q_mat.y, q_mat.z));
    V3 w = v3(w_nod.x, w_nod.y, w_nod.z);
    float al = sv_dot(w, sv_cross(hp, v3(v_.x, v_.y, v_.z)));
    float be = sv_dot(w, sv_cross(v3(u_ar.x, u_ar.y, u_ar.z), hp));
    if (al < 0.0f || al > 1.0f || be < 0.0f || be > 1.0f)
      return;
    const uint32_t node = __float_as_uint(w_nod.w);
    /* exact tie with an earlier candidate: the reference keeps the LATER-visited quad (closed interval,
     * hittable.c:194) and a quad always displaces a sphere at equal t; node ids are its visiting order */
    if (tt == best.t && best.prim != PRIM_NONE && REF_KIND(best.prim) == REF_QUADS && node < best_node)
      return;
    best.t = tt;
    best.prim = MAKE_REF(REF_QUADS, 0, at);
    best.inst = inst;
    best_node = node;
  }
}

/* State of one closest-hit query over a group.  A walker is always in one of three states: it holds an INNER node
 * (wants a two-box visit), it holds a LEAF (one primitive), or it is done (cur == REF_EMPTY).  The scheduling of
 * visit() and leaf() steps across the lanes of a warp is the caller's: trace_group() below for one ray per lane, the
 * megakernel's voted rounds, its warp ray queue for dynamic assignment.
 *
 * The whole group is ONE tree in the group's frame: primitives under Translate / RotateY nodes sit in it with boxes
 * in the group's frame (rtc_bvh.cu), and only their own test moves the ray into their frame, op by op like
 * Translate_hit / RotateY_hit (hittable.c:325-367) -- the hit distance keeps the reference's bits, the walk never
 * changes frame. */
extern __shared__ __align__(16) uint32_t rtc_dyn_smem[]; /* dynamic shared memory of the render kernels */

/* the four 32-byte rows of a WNode (render kernels' inlined loops only: see ldg_node on ptxas and 256-bit loads) */
RTC_DEV void ldg_wnode(const WNode *n, float4 &cx, float4 &hx, float4 &cy, float4 &hy, float4 &cz, float4 &hz,
                       uint4 &ch) {
  uint32_t a[8], b[8], c[8], d[8];
  const char *base = reinterpret_cast<const char *>(n);
#define RTC_LD256(dst, off)                                                                                          \
  asm("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"                                                         \
      : "=r"(dst[0]), "=r"(dst[1]), "=r"(dst[2]), "=r"(dst[3]), "=r"(dst[4]), "=r"(dst[5]), "=r"(dst[6]),            \
        "=r"(dst[7])                                                                                                  \
      : "l"(base + off))
  RTC_LD256(a, 0);
  RTC_LD256(b, 32);
  RTC_LD256(c, 64);
  RTC_LD256(d, 96);
#undef RTC_LD256
#define RTC_F4(v, o) make_float4(__uint_as_float(v[o]), __uint_as_float(v[o + 1]), __uint_as_float(v[o + 2]), __uint_as_float(v[o + 3]))
  cx = RTC_F4(a, 0), hx = RTC_F4(a, 4), cy = RTC_F4(b, 0), hy = RTC_F4(b, 4), cz = RTC_F4(c, 0), hz = RTC_F4(c, 4);
#undef RTC_F4
  ch = make_uint4(d[0], d[1], d[2], d[3]);
}
/* order (da, ra), (db, rb) by distance */
RTC_DEV void wide_cswap(float &da, uint32_t &ra, float &db, uint32_t &rb) {
  const bool sw = db < da;
  const float dl = sw ? db : da, dh = sw ? da : db;
  const uint32_t rl = sw ? rb : ra, rh = sw ? ra : rb;
  da = dl, db = dh, ra = rl, rb = rh;
}

template <int SK, int FEATS = FEAT_ALL, int NW = 2> struct WalkerT {
  static constexpr int kWidth = NW; /* 2: BNode tree (sc.bnodes); 4: WNode tree (sc.wnodes), render kernels only */
  uint32_t *stack;    /* TRACE_STACK entries owned by the caller (a separate local array, so that the scalar members
                         below stay in registers) */
  uint32_t sbase;     /* SK > 0: shared-window address of this lane's column of the CTA's stack (rtc_dyn_smem:
                         [entry][thread]); kept as a 32-bit address so that a push / pop is one LEA + STS / LDS */
  uint32_t sstride;   /*         bytes between two entries = 4 * threads per CTA */
  int sp;
  float d_len2;       /* |d|^2 of the group-frame ray in the reference's operation order: the `a` of its sphere tests */
  SlabRay slab;
  uint32_t cur;       /* subtree to look at next */
  uint32_t best_node; /* node id of the best quad (only consulted on exact ties) */
  TraceHit best;

  /* The ray itself and t_min are not kept here: the caller passes them to every leaf step (`outer` is a callable
   * returning the group-frame ray -- a register copy, or a reload from wherever the caller parked it). */
  RTC_DEV void push(uint32_t v) {
    if (SK > 0) {
      if (sp < SK)
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(sbase + (uint32_t)sp * sstride), "r"(v) : "memory");
      else
        stack[sp - SK] = v;
      sp++;
    } else {
      stack[sp++] = v;
    }
  }
  RTC_DEV void begin(uint32_t root, const Ray &ray, float tmax) {
    sp = 0;
    push(REF_EMPTY); /* the bottom entry ends the walk: pop() needs no emptiness test */
    d_len2 = sv_len2(ray.d), slab = slab_setup(ray);
    cur = root, best_node = 0;
    best.t = tmax, best.prim = PRIM_NONE, best.inst = RT_NONE;
  }
  /* a walk that was suspended (cur, sp, best kept; the stack is the caller's array): the per-ray constants again */
  RTC_DEV void resume(const Ray &ray) { d_len2 = sv_len2(ray.d), slab = slab_setup(ray); }
  RTC_DEV bool at_inner() const { return REF_KIND(cur) == REF_INNER; }
  RTC_DEV bool done() const { return cur == REF_EMPTY; }
  RTC_DEV uint32_t pop() { /* never called once done(): the bottom entry is REF_EMPTY */
    --sp;
    if (SK > 0) {
      if (sp >= SK)
        return stack[sp - SK];
      uint32_t v;
      asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(sbase + (uint32_t)sp * sstride) : "memory");
      return v;
    }
    return stack[sp];
  }

  /* one inner node: both children's boxes, nearer child first, the other one pushed */
  template <bool NONNEG> RTC_DEV void visit(const DScene &sc, float t_min) {
    /* an inner reference is its node index (kind 0, tag 0): no masking, one multiply-add for the address;
     * the 64-byte node arrives in two 256-bit loads */
    float4 xy0, xy1, zz;
    uint2 ch;
    ldg_node<NONNEG && RTC_NODE_WIDE>(sc.bnodes + cur, xy0, xy1, zz, ch); /* NONNEG = the render kernels' inlined loops */
    float e0, e1;
    const bool h0 = slab_test<NONNEG>(slab, xy0.x, xy0.y, xy0.z, xy0.w, zz.x, zz.y, t_min, best.t, e0);
    const bool h1 = slab_test<NONNEG>(slab, xy1.x, xy1.y, xy1.z, xy1.w, zz.z, zz.w, t_min, best.t, e1);
    /* three short independent steps (not an if / else-if chain): ptxas predicates them, the lanes of a warp do not
     * split three ways at every node */
    uint32_t next = h0 ? ch.x : ch.y;
    if (h0 && h1) {
      const bool swap = e1 < e0;
      next = swap ? ch.y : ch.x;
      push(swap ? ch.x : ch.y);
    }
    if (!(h0 || h1))
      next = pop();
    cur = next;
  }
  /* visit() for a caller that parks leaves (walk_round_sync<PARK>): when the child the walk would continue with is a
   * leaf and the parking slot is free, the leaf goes to the slot; if both children were hit the walk then continues
   * with the farther child directly -- no push, no pop -- else with the next stack entry. */
  template <bool NONNEG> RTC_DEV void visit_park(const DScene &sc, float t_min, uint32_t &parked) {
    float4 xy0, xy1, zz;
    uint2 ch;
    ldg_node<NONNEG && RTC_NODE_WIDE>(sc.bnodes + cur, xy0, xy1, zz, ch);
    float e0, e1;
    const bool h0 = slab_test<NONNEG>(slab, xy0.x, xy0.y, xy0.z, xy0.w, zz.x, zz.y, t_min, best.t, e0);
    const bool h1 = slab_test<NONNEG>(slab, xy1.x, xy1.y, xy1.z, xy1.w, zz.z, zz.w, t_min, best.t, e1);
    const bool both = h0 && h1, none = !(h0 || h1);
    const bool swap = both && e1 < e0;
    uint32_t next = (h0 && !swap) ? ch.x : ch.y; /* the only child hit, or the nearer of the two */
    const uint32_t far = swap ? ch.x : ch.y;
    const bool park = parked == REF_EMPTY && !none && REF_KIND(next) != REF_INNER;
    if (park)
      parked = next;
    if (both && !park)
      push(far);
    if (both && park)
      next = far;
    if (none || (park && !both))
      next = pop();
    cur = next;
  }
  /* One 4-wide node (WNode): four box tests, the children that are hit ordered by entry distance (a 5-comparator
   * network; a miss sorts as +inf), the nearest continued with, the others pushed farthest first.  PARK as in
   * visit_park: a nearest child that is a leaf goes to the free parking slot and the walk continues with the next. */
  template <bool NONNEG, bool PARK> RTC_DEV void visit_wide(const DScene &sc, float t_min, uint32_t &parked) {
    float4 cx, hx, cy, hy, cz, hz;
    uint4 ch;
    ldg_wnode(sc.wnodes + cur, cx, hx, cy, hy, cz, hz, ch);
    float d0, d1, d2, d3;
    const bool h0 = slab_test<NONNEG>(slab, cx.x, hx.x, cy.x, hy.x, cz.x, hz.x, t_min, best.t, d0);
    const bool h1 = slab_test<NONNEG>(slab, cx.y, hx.y, cy.y, hy.y, cz.y, hz.y, t_min, best.t, d1);
    const bool h2 = slab_test<NONNEG>(slab, cx.z, hx.z, cy.z, hy.z, cz.z, hz.z, t_min, best.t, d2) && ch.z != REF_EMPTY;
    const bool h3 = slab_test<NONNEG>(slab, cx.w, hx.w, cy.w, hy.w, cz.w, hz.w, t_min, best.t, d3) && ch.w != REF_EMPTY;
    /* (children 0 and 1 always exist: a wide node is made from a binary node's two children) */
    d0 = h0 ? d0 : RTC_INF, d1 = h1 ? d1 : RTC_INF, d2 = h2 ? d2 : RTC_INF, d3 = h3 ? d3 : RTC_INF;
    int n = (int)h0 + (int)h1 + (int)h2 + (int)h3;
    uint32_t r0 = ch.x, r1 = ch.y, r2 = ch.z, r3 = ch.w;
    wide_cswap(d0, r0, d1, r1), wide_cswap(d2, r2, d3, r3);
    wide_cswap(d0, r0, d2, r2), wide_cswap(d1, r1, d3, r3);
    wide_cswap(d1, r1, d2, r2);
    if (PARK) {
      const bool park = parked == REF_EMPTY && n > 0 && REF_KIND(r0) != REF_INNER;
      if (park) {
        parked = r0;
        r0 = r1, r1 = r2, r2 = r3;
        n--;
      }
    }
    if (n > 3)
      push(r3);
    if (n > 2)
      push(r2);
    if (n > 1)
      push(r1);
    uint32_t next = r0;
    if (n <= 0)
      next = pop();
    cur = next;
  }
  /* one leaf: the primitive is tested in its own frame */
  template <class Outer> RTC_DEV void leaf_test(const DScene &sc, uint32_t ref, float t_min, Outer outer) {
    const uint32_t tag = (FEATS & FEAT_INSTANCES) ? REF_TAG(ref) : 0u;
    Ray r = outer();
    float a = d_len2;
    if ((FEATS & FEAT_INSTANCES) && tag != 0u) {
      if (RTC_CONST_CHAINS && tag <= sc.n_const_chains) {
        /* instance 1 or 2, at most two steps: from the parameter bank (selected, not indexed: an indexed access
         * would make the compiler copy the parameters to local memory) */
        const bool first = tag == 1u;
        const float4 op0 = first ? sc.const_op[0][0] : sc.const_op[1][0];
        const float4 op1 = first ? sc.const_op[0][1] : sc.const_op[1][1];
        const uint32_t n_ops = first ? sc.const_count[0] : sc.const_count[1];
        r = chain_step(op0, r);
        if (n_ops > 1u)
          r = chain_step(op1, r);
      } else {
        const uint2 inst = __ldg(reinterpret_cast<const uint2 *>(&sc.insts[tag - 1u]));
        r = chain_apply(sc, inst.x, inst.y, r);
      }
      a = sv_len2(r.d);
    }
    /* a Box leaf is its six quads in list order, each against the running best (closed interval: a later face wins
     * an exact tie, as in HittableList_hit); ONE transform of the ray serves all six */
    uint32_t first = ref, count = 1u;
#ifndef RTC_EXP_NO_BOX
    if ((FEATS & FEAT_QUADS) && REF_KIND(ref) == REF_BOXES)
      first = ref ^ ((REF_BOXES ^ REF_QUADS) << 30), count = 6u;
#endif
    if (FEATS & FEAT_QUADS) {
#pragma unroll 1
      for (uint32_t k = 0; k < count; k++)
        leaf_intersect<FEATS>(sc, first + k, r, a, t_min, tag - 1u /* RT_NONE for tag 0 */, best, best_node);
    } else {
      leaf_intersect<FEATS>(sc, ref, r, a, t_min, RT_NONE, best, best_node);
    }
  }
  template <class Outer> RTC_DEV void leaf(const DScene &sc, float t_min, Outer outer) {
    leaf_test(sc, cur, t_min, outer);
    cur = pop();
  }
};
typedef WalkerT<0> Walker; /* the whole stack in the caller's local array */

/* The world group's oversized primitives (DScene.pre_ref), tested right after begin() of a walk over that group. */
template <class W, class Outer> RTC_DEV void walk_pre(const DScene &sc, W &w, float t_min, Outer outer) {
  if (sc.n_pre > 0u)
    w.leaf_test(sc, sc.pre_ref[0], t_min, outer);
  if (sc.n_pre > 1u)
    w.leaf_test(sc, sc.pre_ref[1], t_min, outer);
}

/* One round of a WARP-SYNCHRONOUS walk (all 32 lanes call it together, a lane without a ray with w.done()):
 * descend while at least `thr8` eighths of the warp's busy lanes hold an inner node, then the lanes that hold a leaf
 * process it.  thr8 = 0 is the classic "while-while" (descend while ANY lane holds an inner node): there one lane on
 * a long descent keeps every lane that already has its leaf waiting -- ncu on Book 1: descents are geometric with
 * mean 3.5 visits, the longest of 32 is 10.7, so 10.6 of 32 lanes are active in the node loop.  Leaving the loop
 * early costs nothing: a lane that is still descending just carries on in the next round. */
template <bool NONNEG, bool PARK, class W, class Outer>
RTC_DEV void walk_round_sync(const DScene &sc, W &w, float t_min, Outer outer, int thr8) {
  const unsigned busy = __ballot_sync(0xffffffffu, !w.done());
  const int thr = max(1, (__popc(busy) * thr8 + 7) >> 3);
  if (PARK) {
    /* PARK (speculative traversal): a lane that reaches a leaf parks it (one slot) and keeps descending towards its
     * next leaf while the loop runs, instead of idling until the vote ends the loop; the leaf phase tests the parked
     * leaf.  The parked leaf has not shrunk t_max yet, so the descent after it can visit nodes the leaf would have
     * culled: measured +3 % on the sphere scenes (Book 1) and on Cornell, -4 % on Book 2 final (media, quads, spheres
     * and a marble texture: the extra visits cost more than the idle lanes there), so the kernels of up to 256 threads
     * per CTA use it and the 512-thread lockstep kernel of the media scenes does not. */
    uint32_t parked = REF_EMPTY;
    if (!w.at_inner() && !w.done()) { /* the leaf the last round's pop left behind */
      parked = w.cur;
      w.cur = w.pop();
    }
    for (;;) {
      const bool in = w.at_inner();
      if (__popc(__ballot_sync(0xffffffffu, in)) < thr)
        break;
      if (in) { /* parking is part of the visit: no branch of its own */
        if (W::kWidth == 4) {
          w.template visit_wide<NONNEG, true>(sc, t_min, parked);
        } else {
          w.template visit_park<NONNEG>(sc, t_min, parked);
          if (RTC_VISITS_PER_VOTE > 1 && w.at_inner()) /* a second visit under the same vote (A/B: -DRTC_VISITS_PER_VOTE=2) */
            w.template visit_park<NONNEG>(sc, t_min, parked);
        }
      }
    }
    if (parked != REF_EMPTY)
      w.leaf_test(sc, parked, t_min, outer);
    return;
  }
  for (;;) {
    const bool in = w.at_inner();
    if (__popc(__ballot_sync(0xffffffffu, in)) < thr)
      break;
    if (in && W::kWidth == 4) {
      uint32_t none = ~REF_EMPTY; /* a parking slot that is never free */
      w.template visit_wide<NONNEG, false>(sc, t_min, none);
    } else if (in) { /* (RTC_VISITS_PER_VOTE = 2 halves the votes but measured slower: code size) */
      w.template visit<NONNEG>(sc, t_min);
      if (RTC_VISITS_PER_VOTE > 1 && w.at_inner())
        w.template visit<NONNEG>(sc, t_min);
    }
  }
  if (!w.at_inner() && !w.done())
    w.leaf(sc, t_min, outer);
}

/* Closest hit in the group rooted at `root` for the ray `r` (given in the group's frame), one ray per lane.
 * "while-while": all lanes of a warp first descend inner nodes together until each holds a leaf (or is done), then
 * all process leaves together -- keeps the 32 lanes in the same code far more often than one if/else per step. */
template <bool WORLD = false> /* WORLD: `root` is the world group's: its oversized primitives are tested first */
RTC_DEV TraceHit trace_group(const DScene &sc, uint32_t root, const Ray &r, float t_min, float t_max) {
  uint32_t stack[TRACE_STACK];
  Walker w;
  w.stack = stack;
  w.begin(root, r, t_max);
  auto outer = [&]() { return r; };
  if (WORLD)
    walk_pre(sc, w, t_min, outer);
  for (;;) {
    while (w.at_inner())
      w.visit<false>(sc, t_min);
    if (w.done())
      break;
    w.leaf(sc, t_min, outer);
  }
  return w.best;
}

static __device__ __noinline__ TraceHit trace_group_call(const DScene &sc, uint32_t root, const Ray &r, float t_min,
                                                   float t_max) {
  return trace_group(sc, root, r, t_min, t_max);
}

/* Sphere_hit (hittable.c:125-146) asked twice the way ConstantMedium_hit asks its boundary (hittable.c:396-401):
 * over (-inf, inf), then over (t_in + 0.0001, inf).  Same arithmetic, same accept/reject tests, evaluated once. */
RTC_DEV bool sphere_boundary(V3 center, float radius, V3 o, V3 d, float &t_in, float &t_out) {
  V3 oc = sv_sub(o, center);
  float a = sv_len2(d);
  float b = sv_dot(oc, d);
  float c = s_sub(sv_len2(oc), s_mul(radius, radius));
  float disc = s_sub(s_mul(b, b), s_mul(a, c));
  if (disc < 0.0f)
    return false;
  float sq = __fsqrt_rn(disc);
  const float r1 = s_div(s_sub(-b, sq), a), r2 = s_div(s_add(-b, sq), a);
  t_in = r1;
  if (r1 <= -RTC_INF || r1 >= RTC_INF) {
    t_in = r2;
    if (r2 <= -RTC_INF || r2 >= RTC_INF)
      return false;
  }
  const float t_min = s_add(t_in, 0.0001f);
  t_out = r1;
  if (r1 <= t_min || r1 >= RTC_INF) {
    t_out = r2;
    if (r2 <= t_min || r2 >= RTC_INF)
      return false;
  }
  return true;
}

/* node id (= the reference's visiting order, DFS pre-order) of the object a TraceHit names */
RTC_DEV uint32_t prim_node(const DScene &sc, uint32_t prim) {
  const uint32_t idx = REF_INDEX(prim);
  if (REF_KIND(prim) == REF_SPHERES)
    return __ldg(&sc.msph_meta[idx].node);
  if (REF_KIND(prim) == REF_QUADS)
    return __float_as_uint(__ldg(&sc.mquads[idx].w_nod).w);
  return sc.media[idx].node;
}

/* Closest surface hit in (t_min, t_max] among the primitives the reference visits BEFORE node `before` (node ids are
 * its visiting order): the t_max a list item at that position is called with (HittableList_hit, hittable.c:74-88).
 * +inf when there is none.  Rare (see trace_media), so out of line. */
static __device__ __noinline__ float trace_early_call(const DScene &sc, const Ray &r, float t_min, float t_max,
                                                       uint32_t before) {
  uint32_t stack[TRACE_STACK];
  Walker w;
  w.stack = stack;
  w.begin(sc.world_root, r, t_max);
  auto outer = [&]() { return r; };
  /* (no oversized primitives to look at here: rtc_bvh.cu only takes them out of the tree in worlds without media) */
  for (;;) {
    while (w.at_inner())
      w.visit<false>(sc, t_min);
    if (w.done())
      break;
    /* (a Box leaf names its first quad; its six faces are one list, all before or all after `before`) */
    if (prim_node(sc, REF_KIND(w.cur) == REF_BOXES ? w.cur ^ ((REF_BOXES ^ REF_QUADS) << 30) : w.cur) < before)
      w.leaf_test(sc, w.cur, t_min, outer);
    w.cur = w.pop();
  }
  return w.best.prim == PRIM_NONE ? RTC_INF : w.best.t;
}

/* ConstantMedium_hit (hittable.c:392-423) for every medium of the scene: two boundary queries, one draw when the
 * clipped interval is not empty.
 *
 * The reference asks a medium AT ITS LIST POSITION: its t_max is the closest hit among the items BEFORE it, and the
 * items after it are then tested against the medium's hit (HittableList_hit, hittable.c:74-88).  Here the surfaces
 * are all traced first (`best` = the closest of ALL of them), so the same decisions -- and the same number of random
 * draws -- are reproduced from that:
 *   - the closest surface is an EARLIER item (node id below the medium's): it is the reference's t_max.  Exact.
 *   - it is a LATER item at t_c: the reference's t_max T (closest earlier item) is only known to be >= t_c.  The draw
 *     happens iff entry < min(exit, T): certain when entry < min(exit, t_c); when the later item lies in front of the
 *     entry it takes one restricted query (is there an earlier item up to the entry?).  The accept test
 *     travel <= (min(exit, T) - entry) * |d| is certain when exit <= t_c or travel fits below t_c already; and when
 *     it is not, the outcome only matters if the medium's hit would beat the later item or another medium follows --
 *     then T is computed (trace_early_call).  The later item finally meets the medium's hit as its t_max: it wins
 *     iff t_c < t_m, or t_c == t_m for a quad (closed interval, hittable.c:194; spheres are open, :134-137).
 * Media under a BVHNode (in_list == 0; none in the reference's scenes) are clipped against `best` as before: the
 * node's box test decides in the reference whether they are asked at all, which a restricted query cannot tell. */
template <bool EXACT>
__device__ __noinline__ void trace_media(const DScene &sc, const Ray &world, float t_min, TraceHit &best, Rng &g) {
  float early_med_t = RTC_INF; /* closest medium hit the reference has accepted so far (even one that a later item beat) */
  /* at most one medium whose acceptance is still open: it has drawn, it cannot win against the later item that holds
   * the hit, and whether the reference accepted it (which needs its exact t_max, a restricted query) only matters if a
   * medium further down the list asks for ITS exact t_max -- resolved then, usually never */
  bool pend = false;
  float pend_tm = 0.0f, pend_in = 0.0f, pend_out = 0.0f, pend_travel = 0.0f, pend_len = 0.0f;
  uint32_t pend_node = 0u;
  auto resolve = [&]() {
    if (!pend)
      return;
    pend = false;
    const float cap = fminf(pend_out, fminf(trace_early_call(sc, world, t_min, RTC_INF, pend_node), early_med_t));
    if (!(pend_travel > s_mul(s_sub(cap, pend_in), pend_len)))
      early_med_t = fminf(early_med_t, pend_tm);
  };
  for (uint32_t m = 0; m < sc.n_media; m++) {
    const MMedium *med = &sc.media[m];
    Ray r = chain_apply_call(sc, med->op_begin, med->op_count, world);
    TraceHit in, out;
    const uint32_t root = med->root;
    if (RTC_MEDIA_FAST && REF_KIND(root) == REF_SPHERES && REF_TAG(root) == 0u) {
      /* boundary = one sphere (the reference's fog and smoke balls): both queries share b, c and the square root */
      const float4 s = __ldg(reinterpret_cast<const float4 *>(&sc.msph[REF_INDEX(root)]));
      if (!sphere_boundary(v3(s.x, s.y, s.z), s.w, r.o, r.d, in.t, out.t))
        continue;
    } else {
      in = trace_group_call(sc, root, r, -RTC_INF, RTC_INF);
      if (in.prim == PRIM_NONE)
        continue;
      out = trace_group_call(sc, root, r, s_add(in.t, 0.0001f), RTC_INF);
      if (out.prim == PRIM_NONE)
        continue;
    }
    /* is the closest surface a LATER list item than this medium?  Asked lazily: it costs a dependent load of the
     * primitive's node id, and the common outcomes below do not depend on it. */
    auto is_late = [&]() {
      return RTC_MEDIA_INLIST && med->in_list != 0u && best.prim != PRIM_NONE && prim_node(sc, best.prim) > med->node;
    };
    float in_t = fmaxf(in.t, t_min);
    float out_t = fminf(out.t, best.t); /* the reference's clipped exit if `best` is an earlier item, else a LOWER bound */
    if (in_t >= out.t)
      continue; /* empty whatever the reference's t_max is: no draw on either side */
    bool late = false, late_known = false, sure = true;
    if (in_t >= out_t) {
      /* `best` ends the interval before the entry.  An earlier item: no draw.  A later item: the reference still
       * draws, unless an EARLIER item (or an earlier medium's hit) ends the interval before the entry too */
      if (med->boundary_listed)
        continue; /* the boundary's own surface is an earlier item: it (or something nearer) ends the interval at in_t */
      late = is_late(), late_known = true;
      if (!late)
        continue;
      if (pend && pend_tm <= in_t)
        resolve();
      /* (the query's upper end is open for spheres, like Sphere_hit's: ask one ulp further and compare afterwards --
       * the commonest occluder is the boundary's own surface, listed as an item before its medium, at exactly in_t) */
      if (fminf(trace_early_call(sc, world, t_min, __uint_as_float(__float_as_uint(in_t) + 1u), med->node),
                early_med_t) <= in_t)
        continue;
      sure = false;
    }
    in_t = in_t > 0.0f ? in_t : 0.0f;
    float ray_len = __fsqrt_rn(sv_len2(r.d));
    float travel = s_mul(med->neg_inv_density, m_log<EXACT>(rng_f32(g)));
    float t_m = s_add(in_t, s_div(travel, ray_len));
    if (sure && travel > s_mul(s_sub(out_t, in_t), ray_len)) {
      /* does not fit below `best`: rejected if best is an earlier item (or the boundary's own exit); if it is a later
       * item at t_c < exit the reference's t_max lies in [t_c, exit] and may still admit the hit */
      if (out.t <= best.t)
        continue;
      if (t_m > best.t && m + 1 >= sc.n_media)
        continue; /* `best` keeps the hit either way, and nobody is asked after this medium */
      late = is_late(), late_known = true;
      if (!late)
        continue;
      sure = false;
    }
    if (!sure) {
      const bool beats = t_m < best.t || (t_m == best.t && REF_KIND(best.prim) == REF_SPHERES);
      if (!beats) {
        /* the later item keeps the hit whatever the reference decided: leave the decision open (see `pend`) */
        if (m + 1 < sc.n_media) {
          resolve();
          pend = true, pend_tm = t_m, pend_in = in_t, pend_out = out.t, pend_travel = travel, pend_len = ray_len;
          pend_node = med->node;
        }
        continue;
      }
      resolve();
      out_t = fminf(out.t, fminf(trace_early_call(sc, world, t_min, RTC_INF, med->node), early_med_t));
      if (travel > s_mul(s_sub(out_t, in_t), ray_len))
        continue;
    }
    /* accepted by the reference */
    early_med_t = fminf(early_med_t, t_m);
    if (t_m >= best.t && best.prim != PRIM_NONE) {
      /* (only by rounding, or behind a later item) an earlier item is overridden by the medium's hit; a later item is
       * tested with t_max = t_m afterwards and wins unless it is a sphere at exactly t_m (hittable.c:80-84, 134-137, 194) */
      if (!late_known)
        late = is_late();
      if (late && !(t_m == best.t && REF_KIND(best.prim) == REF_SPHERES))
        continue;
    }
    best.t = t_m;
    best.prim = MAKE_REF(PRIM_MEDIUM_KIND, 0, m);
    best.inst = RT_NONE;
  }
}

/* Full hit record (world frame) of a TraceHit.  WANT_UV = false skips the sphere's atan2f/acosf. */
template <bool EXACT, int FEATS = FEAT_ALL>
RTC_DEV void trace_finalize(const DScene &sc, const Ray &world, const TraceHit &h, Hit &rec) {
  rec.t = h.t;
  const uint32_t kind = REF_KIND(h.prim), idx = REF_INDEX(h.prim);
  if ((FEATS & FEAT_MEDIA) && kind == PRIM_MEDIUM_KIND) { /* only t, p, material (hittable.c:418-422) */
    const MMedium *med = &sc.media[idx];
    Ray r = chain_apply_call(sc, med->op_begin, med->op_count, world);
    rec.p = sv_at(r.o, r.d, h.t);
    rec.n = v3(0, 0, 0);
    chain_unapply_call(sc, med->op_begin, med->op_count, rec.p, rec.n);
    rec.material = med->material;
    rec.node = med->node;
    rec.u = rec.v = 0.0f;
    rec.front = 0;
    return;
  }
  Ray r = world;
  uint32_t op_begin = 0, op_count = 0;
  if ((FEATS & FEAT_INSTANCES) && h.inst != RT_NONE) {
    const uint2 inst = __ldg(reinterpret_cast<const uint2 *>(&sc.insts[h.inst]));
    op_begin = inst.x, op_count = inst.y;
    r = chain_apply_call(sc, op_begin, op_count, world);
  }
  if (!(FEATS & FEAT_QUADS) || kind == REF_SPHERES) {
    const float4 s = __ldg(reinterpret_cast<const float4 *>(&sc.msph[idx]));
    const uint2 meta = __ldg(reinterpret_cast<const uint2 *>(&sc.msph_meta[idx]));
    rec.material = meta.x;
    rec.node = meta.y;
    rec.u = rec.v = 0.0f;
    sphere_record<EXACT, false>(v3(s.x, s.y, s.z), s.w, r.o, r.d, h.t, rec);
    /* (u, v) cost an atan2 and an acos: only for textures that read them */
    if (EXACT || ((FEATS & FEAT_TEXTURES) && texture_needs_uv(sc, __ldg(&sc.materials[meta.x].texture))))
      sphere_uv<EXACT>(rec.front ? rec.n : v_neg(rec.n), rec);
  } else {
    const MQuad *q = &sc.mquads[idx];
    const float4 n_D = __ldg(&q->n_D), q_mat = __ldg(&q->q_mat), w_nod = __ldg(&q->w_nod), u_ar = __ldg(&q->u_ar),
                 v_ = __ldg(&q->v_);
    V3 nrm = v3(n_D.x, n_D.y, n_D.z), w = v3(w_nod.x, w_nod.y, w_nod.z);
    rec.p = sv_at(r.o, r.d, h.t);
    V3 hp = sv_sub(rec.p, v3(q_mat.x, q_mat.y, q_mat.z));
    rec.u = sv_dot(w, sv_cross(hp, v3(v_.x, v_.y, v_.z)));
    rec.v = sv_dot(w, sv_cross(v3(u_ar.x, u_ar.y, u_ar.z), hp));
    rec.material = __float_as_uint(q_mat.w);
    rec.node = __float_as_uint(w_nod.w);
    rec.front = sv_dot(r.d, nrm) < 0.0f;
    rec.n = rec.front ? nrm : v_neg(nrm);
  }
  if ((FEATS & FEAT_INSTANCES) && op_count)
    chain_unapply_call(sc, op_begin, op_count, rec.p, rec.n);
}

/* The integrator's query (raytracing.c:45): surfaces, then media. */
template <bool EXACT> RTC_DEV TraceHit trace_world(const DScene &sc, const Ray &world, float t_min, float t_max, Rng &g) {
  TraceHit best = trace_group<true>(sc, sc.world_root, world, t_min, t_max);
  if (sc.n_media)
    trace_media<EXACT>(sc, world, t_min, best, g);
  return best;
}

#endif /* RTC_TRACE_CUH */
