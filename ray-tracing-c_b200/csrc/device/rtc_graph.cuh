/*
 * rtc_graph.cuh -- closest hit over the REFERENCE-ORDER graph (the rt_node array exactly as flattened), as an
 * explicit-stack machine.  It visits nodes in the order the reference's recursive vtable calls do
 * (/root/reference/src/hittable.c:74-88 list, :266-277 BVH, :325-334 Translate, :354-367 RotateY, :392-423 medium),
 * so ties, shrinking t_max, the stale fields a medium hit leaves behind and the position of every random draw
 * are the reference's.  Used by the exact kernel (parity anchor); the throughput kernels walk the rebuilt BVH.
 *
 * Limits (checked on the host by rt_scene_create): call depth <= GRAPH_MAX_FRAMES, nested instance transforms
 * <= GRAPH_MAX_RAYS - 1, media nested inside media boundaries <= GRAPH_MAX_RECS - 1.
 */
#ifndef RTC_GRAPH_CUH
#define RTC_GRAPH_CUH

#include "rtc_shade.cuh"

#define GRAPH_MAX_FRAMES 64
#define GRAPH_MAX_RAYS 6
#define GRAPH_MAX_RECS 3

/* hittable.c:38-55, op for op (1/d then multiply; fmaxf/fminf ignore NaNs; early exit per axis) */
RTC_DEV bool aabb_hit_ref(const float *bbox, V3 o, V3 d, float t_min, float t_max) {
  const float oo[3] = {o.x, o.y, o.z}, dd[3] = {d.x, d.y, d.z};
#pragma unroll
  for (int a = 0; a < 3; a++) {
    float inv = s_div(1.0f, dd[a]);
    float t0 = s_mul(s_sub(__ldg(&bbox[2 * a]), oo[a]), inv);
    float t1 = s_mul(s_sub(__ldg(&bbox[2 * a + 1]), oo[a]), inv);
    if (inv < 0.0f) {
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

template <bool EXACT>
__device__ __noinline__ bool graph_hit(const DScene &sc, uint32_t root, V3 ray_o, V3 ray_d, float t_min0, float t_max0,
                                       Hit &out, Rng &g) {
  uint32_t f_node[GRAPH_MAX_FRAMES], f_idx[GRAPH_MAX_FRAMES], f_flags[GRAPH_MAX_FRAMES];
  float f_tmin[GRAPH_MAX_FRAMES], f_tmax[GRAPH_MAX_FRAMES];
  V3 r_o[GRAPH_MAX_RAYS], r_d[GRAPH_MAX_RAYS];
  Hit rec[GRAPH_MAX_RECS];
  float med_in[GRAPH_MAX_RECS];

  /* flags: bit 0 = "some child hit", bits 8-15 ray level, bits 16-23 record level */
#define PUSH(node_, tmin_, tmax_, rl_, cl_)                                                                           \
  do {                                                                                                                \
    sp++;                                                                                                             \
    f_node[sp] = (node_), f_idx[sp] = 0, f_tmin[sp] = (tmin_), f_tmax[sp] = (tmax_);                                  \
    f_flags[sp] = ((uint32_t)(rl_) << 8) | ((uint32_t)(cl_) << 16);                                                   \
  } while (0)

  r_o[0] = ray_o, r_d[0] = ray_d;
  rec[0] = out;
  int sp = -1;
  PUSH(root, t_min0, t_max0, 0, 0);
  bool ret = false;

  while (sp >= 0) {
    const uint32_t id = f_node[sp];
    const rt_node *n = &sc.nodes[id];
    const uint32_t kind = __ldg(&n->kind);
    const uint32_t rl = (f_flags[sp] >> 8) & 255u, cl = (f_flags[sp] >> 16) & 255u;
    const float t_min = f_tmin[sp], t_max = f_tmax[sp];

    if (kind == RT_NODE_SPHERE) {
      ret = sphere_hit<EXACT>(sc.spheres[__ldg(&n->a)], r_o[rl], r_d[rl], t_min, t_max, rec[cl]);
      sp--;
    } else if (kind == RT_NODE_QUAD) {
      ret = quad_hit(sc.quads[__ldg(&n->a)], r_o[rl], r_d[rl], t_min, t_max, rec[cl]);
      sp--;
    } else if (kind == RT_NODE_LIST) {
      uint32_t idx = f_idx[sp];
      if (idx > 0 && ret) { /* back from item idx-1 with a hit: shrink the interval */
        f_tmax[sp] = rec[cl].t;
        f_flags[sp] |= 1u;
      }
      if (idx < __ldg(&n->b)) {
        f_idx[sp] = idx + 1;
        uint32_t child = __ldg(&sc.children[__ldg(&n->a) + idx]);
        float tmax_now = f_tmax[sp];
        PUSH(child, t_min, tmax_now, rl, cl);
      } else {
        ret = (f_flags[sp] & 1u) != 0u;
        sp--;
      }
    } else if (kind == RT_NODE_BVH) {
      uint32_t idx = f_idx[sp];
      if (idx == 0) {
        if (!aabb_hit_ref(n->bbox, r_o[rl], r_d[rl], t_min, t_max)) {
          ret = false;
          sp--;
        } else {
          f_idx[sp] = 1;
          PUSH(__ldg(&n->a), t_min, t_max, rl, cl);
        }
      } else if (idx == 1) {
        if (ret) {
          f_tmax[sp] = rec[cl].t;
          f_flags[sp] |= 1u;
        }
        f_idx[sp] = 2;
        float tmax_now = f_tmax[sp];
        PUSH(__ldg(&n->b), t_min, tmax_now, rl, cl);
      } else {
        ret = ret || (f_flags[sp] & 1u) != 0u;
        sp--;
      }
    } else if (kind == RT_NODE_TRANSLATE) {
      V3 offset = v3(__ldg(&n->f[0]), __ldg(&n->f[1]), __ldg(&n->f[2]));
      if (f_idx[sp] == 0) {
        r_o[rl + 1] = sv_sub(r_o[rl], offset);
        r_d[rl + 1] = r_d[rl];
        f_idx[sp] = 1;
        PUSH(__ldg(&n->a), t_min, t_max, rl + 1, cl);
      } else {
        if (ret)
          rec[cl].p = sv_add(rec[cl].p, offset);
        sp--;
      }
    } else if (kind == RT_NODE_ROTATE_Y) {
      float sn = __ldg(&n->f[0]), cs = __ldg(&n->f[1]);
      if (f_idx[sp] == 0) {
        r_o[rl + 1] = sv_rot_y(r_o[rl], cs, sn);
        r_d[rl + 1] = sv_rot_y(r_d[rl], cs, sn);
        f_idx[sp] = 1;
        PUSH(__ldg(&n->a), t_min, t_max, rl + 1, cl);
      } else {
        if (ret) {
          rec[cl].p = sv_rot_y_inv(rec[cl].p, cs, sn);
          rec[cl].n = sv_rot_y_inv(rec[cl].n, cs, sn);
        }
        sp--;
      }
    } else if (kind == RT_NODE_MEDIUM) {
      uint32_t idx = f_idx[sp];
      if (idx == 0) {
        f_idx[sp] = 1;
        PUSH(__ldg(&n->a), -RTC_INF, RTC_INF, rl, cl + 1);
      } else if (idx == 1) {
        if (!ret) {
          sp--;
        } else {
          med_in[cl] = rec[cl + 1].t;
          f_idx[sp] = 2;
          PUSH(__ldg(&n->a), s_add(med_in[cl], 0.0001f), RTC_INF, rl, cl + 1);
        }
      } else {
        if (ret) {
          ret = false;
          float in_t = fmaxf(med_in[cl], t_min);
          float out_t = fminf(rec[cl + 1].t, t_max);
          if (!(in_t >= out_t)) {
            in_t = in_t > 0.0f ? in_t : 0.0f;
            float ray_len = __fsqrt_rn(sv_len2(r_d[rl]));
            float inside = s_mul(s_sub(out_t, in_t), ray_len);
            float travel = s_mul(__ldg(&n->f[0]), m_log<EXACT>(rng_f32(g)));
            if (!(travel > inside)) {
              rec[cl].t = s_add(in_t, s_div(travel, ray_len));
              rec[cl].p = sv_at(r_o[rl], r_d[rl], rec[cl].t);
              rec[cl].material = __ldg(&n->b);
              rec[cl].node = id;
              ret = true;
            }
          }
        }
        sp--;
      }
    } else {
      ret = false;
      sp--;
    }
  }
#undef PUSH
  out = rec[0];
  return ret;
}

#endif /* RTC_GRAPH_CUH */
