/*
 * rtc_bvh.cu -- validation of an rt_scene_desc and construction of the rebuilt acceleration structure
 * (see rtc_bvh.h).  Host code only.
 */
#include "rtc_bvh.h"
#include "rtc_graph.cuh" /* the graph-machine limits */
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <set>

/* ------------------------------------------------------------------ validation --------------------------- */

std::string rtc_validate_desc(const rt_scene_desc &d) {
  if (d.abi_version != RT_ABI_VERSION)
    return "abi_version mismatch";
  if (d.n_nodes == 0 || d.nodes == nullptr)
    return "scene without nodes";
  if ((d.n_children && !d.children) || (d.n_spheres && !d.spheres) || (d.n_quads && !d.quads) ||
      (d.n_materials && !d.materials) || (d.n_textures && !d.textures) || (d.n_perlins && !d.perlins) ||
      (d.n_images && !d.images) || (d.n_texel_bytes && !d.texels))
    return "NULL array with a non-zero count";
  if (d.objects_root >= d.n_nodes)
    return "objects_root out of range";
  if (d.lights_root != RT_NONE && d.lights_root >= d.n_nodes)
    return "lights_root out of range";
  if (d.n_nodes > 0x1FFFFFFu || d.n_spheres > REF_MAX_INDEX || d.n_quads > REF_MAX_INDEX)
    return "scene too large for 22-bit primitive references";

  for (uint32_t i = 0; i < d.n_textures; i++) {
    const rt_texture &t = d.textures[i];
    switch (t.kind) {
    case RT_TEX_SOLID: break;
    case RT_TEX_CHECKER:
      if (t.a >= d.n_textures || t.b >= d.n_textures)
        return "checker texture child out of range";
      if (t.a <= i || t.b <= i) /* the flattener emits children after parents: also rules out cycles */
        return "checker texture children must follow their parent";
      break;
    case RT_TEX_IMAGE: {
      if (t.a >= d.n_images)
        return "image index out of range";
      const rt_image &im = d.images[t.a];
      if (im.width <= 0 || im.height <= 0 || im.offset + (uint64_t)im.width * im.height * 3 > d.n_texel_bytes)
        return "image texels out of range";
      break;
    }
    case RT_TEX_PERLIN:
      if (t.a >= d.n_perlins)
        return "perlin index out of range";
      if (t.b > 64)
        return "perlin depth out of range";
      for (int k = 0; k < 256; k++) {
        const rt_perlin &p = d.perlins[t.a];
        if ((uint32_t)(p.perm_x[k] ^ p.perm_y[k] ^ p.perm_z[k]) > 255u || (uint32_t)p.perm_x[k] > 255u ||
            (uint32_t)p.perm_y[k] > 255u || (uint32_t)p.perm_z[k] > 255u)
          return "perlin permutation entry out of range";
      }
      break;
    default: return "unknown texture kind";
    }
  }
  for (uint32_t i = 0; i < d.n_materials; i++) {
    const rt_material &m = d.materials[i];
    if (m.tag > RT_MAT_ISOTROPIC)
      return "unknown material tag";
    bool needs_tex = m.tag == RT_MAT_LAMBERTIAN || m.tag == RT_MAT_METAL || m.tag == RT_MAT_DIFFUSE_LIGHT ||
                     m.tag == RT_MAT_ISOTROPIC;
    if (needs_tex && m.texture >= d.n_textures)
      return "material texture out of range";
  }
  for (uint32_t i = 0; i < d.n_spheres; i++)
    if (d.spheres[i].material >= d.n_materials || d.spheres[i].node >= d.n_nodes)
      return "sphere material/node out of range";
  for (uint32_t i = 0; i < d.n_quads; i++)
    if (d.quads[i].material >= d.n_materials || d.quads[i].node >= d.n_nodes)
      return "quad material/node out of range";

  for (uint32_t i = 0; i < d.n_nodes; i++) {
    const rt_node &n = d.nodes[i];
    switch (n.kind) {
    case RT_NODE_LIST:
      if ((uint64_t)n.a + n.b > d.n_children)
        return "list children out of range";
      for (uint32_t k = 0; k < n.b; k++)
        if (d.children[n.a + k] >= d.n_nodes)
          return "list item out of range";
      break;
    case RT_NODE_SPHERE:
      if (n.a >= d.n_spheres)
        return "sphere index out of range";
      break;
    case RT_NODE_QUAD:
      if (n.a >= d.n_quads)
        return "quad index out of range";
      break;
    case RT_NODE_BVH:
      if (n.a >= d.n_nodes || n.b >= d.n_nodes)
        return "bvh child out of range";
      break;
    case RT_NODE_TRANSLATE:
    case RT_NODE_ROTATE_Y:
      if (n.a >= d.n_nodes)
        return "instance child out of range";
      break;
    case RT_NODE_MEDIUM:
      if (n.a >= d.n_nodes || n.b >= d.n_materials)
        return "medium boundary/material out of range";
      break;
    default: return "unknown node kind";
    }
  }

  /* acyclic: iterative DFS with colours */
  std::vector<uint8_t> colour(d.n_nodes, 0);
  std::function<bool(uint32_t, int)> dfs = [&](uint32_t id, int depth) -> bool {
    if (depth > 4096 || colour[id] == 1)
      return false;
    if (colour[id] == 2)
      return true;
    colour[id] = 1;
    const rt_node &n = d.nodes[id];
    bool ok = true;
    switch (n.kind) {
    case RT_NODE_LIST:
      for (uint32_t k = 0; k < n.b && ok; k++)
        ok = dfs(d.children[n.a + k], depth + 1);
      break;
    case RT_NODE_BVH: ok = dfs(n.a, depth + 1) && dfs(n.b, depth + 1); break;
    case RT_NODE_TRANSLATE:
    case RT_NODE_ROTATE_Y:
    case RT_NODE_MEDIUM: ok = dfs(n.a, depth + 1); break;
    default: break;
    }
    colour[id] = 2;
    return ok;
  };
  if (!dfs(d.objects_root, 0))
    return "cycle (or absurd depth) in the Hittable graph";
  if (d.lights_root != RT_NONE && !dfs(d.lights_root, 0))
    return "cycle in the lights graph";
  return "";
}

/* ------------------------------------------------------------------ graph-machine limits ----------------- */

struct Depths {
  int frames, rays, recs;
};

static Depths graph_depths(const rt_scene_desc &d, uint32_t id, std::vector<Depths> &memo) {
  if (memo[id].frames)
    return memo[id];
  const rt_node &n = d.nodes[id];
  Depths r = {1, 0, 0};
  auto absorb = [&](uint32_t child, int extra_ray, int extra_rec) {
    Depths c = graph_depths(d, child, memo);
    r.frames = std::max(r.frames, 1 + c.frames);
    r.rays = std::max(r.rays, extra_ray + c.rays);
    r.recs = std::max(r.recs, extra_rec + c.recs);
  };
  switch (n.kind) {
  case RT_NODE_LIST:
    for (uint32_t k = 0; k < n.b; k++)
      absorb(d.children[n.a + k], 0, 0);
    break;
  case RT_NODE_BVH:
    absorb(n.a, 0, 0);
    absorb(n.b, 0, 0);
    break;
  case RT_NODE_TRANSLATE:
  case RT_NODE_ROTATE_Y: absorb(n.a, 1, 0); break;
  case RT_NODE_MEDIUM: absorb(n.a, 0, 1); break;
  default: break;
  }
  memo[id] = r;
  return r;
}

/* nesting of lists under the lights root; -1 if some reachable list has no item to sample (would never return) */
static int light_nesting(const rt_scene_desc &d, uint32_t id) {
  const rt_node &n = d.nodes[id];
  if (n.kind != RT_NODE_LIST)
    return 0;
  int deepest = 0, usable = 0;
  for (uint32_t k = 0; k < n.b; k++) {
    const rt_node &c = d.nodes[d.children[n.a + k]];
    if (!(c.flags & RT_NODE_HAS_PDF))
      continue;
    usable++;
    int sub = light_nesting(d, d.children[n.a + k]);
    if (sub < 0)
      return -1;
    deepest = std::max(deepest, sub);
  }
  if (usable == 0)
    return -1;
  return 1 + deepest;
}

/* ------------------------------------------------------------------ the rebuilt structure ---------------- */

namespace {

struct Box {
  float lo[3], hi[3];
  void reset() {
    for (int a = 0; a < 3; a++)
      lo[a] = FLT_MAX, hi[a] = -FLT_MAX;
  }
  void grow(const Box &b) {
    for (int a = 0; a < 3; a++)
      lo[a] = std::min(lo[a], b.lo[a]), hi[a] = std::max(hi[a], b.hi[a]);
  }
  float area() const {
    float e[3] = {hi[0] - lo[0], hi[1] - lo[1], hi[2] - lo[2]};
    if (e[0] < 0 || e[1] < 0 || e[2] < 0)
      return 0.0f;
    return 2.0f * (e[0] * e[1] + e[1] * e[2] + e[2] * e[0]);
  }
};

struct Item {
  uint32_t kind;  /* REF_SPHERES / REF_QUADS / REF_BOXES */
  uint32_t index; /* desc sphere / desc quad / first slot in desc.children of the Box's six quads */
  uint32_t tag;   /* 0: in the group's own frame; k: under out.insts[k-1]'s transform chain */
  Box box;        /* in the group's frame */
};

/* SAH constants: one inner visit (two box tests) = 1; tunable through the environment for experiments */
static float env_float(const char *name, float fallback) {
  const char *v = getenv(name);
  return v && *v ? (float)atof(v) : fallback;
}
static const float kCostNode = 1.0f, kCostSphere = env_float("RT_BVH_COST_SPHERE", 0.8f),
                   kCostQuad = env_float("RT_BVH_COST_QUAD", 1.4f), kCostBox = env_float("RT_BVH_COST_BOX", 4.0f);
/* RT_BVH_BOXES=0: every quad its own leaf (the round-1 structure), for A/B measurements; read at every build */
static bool box_leaves_enabled() { return env_float("RT_BVH_BOXES", 0.0f) != 0.0f; }

struct Builder {
  const rt_scene_desc &d;
  BuiltScene &out;
  const bool kBoxLeaves = box_leaves_enabled();
  std::vector<Box> child_boxes; /* 2 per binary node: the exact (lo, hi) boxes the (centre, half) forms were made from */
  Builder(const rt_scene_desc &desc, BuiltScene &o) : d(desc), out(o) {}

  static Box node_box(const rt_node &n) {
    Box b;
    for (int a = 0; a < 3; a++)
      b.lo[a] = n.bbox[2 * a], b.hi[a] = n.bbox[2 * a + 1];
    return b;
  }
  static float item_cost(uint32_t kind) {
    return kind == REF_SPHERES ? kCostSphere : kind == REF_BOXES ? kCostBox : kCostQuad;
  }

  void unsupported(const std::string &why) {
    if (out.mega_supported) {
      out.mega_supported = false;
      out.mega_why = why;
    }
  }

  /* object -> enclosing frame for one point, applied from the innermost op outwards (the inverse of what
   * Translate_hit / RotateY_hit do to the ray, hittable.c:325-367) */
  static void point_to_outer(float p[3], const std::vector<ChainOp> &chain) {
    for (size_t k = chain.size(); k-- > 0;) {
      const ChainOp &op = chain[k];
      if (!op.is_rotate) {
        p[0] += op.f0, p[1] += op.f1, p[2] += op.f2;
      } else {
        const float sn = op.f0, cs = op.f1;
        const float x = p[0], z = p[2];
        p[0] = cs * x + sn * z;
        p[2] = -sn * x + cs * z;
      }
    }
  }
  /* re-derived in float: a hit the reference computes in the object's frame lies within a few ulps of the
   * coordinates of the transformed primitive, the pad is two orders of magnitude above that */
  static void pad_box(Box &b) {
    for (int a = 0; a < 3; a++) {
      float pad = 1e-4f + 1e-5f * std::max(std::fabs(b.lo[a]), std::fabs(b.hi[a]));
      b.lo[a] -= pad, b.hi[a] += pad;
    }
  }
  /* box of an axis-aligned box through a chain (media bounds): its 8 corners, the way the reference bounds a
   * RotateY (hittable.c:369-389) */
  Box chain_to_outer(const Box &b, const std::vector<ChainOp> &chain) const {
    Box r;
    r.reset();
    for (int c = 0; c < 8; c++) {
      float p[3] = {(c & 1) ? b.hi[0] : b.lo[0], (c & 2) ? b.hi[1] : b.lo[1], (c & 4) ? b.hi[2] : b.lo[2]};
      point_to_outer(p, chain);
      Box q;
      for (int a = 0; a < 3; a++)
        q.lo[a] = q.hi[a] = p[a];
      r.grow(q);
    }
    pad_box(r);
    return r;
  }
  /* tight bounds of ONE transformed primitive in the group's frame: a sphere stays a sphere (RotateY and Translate
   * are rigid), a quad is the hull of its four transformed corners */
  Box sphere_box(const rt_sphere &s, const std::vector<ChainOp> &chain) const {
    float c[3] = {s.center[0], s.center[1], s.center[2]};
    point_to_outer(c, chain);
    Box b;
    const float r = std::fabs(s.radius);
    for (int a = 0; a < 3; a++)
      b.lo[a] = c[a] - r, b.hi[a] = c[a] + r;
    pad_box(b);
    return b;
  }
  Box quad_box(const rt_quad &q, const std::vector<ChainOp> &chain) const {
    Box b;
    b.reset();
    for (int c = 0; c < 4; c++) {
      float p[3];
      for (int a = 0; a < 3; a++)
        p[a] = q.Q[a] + ((c & 1) ? q.u[a] : 0.0f) + ((c & 2) ? q.v[a] : 0.0f);
      point_to_outer(p, chain);
      Box pt;
      for (int a = 0; a < 3; a++)
        pt.lo[a] = pt.hi[a] = p[a];
      b.grow(pt);
    }
    pad_box(b);
    return b;
  }

  /* Collects the primitives under `id` into `items` (the group being assembled).  `chain` holds the transforms
   * between the group's frame and the current node and `tag` names that chain (0 = none): a Translate / RotateY node
   * does not become a node of the tree -- its primitives join the group's ONE tree with boxes in the group's frame,
   * each remembering the chain its own test has to apply to the ray.  (The reference never bounds its instances:
   * Translate_hit / RotateY_hit always enter the child, hittable.c:325-367.) */
  static constexpr uint32_t kTagPending = 0xFFFFFFFFu; /* a chain whose Instance is made by its first primitive */
  int bvh_above = 0;               /* BVHNodes between the group's root and the node collect() is looking at */
  std::set<uint32_t> media_in_bvh; /* media met below a BVHNode (see MMedium::in_list) */
  uint32_t new_instance(const std::vector<ChainOp> &chain) {
    if (out.insts.size() >= REF_MAX_TAG)
      unsupported("more than 255 distinct Translate/RotateY chains in the scene");
    Instance inst;
    inst.op_begin = (uint32_t)out.ops.size();
    inst.op_count = (uint32_t)chain.size();
    inst.pad0 = inst.pad1 = 0;
    out.ops.insert(out.ops.end(), chain.begin(), chain.end());
    out.insts.push_back(inst);
    return std::min<uint32_t>((uint32_t)out.insts.size(), REF_MAX_TAG);
  }
  void collect(uint32_t id, std::vector<Item> &items, std::vector<ChainOp> &chain, uint32_t &tag,
               std::set<uint32_t> &seen, bool in_boundary,
               std::vector<std::pair<uint32_t, std::vector<ChainOp>>> &media) {
    const rt_node &n = d.nodes[id];
    switch (n.kind) {
    case RT_NODE_LIST: {
      /* a list of exactly six quads none of which was met before: what Box_new builds (hittable.c:246-264).  ONE leaf
       * (any such list may be treated so: the leaf tests its quads in list order like HittableList_hit). */
      bool six_quads = kBoxLeaves && n.b == 6 && out.mquads.size() + 6 < REF_MAX_INDEX;
      for (uint32_t k = 0; six_quads && k < 6; k++) {
        const uint32_t c = d.children[n.a + k];
        six_quads = d.nodes[c].kind == RT_NODE_QUAD && !seen.count(c);
        for (uint32_t j = 0; six_quads && j < k; j++)
          six_quads = d.children[n.a + j] != c;
      }
      if (!six_quads) {
        for (uint32_t k = 0; k < n.b; k++)
          collect(d.children[n.a + k], items, chain, tag, seen, in_boundary, media);
        break;
      }
      if (tag == kTagPending)
        tag = new_instance(chain);
      Item it;
      it.kind = REF_BOXES, it.index = n.a, it.tag = tag;
      it.box.reset();
      for (uint32_t k = 0; k < 6; k++) {
        const uint32_t c = d.children[n.a + k];
        seen.insert(c);
        it.box.grow(tag == 0 ? node_box(d.nodes[c]) : quad_box(d.quads[d.nodes[c].a], chain));
      }
      items.push_back(it);
      break;
    }
    case RT_NODE_BVH:
      bvh_above++;
      collect(n.a, items, chain, tag, seen, in_boundary, media);
      if (n.b != n.a)
        collect(n.b, items, chain, tag, seen, in_boundary, media);
      bvh_above--;
      break;
    case RT_NODE_SPHERE:
    case RT_NODE_QUAD: {
      if (!seen.insert(id).second)
        break; /* the same object twice in one frame (BVH n==1 stores it as both children) */
      if (tag == kTagPending)
        tag = new_instance(chain);
      Item it;
      it.kind = n.kind == RT_NODE_SPHERE ? REF_SPHERES : REF_QUADS;
      it.index = n.a;
      it.tag = tag;
      if (tag == 0)
        it.box = node_box(n); /* the reference's own box (quads padded, hittable.c:24-37) */
      else
        it.box = n.kind == RT_NODE_SPHERE ? sphere_box(d.spheres[n.a], chain) : quad_box(d.quads[n.a], chain);
      items.push_back(it);
      break;
    }
    case RT_NODE_TRANSLATE:
    case RT_NODE_ROTATE_Y: {
      ChainOp op;
      op.is_rotate = n.kind == RT_NODE_ROTATE_Y;
      op.f0 = n.f[0], op.f1 = n.f[1], op.f2 = n.f[2];
      chain.push_back(op);
      uint32_t my_tag = kTagPending;
      std::set<uint32_t> seen_inside;
      collect(n.a, items, chain, my_tag, seen_inside, in_boundary, media);
      chain.pop_back();
      break;
    }
    case RT_NODE_MEDIUM:
      if (in_boundary)
        unsupported("a ConstantMedium inside the boundary of another ConstantMedium");
      else {
        media.push_back(std::make_pair(id, chain));
        if (bvh_above > 0)
          media_in_bvh.insert(id);
      }
      break;
    default: break;
    }
  }

  uint32_t make_leaf(const Item &it) {
    if (it.kind == REF_SPHERES) {
      const uint32_t at = (uint32_t)out.msph.size();
      const rt_sphere &s = d.spheres[it.index];
      MSphere ms = {s.center[0], s.center[1], s.center[2], s.radius};
      MPrimMeta meta = {s.material, s.node};
      out.msph.push_back(ms);
      out.msph_meta.push_back(meta);
      return MAKE_REF(REF_SPHERES, it.tag, at);
    }
    const uint32_t at = (uint32_t)out.mquads.size();
    const uint32_t count = it.kind == REF_BOXES ? 6u : 1u;
    for (uint32_t k = 0; k < count; k++) {
      const rt_quad &q = d.quads[it.kind == REF_BOXES ? d.nodes[d.children[it.index + k]].a : it.index];
      MQuad mq;
      float mat_bits, node_bits;
      memcpy(&mat_bits, &q.material, 4);
      memcpy(&node_bits, &q.node, 4);
      mq.n_D = make_float4(q.normal[0], q.normal[1], q.normal[2], q.D);
      mq.q_mat = make_float4(q.Q[0], q.Q[1], q.Q[2], mat_bits);
      mq.w_nod = make_float4(q.w[0], q.w[1], q.w[2], node_bits);
      mq.u_ar = make_float4(q.u[0], q.u[1], q.u[2], q.area);
      mq.v_ = make_float4(q.v[0], q.v[1], q.v[2], 0.0f);
      out.mquads.push_back(mq);
    }
    return MAKE_REF(it.kind, it.tag, at);
  }

  /* [lo, hi] as (centre, half extent) with [c - h, c + h] containing it.  The traversal computes the slab distances as
   * fma(-+h, |1/d|, fma(c, 1/d, -o/d)): against the (lo, hi) form that is one more rounding, of the CENTRE's distance,
   * i.e. up to 6e-8 * |c - o| in space; the |o| part is in the ray's own slack (slab_setup), the |c| part and the
   * rounding of c and h themselves are added to h here (4e-7 |c| + 1e-6 h: generous, and nothing next to a box). */
  static void centre_half(float lo, float hi, float &c, float &h) {
    const double cd = 0.5 * ((double)lo + (double)hi);
    c = (float)cd;
    const double need = std::max((double)hi - (double)c, (double)c - (double)lo);
    h = (float)(need * (1.0 + 1e-6) + 4e-7 * std::fabs((double)c));
    h = std::nextafter(h, FLT_MAX);
  }

  /* binned SAH over items[lo, hi); reorders `items` in place; returns the reference of the subtree */
  uint32_t build(std::vector<Item> &items, size_t lo, size_t hi, int &depth) {
    const size_t n = hi - lo;
    if (n == 0)
      return REF_EMPTY;
    Box bounds, cbounds;
    bounds.reset(), cbounds.reset();
    for (size_t k = lo; k < hi; k++) {
      bounds.grow(items[k].box);
      Box c;
      for (int a = 0; a < 3; a++)
        c.lo[a] = c.hi[a] = 0.5f * (items[k].box.lo[a] + items[k].box.hi[a]);
      cbounds.grow(c);
    }
    if (n == 1) /* one primitive per leaf */
      return make_leaf(items[lo]);

    /* best split over 3 axes x 15 planes */
    constexpr int kBins = 16;
    int best_axis = -1, best_plane = -1;
    float best_cost = FLT_MAX;
    const float parent_area = std::max(bounds.area(), 1e-30f);
    for (int axis = 0; axis < 3; axis++) {
      float extent = cbounds.hi[axis] - cbounds.lo[axis];
      if (!(extent > 0.0f))
        continue;
      Box bin_box[kBins];
      float bin_cost[kBins];
      int bin_n[kBins];
      for (int b = 0; b < kBins; b++)
        bin_box[b].reset(), bin_cost[b] = 0.0f, bin_n[b] = 0;
      const float scale = (float)kBins / extent;
      for (size_t k = lo; k < hi; k++) {
        float c = 0.5f * (items[k].box.lo[axis] + items[k].box.hi[axis]);
        int b = std::min(kBins - 1, std::max(0, (int)((c - cbounds.lo[axis]) * scale)));
        bin_box[b].grow(items[k].box);
        bin_cost[b] += item_cost(items[k].kind);
        bin_n[b]++;
      }
      float right_area[kBins], right_cost[kBins];
      Box acc;
      acc.reset();
      float cost = 0.0f;
      for (int b = kBins - 1; b > 0; b--) {
        acc.grow(bin_box[b]);
        cost += bin_cost[b];
        right_area[b] = acc.area(), right_cost[b] = cost;
      }
      acc.reset();
      cost = 0.0f;
      int left_n = 0;
      for (int b = 0; b < kBins - 1; b++) {
        acc.grow(bin_box[b]);
        cost += bin_cost[b];
        left_n += bin_n[b];
        if (left_n == 0 || left_n == (int)n)
          continue;
        float sah = kCostNode + (acc.area() * cost + right_area[b + 1] * right_cost[b + 1]) / parent_area;
        if (sah < best_cost)
          best_cost = sah, best_axis = axis, best_plane = b;
      }
    }

    size_t mid;
    if (best_axis >= 0) {
      const float extent = cbounds.hi[best_axis] - cbounds.lo[best_axis];
      const float scale = (float)kBins / extent;
      const float base = cbounds.lo[best_axis];
      auto first_right = std::partition(items.begin() + lo, items.begin() + hi, [&](const Item &it) {
        float c = 0.5f * (it.box.lo[best_axis] + it.box.hi[best_axis]);
        int b = std::min(kBins - 1, std::max(0, (int)((c - base) * scale)));
        return b <= best_plane;
      });
      mid = (size_t)(first_right - items.begin());
    } else {
      mid = lo + n / 2; /* coincident centroids: split by count */
    }
    if (mid == lo || mid == hi)
      mid = lo + n / 2;

    const uint32_t self = (uint32_t)out.bnodes.size();
    out.bnodes.push_back(BNode());
    int d0 = 0, d1 = 0;
    Box b0, b1;
    b0.reset(), b1.reset();
    for (size_t k = lo; k < mid; k++)
      b0.grow(items[k].box);
    for (size_t k = mid; k < hi; k++)
      b1.grow(items[k].box);
    uint32_t c0 = build(items, lo, mid, d0);
    uint32_t c1 = build(items, mid, hi, d1);
    BNode node;
#if RTC_NODE_CH
    float bc0[3], bh0[3], bc1[3], bh1[3];
    for (int a = 0; a < 3; a++) {
      centre_half(b0.lo[a], b0.hi[a], bc0[a], bh0[a]);
      centre_half(b1.lo[a], b1.hi[a], bc1[a], bh1[a]);
    }
    node.xy0 = make_float4(bc0[0], bh0[0], bc0[1], bh0[1]);
    node.xy1 = make_float4(bc1[0], bh1[0], bc1[1], bh1[1]);
    node.zz = make_float4(bc0[2], bh0[2], bc1[2], bh1[2]);
#else
    node.xy0 = make_float4(b0.lo[0], b0.hi[0], b0.lo[1], b0.hi[1]);
    node.xy1 = make_float4(b1.lo[0], b1.hi[0], b1.lo[1], b1.hi[1]);
    node.zz = make_float4(b0.lo[2], b0.hi[2], b1.lo[2], b1.hi[2]);
#endif
    node.child0 = c0, node.child1 = c1, node.pad0 = 0, node.pad1 = 0;
    out.bnodes[self] = node;
    if (child_boxes.size() < 2 * out.bnodes.size())
      child_boxes.resize(2 * out.bnodes.size());
    child_boxes[2 * self] = b0, child_boxes[2 * self + 1] = b1;
    depth = 1 + std::max(d0, d1);
    return MAKE_REF(REF_INNER, 0, self);
  }

  /* The binary subtree under `ref` as a 4-wide tree: a node's two children are replaced, largest surface first, by
   * their own children until there are four (or only leaves are left).  Returns the wide reference; `stack_need` =
   * the most entries a walk below it can have pushed at once (every visit pushes at most children - 1). */
  uint32_t wide_collapse(uint32_t ref, int &stack_need) {
    stack_need = 0;
    if (ref == REF_EMPTY || REF_KIND(ref) != REF_INNER)
      return ref;
    struct Kid {
      uint32_t ref;
      Box box;
    };
    std::vector<Kid> kids;
    const BNode &root = out.bnodes[ref];
    kids.push_back(Kid{root.child0, child_boxes[2 * ref]});
    kids.push_back(Kid{root.child1, child_boxes[2 * ref + 1]});
    while (kids.size() < 4) {
      int pick = -1;
      for (size_t k = 0; k < kids.size(); k++)
        if (REF_KIND(kids[k].ref) == REF_INNER && (pick < 0 || kids[k].box.area() > kids[(size_t)pick].box.area()))
          pick = (int)k;
      if (pick < 0)
        break;
      const uint32_t inner = kids[(size_t)pick].ref;
      const BNode &n = out.bnodes[inner];
      kids[(size_t)pick] = Kid{n.child0, child_boxes[2 * inner]};
      kids.push_back(Kid{n.child1, child_boxes[2 * inner + 1]});
    }
    const uint32_t self = (uint32_t)out.wnodes.size();
    out.wnodes.push_back(WNode());
    WNode w;
    memset(&w, 0, sizeof(w));
    float c[3][4], h[3][4];
    int deepest = 0;
    for (size_t k = 0; k < 4; k++) {
      if (k < kids.size()) {
        int below = 0;
        w.child[k] = wide_collapse(kids[k].ref, below);
        deepest = std::max(deepest, below);
        for (int a = 0; a < 3; a++)
          centre_half(kids[k].box.lo[a], kids[k].box.hi[a], c[a][k], h[a][k]);
      } else {
        w.child[k] = REF_EMPTY;
        for (int a = 0; a < 3; a++)
          c[a][k] = 0.0f, h[a][k] = 0.0f;
      }
    }
    w.cx = make_float4(c[0][0], c[0][1], c[0][2], c[0][3]), w.hx = make_float4(h[0][0], h[0][1], h[0][2], h[0][3]);
    w.cy = make_float4(c[1][0], c[1][1], c[1][2], c[1][3]), w.hy = make_float4(h[1][0], h[1][1], h[1][2], h[1][3]);
    w.cz = make_float4(c[2][0], c[2][1], c[2][2], c[2][3]), w.hz = make_float4(h[2][0], h[2][1], h[2][2], h[2][3]);
    out.wnodes[self] = w;
    stack_need = (int)kids.size() - 1 + deepest;
    return MAKE_REF(REF_INNER, 0, self);
  }

  /* one group = everything under `root_node` seen from that node's own frame */
  uint32_t build_group(uint32_t root_node, bool in_boundary,
                       std::vector<std::pair<uint32_t, std::vector<ChainOp>>> &media, int &depth) {
    std::vector<Item> items;
    std::vector<ChainOp> chain;
    std::set<uint32_t> seen;
    uint32_t tag = 0;
    collect(root_node, items, chain, tag, seen, in_boundary, media);
    /* Oversized primitives (world group only): a primitive whose box has more than RT_BVH_PRE (default 0.5) of the
     * surface of the whole group's box -- a ground sphere of radius 1000 -- would sit in a leaf right under the root
     * with a box every ray is inside of.  Up to two of them leave the tree and are tested BEFORE the walk, so that the
     * walk starts with the t_max they leave.  The result does not depend on the order candidates are met in. */
    const float pre_frac = env_float("RT_BVH_PRE", 0.5f);
    /* only sphere-only worlds without media (the kernels of the other scene classes do not carry the pre-walk step:
     * measured -1.7 % on the Cornell box and Book 2 final for nothing), and only while two primitives stay in the tree
     * (two equal spheres that both cover the scene -- the checker scene -- are better off as one node) */
    bool spheres_only = !in_boundary && media.empty();
    for (const Item &it : items)
      spheres_only = spheres_only && it.kind == REF_SPHERES;
    if (spheres_only && pre_frac > 0.0f) {
      Box all;
      all.reset();
      for (const Item &it : items)
        all.grow(it.box);
      const float limit = pre_frac * all.area();
      for (size_t k = 0; k < items.size() && out.n_pre < 2 && items.size() > 2;) {
        if (items[k].box.area() > limit) {
          out.pre_ref[out.n_pre++] = make_leaf(items[k]);
          items.erase(items.begin() + (long)k);
        } else {
          k++;
        }
      }
    }
    return build(items, 0, items.size(), depth);
  }
};

} // namespace

void rtc_build_scene(const rt_scene_desc &d, BuiltScene &out) {
  out = BuiltScene();
  std::vector<Depths> memo(d.n_nodes, Depths{0, 0, 0});
  Depths g = graph_depths(d, d.objects_root, memo);
  out.graph_frames = g.frames, out.graph_rays = g.rays + 1, out.graph_recs = g.recs + 1;
  out.light_nesting = d.lights_root == RT_NONE ? 0 : light_nesting(d, d.lights_root);

  Builder b(d, out);
  std::vector<std::pair<uint32_t, std::vector<ChainOp>>> media, nested;
  int depth = 0;
  out.world_root = b.build_group(d.objects_root, false, media, depth);
  out.max_stack = depth;
  if (env_float("RT_BVH_WIDE", 0.0f) != 0.0f && RTC_NODE_CH) { /* the world group again, four children per node */
    int need = 0;
    out.wide_root = b.wide_collapse(out.world_root, need);
    out.wide_max_stack = need;
  }
  for (auto &m : media) {
    const rt_node &n = d.nodes[m.first];
    MMedium mm;
    memset(&mm, 0, sizeof(mm));
    mm.op_begin = (uint32_t)out.ops.size();
    mm.op_count = (uint32_t)m.second.size();
    out.ops.insert(out.ops.end(), m.second.begin(), m.second.end());
    Box nb = Builder::node_box(n);
    Box wb = b.chain_to_outer(nb, m.second);
    for (int a = 0; a < 3; a++)
      mm.bbox[2 * a] = wb.lo[a], mm.bbox[2 * a + 1] = wb.hi[a];
    int bdepth = 0;
    mm.root = b.build_group(n.a, true, nested, bdepth);
    out.max_stack = std::max(out.max_stack, bdepth);
    mm.material = n.b;
    mm.node = m.first;
    mm.neg_inv_density = n.f[0];
    mm.in_list = b.media_in_bvh.count(m.first) ? 0u : 1u;
    mm.boundary_listed = 0;
    const rt_node &top = d.nodes[d.objects_root];
    if (mm.in_list && mm.op_count == 0 && d.nodes[n.a].kind == RT_NODE_SPHERE && top.kind == RT_NODE_LIST)
      for (uint32_t k = 0; k < top.b && d.children[top.a + k] != m.first; k++)
        if (d.children[top.a + k] == n.a)
          mm.boundary_listed = 1;
    out.media.push_back(mm);
  }
  out.n_bnodes = (uint32_t)out.bnodes.size(), out.n_msph = (uint32_t)out.msph.size();
  out.n_mquads = (uint32_t)out.mquads.size(), out.n_insts = (uint32_t)out.insts.size();
  out.n_ops = (uint32_t)out.ops.size();
  /* keep every array non-empty so device pointers are never NULL */
  if (out.bnodes.empty())
    out.bnodes.push_back(BNode());
  if (out.wnodes.empty())
    out.wnodes.push_back(WNode());
  if (out.msph.empty()) {
    out.msph.push_back(MSphere{0, 0, 0, 0});
    out.msph_meta.push_back(MPrimMeta{0, 0});
  }
  if (out.mquads.empty())
    out.mquads.push_back(MQuad());
  if (out.insts.empty())
    out.insts.push_back(Instance{0, 0, 0, 0});
  if (out.ops.empty())
    out.ops.push_back(ChainOp{0, 0, 0, 0});
  if (out.media.empty()) {
    MMedium z;
    memset(&z, 0, sizeof(z));
    z.root = REF_EMPTY;
    out.media.push_back(z);
    out.media.pop_back(); /* n_media stays 0; the upload pads */
  }
}
