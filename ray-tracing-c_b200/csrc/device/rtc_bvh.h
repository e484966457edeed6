/*
 * rtc_bvh.h -- host-side construction of the structure the throughput kernels traverse.
 *
 * The reference builds its BVH by sorting along a RANDOM axis and splitting at the median
 * (/root/reference/src/hittable.c:284-323), tests ~43 boxes per segment on the Book-1 scene, never bounds its
 * Translate/RotateY instances (hittable.c:325-334,354-367) and keeps top-level items in a linear list.  For the
 * megakernel we keep the reference's PRIMITIVES and TRANSFORMS (so hit distances stay bit-identical) but rebuild
 * everything around them: ONE binned-SAH binary BVH per group (the world, or a medium's boundary) with one primitive
 * per leaf; primitives under Translate / RotateY chains are flattened into it with tight boxes in the group's frame
 * and a tag naming their chain; media are evaluated after the surfaces.
 */
#ifndef RTC_BVH_H
#define RTC_BVH_H

#include "rtc_scene.cuh"
#include <string>
#include <vector>

struct BuiltScene {
  std::vector<BNode> bnodes;
  std::vector<WNode> wnodes;         /* world group only, when requested (RT_BVH_WIDE=1) */
  uint32_t wide_root = REF_EMPTY;
  uint32_t n_pre = 0, pre_ref[2] = {REF_EMPTY, REF_EMPTY}; /* oversized world-group primitives tested before the walk */
  int wide_max_stack = 0;
  std::vector<MSphere> msph;
  std::vector<MPrimMeta> msph_meta;
  std::vector<MQuad> mquads;
  std::vector<Instance> insts;
  std::vector<ChainOp> ops;
  std::vector<MMedium> media;
  uint32_t world_root = REF_EMPTY;
  uint32_t n_bnodes = 0, n_msph = 0, n_mquads = 0, n_insts = 0, n_ops = 0; /* before padding to non-empty */
  int max_stack = 0; /* deepest traversal stack the kernels may need (entries) */
  /* limits of the reference-order graph machine (rtc_graph.cuh) */
  int graph_frames = 0, graph_rays = 0, graph_recs = 0, light_nesting = 0;
  bool mega_supported = true;
  std::string mega_why; /* why the rebuilt structure cannot express this graph (RT_ERR_UNSUPPORTED for mega/wave) */
};

/* Returns an empty string on success, else the reason the description is malformed (RT_ERR_INVALID). */
std::string rtc_validate_desc(const rt_scene_desc &d);
void rtc_build_scene(const rt_scene_desc &d, BuiltScene &out);

#endif /* RTC_BVH_H */
