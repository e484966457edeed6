/*
 * rtc_kernels.h -- internal launch interface between the C-ABI layer (rtc_api.cu) and the kernel translation
 * units.  Not installed; the public surface is include/rtc_b200.h.
 */
#ifndef RTC_KERNELS_H
#define RTC_KERNELS_H

#include "rtc_scene.cuh"

/* device-side counters of one render (reduced per warp, then atomically) */
struct RenderCounters {
  unsigned long long segments, light_pdf_evals, samples, nan_samples;
  unsigned int next_item; /* dynamic work fetch of the persistent kernels */
  unsigned int pad;
};

struct RenderJob {
  rt_camera cam;
  int sample_begin, sample_count;
  uint32_t rng_mode, flags;
  float *radiance;          /* float[h][w][3], overwritten with the per-pixel SUM over the rendered samples */
  RenderCounters *counters; /* zeroed by the caller */
};

struct HitBatch {
  size_t n;
  const float *origins, *dirs;
  float t_min, t_max;
  uint64_t *rng_io;
  float *out_t;
  uint32_t *out_node;
  float *out_p, *out_normal, *out_uv;
  uint8_t *out_front;
  uint32_t *out_material;
};

/* rtc_exact.cu: reference-order graph, portable math, recursion unwound exactly (bit-identical to the oracle) */
size_t exact_scratch_bytes(const rt_camera &cam, int *n_threads);
cudaError_t exact_render(const DScene &sc, const RenderJob &job, void *scratch, int n_threads, cudaStream_t stream);
cudaError_t exact_hit_batch(const DScene &sc, const HitBatch &b, cudaStream_t stream);

/* rtc_mega.cu: persistent megakernel over the rebuilt BVH */
struct MegaPlan {
  int blocks, threads;   /* persistent grid */
  int minb;              /* __launch_bounds__ variant: resident CTAs per SM the registers allow */
  int nv;                /* paths per lane: 0 = register kernel, else the warp-ray-queue kernel (rtc_mega.cu) */
  int variant;           /* index into rtc_mega.cu's table of built kernels */
  int thr8;              /* node loop runs while >= thr8/8 of the busy lanes want it (walk_round_sync) */
  int resume8;           /* RESUME kernels: a trip's traversal ends once fewer than resume8/8 of its walkers remain */
  int cta8;              /* lockstep RESUME kernels: ... or once cta8/8 of the CTA's warps have finished theirs (0: off) */
  int chunk;             /* samples of a pixel's first (largest) work item */
  int n_chunks;          /* work items per pixel (<= MEGA_MAX_CHUNKS) */
  int begin[65];         /* first sample (relative to the job's sample_begin) of item c; begin[n_chunks] = sample count.
                            Item sizes DECREASE (guided self-scheduling, see mega_plan): items are handed out chunk by
                            chunk, and an item on an expensive pixel must not outlast everything handed out after it */
  size_t partial_bytes;  /* per-chunk partial sums when n_chunks > 1 */
};
#define MEGA_MAX_CHUNKS 64
/* first sample (relative to the job's sample_begin) of a pixel's work item c; item c ends where item c + 1 begins */
__host__ __device__ inline int plan_chunk_begin(const MegaPlan &p, int c) { return p.begin[c]; }
/* equal items of `chunk` samples (the wavefront kernel, RT_MEGA_GUIDED=0, RT_MEGA_CHUNKS) */
inline void plan_equal_chunks(MegaPlan *p, int sample_count, int n_chunks) {
  if (n_chunks > MEGA_MAX_CHUNKS)
    n_chunks = MEGA_MAX_CHUNKS;
  if (n_chunks < 1)
    n_chunks = 1;
  p->chunk = (sample_count + n_chunks - 1) / n_chunks;
  p->n_chunks = p->chunk > 0 ? (sample_count + p->chunk - 1) / p->chunk : 1;
  for (int c = 0; c <= p->n_chunks; c++)
    p->begin[c] = c * p->chunk < sample_count ? c * p->chunk : sample_count;
}
/* what the scene contains, as far as the choice of kernel variant goes (rtc_mega.cu: mega_variant) */
enum { MEGA_SCENE_SPHERES = 0, MEGA_SCENE_QUADS = 1, MEGA_SCENE_MEDIA = 2 };
#define MEGA_FEATS_WIDE_TREE (1 << 30) /* or'ed into scene_feats: walk the scene's 4-wide tree (DScene.wnodes) */
/* scene_feats: FEAT_* mask of the scene (rtc_scene.cuh): a kernel without the code of absent features is preferred */
cudaError_t mega_plan(const rt_camera &cam, int sample_count, uint32_t rng_mode, int device, int scene_class,
                      int scene_feats, MegaPlan *plan);
cudaError_t mega_render(const DScene &sc, const RenderJob &job, const MegaPlan &plan, float *partial,
                        cudaStream_t stream, int *launches);
cudaError_t mega_hit_batch(const DScene &sc, const HitBatch &b, cudaStream_t stream);

/* rtc_wave.cu: wavefront variant (generate / extend+classify / shade per material queue) over a path-state pool */
struct WavePlan {
  MegaPlan mega;        /* chunking of the sample range and the partial-sum buffer (same scheme as the megakernel) */
  uint32_t slots;       /* path states in the pool */
  int shade_blocks;     /* grid of each shade launch (grid-stride over its queue) */
  size_t pool_bytes;    /* SoA pool + queues */
  size_t scratch_bytes; /* partial sums + pool */
};
cudaError_t wave_plan(const rt_camera &cam, int sample_count, uint32_t rng_mode, int device, WavePlan *plan);
/* Blocks the host between batches of iterations (it polls the live-slot count); `stream` is idle on return. */
cudaError_t wave_render(const DScene &sc, const RenderJob &job, const WavePlan &plan, void *scratch,
                        cudaStream_t stream, int *launches);

#endif /* RTC_KERNELS_H */
