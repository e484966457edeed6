/*
 * rtc_mega.cu -- RT_KERNEL_MEGA: the persistent megakernel (throughput path).
 *
 * Replaces the loop nest of Camera_render + Camera_ray_color (/root/reference/src/raytracing.c:39-133).
 * One CTA set sized to the machine (SM count x resident CTAs) stays resident; a path that ends is regenerated on the
 * spot: next sample of its work item, or a new work item fetched from a global counter.  So every trip of the main
 * loop is "one segment for every path" -- nothing idles on a finished path, which is what depth-50 tails
 * (dielectrics) would otherwise cost.
 *
 * Work item = (pixel, sample chunk).  Its samples are summed privately and the partial sum is stored once; chunks
 * of a pixel are added up in chunk order by a second tiny kernel, so the image is deterministic (no float atomics).
 * The recursion  colour = emitted + albedo * colour(next) * ratio  is run forwards as  L += beta * emitted;
 * beta *= albedo * ratio  (raytracing.c:57-58,69-71).
 *
 * Two kernels share the per-path code (camera ray, shading):
 *   mega_render_kernel  one path per lane, state in registers, one ray per lane per trip.
 *   mega_queue_kernel   NV paths per lane ("virtual lanes"), state in shared memory, and the trip's NV*32 rays of a
 *                       warp are handed to its 32 lanes DYNAMICALLY: a lane that finishes a short ray takes the next
 *                       ray of the warp's queue instead of waiting for the longest ray of the trip.  ncu on the
 *                       one-path kernel (Book 1): 31 node-loop iterations per trip against a mean of 10.8 visits
 *                       per ray -- two thirds of the lane-slots of the dominant loop idle behind the longest ray.
 */
#define RTC_FAST_SHADING /* approximate div/sqrt in the shading code only; see rtc_math.cuh */
#include "rtc_kernels.h"
#include "rtc_trace.cuh"
#include <cstdio>
#include <cstdlib>

#define MEGA_BLOCK 128
#define MEGA_THR8_DEFAULT 2 /* measured: see DESIGN.md */
#define MEGA_CTA8_DEFAULT 3 /* measured on Book 2 final: 3 +6 %, 4 +5.6 %, 6 +3.3 %, 7 +1.9 % (0 = off) */
#define MEGA_GUIDED_DEFAULT 4 /* a work item = a quarter of the samples its pixel has left (mega_plan) */
#define MEGA_RESUME8_DEFAULT 1 /* measured on Book 2 final: 1 +1.7 %, 2 +0.6 %, 3 -3 %, 4 -7 % */

RTC_DEV bool has_nan3(V3 a) { return a.x != a.x || a.y != a.y || a.z != a.z; }

/* pixel order: 8x4 tiles, row-major inside a tile, so the 32 lanes of a warp start on one tile */
RTC_DEV void item_to_pixel(unsigned item_pix, int w, int h, int &i, int &j) {
  const int tiles_x = (w + 7) >> 3;
  unsigned tile = item_pix >> 5, lane = item_pix & 31u;
  int ty = (int)(tile / (unsigned)tiles_x), tx = (int)(tile - (unsigned)ty * (unsigned)tiles_x);
  i = tx * 8 + (int)(lane & 7u);
  j = ty * 4 + (int)(lane >> 3);
}

/* what every path of a render shares (kernel-uniform) */
struct Frame {
  int w, h, max_depth;
  unsigned tiles, n_items;
  V3 px00, du, dv, eye, lens_u, lens_v, background;
  float prob, dof_angle;
  bool use_lights;
};
RTC_DEV Frame frame_of(const DScene &sc, const RenderJob &job, const MegaPlan &plan) {
  const rt_camera &cam = job.cam;
  Frame f;
  f.w = cam.img_width, f.h = cam.img_height, f.max_depth = cam.max_depth;
  f.tiles = (unsigned)(((f.w + 7) >> 3) * ((f.h + 3) >> 2));
  f.n_items = f.tiles * 32u * (unsigned)plan.n_chunks;
  f.px00 = v_load(cam.pixel00), f.du = v_load(cam.pixel_delta_u), f.dv = v_load(cam.pixel_delta_v);
  f.eye = v_load(cam.origin), f.lens_u = v_load(cam.dof_disc_u), f.lens_v = v_load(cam.dof_disc_v);
  f.background = v_load(cam.background);
  f.prob = cam.lights_sampling_prob, f.dof_angle = cam.dof_angle;
  f.use_lights = !(f.prob == 0.0f || sc.lights_root == RT_NONE);
  return f;
}

#ifdef RTC_TIMELINE
#define TL_MAX_WARPS 16384
__device__ unsigned long long tl_records[TL_MAX_WARPS][8];
#endif

/* A word of per-lane state: a register, or (COLD) one slot of shared memory -- for what no traversal touches (the
 * sample sums of the work item, the launch counters), so that the node loop has the registers to itself. */
template <bool COLD, typename T> struct Kept {
  T reg;
  volatile T *mem;
  RTC_DEV explicit Kept(T *slot) : reg(0), mem(slot) {
    if (COLD)
      *mem = 0;
  }
  RTC_DEV T get() const { return COLD ? *mem : reg; }
  RTC_DEV void set(T v) {
    if (COLD)
      *mem = v;
    else
      reg = v;
  }
  RTC_DEV void add(T v) { set(get() + v); } /* float: one rounded addition either way (no product to contract) */
};

/* Fetches the next work item that lies inside the image.  false = no items left.  (The kernels built with POOL draw
 * 32 items per warp instead, mega_render_kernel.) */
/* RTC_ITEM_POOL: -1 = as the variant table says, 0 / 1 = every kernel without / with the warp pool (A/B builds) */
#ifndef RTC_ITEM_POOL
#define RTC_ITEM_POOL -1
#endif
RTC_DEV bool fetch_item(const Frame &f, const RenderJob &job, int &pi, int &pj, int &chunk_id) {
  for (;;) {
    unsigned item = atomicAdd(&job.counters->next_item, 1u);
    if (item >= f.n_items)
      return false;
    unsigned c = item / (f.tiles * 32u), pix = item - c * (f.tiles * 32u);
#ifdef RTC_SCRAMBLE /* measurement only: consecutive items are pixels of 32 different tiles (how much is coherence worth?) */
    pix = (pix % 32u) * f.tiles + pix / 32u;
#endif
    item_to_pixel(pix, f.w, f.h, pi, pj);
    if (pi >= f.w || pj >= f.h)
      continue; /* padding of the last tile column / row */
    chunk_id = (int)c;
    return true;
  }
}

/* camera ray of one sample (raytracing.c:100-122) */
RTC_DEV Ray camera_ray(const Frame &f, int pi, int pj, Rng &rng) {
  const V3 pixel = v_add(v_add(f.px00, v_scale(f.du, (float)pi)), v_scale(f.dv, (float)pj));
  float jx = rng_between(rng, -0.5f, 0.5f);
  float jy = rng_between(rng, -0.5f, 0.5f);
  Ray ray;
  ray.o = f.eye;
  if (f.dof_angle > 0.0f) {
    float a, b;
    for (;;) {
      a = rng_between(rng, -1.0f, 1.0f);
      b = rng_between(rng, -1.0f, 1.0f);
      if (a * a + b * b < 1.0f)
        break;
    }
    ray.o = v_add(v_add(f.eye, v_scale(f.lens_u, a)), v_scale(f.lens_v, b));
  }
  ray.d = v_add(v_add(v_add(pixel, v_scale(f.du, jx)), v_scale(f.dv, jy)), v_neg(ray.o));
  return ray;
}

/* One step of Camera_ray_color (raytracing.c:39-72) for a path whose closest hit is `th`: adds what the hit emits,
 * scatters, updates the throughput.  Returns true when the path ended (its radiance L is final). */
template <bool DEFER, int FEATS = FEAT_ALL>
RTC_DEV bool shade_segment(const DScene &sc, const Frame &f, const TraceHit &th, Ray &ray, V3 &beta, V3 &L, int &depth,
                           Rng &rng, unsigned &n_pdf, PendingTex &pend) {
  if (th.prim == PRIM_NONE) {
    L = v_add(L, v_mul(beta, f.background));
    return true;
  }
  Hit rec;
  trace_finalize<false, FEATS>(sc, ray, th, rec);
  V3 r_out, albedo;
  bool skip_pdf;
  const uint32_t tag = __ldg(&sc.materials[rec.material].tag);
  if (tag == RT_MAT_DIFFUSE_LIGHT || tag == RT_MAT_SURFACE_NORMAL) {
    L = v_add(L, v_mul(beta, material_emit<false>(sc, rec)));
    return true;
  }
  if (!material_scatter<false, DEFER, FEATS>(sc, rec, ray.d, r_out, albedo, skip_pdf, rng, pend))
    return true;
  if ((FEATS & FEAT_LIGHTS) && !skip_pdf && f.use_lights) { /* Book-3 mixture (raytracing.c:60-71) */
    /* world.lights = one quad (Cornell box, Book 2 final): HittableList_rand's index draw (`u32 % 1`, hittable.c:101-107)
     * is still made, HittableList_pdf's mean over one item is the item's pdf (hittable.c:89-100) */
    if (rng_f32(rng) < f.prob) {
#ifdef RTC_EXP_NO_ONE
      if (false) {
#else
      if (sc.one_quad_light) {
#endif
        rng_u32(rng);
        r_out = quad_rand(sc.light, rec.p, rng);
      } else {
        r_out = lights_rand<false>(sc, sc.lights_root, rec.p, rng);
      }
    }
    float scatter_pdf = material_scatter_pdf(sc, rec.material, rec.n, r_out);
    n_pdf++;
    const float light_pdf =
#ifndef RTC_EXP_NO_ONE
        sc.one_quad_light ? quad_pdf(sc.light, rec.p, r_out) :
#endif
        lights_pdf<false>(sc, sc.lights_root, rec.p, r_out);
    float sampling_pdf = (1.0f - f.prob) * scatter_pdf + f.prob * light_pdf;
    albedo = v_scale(albedo, f_div(scatter_pdf, sampling_pdf));
  }
  beta = v_mul(beta, albedo);
  ray.o = rec.p;
  ray.d = r_out;
  depth++;
  return depth >= f.max_depth; /* the next call would return 0 (raytracing.c:40-41) */
}

/* warp-reduces the four counters of a render and adds them to the job's */
RTC_DEV void flush_counters(const RenderJob &job, unsigned n_segments, unsigned n_pdf, unsigned n_samples,
                            unsigned n_nan) {
  unsigned long long v[4] = {n_segments, n_pdf, n_samples, n_nan};
  __syncwarp();
#pragma unroll
  for (int k = 0; k < 4; k++)
    for (int off = 16; off > 0; off >>= 1)
      v[k] += __shfl_down_sync(0xffffffffu, v[k], off);
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(&job.counters->segments, v[0]);
    atomicAdd(&job.counters->light_pdf_evals, v[1]);
    atomicAdd(&job.counters->samples, v[2]);
    atomicAdd(&job.counters->nan_samples, v[3]);
  }
}

/* ------------------------------------------------------------------ one path per lane -------------------- */

/* MINB = resident CTAs per SM the register allocation must allow (occupancy vs spills; chosen from measurements,
 * RT_MEGA_MINB overrides for tuning) */
/* BLOCK / LOCKSTEP: the default is 128-thread CTAs whose warps run freely.  LOCKSTEP puts a CTA-wide barrier between
 * the phases, so that all warps of a (large) CTA execute the same phase at the same time: on scenes that use every
 * code path (Book 2 final) the kernel is instruction-cache bound, and one phase's code fits the cache where the whole
 * loop body does not. */
/* RESUME: a trip's traversal does not wait for its longest rays.  Once fewer than plan.resume8/8 of the lanes that
 * entered it are still walking it ends; the finished lanes shade, regenerate and start their next ray, the stragglers
 * keep their walk (cursor, stack pointer, best hit: six registers; the stack is already per-thread local memory) and
 * carry on in the next trip.  Pays where ray lengths spread widely (Book 2 final: a few rays cross the box field or
 * the sphere cluster while most leave after a handful of nodes).
 * In the LOCKSTEP kernels the same machinery bounds the wait at the CTA barrier (ncu: 3.5 of the 14 cycles between two
 * issues of a warp there): once plan.cta8/8 of the CTA's warps have finished the trip's traversal, the others suspend
 * theirs too (after at least one round, so every trip makes progress; warps without work do not count, so nobody is
 * suspended when few warps have work left). */
/* POOL: the warp draws 32 consecutive items at a time -- one chunk of the 32 pixels of ONE tile -- and hands them to its
 * lanes as they come free.  With one draw per lane the lanes of a warp drift to 32 unrelated pixels after their first
 * item; with the pool they stay on one or two tiles (coherent camera rays, shared nodes), and the counter is hit once per
 * 32 items.  Measured (tools/gpu_run27.sh, gpu_run28.sh): items scattered over 32 tiles from the start
 * cost Book 1 final +7 %; the pool gains it 6.2 % at 125 and at 1000 spp, the Cornell box 1.6-2.0 %, scene 0 2.5 %; the
 * lockstep kernels lose 0.3-1.9 % (Book 2 final, the quad scene), so they keep the per-lane draw.
 * COLD: the sample sums and the launch counters live in shared memory (Kept).  tools/gpu_run26.sh: the Cornell box
 * -1.8 %, the free-running all-features kernel -1.4..-2.0 % (scenes 2, 5); Book 1 final +3.6 %, Book 2 final +1.0 %,
 * the 256-thread lockstep kernel +-0: built where it pays. */
template <int MINB, int BLOCK, bool LOCKSTEP, bool RESUME, int FEATS, int NW = 2, bool COLD = false, bool POOL_ = false>
__global__ void __launch_bounds__(BLOCK, MINB)
    mega_render_kernel(const __grid_constant__ DScene sc, const __grid_constant__ RenderJob job,
                       const __grid_constant__ MegaPlan plan, float *partial) {
  constexpr bool POOL = RTC_ITEM_POOL < 0 ? POOL_ : RTC_ITEM_POOL != 0;
  const Frame f = frame_of(sc, job, plan);
  /* per-lane counts of one launch: a lane would need days to overflow 32 bits */
  /* COLD: these seven words live in shared memory, [field][thread] (every access one conflict-free ld/st.shared) */
  __shared__ unsigned cold_u[COLD ? 4 : 1][COLD ? BLOCK : 1];
  __shared__ float cold_f[COLD ? 3 : 1][COLD ? BLOCK : 1];
  const int cold_t = COLD ? (int)threadIdx.x : 0;
  /* POOL: the warp's drawn and not yet handed-out items [next, end), and their chunk and tile */
  __shared__ volatile unsigned pool_w[5][POOL ? BLOCK / 32 : 1];
  if (POOL) {
    if ((threadIdx.x & 31u) == 0)
      pool_w[0][threadIdx.x >> 5] = pool_w[1][threadIdx.x >> 5] = 0;
    __syncwarp();
  }
  Kept<COLD, unsigned> n_segments(&cold_u[0][cold_t]), n_pdf(&cold_u[COLD ? 1 : 0][cold_t]),
      n_samples(&cold_u[COLD ? 2 : 0][cold_t]), n_nan(&cold_u[COLD ? 3 : 0][cold_t]);
  Kept<COLD, float> sum_x(&cold_f[0][cold_t]), sum_y(&cold_f[COLD ? 1 : 0][cold_t]), sum_z(&cold_f[COLD ? 2 : 0][cold_t]);
#ifdef RTC_TIMELINE /* development build: when does every CTA start, run out of items, and end? */
  unsigned long long tl_in = 0, tl_dry = 0, tl_last = 0; /* tl_last: the warp's last draw from the queue (POOL) */
  unsigned tl_last_chunk = 0, tl_draws = 0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tl_in));
#endif

  /* per-lane path state */
  Rng rng;
  rng.state = 0, rng.inc = 1;
  Ray ray;
  ray.o = ray.d = v3(0, 0, 0);
  V3 beta = v3(1, 1, 1);
  int pi = 0, pj = 0, s = 0, depth = 0, chunk_id = 0;
  bool have_item = false, alive = false;
  /* the traversal state outlives a trip only in the RESUME kernels */
  uint32_t stack[TRACE_STACK - RTC_SMEM_STACK];
  WalkerT<RTC_SMEM_STACK, FEATS, NW> wk; /* NW = 4: the world group's 4-wide tree (sc.wnodes) */
  wk.stack = stack, wk.cur = REF_EMPTY, wk.sp = 1;
  /* dynamic shared memory: RTC_SMEM_STACK * BLOCK words (mega_variants) */
  wk.sbase = (uint32_t)__cvta_generic_to_shared(rtc_dyn_smem + threadIdx.x), wk.sstride = 4u * BLOCK;
  bool walking = false;

  /* lockstep RESUME kernels: how many warps of the CTA have finished this trip's traversal (one counter per trip
   * parity; the idle one is cleared between the barriers) */
  __shared__ int warps_finished[2];
  if (LOCKSTEP && RESUME) {
    if (threadIdx.x < 2)
      warps_finished[threadIdx.x] = 0;
    __syncthreads();
  }
  const int stop_warps = (LOCKSTEP && RESUME && plan.cta8 > 0) ? max(1, ((BLOCK / 32) * plan.cta8) >> 3) : 0x7fffffff;
  unsigned trip = 0;

  bool done = false; /* this lane has no work left; it stays in the loop (idle) until its whole warp is done */
  for (;; trip++) {
    /* ---- phase 1: regenerate.  Lanes whose path ended take the next sample of their item or fetch a new item. */
    if (POOL) {
      /* finished chunks are stored, then the lanes without an item are served from the warp's pool, lowest lane first */
      bool need = !alive && !done;
      if (need && have_item) {
        if (s == job.sample_begin + min(plan_chunk_begin(plan, chunk_id + 1), job.sample_count)) {
          float *dst = (plan.n_chunks > 1 ? partial + (size_t)chunk_id * (size_t)f.w * f.h * 3 : job.radiance) +
                       ((size_t)pj * f.w + pi) * 3;
          dst[0] = sum_x.get(), dst[1] = sum_y.get(), dst[2] = sum_z.get();
          have_item = false;
        }
      }
      need = need && !have_item;
      const unsigned lane_id = threadIdx.x & 31u, warp_id = threadIdx.x >> 5;
      unsigned need_mask = __ballot_sync(0xffffffffu, need);
      while (need_mask) {
        unsigned next = pool_w[0][warp_id], end = pool_w[1][warp_id];
        /* the batch is one chunk of one tile: decoded once per draw, not once per item */
        int batch_chunk = (int)pool_w[2][warp_id], batch_x = (int)pool_w[3][warp_id], batch_y = (int)pool_w[4][warp_id];
        if (next == end) {
          unsigned base = 0;
          if (lane_id == 0)
            base = atomicAdd(&job.counters->next_item, 32u);
          base = __shfl_sync(0xffffffffu, base, 0);
          if (base >= f.n_items) { /* n_items is a multiple of 32: nothing left for anybody */
            if (need)
              done = true;
#ifdef RTC_TIMELINE
            if (tl_dry == 0)
              asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tl_dry));
#endif
            break;
          }
          next = base, end = base + 32u;
          const unsigned c = base / (f.tiles * 32u);
          batch_chunk = (int)c;
          item_to_pixel(base - c * (f.tiles * 32u), f.w, f.h, batch_x, batch_y); /* the tile's first pixel */
#ifdef RTC_TIMELINE
          asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tl_last));
          tl_last_chunk = c, tl_draws++;
#endif
        }
        const unsigned rank = __popc(need_mask & ((1u << lane_id) - 1u));
        if (need && rank < end - next) {
          const unsigned k = (next + rank) & 31u; /* 8x4 tile, row-major (item_to_pixel) */
          pi = batch_x + (int)(k & 7u), pj = batch_y + (int)(k >> 3);
          if (pi < f.w && pj < f.h) { /* else: padding of the last tile column / row, the lane asks again */
            chunk_id = batch_chunk;
            s = job.sample_begin + plan_chunk_begin(plan, chunk_id);
            sum_x.set(0.0f), sum_y.set(0.0f), sum_z.set(0.0f);
            if (job.rng_mode == RT_RNG_PIXEL)
              rng_seed_pixel(rng, pi, pj);
            have_item = true, need = false;
          }
        }
        next += min(end - next, (unsigned)__popc(need_mask));
        __syncwarp();
        if (lane_id == 0) {
          pool_w[0][warp_id] = next, pool_w[1][warp_id] = end;
          pool_w[2][warp_id] = (unsigned)batch_chunk, pool_w[3][warp_id] = (unsigned)batch_x, pool_w[4][warp_id] = (unsigned)batch_y;
        }
        __syncwarp();
        need_mask = __ballot_sync(0xffffffffu, need);
      }
    }
    if (!alive && !done) {
      if (!POOL && have_item && s == job.sample_begin + min(plan_chunk_begin(plan, chunk_id + 1), job.sample_count)) {
        /* chunk finished: one store per item */
        float *dst = (plan.n_chunks > 1 ? partial + (size_t)chunk_id * (size_t)f.w * f.h * 3 : job.radiance) +
                     ((size_t)pj * f.w + pi) * 3;
        dst[0] = sum_x.get(), dst[1] = sum_y.get(), dst[2] = sum_z.get();
        have_item = false;
      }
      if (!POOL && !have_item) {
        if (fetch_item(f, job, pi, pj, chunk_id)) {
          s = job.sample_begin + plan_chunk_begin(plan, chunk_id);
          sum_x.set(0.0f), sum_y.set(0.0f), sum_z.set(0.0f);
          if (job.rng_mode == RT_RNG_PIXEL)
            rng_seed_pixel(rng, pi, pj);
          have_item = true;
        } else {
          done = true;
#ifdef RTC_TIMELINE
          if (tl_dry == 0)
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tl_dry));
#endif
        }
      }
      if (!done) {
        if (job.rng_mode == RT_RNG_SAMPLE)
          rng_seed_sample(rng, pi, pj, s);
        ray = camera_ray(f, pi, pj, rng);
        beta = v3(1, 1, 1);
        depth = 0;
        alive = f.max_depth > 0;
        n_samples.add(1u);
        if (!alive)
          s++; /* depth 0: the sample contributes nothing (raytracing.c:40-41) */
      }
    }
    /* warp-uniform (CTA-uniform) exit, and the point where all lanes meet again before the traversal */
    if (LOCKSTEP ? __syncthreads_and(done) : __all_sync(0xffffffffu, done))
      break;

    /* ---- phase 2: one closest-hit query for every live lane, entered together */
    TraceHit th;
    {
      constexpr bool PARK = (RTC_POSTPONE == 1 && BLOCK <= 256) || RTC_POSTPONE == 2;
      if (RESUME && walking)
        wk.resume(ray);
      else
        wk.begin(alive ? (NW == 4 ? sc.wide_root : sc.world_root) : REF_EMPTY, ray, RTC_INF);
      auto outer = [&]() { return ray; };
      if ((FEATS & FEAT_PRE) && !(RESUME && walking) && alive)
        walk_pre(sc, wk, 1e-3f, outer);
      if (RESUME) {
        const int entered = __popc(__ballot_sync(0xffffffffu, !wk.done()));
        const int stop_below = (entered * plan.resume8 + 7) >> 3;
        volatile int *finished = &warps_finished[trip & 1u];
        if (LOCKSTEP && threadIdx.x == 0)
          warps_finished[(trip & 1u) ^ 1u] = 0; /* last used before the previous trip's second barrier */
        bool natural = true;
        for (int round = 0;; round++) {
          const int left = __popc(__ballot_sync(0xffffffffu, !wk.done()));
          if (left == 0 || left < stop_below)
            break;
          /* most of the CTA's working warps wait at the barrier: suspend this warp's walks (never before its first
           * round of the trip, so every trip makes progress) */
          /* (other warps add to the counter concurrently: the decision is made from ONE read, lane 0's, so that all
           * 32 lanes leave together -- the next statement after the loop is a full-mask ballot) */
          if (LOCKSTEP && round > 0 && __shfl_sync(0xffffffffu, *finished, 0) >= stop_warps) {
            natural = false;
            break;
          }
          walk_round_sync<true, PARK>(sc, wk, 1e-3f, outer, plan.thr8);
        }
        /* a warp that had nothing to trace does not count: when few warps have work left nobody is suspended */
        if (LOCKSTEP && natural && entered > 0 && (threadIdx.x & 31) == 0)
          atomicAdd(&warps_finished[trip & 1u], 1);
        walking = !wk.done();
      } else {
        while (__any_sync(0xffffffffu, !wk.done()))
          walk_round_sync<true, PARK>(sc, wk, 1e-3f, outer, plan.thr8);
      }
      th = wk.best;
    }
    if (alive && !(RESUME && walking)) {
      n_segments.add(1u);
      if ((FEATS & FEAT_MEDIA) && sc.n_media)
        trace_media<false>(sc, ray, 1e-3f, th, rng);
    }
    if (LOCKSTEP)
      __syncthreads();
    else
      __syncwarp();

    /* ---- phase 3: shade.  The radiance of a path stays 0 until the step that ends it (emission or background),
     * so it is not part of the state a lane carries through the walk. */
    V3 L = v3(0, 0, 0);
    PendingTex pend;
    pend.tex = RT_NONE;
    unsigned pdf_now = 0;
    if (alive && !(RESUME && walking) &&
        shade_segment<true, FEATS>(sc, f, th, ray, beta, L, depth, rng, pdf_now, pend)) {
      alive = false;
      s++;
      if (has_nan3(L)) {
        n_nan.add(1u);
        if (!(job.flags & RT_FLAG_DROP_NAN))
          sum_x.add(L.x), sum_y.add(L.y), sum_z.add(L.z);
      } else {
        sum_x.add(L.x), sum_y.add(L.y), sum_z.add(L.z);
      }
    }
    if (pdf_now)
      n_pdf.add(pdf_now);
    /* Perlin albedos owed by this trip's hits: evaluated by the whole warp, one lookup after the other (the point is the
     * scattered ray's origin; a path that ended at max_depth owes one too, its throughput is not used again) */
    if ((FEATS & FEAT_TEXTURES) && sc.n_perlins)
      perlin_settle_warp(sc, pend.tex, ray.o, beta);
  }
  /* a lane that ran out of items while holding a finished chunk already stored it in phase 1 */
  flush_counters(job, n_segments.get(), n_pdf.get(), n_samples.get(), n_nan.get());
#ifdef RTC_TIMELINE
  if ((threadIdx.x & 31u) == 0) { /* one line per warp */
    unsigned long long tl_out;
    unsigned smid;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tl_out));
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    /* into a device array the host prints after the launch (printf from 4 736 warps stretches the kernel by ms) */
    const unsigned w = blockIdx.x * (BLOCK / 32) + (threadIdx.x >> 5);
    if (w < TL_MAX_WARPS) {
      unsigned long long *r = tl_records[w];
      r[0] = tl_in, r[1] = tl_dry, r[2] = tl_out, r[3] = trip, r[4] = tl_last, r[5] = tl_last_chunk, r[6] = tl_draws, r[7] = smid;
    }
  }
#endif
}

/* ------------------------------------------------------------------ NV paths per lane, warp ray queue ----- */

/* Path state of one "virtual lane" v = k*32 + lane (k < NV), one 32-bit word per field, laid out [field][k][lane] in
 * the warp's slice of shared memory: a lane working on its own paths touches its own bank.  The radiance of a path
 * needs no slot: it stays 0 until the step that ends the path (emission or background). */
enum {
  F_OX, F_OY, F_OZ, F_DX, F_DY, F_DZ, /* ray of the next / current segment */
  F_BX, F_BY, F_BZ,                   /* throughput */
  F_SX, F_SY, F_SZ,                   /* sum over the samples of the work item so far */
  F_RNG_LO, F_RNG_HI,                 /* pcg32 state; its increment follows from the pixel column (pcg32.c:14) */
  F_PIX,                              /* pixel: column | row << 16 */
  F_S,                                /* sample index */
  F_META,                             /* depth | chunk << 16 | flags */
  F_T, F_PRIM, F_INST,                /* closest hit of the current segment, written by the warp's traversal */
  F_COUNT
};
#define META_ALIVE (1u << 24)
#define META_ITEM (1u << 25)
#define META_DONE (1u << 26)

#define mega_smem rtc_dyn_smem
template <int NV> struct WarpSlots {
  unsigned base; /* word offset of the warp's [F_COUNT][NV][32] words (an offset, not a pointer: the accesses stay
                    plain shared-memory loads and stores) */
  RTC_DEV uint32_t &u(int field, int v) const { return mega_smem[base + field * NV * 32 + v]; }
  RTC_DEV float &f(int field, int v) const { return reinterpret_cast<float *>(mega_smem)[base + field * NV * 32 + v]; }
  RTC_DEV V3 v3f(int field, int v) const { return v3(f(field, v), f(field + 1, v), f(field + 2, v)); }
  RTC_DEV void put(int field, int v, V3 a) const { f(field, v) = a.x, f(field + 1, v) = a.y, f(field + 2, v) = a.z; }
};
template <int NV> constexpr size_t mega_queue_smem() {
  return (size_t)(MEGA_BLOCK / 32) * ((size_t)F_COUNT * NV * 32 * 4 + (size_t)NV * 32);
}
RTC_DEV uint64_t rng_inc_of_column(int pi) { return ((23u + (uint64_t)pi) << 1u) | 1u; }

template <int MINB, int NV>
__global__ void __launch_bounds__(MEGA_BLOCK, MINB)
    mega_queue_kernel(const __grid_constant__ DScene sc, const __grid_constant__ RenderJob job,
                      const __grid_constant__ MegaPlan plan, float *partial) {
  const Frame f = frame_of(sc, job, plan);
  const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, lt_mask = (1u << lane) - 1u;
  WarpSlots<NV> S;
  S.base = warp * F_COUNT * NV * 32;
  const unsigned list_base = (MEGA_BLOCK / 32) * F_COUNT * NV * 32 * 4 + warp * NV * 32; /* byte offset */
  uint8_t *const smem_bytes = reinterpret_cast<uint8_t *>(mega_smem);
  /* per-lane counts of one launch: a lane would need days to overflow 32 bits */
  unsigned n_segments = 0, n_pdf = 0, n_samples = 0, n_nan = 0;
  const int s_last = job.sample_begin + job.sample_count;

#pragma unroll
  for (int k = 0; k < NV; k++)
    S.u(F_META, k * 32 + lane) = 0u;

  for (;;) {
    /* ---- phase 1: regenerate the paths that ended; queue the live ones */
    int n_list = 0;
    bool all_done = true;
#pragma unroll 1
    for (int k = 0; k < NV; k++) {
      const int v = k * 32 + (int)lane;
      uint32_t meta = S.u(F_META, v);
      if (!(meta & (META_ALIVE | META_DONE))) {
        uint32_t pix = S.u(F_PIX, v);
        int pi = (int)(pix & 0xffffu), pj = (int)(pix >> 16), s = (int)S.u(F_S, v), chunk_id = (int)((meta >> 16) & 0xffu);
        if (meta & META_ITEM) {
          const int s_end = min(job.sample_begin + plan_chunk_begin(plan, chunk_id + 1), s_last);
          if (s == s_end) { /* chunk finished: one store per item */
            float *dst = (plan.n_chunks > 1 ? partial + (size_t)chunk_id * (size_t)f.w * f.h * 3 : job.radiance) +
                         ((size_t)pj * f.w + pi) * 3;
            dst[0] = S.f(F_SX, v), dst[1] = S.f(F_SY, v), dst[2] = S.f(F_SZ, v);
            meta &= ~META_ITEM;
          }
        }
        Rng rng;
        bool seeded = false;
        if (!(meta & META_ITEM)) {
          if (fetch_item(f, job, pi, pj, chunk_id)) {
            s = job.sample_begin + plan_chunk_begin(plan, chunk_id);
            S.put(F_SX, v, v3(0, 0, 0));
            S.u(F_PIX, v) = (uint32_t)pi | ((uint32_t)pj << 16);
            if (job.rng_mode == RT_RNG_PIXEL) {
              rng_seed_pixel(rng, pi, pj);
              seeded = true;
            }
            meta = META_ITEM | ((uint32_t)chunk_id << 16);
          } else {
            meta = META_DONE;
          }
        }
        if (!(meta & META_DONE)) {
          if (job.rng_mode == RT_RNG_SAMPLE) {
            rng_seed_sample(rng, pi, pj, s);
          } else if (!seeded) {
            rng.state = (uint64_t)S.u(F_RNG_LO, v) | ((uint64_t)S.u(F_RNG_HI, v) << 32);
            rng.inc = rng_inc_of_column(pi);
          }
          const Ray ray = camera_ray(f, pi, pj, rng);
          S.put(F_OX, v, ray.o);
          S.put(F_DX, v, ray.d);
          S.put(F_BX, v, v3(1, 1, 1));
          S.u(F_RNG_LO, v) = (uint32_t)rng.state, S.u(F_RNG_HI, v) = (uint32_t)(rng.state >> 32);
          meta &= ~0xffffu; /* depth 0 */
          n_samples++;
          if (f.max_depth > 0)
            meta |= META_ALIVE;
          else
            s++; /* depth 0: the sample contributes nothing (raytracing.c:40-41) */
          S.u(F_S, v) = (uint32_t)s;
        }
        S.u(F_META, v) = meta;
      }
      const bool alive = (meta & META_ALIVE) != 0u;
      const unsigned m = __ballot_sync(0xffffffffu, alive);
      if (alive) {
        smem_bytes[list_base + n_list + __popc(m & lt_mask)] = (uint8_t)v;
        n_segments++;
      }
      n_list += __popc(m);
      all_done = all_done && __all_sync(0xffffffffu, (meta & META_DONE) != 0u);
    }
    if (all_done)
      break;
    __syncwarp();

    /* ---- phase 2: the warp's n_list rays, handed to its lanes as they become free */
    {
      uint32_t stack[TRACE_STACK];
      Walker wk;
      wk.stack = stack, wk.cur = REF_EMPTY, wk.sp = 0;
      int mine = -1, next = 0;
      auto outer = [&]() { /* the ray as queued (only needed again when the walk leaves an instance) */
        Ray q;
        q.o = S.v3f(F_OX, mine), q.d = S.v3f(F_DX, mine);
        return q;
      };
      for (;;) {
        const bool idle = wk.done();
        const unsigned m = __ballot_sync(0xffffffffu, idle);
        if (m != 0u) { /* some lane finished its ray (or has none yet): park the hit, take the next ray */
          if (idle && mine >= 0) {
            S.f(F_T, mine) = wk.best.t, S.u(F_PRIM, mine) = wk.best.prim, S.u(F_INST, mine) = wk.best.inst;
            mine = -1;
          }
          if (next < n_list) {
            const int idx = next + __popc(m & lt_mask);
            if (idle && idx < n_list) {
              mine = smem_bytes[list_base + idx];
              wk.begin(sc.world_root, outer(), RTC_INF);
              walk_pre(sc, wk, 1e-3f, outer);
            }
            next += __popc(m);
          } else if (m == 0xffffffffu) {
            break;
          }
        }
        walk_round_sync<true, false>(sc, wk, 1e-3f, outer, plan.thr8);
      }
    }
    __syncwarp();

    /* ---- phase 3: shade */
#pragma unroll 1
    for (int k = 0; k < NV; k++) {
      const int v = k * 32 + (int)lane;
      uint32_t meta = S.u(F_META, v);
      PendingTex pend;
      pend.tex = RT_NONE;
      if (meta & META_ALIVE) {
        Ray ray;
        ray.o = S.v3f(F_OX, v), ray.d = S.v3f(F_DX, v);
        TraceHit th;
        th.t = S.f(F_T, v), th.prim = S.u(F_PRIM, v), th.inst = S.u(F_INST, v);
        Rng rng;
        rng.state = (uint64_t)S.u(F_RNG_LO, v) | ((uint64_t)S.u(F_RNG_HI, v) << 32);
        rng.inc = rng_inc_of_column((int)(S.u(F_PIX, v) & 0xffffu));
        V3 beta = S.v3f(F_BX, v), L = v3(0, 0, 0);
        int depth = (int)(meta & 0xffffu);
        if (sc.n_media)
          trace_media<false>(sc, ray, 1e-3f, th, rng);
        if (shade_segment<true>(sc, f, th, ray, beta, L, depth, rng, n_pdf, pend)) {
          pend.tex = RT_NONE; /* the path ended: its throughput is not used again */
          meta &= ~META_ALIVE;
          S.u(F_S, v) += 1u;
          bool add = true;
          if (has_nan3(L)) {
            n_nan++;
            add = !(job.flags & RT_FLAG_DROP_NAN);
          }
          if (add)
            S.put(F_SX, v, v_add(S.v3f(F_SX, v), L));
        } else {
          S.put(F_OX, v, ray.o);
          S.put(F_DX, v, ray.d);
          S.put(F_BX, v, beta);
          meta = (meta & ~0xffffu) | (uint32_t)depth;
        }
        S.u(F_RNG_LO, v) = (uint32_t)rng.state, S.u(F_RNG_HI, v) = (uint32_t)(rng.state >> 32);
        S.u(F_META, v) = meta;
      }
      if (sc.n_perlins) { /* Perlin albedos owed by this slot's hits, evaluated by the whole warp */
        V3 beta = S.v3f(F_BX, v);
        perlin_settle_warp(sc, pend.tex, S.v3f(F_OX, v), beta);
        if (pend.tex != RT_NONE)
          S.put(F_BX, v, beta);
      }
    }
    __syncwarp();
  }
  flush_counters(job, n_segments, n_pdf, n_samples, n_nan);
}

/* radiance[p] = sum over chunks, in chunk order (deterministic) */
__global__ void mega_sum_chunks_kernel(const float *__restrict__ partial, float *__restrict__ radiance, size_t n_values,
                                       int n_chunks) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_values)
    return;
  float acc = partial[i];
  for (int c = 1; c < n_chunks; c++)
    acc += partial[(size_t)c * n_values + i];
  radiance[i] = acc;
}

/* ------------------------------------------------------------------ launch ------------------------------- */

typedef void (*MegaKernel)(const DScene, const RenderJob, const MegaPlan, float *);
struct MegaVariant {
  int nv, minb; /* nv = 0: one path per lane (registers); nv = -1: the same in lockstep CTAs */
  MegaKernel fn;
  size_t smem;
  int block;
  int feats; /* FEAT_* mask the kernel was instantiated with: it renders scenes whose mask is a subset */
  int nw;    /* 2: walks the binary tree; 4: the 4-wide tree (only scenes built with RT_BVH_WIDE=1 have one) */
};
#define STACK_SMEM(block) ((size_t)RTC_SMEM_STACK * (block) * sizeof(uint32_t)) /* WalkerT's shared-memory entries */
static const MegaVariant mega_variants[] = {
    {2, 6, mega_queue_kernel<6, 2>, mega_queue_smem<2>(), MEGA_BLOCK, FEAT_ALL, 2}, /* first entry of an nv = its default */
    {4, 4, mega_queue_kernel<4, 4>, mega_queue_smem<4>(), MEGA_BLOCK, FEAT_ALL, 2},
    {0, 8, mega_render_kernel<8, MEGA_BLOCK, false, false, FEAT_ALL, 2, true>, STACK_SMEM(MEGA_BLOCK), MEGA_BLOCK,
     FEAT_ALL, 2},
    {0, 6, mega_render_kernel<6, MEGA_BLOCK, false, false, FEAT_ALL>, STACK_SMEM(MEGA_BLOCK), MEGA_BLOCK, FEAT_ALL, 2},
    {0, 4, mega_render_kernel<4, MEGA_BLOCK, false, false, FEAT_ALL>, STACK_SMEM(MEGA_BLOCK), MEGA_BLOCK, FEAT_ALL, 2},
    {-1, 4, mega_render_kernel<4, 256, true, false, FEAT_ALL>, STACK_SMEM(256), 256, FEAT_ALL, 2},
    {-1, 2, mega_render_kernel<2, 512, true, false, FEAT_ALL>, STACK_SMEM(512), 512, FEAT_ALL, 2},
    {-1, 1, mega_render_kernel<1, 1024, true, false, FEAT_ALL>, STACK_SMEM(1024), 1024, FEAT_ALL, 2},
    /* nv = -2: lockstep CTAs whose trips do not wait for their longest rays (RESUME); nv = -3: the same, free-running */
    {-2, 2, mega_render_kernel<2, 512, true, true, FEAT_ALL>, STACK_SMEM(512), 512, FEAT_ALL, 2},
    {-2, 4, mega_render_kernel<4, 256, true, true, FEAT_ALL>, STACK_SMEM(256), 256, FEAT_ALL, 2},
    {-2, 3, mega_render_kernel<2, 384, true, true, FEAT_ALL>, STACK_SMEM(384), 384, FEAT_ALL, 2}, /* 80 registers */
    {-3, 8, mega_render_kernel<8, MEGA_BLOCK, false, true, FEAT_ALL>, STACK_SMEM(MEGA_BLOCK), MEGA_BLOCK, FEAT_ALL, 2},
    /* LEAN instantiations: the same kernels without the code of what a scene does not contain.  FEATS = 0: spheres
     * with solid-colour lambertian / metal / dielectric / emissive materials (scene 0, Book 1 final); quads + instances
     * + light sampling, no media, no textures (the Cornell box). */
    {0, 8, mega_render_kernel<8, MEGA_BLOCK, false, false, 0, 2, false, true>, STACK_SMEM(MEGA_BLOCK), MEGA_BLOCK, 0, 2},
    {0, 10, mega_render_kernel<10, MEGA_BLOCK, false, false, 0>, STACK_SMEM(MEGA_BLOCK), MEGA_BLOCK, 0, 2},
    {0, 12, mega_render_kernel<12, MEGA_BLOCK, false, false, 0>, STACK_SMEM(MEGA_BLOCK), MEGA_BLOCK, 0, 2},
    {-1, 4, mega_render_kernel<4, 256, true, false, FEAT_QUADS | FEAT_INSTANCES | FEAT_LIGHTS>, STACK_SMEM(256), 256,
     FEAT_QUADS | FEAT_INSTANCES | FEAT_LIGHTS, 2},
    {0, 8, mega_render_kernel<8, MEGA_BLOCK, false, false, FEAT_QUADS | FEAT_INSTANCES | FEAT_LIGHTS, 2, true, true>,
     STACK_SMEM(MEGA_BLOCK), MEGA_BLOCK, FEAT_QUADS | FEAT_INSTANCES | FEAT_LIGHTS, 2},
    /* sphere-only worlds with an oversized primitive tested before the walk (FEAT_PRE): lean, and with everything */
    {0, 8, mega_render_kernel<8, MEGA_BLOCK, false, false, FEAT_PRE, 2, false, true>, STACK_SMEM(MEGA_BLOCK), MEGA_BLOCK, FEAT_PRE,
     2},
    {0, 8, mega_render_kernel<8, MEGA_BLOCK, false, false, FEAT_ALL | FEAT_PRE>, STACK_SMEM(MEGA_BLOCK), MEGA_BLOCK,
     FEAT_ALL | FEAT_PRE, 2},
    /* the three default shapes again, walking the 4-wide tree */
    {0, 8, mega_render_kernel<8, MEGA_BLOCK, false, false, 0, 4>, STACK_SMEM(MEGA_BLOCK), MEGA_BLOCK, 0, 4},
    {0, 8, mega_render_kernel<8, MEGA_BLOCK, false, false, FEAT_QUADS | FEAT_INSTANCES | FEAT_LIGHTS, 4>,
     STACK_SMEM(MEGA_BLOCK), MEGA_BLOCK, FEAT_QUADS | FEAT_INSTANCES | FEAT_LIGHTS, 4},
    {0, 8, mega_render_kernel<8, MEGA_BLOCK, false, false, FEAT_ALL, 4>, STACK_SMEM(MEGA_BLOCK), MEGA_BLOCK, FEAT_ALL, 4},
    {-2, 2, mega_render_kernel<2, 512, true, true, FEAT_ALL, 4>, STACK_SMEM(512), 512, FEAT_ALL, 4},
};
static int env_int(const char *name, int fallback) {
  const char *v = getenv(name);
  return v ? atoi(v) : fallback;
}
/* Which kernel renders a scene.  Default, from measurements on the eight reference scenes: free-running
 * 128-thread CTAs for sphere-only scenes; lockstep CTAs of 256 threads once the scene has quads (Cornell +5 %, the
 * quad scene +12 %), of 512 threads once it has media (Book 2 final +34 %: instruction cache; there also RESUME,
 * +1.7 %).  RT_MEGA_NV (paths per lane: 0 = free-running register kernel, -1 = lockstep, -2 = lockstep + RESUME,
 * -3 = free-running + RESUME, 2/4 = warp ray queue) and RT_MEGA_MINB (CTAs per SM the
 * registers must allow) override; a combination that is not built falls back to the nv's first entry. */
static const MegaVariant *mega_variant(const rt_camera &cam, int n_chunks, int scene_class, int scene_feats) {
  /* bit 30 of scene_feats: the scene has a 4-wide tree (built with RT_BVH_WIDE=1) and the walk should use it */
  const int want_nw = (scene_feats & MEGA_FEATS_WIDE_TREE) ? 4 : 2;
  scene_feats &= ~MEGA_FEATS_WIDE_TREE;
  int nv = env_int("RT_MEGA_NV", -99), minb = env_int("RT_MEGA_MINB", 0);
  if (nv == -99) {
    /* quads without media or textures (the Cornell box): the lean free-running kernel (9 318 Mseg/s) beats the lean
     * lockstep one (8 583) -- what lockstep bought there was instruction-cache room the lean instantiation no longer
     * needs */
    const bool cornell_like = scene_class == MEGA_SCENE_QUADS &&
                              (scene_feats & ~(FEAT_QUADS | FEAT_INSTANCES | FEAT_LIGHTS)) == 0 &&
                              env_int("RT_MEGA_LEAN", 1) != 0;
    nv = (scene_class == MEGA_SCENE_SPHERES || cornell_like) ? 0 : scene_class == MEGA_SCENE_MEDIA ? -2 : -1;
    if (minb == 0 && nv < 0)
      minb = scene_class == MEGA_SCENE_MEDIA ? 2 : 4;
  }
  if (env_int("RT_MEGA_LEAN", 1) == 0)
    scene_feats = FEAT_ALL | (scene_feats & FEAT_PRE); /* A/B: only the kernels that carry everything */
  int want_nv = nv;
  /* the slot layout packs the pixel into 16+16 bits, the depth into 16 and the chunk into 8 */
  if (want_nv > 0 && (cam.img_width > 65535 || cam.img_height > 65535 || cam.max_depth > 65535 || n_chunks > 256))
    want_nv = 0;
  const int n = (int)(sizeof(mega_variants) / sizeof(mega_variants[0]));
  /* among the kernels of that shape that cover the scene's features: the one that carries the least besides
   * (pass 0: the requested register budget too; pass 1: any) */
  for (int pass = 0; pass < 2; pass++) {
    const MegaVariant *pick = nullptr;
    for (int i = 0; i < n; i++) {
      const MegaVariant &v = mega_variants[i];
      if (v.nv != want_nv || v.nw != want_nw || (pass == 0 && v.minb != minb) || (scene_feats & ~v.feats) != 0)
        continue;
      if (!pick || __builtin_popcount(v.feats) < __builtin_popcount(pick->feats))
        pick = &v;
    }
    if (pick)
      return pick;
  }
  if (want_nw == 4) /* no wide kernel of that shape is built: the binary walk of the same scene */
    return mega_variant(cam, n_chunks, scene_class, scene_feats);
  /* no kernel of the requested shape covers the scene's features (a forced RT_MEGA_NV on a FEAT_PRE scene): any
   * binary-walk kernel that does -- never one that would skip part of the scene */
  for (int i = 0; i < n; i++)
    if (mega_variants[i].nw == 2 && mega_variants[i].nv <= 0 && (scene_feats & ~mega_variants[i].feats) == 0)
      return &mega_variants[i];
  for (int i = 0; i < n; i++)
    if (mega_variants[i].nv == 0 && mega_variants[i].feats == FEAT_ALL && mega_variants[i].nw == 2)
      return &mega_variants[i];
  return &mega_variants[0];
}

cudaError_t mega_plan(const rt_camera &cam, int sample_count, uint32_t rng_mode, int device, int scene_class,
                      int scene_feats, MegaPlan *plan) {
  int sms = 0, per_sm = 0;
  cudaError_t err = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  if (err != cudaSuccess)
    return err;
  /* chunking first (the variant depends on it only through a capacity check); the grid size is needed for the
   * chunk count, so take the default variant's occupancy for that estimate */
  const MegaVariant *var = mega_variant(cam, 1, scene_class, scene_feats);
  if (var->smem > 48 * 1024) {
    err = cudaFuncSetAttribute(var->fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)var->smem);
    if (err != cudaSuccess)
      return err;
  }
  err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, var->fn, var->block, var->smem);
  if (err != cudaSuccess)
    return err;
  if (per_sm < 1)
    per_sm = 1;
  plan->blocks = sms * per_sm;
  plan->threads = var->block;
  const long long paths = (long long)plan->blocks * var->block * (var->nv > 0 ? var->nv : 1);
  const long long pixels = (long long)cam.img_width * cam.img_height;
  int n_chunks = 1;
  if (rng_mode == RT_RNG_SAMPLE && pixels > 0) {
    /* enough items that the end-of-kernel tail (paths waiting for the last items) stays small, but chunks of at
     * least 8 samples so the per-item store and fetch are noise */
    long long want = (64 * paths + pixels - 1) / pixels;
    long long most = sample_count / 8;
    n_chunks = (int)(want < 1 ? 1 : want);
    if (n_chunks > most)
      n_chunks = (int)(most < 1 ? 1 : most);
    if (n_chunks > MEGA_MAX_CHUNKS)
      n_chunks = MEGA_MAX_CHUNKS;
    int env = env_int("RT_MEGA_CHUNKS", 0); /* tuning: that many EQUAL items */
    if (env > 0 && env <= sample_count)
      n_chunks = env;
  }
  plan_equal_chunks(plan, sample_count, n_chunks);
  /* Guided self-scheduling.  Work items are handed out chunk by chunk (all pixels' item 0, then all pixels' item 1,
   * ...) and the kernel ends when the LAST item ends.  Items cost very differently: a pixel on a glass sphere runs
   * depth-50 paths, ~18x the average sample.  With equal items one such item handed out near the end outlasts
   * everything else: timestamps (tools/timeline.py, Book 1 at 125 spp per GPU) show the queue empty at 25.1 ms, the
   * median CTA done at 25.2 ms and the last one at 27.3 ms.  So an item takes a QUARTER of the samples its pixel has
   * left (at least one): whatever an expensive pixel's item costs, three times as much work is still to be handed out
   * behind it.  The chunks are still summed in order (mega_sum_chunks_kernel): images stay deterministic. */
  int guided = env_int("RT_MEGA_GUIDED", MEGA_GUIDED_DEFAULT); /* 0: equal items; d >= 2: an item = 1/d of the rest */
  if (guided == 1)
    guided = MEGA_GUIDED_DEFAULT;
  if (n_chunks > 1 && guided >= 2 && env_int("RT_MEGA_CHUNKS", 0) <= 0) {
    int c = 0, at = 0;
    while (at < sample_count && c < MEGA_MAX_CHUNKS - 1) {
      plan->begin[c++] = at;
      const int left = sample_count - at;
      at += left / guided > 1 ? left / guided : 1;
    }
    if (at < sample_count) /* (more than 64 items are never needed below ~1e8 samples; the last one takes the rest) */
      plan->begin[c++] = at;
    plan->begin[c] = sample_count;
    plan->n_chunks = c;
    plan->chunk = plan->begin[1] - plan->begin[0];
  }
  /* every work item of a pixel owns one partial-sum frame: never more than 8 GiB of them (a 4K frame at 4096 spp
   * uses 33 = 3.3 GB; an 8K x 8K frame would ask for 26 GB) -- the trailing items are merged into the last one */
  if (plan->n_chunks > 2 && pixels > 0) {
    const long long frame_bytes = pixels * 3 * (long long)sizeof(float);
    long long fit = (8ll << 30) / frame_bytes;
    if (fit < 2)
      fit = 2;
    if ((long long)plan->n_chunks > fit) {
      plan->n_chunks = (int)fit;
      plan->begin[plan->n_chunks] = sample_count;
    }
  }
  plan->partial_bytes = plan->n_chunks > 1 ? (size_t)plan->n_chunks * (size_t)pixels * 3 * sizeof(float) : 0;
  if (mega_variant(cam, plan->n_chunks, scene_class, scene_feats) != var) { /* capacity check failed: the register kernel */
    var = mega_variant(cam, plan->n_chunks, scene_class, scene_feats);
    err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, var->fn, var->block, var->smem);
    if (err != cudaSuccess)
      return err;
    plan->blocks = sms * (per_sm < 1 ? 1 : per_sm);
    plan->threads = var->block;
  }
  plan->variant = (int)(var - mega_variants);
  {
    int t = env_int("RT_MEGA_THR8", MEGA_THR8_DEFAULT); /* see walk_round_sync */
    plan->thr8 = t < 0 ? 0 : t > 8 ? 8 : t;
    int r = env_int("RT_MEGA_RESUME8", MEGA_RESUME8_DEFAULT); /* see mega_render_kernel, RESUME */
    plan->resume8 = r < 0 ? 0 : r > 8 ? 8 : r;
    int c = env_int("RT_MEGA_CTA8", MEGA_CTA8_DEFAULT);
    plan->cta8 = c < 0 ? 0 : c > 8 ? 8 : c;
  }
  plan->nv = var->nv;
  plan->minb = var->minb;
  return cudaSuccess;
}

cudaError_t mega_render(const DScene &sc, const RenderJob &job, const MegaPlan &plan, float *partial,
                        cudaStream_t stream, int *launches) {
  *launches = 0;
  if (plan.variant < 0 || plan.variant >= (int)(sizeof(mega_variants) / sizeof(mega_variants[0])))
    return cudaErrorInvalidValue;
  const MegaVariant *var = &mega_variants[plan.variant];
  var->fn<<<plan.blocks, plan.threads, var->smem, stream>>>(sc, job, plan, partial);
  cudaError_t err = cudaGetLastError();
  *launches = 1;
  if (err != cudaSuccess)
    return err;
#ifdef RTC_TIMELINE
  {
    static unsigned long long host_records[TL_MAX_WARPS][8];
    cudaStreamSynchronize(stream);
    cudaMemcpyFromSymbol(host_records, tl_records, sizeof(host_records));
    const unsigned n_warps = plan.blocks * (plan.threads / 32);
    for (unsigned w = 0; w < n_warps && w < TL_MAX_WARPS; w++) {
      const unsigned long long *r = host_records[w];
      printf("TL %u %llu %llu %llu %llu %llu %llu %llu %llu\n", w, r[0], r[1], r[2], r[3], r[4], r[5], r[6], r[7]);
    }
    fflush(stdout);
  }
#endif
  if (plan.n_chunks > 1) {
    size_t n_values = (size_t)job.cam.img_width * job.cam.img_height * 3;
    unsigned blocks = (unsigned)((n_values + 255) / 256);
    mega_sum_chunks_kernel<<<blocks, 256, 0, stream>>>(partial, job.radiance, n_values, plan.n_chunks);
    err = cudaGetLastError();
    *launches = 2;
  }
  return err;
}

