/*
 * rtc_wave.cu -- RT_KERNEL_WAVE: the wavefront variant of the render loop.
 *
 * The recursion of Camera_ray_color (/root/reference/src/raytracing.c:39-84) is cut at its segment boundary into
 * two kernels per iteration that communicate through a structure-of-arrays pool of path states in HBM:
 *
 *   wave_extend   one closest-hit query per live slot over the rebuilt BVH (raytracing.c:45) -- a small kernel, 64
 *                 registers, 32 warps per SM -- then classifies the hit by material and appends the slot to that
 *                 material's queue (ballot-aggregated atomics: one atomic per warp and class)
 *   wave_shade    walks the six queues back to back (miss, emitter, lambertian, metal, dielectric, isotropic), so a
 *                 warp runs ONE material's code (material.c:103-142, the Book-3 mixture raytracing.c:60-71); a path
 *                 that ends regenerates in place: next sample of its work item, or a new item (raytracing.c:93-122)
 *
 * (wave_fill seeds the pool once.)  Against the megakernel this trades ~250 bytes of HBM traffic per segment for
 * (a) no divergence between materials and no shading registers in the traversal, (b) full occupancy and a loop that
 * fits the instruction cache.  Work items and per-item partial sums are the megakernel's, so images are
 * deterministic and the two variants agree sample for sample up to fast-math rounding.  The host enqueues
 * WAVE_BATCH iterations (2 launches each) at a time and reads one counter between batches to know when to stop.
 */
#define RTC_FAST_SHADING
#include "rtc_kernels.h"
#include "rtc_trace.cuh"
#include <cstdlib>

#define WAVE_BLOCK 256
#define WAVE_CLASSES 6 /* 0 miss, 1 emitter (DIFFUSE_LIGHT / SURFACE_NORMAL), 2 lambertian, 3 metal, 4 dielectric, 5 isotropic */
#define WAVE_BATCH 32  /* iterations per graph launch; the host looks at the done-counter between batches */

enum { SLOT_NEED = 0u, SLOT_ALIVE = 1u, SLOT_DONE = 2u };

struct WaveCtl {
  uint32_t queue_count[2][8]; /* per parity: WAVE_CLASSES counters (+ padding) */
  uint32_t done;              /* slots that ran out of work items */
  uint32_t pad[15];
};

struct WavePool {
  float4 *o_bx;     /* ray origin, beta.x */
  float4 *d_by;     /* ray direction, beta.y */
  float4 *L_bz;     /* radiance of the current sample, beta.z */
  float4 *sum;      /* sum over the finished samples of the item */
  uint4 *rng;       /* PCG32 state, inc */
  uint4 *item;      /* pixel index, next sample, end sample, chunk id */
  uint4 *hit;       /* t bits, primitive, instance, - */
  uint32_t *status; /* SLOT_* | depth << 8 */
  uint32_t *queue;  /* WAVE_CLASSES queues of `slots` entries */
  WaveCtl *ctl;
  uint32_t slots;
};

RTC_DEV bool wave_has_nan(V3 a) { return a.x != a.x || a.y != a.y || a.z != a.z; }

RTC_DEV Rng rng_unpack(uint4 v) {
  Rng g;
  g.state = ((uint64_t)v.y << 32) | v.x, g.inc = ((uint64_t)v.w << 32) | v.z;
  return g;
}
RTC_DEV uint4 rng_pack(const Rng &g) {
  return make_uint4((uint32_t)g.state, (uint32_t)(g.state >> 32), (uint32_t)g.inc, (uint32_t)(g.inc >> 32));
}

/* ------------------------------------------------------------------ regeneration ------------------------- */

/* The slot's path has ended (or it never had one): store a finished item, fetch the next one if needed, and build
 * the camera ray of the next sample (raytracing.c:93-122).  Leaves the slot ALIVE, NEED (max_depth == 0) or DONE. */
__device__ __noinline__ unsigned wave_regenerate(const RenderJob &job, const MegaPlan &plan, float *partial,
                                             const WavePool &pool, uint32_t slot, uint4 it, float4 sum, Rng rng) {
  const rt_camera &cam = job.cam;
  const int w = cam.img_width, h = cam.img_height;
  unsigned started = 0; /* camera samples begun here (counted per warp by the caller: no same-address atomics) */
  for (;;) {
    bool have = it.z != 0u; /* s_end == 0 marks "no item yet" */
    if (have && it.y == it.z) { /* item finished: one store */
      float *dst = (plan.n_chunks > 1 ? partial + (size_t)it.w * (size_t)w * h * 3 : job.radiance) + (size_t)it.x * 3;
      dst[0] = sum.x, dst[1] = sum.y, dst[2] = sum.z;
      have = false;
    }
    if (!have) {
      const unsigned tiles = (unsigned)(((w + 7) >> 3) * ((h + 3) >> 2));
      const unsigned n_items = tiles * 32u * (unsigned)plan.n_chunks;
      const int tiles_x = (w + 7) >> 3;
      for (;;) {
        unsigned id = atomicAdd(&job.counters->next_item, 1u);
        if (id >= n_items)
          break;
        unsigned c = id / (tiles * 32u), pix = id - c * (tiles * 32u);
        unsigned tile = pix >> 5, lane = pix & 31u;
        int ty = (int)(tile / (unsigned)tiles_x), tx = (int)(tile - (unsigned)ty * (unsigned)tiles_x);
        int pi = tx * 8 + (int)(lane & 7u), pj = ty * 4 + (int)(lane >> 3);
        if (pi >= w || pj >= h)
          continue; /* padding of the last tile column / row */
        it.x = (uint32_t)(pj * w + pi);
        it.y = (uint32_t)(job.sample_begin + plan_chunk_begin(plan, (int)c));
        it.z = (uint32_t)(job.sample_begin + min(plan_chunk_begin(plan, (int)c + 1), job.sample_count));
        it.w = c;
        sum = make_float4(0, 0, 0, 0);
        have = true;
        break;
      }
    }
    if (!have) {
      pool.status[slot] = SLOT_DONE;
      atomicAdd(&pool.ctl->done, 1u);
      return started;
    }
    const int pj = (int)(it.x / (uint32_t)w), pi = (int)(it.x - (uint32_t)pj * (uint32_t)w);
    const int s = (int)it.y;
    if (job.rng_mode == RT_RNG_SAMPLE)
      rng_seed_sample(rng, pi, pj, s);
    else if (s == job.sample_begin)
      rng_seed_pixel(rng, pi, pj);
    const V3 du = v_load(cam.pixel_delta_u), dv = v_load(cam.pixel_delta_v), eye = v_load(cam.origin);
    V3 pixel = v_add(v_add(v_load(cam.pixel00), v_scale(du, (float)pi)), v_scale(dv, (float)pj));
    float jx = rng_between(rng, -0.5f, 0.5f);
    float jy = rng_between(rng, -0.5f, 0.5f);
    V3 o = eye;
    if (cam.dof_angle > 0.0f) {
      float a, b;
      for (;;) {
        a = rng_between(rng, -1.0f, 1.0f);
        b = rng_between(rng, -1.0f, 1.0f);
        if (a * a + b * b < 1.0f)
          break;
      }
      o = v_add(v_add(eye, v_scale(v_load(cam.dof_disc_u), a)), v_scale(v_load(cam.dof_disc_v), b));
    }
    V3 d = v_add(v_add(v_add(pixel, v_scale(du, jx)), v_scale(dv, jy)), v_neg(o));
    started++;
    if (cam.max_depth <= 0) { /* raytracing.c:40-41: nothing to trace, the sample contributes 0 */
      it.y++;
      continue;
    }
    pool.o_bx[slot] = make_float4(o.x, o.y, o.z, 1.0f);
    pool.d_by[slot] = make_float4(d.x, d.y, d.z, 1.0f);
    pool.L_bz[slot] = make_float4(0, 0, 0, 1.0f);
    pool.sum[slot] = sum;
    pool.rng[slot] = rng_pack(rng);
    pool.item[slot] = it;
    pool.status[slot] = SLOT_ALIVE; /* depth 0 */
    return started;
  }
}

/* adds per-lane event counts to the global counters with one atomic per warp (all 32 lanes must call) */
RTC_DEV void wave_count(unsigned long long *counter, unsigned mine) {
  unsigned total = __reduce_add_sync(0xffffffffu, mine);
  if ((threadIdx.x & 31u) == 0u && total)
    atomicAdd(counter, (unsigned long long)total);
}

__global__ void __launch_bounds__(WAVE_BLOCK)
    wave_fill_kernel(const RenderJob job, const __grid_constant__ MegaPlan plan, float *partial, WavePool pool) {
  const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
  unsigned started = 0;
  if (slot < pool.slots) {
    Rng rng;
    rng.state = 0, rng.inc = 1;
    started = wave_regenerate(job, plan, partial, pool, slot, make_uint4(0, 0, 0, 0), make_float4(0, 0, 0, 0), rng);
  }
  wave_count(&job.counters->samples, started);
}

/* ------------------------------------------------------------------ extend + classify -------------------- */

__global__ void __launch_bounds__(WAVE_BLOCK)
    wave_extend_kernel(const DScene sc, const RenderJob job, WavePool pool, int parity) {
  const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
  const bool alive = slot < pool.slots && (pool.status[slot] & 255u) == SLOT_ALIVE;
  int cls = -1;
  if (alive) {
    const float4 o = pool.o_bx[slot], d = pool.d_by[slot];
    Ray ray;
    ray.o = v3(o.x, o.y, o.z), ray.d = v3(d.x, d.y, d.z);
    TraceHit th = trace_group<true>(sc, sc.world_root, ray, 1e-3f, RTC_INF);
    if (sc.n_media) { /* the medium's free-path draw comes from the path's own stream */
      const uint4 rs = pool.rng[slot];
      Rng rng = rng_unpack(rs);
      trace_media<false>(sc, ray, 1e-3f, th, rng);
      pool.rng[slot] = rng_pack(rng);
    }
    pool.hit[slot] = make_uint4(__float_as_uint(th.t), th.prim, th.inst, 0u);
    cls = 0;
    if (th.prim != PRIM_NONE) {
      uint32_t material;
      const uint32_t kind = REF_KIND(th.prim), idx = REF_INDEX(th.prim);
      if (kind == PRIM_MEDIUM_KIND)
        material = sc.media[idx].material;
      else if (kind == REF_SPHERES)
        material = __ldg(&sc.msph_meta[idx].material);
      else
        material = __float_as_uint(__ldg(&sc.mquads[idx].q_mat).w);
      const uint32_t tag = __ldg(&sc.materials[material].tag);
      cls = tag == RT_MAT_LAMBERTIAN ? 2 : tag == RT_MAT_METAL ? 3 : tag == RT_MAT_DIELECTRIC ? 4
            : tag == RT_MAT_ISOTROPIC ? 5 : 1;
    }
  }
  /* per-material queues: one atomic per warp and class */
  const unsigned lane = threadIdx.x & 31u;
  uint32_t *count = pool.ctl->queue_count[parity];
#pragma unroll
  for (int c = 0; c < WAVE_CLASSES; c++) {
    const unsigned mask = __ballot_sync(0xffffffffu, cls == c);
    if (mask == 0u)
      continue;
    unsigned base = 0;
    const int leader = __ffs(mask) - 1;
    if ((int)lane == leader)
      base = atomicAdd(&count[c], (unsigned)__popc(mask));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (cls == c)
      pool.queue[(size_t)c * pool.slots + base + __popc(mask & ((1u << lane) - 1u))] = slot;
  }
  const unsigned n_alive = __popc(__ballot_sync(0xffffffffu, alive));
  if (lane == 0 && n_alive)
    atomicAdd(&job.counters->segments, (unsigned long long)n_alive);
}

/* ------------------------------------------------------------------ shade -------------------------------- */

__global__ void __launch_bounds__(WAVE_BLOCK)
    wave_shade_kernel(const DScene sc, const RenderJob job, const __grid_constant__ MegaPlan plan, float *partial, WavePool pool,
                      int parity) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i == 0) { /* the other parity's counters are read by nobody now: reset them for the next extend */
#pragma unroll
    for (int c = 0; c < WAVE_CLASSES; c++)
      pool.ctl->queue_count[parity ^ 1][c] = 0;
  }
  /* position i of the concatenated queues -> (class, entry) */
  const uint32_t *count = pool.ctl->queue_count[parity];
  int cls = -1;
  uint32_t entry = i;
#pragma unroll
  for (int c = 0; c < WAVE_CLASSES; c++) {
    const uint32_t n = count[c];
    if (cls < 0) {
      if (entry < n)
        cls = c;
      else
        entry -= n;
    }
  }
  unsigned n_started = 0, n_pdf = 0, n_nan = 0;
  if (cls >= 0) {
  const rt_camera &cam = job.cam;
  const float prob = cam.lights_sampling_prob;
  const bool use_lights = !(prob == 0.0f || sc.lights_root == RT_NONE);
  const uint32_t slot = pool.queue[(size_t)cls * pool.slots + entry];
  float4 o = pool.o_bx[slot], d = pool.d_by[slot], Lb = pool.L_bz[slot];
  V3 beta = v3(o.w, d.w, Lb.w), L = v3(Lb.x, Lb.y, Lb.z);
  int depth = (int)(pool.status[slot] >> 8);
  Rng rng = rng_unpack(pool.rng[slot]);
  bool ended = false;
  if (cls == 0) {
    L = v_add(L, v_mul(beta, v_load(cam.background)));
    ended = true;
  } else {
    Ray ray;
    ray.o = v3(o.x, o.y, o.z), ray.d = v3(d.x, d.y, d.z);
    const uint4 hv = pool.hit[slot];
    TraceHit th;
    th.t = __uint_as_float(hv.x), th.prim = hv.y, th.inst = hv.z;
    Hit rec;
    trace_finalize<false>(sc, ray, th, rec);
    if (cls == 1) {
      L = v_add(L, v_mul(beta, material_emit<false>(sc, rec)));
      ended = true;
    } else {
      V3 r_out, albedo;
      bool skip_pdf;
      if (!material_scatter<false>(sc, rec, ray.d, r_out, albedo, skip_pdf, rng)) {
        ended = true;
      } else {
        if (!skip_pdf && use_lights) { /* Book-3 mixture (raytracing.c:60-71) */
          if (rng_f32(rng) < prob)
            r_out = lights_rand<false>(sc, sc.lights_root, rec.p, rng);
          float scatter_pdf = material_scatter_pdf(sc, rec.material, rec.n, r_out);
          n_pdf++;
          float sampling_pdf = (1.0f - prob) * scatter_pdf + prob * lights_pdf<false>(sc, sc.lights_root, rec.p, r_out);
          albedo = v_scale(albedo, f_div(scatter_pdf, sampling_pdf));
        }
        beta = v_mul(beta, albedo);
        o = make_float4(rec.p.x, rec.p.y, rec.p.z, beta.x);
        d = make_float4(r_out.x, r_out.y, r_out.z, beta.y);
        depth++;
        ended = depth >= cam.max_depth;
      }
    }
  }
  if (!ended) {
    pool.o_bx[slot] = o;
    pool.d_by[slot] = d;
    pool.L_bz[slot] = make_float4(L.x, L.y, L.z, beta.z);
    pool.rng[slot] = rng_pack(rng);
    pool.status[slot] = SLOT_ALIVE | ((uint32_t)depth << 8);
  } else {
  uint4 it = pool.item[slot];
  it.y++;
  float4 sum = pool.sum[slot];
  if (wave_has_nan(L)) {
    n_nan++;
    if (!(job.flags & RT_FLAG_DROP_NAN))
      sum.x += L.x, sum.y += L.y, sum.z += L.z;
  } else {
    sum.x += L.x, sum.y += L.y, sum.z += L.z;
  }
  n_started = wave_regenerate(job, plan, partial, pool, slot, it, sum, rng);
  }
  }
  wave_count(&job.counters->samples, n_started);
  wave_count(&job.counters->light_pdf_evals, n_pdf);
  wave_count(&job.counters->nan_samples, n_nan);
}

__global__ void wave_sum_chunks_kernel(const float *__restrict__ partial, float *__restrict__ radiance, size_t n_values,
                                       int n_chunks) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_values)
    return;
  float acc = partial[i];
  for (int c = 1; c < n_chunks; c++)
    acc += partial[(size_t)c * n_values + i];
  radiance[i] = acc;
}

/* ------------------------------------------------------------------ host side ---------------------------- */

static size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

cudaError_t wave_plan(const rt_camera &cam, int sample_count, uint32_t rng_mode, int device, WavePlan *plan) {
  int sms = 0;
  cudaError_t err = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  if (err != cudaSuccess)
    return err;
  const long long pixels = (long long)cam.img_width * cam.img_height;
  int n_chunks = 1;
  const char *chunk_env = getenv("RT_WAVE_CHUNKS");
  if (rng_mode == RT_RNG_SAMPLE) {
    /* items small enough that the drain at the end (slots running out of items) stays short */
    n_chunks = sample_count / 16;
    if (n_chunks < 1)
      n_chunks = 1;
    if (n_chunks > 64)
      n_chunks = 64;
    if (chunk_env && atoi(chunk_env) > 0 && atoi(chunk_env) <= sample_count)
      n_chunks = atoi(chunk_env);
  }
  plan->mega.blocks = 0, plan->mega.threads = 0, plan->mega.minb = 0, plan->mega.nv = 0;
  plan_equal_chunks(&plan->mega, sample_count, n_chunks);
  plan->mega.partial_bytes =
      plan->mega.n_chunks > 1 ? (size_t)plan->mega.n_chunks * (size_t)pixels * 3 * sizeof(float) : 0;
  /* pool: many full waves of every SM per launch (the two launches of an iteration cost a few microseconds of
   * fixed latency each), never more slots than there are work items */
  const char *waves_env = getenv("RT_WAVE_WAVES"); /* tuning: pool size in full waves of the machine */
  const int waves = waves_env && atoi(waves_env) > 0 ? atoi(waves_env) : 4;
  long long slots = (long long)sms * 2048 * waves;
  const long long items = pixels * plan->mega.n_chunks;
  if (slots > items)
    slots = items;
  slots = (slots + WAVE_BLOCK - 1) / WAVE_BLOCK * WAVE_BLOCK;
  plan->slots = (uint32_t)slots;
  plan->shade_blocks = (int)(slots / WAVE_BLOCK);
  const size_t s = (size_t)slots;
  plan->pool_bytes = 7 * align256(s * 16) + align256(s * 4) + align256(s * 4 * WAVE_CLASSES) + align256(sizeof(WaveCtl));
  plan->scratch_bytes = align256(plan->mega.partial_bytes) + plan->pool_bytes;
  return cudaSuccess;
}

cudaError_t wave_render(const DScene &sc, const RenderJob &job, const WavePlan &plan, void *scratch,
                        cudaStream_t stream, int *launches) {
  char *base = (char *)scratch;
  float *partial = (float *)base;
  char *p = base + align256(plan.mega.partial_bytes);
  const size_t s = plan.slots;
  WavePool pool;
  pool.slots = plan.slots;
  auto take = [&](size_t bytes) {
    char *q = p;
    p += align256(bytes);
    return q;
  };
  pool.o_bx = (float4 *)take(s * 16), pool.d_by = (float4 *)take(s * 16), pool.L_bz = (float4 *)take(s * 16);
  pool.sum = (float4 *)take(s * 16), pool.rng = (uint4 *)take(s * 16), pool.item = (uint4 *)take(s * 16);
  pool.hit = (uint4 *)take(s * 16);
  pool.status = (uint32_t *)take(s * 4);
  pool.queue = (uint32_t *)take(s * 4 * WAVE_CLASSES);
  pool.ctl = (WaveCtl *)take(sizeof(WaveCtl));

  cudaError_t err;
  if ((err = cudaMemsetAsync(pool.ctl, 0, sizeof(WaveCtl), stream)) != cudaSuccess)
    return err;
  const unsigned blocks = (unsigned)(s / WAVE_BLOCK);
  int n_launch = 1;
  wave_fill_kernel<<<blocks, WAVE_BLOCK, 0, stream>>>(job, plan.mega, partial, pool);

  /* generous bound on the number of iterations, so that a bug cannot spin forever */
  const long long items_per_slot =
      ((long long)job.cam.img_width * job.cam.img_height * plan.mega.n_chunks + (long long)s - 1) / (long long)s + 1;
  const long long max_iters = 64 + items_per_slot * (long long)plan.mega.chunk * ((long long)job.cam.max_depth + 1);
  bool finished = false;
  for (long long iter = 0; iter < max_iters;) {
    for (int k = 0; k < WAVE_BATCH; k++) {
      wave_extend_kernel<<<blocks, WAVE_BLOCK, 0, stream>>>(sc, job, pool, k & 1);
      wave_shade_kernel<<<blocks, WAVE_BLOCK, 0, stream>>>(sc, job, plan.mega, partial, pool, k & 1);
    }
    n_launch += 2 * WAVE_BATCH;
    iter += WAVE_BATCH;
    if ((err = cudaGetLastError()) != cudaSuccess)
      return err;
    uint32_t done = 0;
    if ((err = cudaMemcpyAsync(&done, &pool.ctl->done, 4, cudaMemcpyDeviceToHost, stream)) != cudaSuccess ||
        (err = cudaStreamSynchronize(stream)) != cudaSuccess)
      return err;
    if (done >= plan.slots) {
      finished = true;
      break;
    }
  }
  if (!finished) /* the iteration bound ran out with live paths: an incomplete frame is an error, not a result */
    return cudaErrorLaunchTimeout;
  if (plan.mega.n_chunks > 1) {
    size_t n_values = (size_t)job.cam.img_width * job.cam.img_height * 3;
    wave_sum_chunks_kernel<<<(unsigned)((n_values + 255) / 256), 256, 0, stream>>>(partial, job.radiance, n_values,
                                                                                  plan.mega.n_chunks);
    n_launch++;
  }
  *launches = n_launch;
  return cudaGetLastError();
}
