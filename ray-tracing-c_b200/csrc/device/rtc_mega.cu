/*
 * rtc_mega.cu -- RT_KERNEL_MEGA: the persistent megakernel (throughput path).
 *
 * Replaces the loop nest of Camera_render + Camera_ray_color (/root/reference/src/raytracing.c:39-133).
 * One CTA set sized to the machine (SM count x resident CTAs) stays resident; every lane owns one path at a time
 * and, the moment its path ends, regenerates: next sample of its work item, or a new work item fetched from a
 * global counter.  So every trip of the main loop is "one segment for 32 lanes" -- no lane idles on a finished
 * path, which is what depth-50 tails (dielectrics) would otherwise cost.
 *
 * Work item = (pixel, sample chunk).  A lane sums its chunk in registers and stores the partial sum once; chunks
 * of a pixel are added up in chunk order by a second tiny kernel, so the image is deterministic (no float atomics).
 * The recursion  colour = emitted + albedo * colour(next) * ratio  is run forwards as  L += beta * emitted;
 * beta *= albedo * ratio  (raytracing.c:57-58,69-71).
 */
#define RTC_FAST_SHADING /* approximate div/sqrt in the shading code only; see rtc_math.cuh */
#include "rtc_kernels.h"
#include "rtc_trace.cuh"
#include <cstdlib>

#define MEGA_BLOCK 128

RTC_DEV bool has_nan3(V3 a) { return a.x != a.x || a.y != a.y || a.z != a.z; }

/* pixel order: 8x4 tiles, row-major inside a tile, so the 32 lanes of a warp start on one tile */
RTC_DEV void item_to_pixel(unsigned item_pix, int w, int h, int &i, int &j) {
  const int tiles_x = (w + 7) >> 3;
  unsigned tile = item_pix >> 5, lane = item_pix & 31u;
  int ty = (int)(tile / (unsigned)tiles_x), tx = (int)(tile - (unsigned)ty * (unsigned)tiles_x);
  i = tx * 8 + (int)(lane & 7u);
  j = ty * 4 + (int)(lane >> 3);
}

/* MINB = resident CTAs per SM the register allocation must allow (occupancy vs spills; chosen from measurements,
 * RT_MEGA_MINB overrides for tuning) */
template <int MINB, int TRAV>
__global__ void __launch_bounds__(MEGA_BLOCK, MINB)
    mega_render_kernel(const DScene sc, const RenderJob job, const MegaPlan plan, float *partial) {
  const rt_camera &cam = job.cam;
  const int w = cam.img_width, h = cam.img_height;
  const unsigned tiles = (unsigned)(((w + 7) >> 3) * ((h + 3) >> 2));
  const unsigned n_items = tiles * 32u * (unsigned)plan.n_chunks;
  const V3 px00 = v_load(cam.pixel00), du = v_load(cam.pixel_delta_u), dv = v_load(cam.pixel_delta_v);
  const V3 eye = v_load(cam.origin), lens_u = v_load(cam.dof_disc_u), lens_v = v_load(cam.dof_disc_v);
  const V3 background = v_load(cam.background);
  const float prob = cam.lights_sampling_prob;
  const bool use_lights = !(prob == 0.0f || sc.lights_root == RT_NONE);
  const int max_depth = cam.max_depth;

  unsigned long long n_segments = 0, n_pdf = 0, n_samples = 0, n_nan = 0;

  /* per-lane path state */
  Rng rng;
  rng.state = 0, rng.inc = 1;
  Ray ray;
  ray.o = ray.d = v3(0, 0, 0);
  V3 beta = v3(1, 1, 1), L = v3(0, 0, 0), sum = v3(0, 0, 0), pixel = v3(0, 0, 0);
  int pi = 0, pj = 0, s = 0, s_end = 0, depth = 0, chunk_id = 0;
  bool have_item = false, alive = false;

  bool done = false; /* this lane has no work left; it stays in the loop (idle) until its whole warp is done */
  for (;;) {
    /* ---- phase 1: regenerate.  Lanes whose path ended take the next sample of their item or fetch a new item. */
    if (!alive && !done) {
      if (have_item && s == s_end) { /* chunk finished: one store per item */
        float *dst = (plan.n_chunks > 1 ? partial + (size_t)chunk_id * (size_t)w * h * 3 : job.radiance) +
                     ((size_t)pj * w + pi) * 3;
        dst[0] = sum.x, dst[1] = sum.y, dst[2] = sum.z;
        have_item = false;
      }
      if (!have_item) {
        for (;;) {
          unsigned item = atomicAdd(&job.counters->next_item, 1u);
          if (item >= n_items)
            break;
          unsigned c = item / (tiles * 32u), pix = item - c * (tiles * 32u);
          item_to_pixel(pix, w, h, pi, pj);
          if (pi >= w || pj >= h)
            continue; /* padding of the last tile column / row */
          chunk_id = (int)c;
          s = job.sample_begin + (int)c * plan.chunk;
          s_end = min(s + plan.chunk, job.sample_begin + job.sample_count);
          sum = v3(0, 0, 0);
          pixel = v_add(v_add(px00, v_scale(du, (float)pi)), v_scale(dv, (float)pj));
          if (job.rng_mode == RT_RNG_PIXEL)
            rng_seed_pixel(rng, pi, pj);
          have_item = true;
          break;
        }
        if (!have_item)
          done = true;
      }
      if (!done) { /* camera ray of sample s (raytracing.c:100-122) */
        if (job.rng_mode == RT_RNG_SAMPLE)
          rng_seed_sample(rng, pi, pj, s);
        float jx = rng_between(rng, -0.5f, 0.5f);
        float jy = rng_between(rng, -0.5f, 0.5f);
        ray.o = eye;
        if (cam.dof_angle > 0.0f) {
          float a, b;
          for (;;) {
            a = rng_between(rng, -1.0f, 1.0f);
            b = rng_between(rng, -1.0f, 1.0f);
            if (a * a + b * b < 1.0f)
              break;
          }
          ray.o = v_add(v_add(eye, v_scale(lens_u, a)), v_scale(lens_v, b));
        }
        ray.d = v_add(v_add(v_add(pixel, v_scale(du, jx)), v_scale(dv, jy)), v_neg(ray.o));
        beta = v3(1, 1, 1);
        L = v3(0, 0, 0);
        depth = 0;
        alive = max_depth > 0;
        n_samples++;
        if (!alive)
          s++; /* depth 0: the sample contributes nothing (raytracing.c:40-41) */
      }
    }
    /* warp-uniform exit, and the point where all 32 lanes meet again before the traversal */
    if (__all_sync(0xffffffffu, done))
      break;

    /* ---- phase 2: one closest-hit query for every live lane, entered together */
    TraceHit th;
    th.t = RTC_INF, th.prim = PRIM_NONE, th.inst = RT_NONE;
    if (alive) {
      n_segments++;
      th = trace_world<false, TRAV>(sc, ray, 1e-3f, RTC_INF, rng);
    }
    __syncwarp();

    /* ---- phase 3: shade */
    if (alive) {
      bool ended = false;
      if (th.prim == PRIM_NONE) {
        L = v_add(L, v_mul(beta, background));
        ended = true;
      } else {
        Hit rec;
        trace_finalize<false>(sc, ray, th, rec);
        V3 r_out, albedo;
        bool skip_pdf;
        const uint32_t tag = __ldg(&sc.materials[rec.material].tag);
        if (tag == RT_MAT_DIFFUSE_LIGHT || tag == RT_MAT_SURFACE_NORMAL) {
          L = v_add(L, v_mul(beta, material_emit<false>(sc, rec)));
          ended = true;
        } else if (!material_scatter<false>(sc, rec, ray.d, r_out, albedo, skip_pdf, rng)) {
          ended = true;
        } else {
          if (!skip_pdf && use_lights) { /* Book-3 mixture (raytracing.c:60-71) */
            if (rng_f32(rng) < prob)
              r_out = lights_rand<false>(sc, sc.lights_root, rec.p, rng);
            float scatter_pdf = material_scatter_pdf(sc, rec.material, rec.n, r_out);
            n_pdf++;
            float sampling_pdf =
                (1.0f - prob) * scatter_pdf + prob * lights_pdf<false>(sc, sc.lights_root, rec.p, r_out);
            albedo = v_scale(albedo, f_div(scatter_pdf, sampling_pdf));
          }
          beta = v_mul(beta, albedo);
          ray.o = rec.p;
          ray.d = r_out;
          depth++;
          ended = depth >= max_depth; /* the next call would return 0 (raytracing.c:40-41) */
        }
      }
      if (ended) {
        alive = false;
        s++;
        if (has_nan3(L)) {
          n_nan++;
          if (!(job.flags & RT_FLAG_DROP_NAN))
            sum = v_add(sum, L);
        } else {
          sum = v_add(sum, L);
        }
      }
    }
  }
  /* a lane that ran out of items while holding a finished chunk already stored it in phase 1 */

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

typedef void (*MegaKernel)(const DScene, const RenderJob, const MegaPlan, float *);
static MegaKernel mega_kernel_for(int minb, int trav) {
  if (trav == 1)
    return minb == 4 ? mega_render_kernel<4, 1> : minb == 6 ? mega_render_kernel<6, 1> : mega_render_kernel<8, 1>;
  return minb == 4 ? mega_render_kernel<4, 0> : minb == 6 ? mega_render_kernel<6, 0> : mega_render_kernel<8, 0>;
}
static int mega_trav() { /* 0 = while-while, 1 = leaves intersected inside the node loop */
  static int cached = -1;
  if (cached < 0) {
    const char *v = getenv("RT_MEGA_TRAV");
    cached = v ? (atoi(v) == 1 ? 1 : 0) : 0;
  }
  return cached;
}
static int mega_minb() {
  static int cached = 0;
  if (cached == 0) {
    const char *v = getenv("RT_MEGA_MINB");
    int m = v ? atoi(v) : 8;
    cached = (m == 4 || m == 6 || m == 8) ? m : 8;
  }
  return cached;
}

cudaError_t mega_plan(const rt_camera &cam, int sample_count, uint32_t rng_mode, int device, MegaPlan *plan) {
  int sms = 0, per_sm = 0;
  cudaError_t err = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  if (err != cudaSuccess)
    return err;
  plan->minb = mega_minb();
  plan->trav = mega_trav();
  err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, mega_kernel_for(plan->minb, plan->trav), MEGA_BLOCK, 0);
  if (err != cudaSuccess)
    return err;
  if (per_sm < 1)
    per_sm = 1;
  plan->blocks = sms * per_sm;
  plan->threads = MEGA_BLOCK;
  const long long lanes = (long long)plan->blocks * MEGA_BLOCK;
  const long long pixels = (long long)cam.img_width * cam.img_height;
  int n_chunks = 1;
  if (rng_mode == RT_RNG_SAMPLE && pixels > 0) {
    /* enough items that the end-of-kernel tail (lanes waiting for the last items) stays small, but chunks of at
     * least 8 samples so the per-item store and fetch are noise */
    long long want = (64 * lanes + pixels - 1) / pixels;
    long long most = sample_count / 8;
    n_chunks = (int)(want < 1 ? 1 : want);
    if (n_chunks > most)
      n_chunks = (int)(most < 1 ? 1 : most);
    if (n_chunks > 64)
      n_chunks = 64;
    const char *env = getenv("RT_MEGA_CHUNKS"); /* tuning */
    if (env && atoi(env) > 0 && atoi(env) <= sample_count)
      n_chunks = atoi(env);
  }
  plan->chunk = (sample_count + n_chunks - 1) / n_chunks;
  plan->n_chunks = plan->chunk > 0 ? (sample_count + plan->chunk - 1) / plan->chunk : 1;
  plan->partial_bytes = plan->n_chunks > 1 ? (size_t)plan->n_chunks * (size_t)pixels * 3 * sizeof(float) : 0;
  return cudaSuccess;
}

cudaError_t mega_render(const DScene &sc, const RenderJob &job, const MegaPlan &plan, float *partial,
                        cudaStream_t stream, int *launches) {
  mega_kernel_for(plan.minb, plan.trav)<<<plan.blocks, plan.threads, 0, stream>>>(sc, job, plan, partial);
  cudaError_t err = cudaGetLastError();
  *launches = 1;
  if (err != cudaSuccess)
    return err;
  if (plan.n_chunks > 1) {
    size_t n_values = (size_t)job.cam.img_width * job.cam.img_height * 3;
    unsigned blocks = (unsigned)((n_values + 255) / 256);
    mega_sum_chunks_kernel<<<blocks, 256, 0, stream>>>(partial, job.radiance, n_values, plan.n_chunks);
    err = cudaGetLastError();
    *launches = 2;
  }
  return err;
}

/* ------------------------------------------------------------------ rt_hit_batch over the rebuilt BVH ---- */

template <int TRAV> __global__ void __launch_bounds__(MEGA_BLOCK) mega_hit_kernel(const DScene sc, const HitBatch b) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= b.n)
    return;
  Ray ray;
  ray.o = v_load(b.origins + 3 * i), ray.d = v_load(b.dirs + 3 * i);
  Rng g;
  g.state = 0, g.inc = 1;
  if (b.rng_io) {
    g.state = b.rng_io[2 * i];
    g.inc = b.rng_io[2 * i + 1];
  }
  TraceHit th = trace_world<true, TRAV>(sc, ray, b.t_min, b.t_max, g);
  if (b.rng_io)
    b.rng_io[2 * i] = g.state;
  const bool hit = th.prim != PRIM_NONE;
  Hit rec;
  rec.p = v3(0, 0, 0), rec.n = v3(0, 0, 0), rec.material = RT_NONE, rec.t = 0, rec.u = 0, rec.v = 0, rec.front = 0,
  rec.node = RT_NONE;
  if (hit)
    trace_finalize<true>(sc, ray, th, rec);
  b.out_t[i] = hit ? rec.t : RTC_INF;
  b.out_node[i] = hit ? rec.node : RT_NONE;
  if (b.out_p)
    b.out_p[3 * i] = rec.p.x, b.out_p[3 * i + 1] = rec.p.y, b.out_p[3 * i + 2] = rec.p.z;
  if (b.out_normal)
    b.out_normal[3 * i] = rec.n.x, b.out_normal[3 * i + 1] = rec.n.y, b.out_normal[3 * i + 2] = rec.n.z;
  if (b.out_uv)
    b.out_uv[2 * i] = rec.u, b.out_uv[2 * i + 1] = rec.v;
  if (b.out_front)
    b.out_front[i] = (uint8_t)rec.front;
  if (b.out_material)
    b.out_material[i] = hit ? rec.material : RT_NONE;
}

cudaError_t mega_hit_batch(const DScene &sc, const HitBatch &b, cudaStream_t stream) {
  if (b.n == 0)
    return cudaSuccess;
  unsigned blocks = (unsigned)((b.n + MEGA_BLOCK - 1) / MEGA_BLOCK);
  if (mega_trav() == 1)
    mega_hit_kernel<1><<<blocks, MEGA_BLOCK, 0, stream>>>(sc, b);
  else
    mega_hit_kernel<0><<<blocks, MEGA_BLOCK, 0, stream>>>(sc, b);
  return cudaGetLastError();
}
