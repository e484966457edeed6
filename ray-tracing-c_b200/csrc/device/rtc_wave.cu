/*
 * rtc_wave.cu -- RT_KERNEL_WAVE: the wavefront variant of the render loop.
 *
 * The recursion of Camera_ray_color (/root/reference/src/raytracing.c:39-84) is cut at its segment boundary into
 * three kinds of kernels that communicate through a structure-of-arrays pool of path states in HBM:
 *
 *   wave_generate   slots whose path ended: finish / fetch a work item (pixel, sample chunk), seed the PCG32 stream,
 *                   build the camera ray (raytracing.c:93-122)
 *   wave_extend     one closest-hit query per live slot over the rebuilt BVH (raytracing.c:45), then classify the
 *                   hit by material and append the slot to that material's queue (ballot-aggregated atomics)
 *   wave_shade      one launch per queue -- miss, emitter, lambertian, metal, dielectric, isotropic -- so every warp
 *                   runs ONE material's code (material.c:103-142, the Book-3 mixture raytracing.c:60-71)
 *
 * Against the megakernel this trades ~230 bytes of HBM traffic per segment for (a) no divergence between materials,
 * (b) a traversal kernel with few registers and full occupancy.  Work items and per-item partial sums are the
 * megakernel's, so images are deterministic and the two variants agree sample for sample up to fast-math rounding.
 */
#define RTC_FAST_SHADING
#include "rtc_kernels.h"
#include "rtc_trace.cuh"

#define WAVE_BLOCK 256
#define WAVE_CLASSES 6 /* 0 miss, 1 emitter (DIFFUSE_LIGHT / SURFACE_NORMAL), 2 lambertian, 3 metal, 4 dielectric, 5 isotropic */

enum { SLOT_NEED = 0u, SLOT_ALIVE = 1u, SLOT_DONE = 2u };

struct WavePool {
  float4 *o_bx;   /* ray origin, beta.x */
  float4 *d_by;   /* ray direction, beta.y */
  float4 *L_bz;   /* radiance of the current sample, beta.z */
  float4 *sum;    /* sum over the finished samples of the item */
  uint4 *rng;     /* PCG32 state, inc */
  uint4 *item;    /* pixel index, next sample, end sample, chunk id */
  uint4 *hit;     /* t bits, primitive, instance, - */
  uint32_t *status; /* SLOT_* | depth << 8 */
  uint32_t *queue;  /* WAVE_CLASSES queues of `slots` entries */
  uint32_t *queue_count; /* WAVE_CLASSES counters */
  uint32_t *alive_hist;  /* live slots after each generate of the current batch */
  uint32_t slots;
};

RTC_DEV bool wave_has_nan(V3 a) { return a.x != a.x || a.y != a.y || a.z != a.z; }

/* ------------------------------------------------------------------ generate ----------------------------- */

__global__ void __launch_bounds__(WAVE_BLOCK)
    wave_generate_kernel(const RenderJob job, const MegaPlan plan, float *partial, WavePool pool, uint32_t iter) {
  const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
  if (slot < WAVE_CLASSES)
    pool.queue_count[slot] = 0; /* the queues of this iteration start empty (extend runs after this kernel) */
  bool live = false;
  if (slot < pool.slots) {
    const rt_camera &cam = job.cam;
    const int w = cam.img_width, h = cam.img_height;
    uint32_t status = pool.status[slot];
    if ((status & 255u) == SLOT_ALIVE) {
      live = true;
    } else if ((status & 255u) == SLOT_NEED) {
      uint4 it = pool.item[slot];
      float4 sum = pool.sum[slot];
      bool have = it.z != 0u; /* s_end == 0 marks "no item yet" */
      if (have && it.y == it.z) { /* item finished: one store */
        float *dst = (plan.n_chunks > 1 ? partial + (size_t)it.w * (size_t)w * h * 3 : job.radiance) + (size_t)it.x * 3;
        dst[0] = sum.x, dst[1] = sum.y, dst[2] = sum.z;
        have = false;
      }
      int pi = 0, pj = 0;
      if (!have) {
        const unsigned tiles = (unsigned)(((w + 7) >> 3) * ((h + 3) >> 2));
        const unsigned n_items = tiles * 32u * (unsigned)plan.n_chunks;
        for (;;) {
          unsigned id = atomicAdd(&job.counters->next_item, 1u);
          if (id >= n_items)
            break;
          unsigned c = id / (tiles * 32u), pix = id - c * (tiles * 32u);
          const int tiles_x = (w + 7) >> 3;
          unsigned tile = pix >> 5, lane = pix & 31u;
          int ty = (int)(tile / (unsigned)tiles_x), tx = (int)(tile - (unsigned)ty * (unsigned)tiles_x);
          pi = tx * 8 + (int)(lane & 7u);
          pj = ty * 4 + (int)(lane >> 3);
          if (pi >= w || pj >= h)
            continue;
          it.x = (uint32_t)(pj * w + pi);
          it.y = (uint32_t)(job.sample_begin + (int)c * plan.chunk);
          it.z = (uint32_t)min((int)it.y + plan.chunk, job.sample_begin + job.sample_count);
          it.w = c;
          sum = make_float4(0, 0, 0, 0);
          have = true;
          break;
        }
      } else {
        pj = (int)(it.x / (uint32_t)w), pi = (int)(it.x - (uint32_t)pj * (uint32_t)w);
      }
      if (!have) {
        pool.status[slot] = SLOT_DONE;
      } else {
        /* camera ray of sample it.y (raytracing.c:100-122) */
        uint4 rs = pool.rng[slot];
        Rng rng;
        rng.state = ((uint64_t)rs.y << 32) | rs.x, rng.inc = ((uint64_t)rs.w << 32) | rs.z;
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
        pool.o_bx[slot] = make_float4(o.x, o.y, o.z, 1.0f);
        pool.d_by[slot] = make_float4(d.x, d.y, d.z, 1.0f);
        pool.L_bz[slot] = make_float4(0, 0, 0, 1.0f);
        pool.sum[slot] = sum;
        pool.rng[slot] = make_uint4((uint32_t)rng.state, (uint32_t)(rng.state >> 32), (uint32_t)rng.inc,
                                    (uint32_t)(rng.inc >> 32));
        atomicAdd(&job.counters->samples, 1ull);
        if (cam.max_depth > 0) {
          pool.item[slot] = it;
          pool.status[slot] = SLOT_ALIVE; /* depth 0 */
          live = true;
        } else { /* raytracing.c:40-41: nothing to trace; the sample contributes 0 */
          it.y++;
          pool.item[slot] = it;
          live = true; /* still has work next iteration */
        }
      }
    }
  }
  unsigned n_live = __popc(__ballot_sync(0xffffffffu, live));
  if ((threadIdx.x & 31) == 0 && n_live)
    atomicAdd(&pool.alive_hist[iter], n_live);
}

/* ------------------------------------------------------------------ extend + classify -------------------- */

__global__ void __launch_bounds__(WAVE_BLOCK) wave_extend_kernel(const DScene sc, const RenderJob job, WavePool pool) {
  const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
  const bool alive = slot < pool.slots && (pool.status[slot] & 255u) == SLOT_ALIVE;
  int cls = -1;
  if (alive) {
    const float4 o = pool.o_bx[slot], d = pool.d_by[slot];
    Ray ray;
    ray.o = v3(o.x, o.y, o.z), ray.d = v3(d.x, d.y, d.z);
    TraceHit th = trace_group(sc, sc.world_root, ray, 1e-3f, RTC_INF);
    if (sc.n_media) { /* the medium's free-path draw comes from the path's own stream */
      uint4 rs = pool.rng[slot];
      Rng rng;
      rng.state = ((uint64_t)rs.y << 32) | rs.x, rng.inc = ((uint64_t)rs.w << 32) | rs.z;
      trace_media<false>(sc, ray, 1e-3f, th, rng);
      pool.rng[slot] = make_uint4((uint32_t)rng.state, (uint32_t)(rng.state >> 32), rs.z, rs.w);
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
#pragma unroll
  for (int c = 0; c < WAVE_CLASSES; c++) {
    const unsigned mask = __ballot_sync(0xffffffffu, cls == c);
    if (mask == 0u)
      continue;
    unsigned base = 0;
    const int leader = __ffs(mask) - 1;
    if ((int)lane == leader)
      base = atomicAdd(&pool.queue_count[c], (unsigned)__popc(mask));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (cls == c)
      pool.queue[(size_t)c * pool.slots + base + __popc(mask & ((1u << lane) - 1u))] = slot;
  }
  const unsigned n_alive = __popc(__ballot_sync(0xffffffffu, alive));
  if (lane == 0 && n_alive)
    atomicAdd(&job.counters->segments, (unsigned long long)n_alive);
}

/* ------------------------------------------------------------------ shade, one launch per queue ---------- */

__global__ void __launch_bounds__(WAVE_BLOCK)
    wave_shade_kernel(const DScene sc, const RenderJob job, WavePool pool, int cls) {
  const uint32_t count = pool.queue_count[cls];
  const rt_camera &cam = job.cam;
  const float prob = cam.lights_sampling_prob;
  const bool use_lights = !(prob == 0.0f || sc.lights_root == RT_NONE);
  unsigned long long n_pdf = 0, n_nan = 0;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
    const uint32_t slot = pool.queue[(size_t)cls * pool.slots + i];
    float4 o = pool.o_bx[slot], d = pool.d_by[slot], Lb = pool.L_bz[slot];
    V3 beta = v3(o.w, d.w, Lb.w), L = v3(Lb.x, Lb.y, Lb.z);
    uint32_t status = pool.status[slot];
    int depth = (int)(status >> 8);
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
        uint4 rs = pool.rng[slot];
        Rng rng;
        rng.state = ((uint64_t)rs.y << 32) | rs.x, rng.inc = ((uint64_t)rs.w << 32) | rs.z;
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
            float sampling_pdf =
                (1.0f - prob) * scatter_pdf + prob * lights_pdf<false>(sc, sc.lights_root, rec.p, r_out);
            albedo = v_scale(albedo, f_div(scatter_pdf, sampling_pdf));
          }
          beta = v_mul(beta, albedo);
          o = make_float4(rec.p.x, rec.p.y, rec.p.z, beta.x);
          d = make_float4(r_out.x, r_out.y, r_out.z, beta.y);
          depth++;
          ended = depth >= cam.max_depth;
        }
        pool.rng[slot] = make_uint4((uint32_t)rng.state, (uint32_t)(rng.state >> 32), rs.z, rs.w);
      }
    }
    if (ended) {
      uint4 it = pool.item[slot];
      it.y++;
      pool.item[slot] = it;
      float4 sum = pool.sum[slot];
      if (wave_has_nan(L)) {
        n_nan++;
        if (!(job.flags & RT_FLAG_DROP_NAN))
          sum.x += L.x, sum.y += L.y, sum.z += L.z;
      } else {
        sum.x += L.x, sum.y += L.y, sum.z += L.z;
      }
      pool.sum[slot] = sum;
      pool.status[slot] = SLOT_NEED;
    } else {
      pool.o_bx[slot] = o;
      pool.d_by[slot] = d;
      pool.L_bz[slot] = make_float4(L.x, L.y, L.z, beta.z);
      pool.status[slot] = SLOT_ALIVE | ((uint32_t)depth << 8);
    }
  }
  if (n_pdf)
    atomicAdd(&job.counters->light_pdf_evals, n_pdf);
  if (n_nan)
    atomicAdd(&job.counters->nan_samples, n_nan);
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

#define WAVE_BATCH 32 /* iterations enqueued between two looks at the live-slot count */

static size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

cudaError_t wave_plan(const rt_camera &cam, int sample_count, uint32_t rng_mode, int device, WavePlan *plan) {
  int sms = 0;
  cudaError_t err = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  if (err != cudaSuccess)
    return err;
  const long long pixels = (long long)cam.img_width * cam.img_height;
  /* pool: enough slots for several full waves of every SM, never more than there are pixels */
  long long slots = (long long)sms * 2048 * 4;
  if (slots > pixels)
    slots = pixels;
  slots = (slots + WAVE_BLOCK - 1) / WAVE_BLOCK * WAVE_BLOCK;
  plan->slots = (uint32_t)slots;
  plan->shade_blocks = sms * 8;
  int n_chunks = 1;
  if (rng_mode == RT_RNG_SAMPLE) {
    /* items small enough that the drain at the end (slots running out of items) stays short */
    n_chunks = sample_count / 64;
    if (n_chunks < 1)
      n_chunks = 1;
    if (n_chunks > 16)
      n_chunks = 16;
  }
  plan->mega.blocks = 0, plan->mega.threads = 0, plan->mega.minb = 0;
  plan->mega.chunk = (sample_count + n_chunks - 1) / n_chunks;
  plan->mega.n_chunks = (sample_count + plan->mega.chunk - 1) / plan->mega.chunk;
  plan->mega.partial_bytes =
      plan->mega.n_chunks > 1 ? (size_t)plan->mega.n_chunks * (size_t)pixels * 3 * sizeof(float) : 0;
  const size_t s = (size_t)slots;
  plan->pool_bytes = 7 * align256(s * 16) + align256(s * 4) + align256(s * 4 * WAVE_CLASSES) + align256(64) +
                     align256(WAVE_BATCH * 4);
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
  pool.queue_count = (uint32_t *)take(64);
  pool.alive_hist = (uint32_t *)take(WAVE_BATCH * 4);

  cudaError_t err;
  /* every slot starts as "needs a sample, holds no item" */
  if ((err = cudaMemsetAsync(pool.status, 0, s * 4, stream)) != cudaSuccess ||
      (err = cudaMemsetAsync(pool.item, 0, s * 16, stream)) != cudaSuccess ||
      (err = cudaMemsetAsync(pool.rng, 0, s * 16, stream)) != cudaSuccess ||
      (err = cudaMemsetAsync(pool.sum, 0, s * 16, stream)) != cudaSuccess)
    return err;

  const unsigned slot_blocks = (unsigned)(s / WAVE_BLOCK);
  int n_launch = 0;
  /* upper bound on iterations: every item runs its samples' segments one per iteration */
  const long long max_iters = 64ll + (long long)plan.mega.n_chunks * 0 +
                              ((long long)job.cam.img_width * job.cam.img_height * plan.mega.n_chunks / (long long)s + 2) *
                                  (long long)plan.mega.chunk * ((long long)job.cam.max_depth + 1);
  for (long long done_iters = 0; done_iters < max_iters;) {
    if ((err = cudaMemsetAsync(pool.alive_hist, 0, WAVE_BATCH * 4, stream)) != cudaSuccess)
      return err;
    for (uint32_t k = 0; k < WAVE_BATCH; k++) {
      wave_generate_kernel<<<slot_blocks, WAVE_BLOCK, 0, stream>>>(job, plan.mega, partial, pool, k);
      wave_extend_kernel<<<slot_blocks, WAVE_BLOCK, 0, stream>>>(sc, job, pool);
      for (int c = 0; c < WAVE_CLASSES; c++)
        wave_shade_kernel<<<plan.shade_blocks, WAVE_BLOCK, 0, stream>>>(sc, job, pool, c);
      n_launch += 2 + WAVE_CLASSES;
    }
    if ((err = cudaGetLastError()) != cudaSuccess)
      return err;
    done_iters += WAVE_BATCH;
    uint32_t alive_last = 0;
    if ((err = cudaMemcpyAsync(&alive_last, pool.alive_hist + (WAVE_BATCH - 1), 4, cudaMemcpyDeviceToHost, stream)) !=
            cudaSuccess ||
        (err = cudaStreamSynchronize(stream)) != cudaSuccess)
      return err;
    if (alive_last == 0)
      break;
  }
  if (plan.mega.n_chunks > 1) {
    size_t n_values = (size_t)job.cam.img_width * job.cam.img_height * 3;
    wave_sum_chunks_kernel<<<(unsigned)((n_values + 255) / 256), 256, 0, stream>>>(partial, job.radiance, n_values,
                                                                                  plan.mega.n_chunks);
    n_launch++;
  }
  *launches = n_launch;
  return cudaGetLastError();
}
