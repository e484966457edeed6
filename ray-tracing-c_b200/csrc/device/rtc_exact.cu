/*
 * rtc_exact.cu -- RT_KERNEL_EXACT: the parity anchor.  COMPILED WITH -fmad=false.
 *
 * One thread per pixel walks the reference-order graph (rtc_graph.cuh) with the portable transcendentals, draws
 * from the pixel's PCG32 stream at exactly the reference's program points, and un-winds the recursion of
 * Camera_ray_color (/root/reference/src/raytracing.c:39-84) in the reference's order:
 *     colour(depth) = emitted + albedo * colour(depth-1) [* scatter_pdf / sampling_pdf]
 * is evaluated from the LAST segment back to the first using a per-thread stack of (albedo, ratio) kept in a
 * global scratch array, because float products do not associate.  The result equals the oracle's PORTABLE mode
 * bit for bit (tests/test_exact_parity.py), which in turn equals the compiled reference up to the libm swap.
 * It is not the throughput path.
 */
#include "rtc_graph.cuh"
#include "rtc_kernels.h"

#define EXACT_BLOCK 128
#define EXACT_MAX_DEPTH 128 /* two 64-bit masks of "this level used the mixture form" */

RTC_DEV bool has_nan(V3 a) { return a.x != a.x || a.y != a.y || a.z != a.z; }

struct ExactCounts {
  unsigned long long segments, pdf_evals, nan_samples;
};

/* raytracing.c:39-84 */
__device__ V3 exact_ray_color(const DScene &sc, const rt_camera &cam, V3 o, V3 d, Rng &g, float4 *stack,
                              size_t stride, ExactCounts &cnt) {
  unsigned long long mixture_mask[2] = {0ull, 0ull};
  int level = 0;
  V3 result;
  const float prob = cam.lights_sampling_prob;
  for (;;) {
    if (cam.max_depth - level <= 0) {
      result = v3(0, 0, 0);
      break;
    }
    cnt.segments++;
    Hit rec;
    rec.p = v3(0, 0, 0), rec.n = v3(0, 0, 0), rec.material = 0, rec.t = 0, rec.u = 0, rec.v = 0, rec.front = 0,
    rec.node = 0;
    if (!graph_hit<true>(sc, sc.objects_root, o, d, 1e-3f, RTC_INF, rec, g)) {
      result = v_load(cam.background);
      break;
    }
    V3 r_out, albedo;
    bool skip_pdf;
    V3 emitted = material_emit<true>(sc, rec);
    if (!material_scatter<true>(sc, rec, d, r_out, albedo, skip_pdf, g)) {
      result = emitted;
      break;
    }
    /* every scattering material emits (0,0,0); the reference still adds it, which only matters for -0 */
    float ratio = 1.0f;
    if (!(skip_pdf || prob == 0.0f || sc.lights_root == RT_NONE)) {
      if (rng_f32(g) < prob)
        r_out = lights_rand<true>(sc, sc.lights_root, rec.p, g);
      float scatter_pdf = material_scatter_pdf(sc, rec.material, rec.n, r_out);
      cnt.pdf_evals++;
      float sampling_pdf = (1.0f - prob) * scatter_pdf + prob * lights_pdf<true>(sc, sc.lights_root, rec.p, r_out);
      ratio = __fdiv_rn(scatter_pdf, sampling_pdf);
      mixture_mask[level >> 6] |= 1ull << (level & 63);
    }
    stack[(size_t)level * stride] = make_float4(albedo.x, albedo.y, albedo.z, ratio);
    level++;
    o = rec.p;
    d = r_out;
  }
  while (level > 0) {
    level--;
    float4 s = stack[(size_t)level * stride];
    V3 through = v_mul(v3(s.x, s.y, s.z), result);
    if ((mixture_mask[level >> 6] >> (level & 63)) & 1ull)
      through = v_scale(through, s.w);
    result = v_add(v3(0, 0, 0), through);
  }
  return result;
}

/* raytracing.c:86-133 without the tonemap (rt_tonemap_device is the epilogue) */
__global__ void __launch_bounds__(EXACT_BLOCK)
    exact_render_kernel(const DScene sc, const RenderJob job, float4 *scratch, int n_threads) {
  const rt_camera &cam = job.cam;
  const int w = cam.img_width, h = cam.img_height;
  const int tid = blockIdx.x * blockDim.x + threadIdx.x;
  ExactCounts cnt = {0ull, 0ull, 0ull};
  unsigned long long samples = 0;
  const V3 px00 = v_load(cam.pixel00), du = v_load(cam.pixel_delta_u), dv = v_load(cam.pixel_delta_v);
  const V3 eye = v_load(cam.origin), lens_u = v_load(cam.dof_disc_u), lens_v = v_load(cam.dof_disc_v);

  for (int pix = tid; pix < w * h; pix += n_threads) {
    const int j = pix / w, i = pix - j * w;
    Rng rng;
    rng_seed_pixel(rng, i, j);
    V3 pixel = v_add(v_add(px00, v_scale(du, (float)i)), v_scale(dv, (float)j));
    V3 sum = v3(0, 0, 0);
    for (int s = job.sample_begin; s < job.sample_begin + job.sample_count; s++) {
      if (job.rng_mode == RT_RNG_SAMPLE)
        rng_seed_sample(rng, i, j, s);
      float jx = rng_between(rng, -0.5f, 0.5f);
      float jy = rng_between(rng, -0.5f, 0.5f);
      V3 o = eye;
      if (cam.dof_angle > 0.0f) { /* raytracing.c:108-117 */
        float a, b;
        for (;;) {
          a = rng_between(rng, -1.0f, 1.0f);
          b = rng_between(rng, -1.0f, 1.0f);
          if (a * a + b * b < 1.0f)
            break;
        }
        o = v_add(v_add(eye, v_scale(lens_u, a)), v_scale(lens_v, b));
      }
      V3 d = v_add(v_add(v_add(pixel, v_scale(du, jx)), v_scale(dv, jy)), v_neg(o));
      V3 L = exact_ray_color(sc, cam, o, d, rng, scratch + tid, (size_t)n_threads, cnt);
      samples++;
      if (has_nan(L)) {
        cnt.nan_samples++;
        if (job.flags & RT_FLAG_DROP_NAN)
          continue;
      }
      sum = v_add(sum, L);
    }
    job.radiance[(size_t)pix * 3 + 0] = sum.x;
    job.radiance[(size_t)pix * 3 + 1] = sum.y;
    job.radiance[(size_t)pix * 3 + 2] = sum.z;
  }

  /* counters: warp reduce, one atomic per warp */
  unsigned long long v[4] = {cnt.segments, cnt.pdf_evals, samples, cnt.nan_samples};
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

size_t exact_scratch_bytes(const rt_camera &cam, int *n_threads) {
  long long pixels = (long long)cam.img_width * cam.img_height;
  long long cap = 148ll * 8 * EXACT_BLOCK; /* enough warps to hide latency; bounds the scratch */
  long long t = pixels < cap ? pixels : cap;
  t = (t + EXACT_BLOCK - 1) / EXACT_BLOCK * EXACT_BLOCK;
  if (t < EXACT_BLOCK)
    t = EXACT_BLOCK;
  *n_threads = (int)t;
  int depth = cam.max_depth > 0 ? cam.max_depth : 1;
  return (size_t)t * (size_t)depth * sizeof(float4);
}

cudaError_t exact_render(const DScene &sc, const RenderJob &job, void *scratch, int n_threads, cudaStream_t stream) {
  if (job.cam.max_depth > EXACT_MAX_DEPTH)
    return cudaErrorInvalidValue;
  exact_render_kernel<<<n_threads / EXACT_BLOCK, EXACT_BLOCK, 0, stream>>>(sc, job, (float4 *)scratch, n_threads);
  return cudaGetLastError();
}

/* ------------------------------------------------------------------ rt_hit_batch, exact ------------------ */

__global__ void __launch_bounds__(EXACT_BLOCK) exact_hit_kernel(const DScene sc, const HitBatch b) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= b.n)
    return;
  V3 o = v_load(b.origins + 3 * i), d = v_load(b.dirs + 3 * i);
  Rng g;
  g.state = 0, g.inc = 1;
  if (b.rng_io) {
    g.state = b.rng_io[2 * i];
    g.inc = b.rng_io[2 * i + 1];
  }
  Hit rec;
  rec.p = v3(0, 0, 0), rec.n = v3(0, 0, 0), rec.material = RT_NONE, rec.t = 0, rec.u = 0, rec.v = 0, rec.front = 0,
  rec.node = RT_NONE;
  bool hit = graph_hit<true>(sc, sc.objects_root, o, d, b.t_min, b.t_max, rec, g);
  if (b.rng_io)
    b.rng_io[2 * i] = g.state;
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

cudaError_t exact_hit_batch(const DScene &sc, const HitBatch &b, cudaStream_t stream) {
  if (b.n == 0)
    return cudaSuccess;
  unsigned blocks = (unsigned)((b.n + EXACT_BLOCK - 1) / EXACT_BLOCK);
  exact_hit_kernel<<<blocks, EXACT_BLOCK, 0, stream>>>(sc, b);
  return cudaGetLastError();
}
