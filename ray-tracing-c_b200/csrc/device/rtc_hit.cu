/*
 * rtc_hit.cu -- rt_hit_batch over the REBUILT structure (RT_KERNEL_MEGA / RT_KERNEL_WAVE): the HittableVTable.hit slot
 * (/root/reference/include/hittable.h:30) on ray batches, the harness of the hit-parity tests.
 *
 * Its own translation unit because it is compiled -fmad=false (csrc/Makefile), like rtc_exact.cu: a medium's free
 * path is neg_inv_density * logf(u) (hittable.c:412), and the portable logf (rtc_portable_math.h) only gives the
 * oracle's bits when the compiler contracts nothing -- with contraction one medium hit in 300 came out an ulp off.
 * The traversal itself is the render kernels' (rtc_trace.cuh); its box tests merely lose their fused multiply-adds
 * here, which no result depends on.
 */
#include "rtc_kernels.h"
#include "rtc_trace.cuh"

#define MEGA_BLOCK 128


__global__ void __launch_bounds__(MEGA_BLOCK) mega_hit_kernel(const DScene sc, const HitBatch b) {
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
  TraceHit th = trace_world<true>(sc, ray, b.t_min, b.t_max, g);
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
  mega_hit_kernel<<<blocks, MEGA_BLOCK, 0, stream>>>(sc, b);
  return cudaGetLastError();
}
