/*
 * rtc_api.cu -- the C ABI of librtc_b200.so (include/rtc_b200.h): scene upload, render entry points, tonemap,
 * hit batches, single-process multi-GPU render.  No torch, no C++ types across the boundary, no CPU fallback:
 * every compute entry point fails with RT_ERR_NO_DEVICE when there is no CUDA device.
 */
#include "rtc_bvh.h"
#include "rtc_kernels.h"
#include "rtc_graph.cuh"
#include "rtc_trace.cuh"
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <dlfcn.h>
#include <mutex>
#include <nccl.h> /* types and prototypes only: the library itself is dlopen'ed (rt_render_multi) */
#include <new>
#include <vector>

/* ------------------------------------------------------------------ errors ------------------------------- */

static thread_local char g_error[512] = "";

static int fail(int status, const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
  return status;
}
#define CUDA_TRY(expr)                                                                                               \
  do {                                                                                                               \
    cudaError_t err__ = (expr);                                                                                      \
    if (err__ != cudaSuccess)                                                                                        \
      return fail(err__ == cudaErrorMemoryAllocation ? RT_ERR_NOMEM : RT_ERR_CUDA, "%s: %s", #expr,                  \
                  cudaGetErrorString(err__));                                                                        \
  } while (0)

extern "C" const char *rt_last_error(void) { return g_error; }
extern "C" uint32_t rt_abi_version(void) { return RT_ABI_VERSION; }
extern "C" int rt_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

/* ------------------------------------------------------------------ scene -------------------------------- */

struct rt_scene {
  int device = 0;
  std::vector<void *> allocations;
  DScene view;
  BuiltScene built_info; /* limits only; the arrays are cleared after upload */
  bool mega_ok = true;
  std::string mega_why;
  /* per-scene scratch reused across renders */
  RenderCounters *counters = nullptr;
  void *scratch = nullptr;
  size_t scratch_bytes = 0;
  float *radiance = nullptr;
  size_t radiance_bytes = 0;
  uint8_t *rgb8 = nullptr;
  size_t rgb8_bytes = 0;
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
  rt_stats last = {};
  int last_launches = 0;
  /* scene cache (rt_scene_acquire / rt_scene_release) */
  bool cached = false, in_use = false;
  uint64_t key[2] = {0, 0};
  uint64_t last_use = 0;
};

namespace {

inline float __uint_as_float_host(uint32_t u) {
  float f;
  memcpy(&f, &u, 4);
  return f;
}

struct DeviceGuard {
  int previous = -1;
  explicit DeviceGuard(int device) {
    cudaGetDevice(&previous);
    if (device != previous)
      cudaSetDevice(device);
  }
  ~DeviceGuard() {
    if (previous >= 0)
      cudaSetDevice(previous);
  }
};

/* Device memory comes from the device's stream-ordered pool with its release threshold raised, so that the
 * create / render / destroy cycle of Camera_render (one scene per call, like the reference) reuses the same pages:
 * with plain cudaMalloc / cudaFree the driver hands the ~100 MB of per-chunk partial sums back to the OS at every
 * destroy, and one call in three then stalled for 30-400 ms in the next allocation (measured, Book 1 final). */
static cudaError_t dev_alloc(void **ptr, size_t bytes) {
  static bool pool_ready[64] = {false};
  int dev = 0;
  cudaError_t err = cudaGetDevice(&dev);
  if (err != cudaSuccess)
    return err;
  if (dev >= 0 && dev < 64 && !pool_ready[dev]) {
    cudaMemPool_t pool;
    err = cudaDeviceGetDefaultMemPool(&pool, dev);
    if (err != cudaSuccess)
      return err;
    unsigned long long keep = ~0ull;
    err = cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    if (err != cudaSuccess)
      return err;
    pool_ready[dev] = true;
  }
  err = cudaMallocAsync(ptr, bytes, (cudaStream_t)0);
  if (err != cudaSuccess)
    return err;
  return cudaStreamSynchronize((cudaStream_t)0); /* callers use the memory from other streams right away */
}
static void dev_free(void *ptr) {
  if (ptr)
    cudaFreeAsync(ptr, (cudaStream_t)0); /* the legacy default stream orders this after all earlier work */
}

template <typename T> int upload(rt_scene *s, const T *host, size_t count, const T **out) {
  size_t bytes = (count ? count : 1) * sizeof(T);
  void *dev = nullptr;
  CUDA_TRY(dev_alloc(&dev, bytes));
  s->allocations.push_back(dev);
  if (count)
    CUDA_TRY(cudaMemcpy(dev, host, count * sizeof(T), cudaMemcpyHostToDevice));
  else
    CUDA_TRY(cudaMemset(dev, 0, bytes));
  *out = static_cast<const T *>(dev);
  return RT_OK;
}

int ensure(void **ptr, size_t *have, size_t need) {
  if (*have >= need && *ptr)
    return RT_OK;
  if (*ptr) {
    cudaDeviceSynchronize(); /* the old buffer may still be in use on a caller's stream */
    dev_free(*ptr);
  }
  *ptr = nullptr;
  *have = 0;
  CUDA_TRY(dev_alloc(ptr, need ? need : 16));
  *have = need;
  return RT_OK;
}

} // namespace

static void scene_free(rt_scene *scene) {
  DeviceGuard guard(scene->device);
  cudaDeviceSynchronize(); /* renders enqueued on caller streams (rt_render_device) may still be running */
  for (void *p : scene->allocations)
    dev_free(p);
  dev_free(scene->counters);
  dev_free(scene->scratch);
  dev_free(scene->radiance);
  dev_free(scene->rgb8);
  for (cudaEvent_t e : scene->ev)
    if (e)
      cudaEventDestroy(e);
  delete scene;
}

/* ------------------------------------------------------------------ scene cache -------------------------- */

/* Camera_render builds, uploads and destroys one scene per call, like the reference (which holds nothing between
 * calls).  For a frame that takes 30 ms on eight GPUs the flatten + SAH build + upload is 1 ms of it, so uploaded
 * scenes are kept, keyed by a 128-bit hash of the CONTENT of the description (no pointer of the caller is retained):
 * a caller that renders the same World again gets the resident copy. */
namespace {
std::mutex g_cache_mutex;
std::vector<rt_scene *> g_cache;
uint64_t g_cache_clock = 0;
constexpr size_t kCachePerDevice = 4;

inline uint64_t mix64(uint64_t h, uint64_t v, uint64_t mul) {
  h ^= v;
  h *= mul;
  return h ^ (h >> 29);
}
void hash_bytes(const void *data, size_t bytes, uint64_t key[2]) {
  if (data == nullptr)
    bytes = 0; /* an empty array of the description (its pointer may be NULL): only the length goes in */
  static const uint8_t nothing[32] = {0};
  if (bytes == 0)
    data = nothing;
  /* four independent lanes over 32-byte blocks (a single multiply chain would be latency-bound: the image texture
   * of the Book 2 scene alone is 1.5 MB), folded into two 64-bit words */
  const uint8_t *p = static_cast<const uint8_t *>(data);
  uint64_t a = key[0] ^ 0x9E3779B97F4A7C15ull, b = key[1] ^ 0xC2B2AE3D27D4EB4Full, c = bytes, d = ~bytes;
  size_t n = bytes / 32;
  for (size_t i = 0; i < n; i++, p += 32) {
    uint64_t w[4];
    memcpy(w, p, 32);
    a = mix64(a, w[0], 0xFF51AFD7ED558CCDull), b = mix64(b, w[1], 0xC4CEB9FE1A85EC53ull);
    c = mix64(c, w[2], 0x9FB21C651E98DF25ull), d = mix64(d, w[3], 0xD6E8FEB86659FD93ull);
  }
  uint64_t tail[4] = {0, 0, 0, 0};
  memcpy(tail, p, bytes % 32);
  a = mix64(a, tail[0], 0xFF51AFD7ED558CCDull), b = mix64(b, tail[1], 0xC4CEB9FE1A85EC53ull);
  c = mix64(c, tail[2], 0x9FB21C651E98DF25ull), d = mix64(d, tail[3], 0xD6E8FEB86659FD93ull);
  key[0] = mix64(mix64(a, c, 0x94D049BB133111EBull), b, 0xBF58476D1CE4E5B9ull);
  key[1] = mix64(mix64(b, d, 0xBF58476D1CE4E5B9ull), a, 0x94D049BB133111EBull);
}
void hash_desc(const rt_scene_desc &d, uint64_t key[2]) {
  key[0] = key[1] = 0;
  const uint32_t head[12] = {d.abi_version, d.n_nodes,  d.n_children, d.n_spheres,    d.n_quads,    d.n_materials,
                             d.n_textures,  d.n_perlins, d.n_images,   d.objects_root, d.lights_root, 0};
  hash_bytes(head, sizeof(head), key);
  hash_bytes(&d.n_texel_bytes, sizeof(d.n_texel_bytes), key);
  hash_bytes(d.nodes, sizeof(rt_node) * d.n_nodes, key);
  hash_bytes(d.children, 4ull * d.n_children, key);
  hash_bytes(d.spheres, sizeof(rt_sphere) * d.n_spheres, key);
  hash_bytes(d.quads, sizeof(rt_quad) * d.n_quads, key);
  hash_bytes(d.materials, sizeof(rt_material) * d.n_materials, key);
  hash_bytes(d.textures, sizeof(rt_texture) * d.n_textures, key);
  hash_bytes(d.perlins, sizeof(rt_perlin) * d.n_perlins, key);
  hash_bytes(d.images, sizeof(rt_image) * d.n_images, key);
  hash_bytes(d.texels, d.n_texel_bytes, key);
  /* what rt_scene_create builds also depends on these build-time knobs: a scene built under one setting must not be
   * handed out under another */
  static const char *const knobs[] = {"RT_BVH_BOXES", "RT_BVH_WIDE", "RT_BVH_PRE", "RT_BVH_COST_SPHERE",
                                      "RT_BVH_COST_QUAD", "RT_BVH_COST_BOX", "RT_ONE_LIGHT_OFF", "RT_CONST_CHAINS_OFF"};
  for (const char *name : knobs) {
    const char *v = getenv(name);
    hash_bytes(name, strlen(name), key);
    hash_bytes(v ? v : "", v ? strlen(v) : 0, key);
  }
}
} // namespace

extern "C" void rt_scene_destroy(rt_scene *scene) {
  if (!scene)
    return;
  if (scene->cached) { /* destroying a cached scene takes it out of the cache */
    std::lock_guard<std::mutex> lock(g_cache_mutex);
    for (size_t i = 0; i < g_cache.size(); i++)
      if (g_cache[i] == scene) {
        g_cache.erase(g_cache.begin() + (long)i);
        break;
      }
  }
  scene_free(scene);
}

extern "C" int rt_scene_acquire(const rt_scene_desc *desc, int device, rt_scene **out) {
  if (!desc || !out)
    return fail(RT_ERR_INVALID, "rt_scene_acquire: NULL argument");
  *out = nullptr;
  std::string why = rtc_validate_desc(*desc);
  if (!why.empty())
    return fail(RT_ERR_INVALID, "rt_scene_acquire: %s", why.c_str());
  if (rt_device_count() <= 0)
    return fail(RT_ERR_NO_DEVICE, "rt_scene_acquire: no CUDA device (this library does not compute on the host)");
  if (device < 0)
    CUDA_TRY(cudaGetDevice(&device));
  uint64_t key[2];
  hash_desc(*desc, key);
  {
    std::lock_guard<std::mutex> lock(g_cache_mutex);
    for (rt_scene *s : g_cache)
      if (!s->in_use && s->device == device && s->key[0] == key[0] && s->key[1] == key[1]) {
        s->in_use = true;
        s->last_use = ++g_cache_clock;
        *out = s;
        return RT_OK;
      }
  }
  rt_scene *s = nullptr;
  int rc = rt_scene_create(desc, device, &s);
  if (rc != RT_OK)
    return rc;
  s->key[0] = key[0], s->key[1] = key[1];
  s->cached = true, s->in_use = true;
  std::vector<rt_scene *> evict;
  {
    std::lock_guard<std::mutex> lock(g_cache_mutex);
    s->last_use = ++g_cache_clock;
    g_cache.push_back(s);
    for (;;) { /* least recently used idle scenes of this device beyond the capacity */
      size_t count = 0, oldest = g_cache.size();
      for (size_t i = 0; i < g_cache.size(); i++)
        if (g_cache[i]->device == device) {
          count++;
          if (!g_cache[i]->in_use && (oldest == g_cache.size() || g_cache[i]->last_use < g_cache[oldest]->last_use))
            oldest = i;
        }
      if (count <= kCachePerDevice || oldest == g_cache.size())
        break;
      evict.push_back(g_cache[oldest]);
      g_cache.erase(g_cache.begin() + (long)oldest);
    }
  }
  for (rt_scene *e : evict)
    scene_free(e);
  *out = s;
  return RT_OK;
}

extern "C" void rt_scene_release(rt_scene *scene) {
  if (!scene)
    return;
  if (!scene->cached) {
    scene_free(scene);
    return;
  }
  std::lock_guard<std::mutex> lock(g_cache_mutex);
  scene->in_use = false;
}

extern "C" int rt_scene_create(const rt_scene_desc *desc, int device, rt_scene **out) {
  if (!desc || !out)
    return fail(RT_ERR_INVALID, "rt_scene_create: NULL argument");
  *out = nullptr;
  std::string why = rtc_validate_desc(*desc);
  if (!why.empty())
    return fail(RT_ERR_INVALID, "rt_scene_create: %s", why.c_str());
  if (rt_device_count() <= 0)
    return fail(RT_ERR_NO_DEVICE, "rt_scene_create: no CUDA device (this library does not compute on the host)");
  if (device < 0)
    CUDA_TRY(cudaGetDevice(&device));

  BuiltScene built;
  rtc_build_scene(*desc, built);
  if (built.graph_frames > GRAPH_MAX_FRAMES || built.graph_rays > GRAPH_MAX_RAYS || built.graph_recs > GRAPH_MAX_RECS)
    return fail(RT_ERR_UNSUPPORTED, "rt_scene_create: graph nesting (calls %d, instance levels %d, medium levels %d) "
                                    "exceeds the limits %d/%d/%d",
                built.graph_frames, built.graph_rays, built.graph_recs, GRAPH_MAX_FRAMES, GRAPH_MAX_RAYS,
                GRAPH_MAX_RECS);
  if (built.light_nesting < 0)
    return fail(RT_ERR_INVALID, "rt_scene_create: a list under world.lights has no item that can be sampled");
  if (built.light_nesting > RTC_LIGHT_NESTING)
    return fail(RT_ERR_UNSUPPORTED, "rt_scene_create: lists nested %d deep under world.lights (limit %d)",
                built.light_nesting, RTC_LIGHT_NESTING);

  rt_scene *s = new (std::nothrow) rt_scene();
  if (!s)
    return fail(RT_ERR_NOMEM, "rt_scene_create: out of host memory");
  s->device = device;
  DeviceGuard guard(device);
  s->mega_ok = built.mega_supported;
  s->mega_why = built.mega_why;
  if (s->mega_ok && built.max_stack + 2 > TRACE_STACK) {
    s->mega_ok = false;
    s->mega_why = "rebuilt BVH deeper than the traversal stack";
  }

  DScene &v = s->view;
  memset(&v, 0, sizeof(v));
  int rc = RT_OK;
#define UP(field, host, count)                                                                                       \
  if (rc == RT_OK)                                                                                                   \
  rc = upload(s, host, (size_t)(count), &v.field)
  UP(nodes, desc->nodes, desc->n_nodes);
  UP(children, desc->children, desc->n_children);
  UP(spheres, desc->spheres, desc->n_spheres);
  UP(quads, desc->quads, desc->n_quads);
  UP(materials, desc->materials, desc->n_materials);
  UP(textures, desc->textures, desc->n_textures);
  UP(perlins, desc->perlins, desc->n_perlins);
  UP(images, desc->images, desc->n_images);
  UP(texels, desc->texels, desc->n_texel_bytes);
  UP(bnodes, built.bnodes.data(), built.bnodes.size());
  UP(wnodes, built.wnodes.data(), built.wnodes.size());
  UP(msph, built.msph.data(), built.msph.size());
  UP(msph_meta, built.msph_meta.data(), built.msph_meta.size());
  UP(mquads, built.mquads.data(), built.mquads.size());
  UP(insts, built.insts.data(), built.insts.size());
  UP(ops, built.ops.data(), built.ops.size());
  UP(media, built.media.data(), built.media.size());
#undef UP
  v.objects_root = desc->objects_root;
  v.lights_root = desc->lights_root;
  v.world_root = built.world_root;
  v.n_pre = built.n_pre, v.pre_ref[0] = built.pre_ref[0], v.pre_ref[1] = built.pre_ref[1];
  v.wide_root = (built.wide_root != REF_EMPTY && built.wide_max_stack + 2 <= TRACE_STACK) ? built.wide_root : REF_EMPTY;
  v.n_media = (uint32_t)built.media.size();
  v.n_perlins = desc->n_perlins;
  v.feats = (built.n_mquads ? FEAT_QUADS : 0) | (built.n_insts ? FEAT_INSTANCES : 0) |
            (desc->lights_root != RT_NONE ? FEAT_LIGHTS : 0) | (built.media.empty() ? 0 : FEAT_MEDIA);
  if (built.n_pre)
    v.feats |= FEAT_PRE;
  for (uint32_t k = 0; k < desc->n_textures; k++)
    if (desc->textures[k].kind != RT_TEX_SOLID)
      v.feats |= FEAT_TEXTURES;
  for (uint32_t k = 0; k < desc->n_materials; k++)
    if (desc->materials[k].tag == RT_MAT_ISOTROPIC)
      v.feats |= FEAT_MEDIA;
  if (getenv("RT_CONST_CHAINS_OFF") == nullptr) {
    for (uint32_t k = 0; k < 2 && k < built.n_insts; k++) {
      const Instance &in = built.insts[k];
      if (in.op_count < 1 || in.op_count > 2)
        break; /* (instance k + 1 needs instance k to be there too: the leaf test compares tag <= n_const_chains) */
      v.const_count[k] = in.op_count;
      for (uint32_t j = 0; j < in.op_count; j++) {
        const ChainOp &op = built.ops[in.op_begin + j];
        v.const_op[k][j] = make_float4(op.f0, op.f1, op.f2, __uint_as_float_host(op.is_rotate));
      }
      v.n_const_chains = k + 1;
    }
  }
  if (desc->lights_root != RT_NONE && getenv("RT_ONE_LIGHT_OFF") == nullptr) {
    const rt_node &l = desc->nodes[desc->lights_root];
    if (l.kind == RT_NODE_LIST && l.b == 1 && desc->nodes[desc->children[l.a]].kind == RT_NODE_QUAD) {
      v.one_quad_light = 1;
      v.light = desc->quads[desc->nodes[desc->children[l.a]].a];
    }
  }
  if (rc == RT_OK) {
    cudaError_t err = dev_alloc((void **)&s->counters, sizeof(RenderCounters));
    for (int k = 0; k < 4 && err == cudaSuccess; k++)
      err = cudaEventCreate(&s->ev[k]);
    if (err != cudaSuccess)
      rc = fail(RT_ERR_CUDA, "rt_scene_create: %s", cudaGetErrorString(err));
  }
  if (rc != RT_OK) {
    rt_scene_destroy(s);
    return rc;
  }
  built.bnodes.clear(), built.wnodes.clear(), built.msph.clear(), built.msph_meta.clear(), built.mquads.clear();
  s->built_info = built;
  *out = s;
  return RT_OK;
}

extern "C" int rt_scene_inspect(const rt_scene_desc *desc, rt_scene_info *info) {
  if (!desc || !info)
    return fail(RT_ERR_INVALID, "rt_scene_inspect: NULL argument");
  std::string why = rtc_validate_desc(*desc);
  if (!why.empty())
    return fail(RT_ERR_INVALID, "rt_scene_inspect: %s", why.c_str());
  BuiltScene b;
  rtc_build_scene(*desc, b);
  memset(info, 0, sizeof(*info));
  info->bvh_nodes = b.n_bnodes;
  info->bvh_depth = (uint32_t)b.max_stack;
  info->spheres = b.n_msph, info->quads = b.n_mquads;
  info->instances = b.n_insts, info->media = (uint32_t)b.media.size();
  info->chain_ops = b.n_ops;
  info->graph_frames = (uint32_t)b.graph_frames, info->graph_rays = (uint32_t)b.graph_rays;
  info->graph_recs = (uint32_t)b.graph_recs;
  info->light_nesting = b.light_nesting;
  info->mega_supported = b.mega_supported && b.max_stack + 2 <= TRACE_STACK;
  info->pre_prims = b.n_pre;
  info->wide_nodes = b.wide_root != REF_EMPTY && REF_KIND(b.wide_root) == REF_INNER ? (uint32_t)b.wnodes.size() : 0u;
  info->device_bytes = sizeof(rt_node) * desc->n_nodes + 4ull * desc->n_children + sizeof(rt_sphere) * desc->n_spheres +
                       sizeof(rt_quad) * desc->n_quads + sizeof(rt_material) * desc->n_materials +
                       sizeof(rt_texture) * desc->n_textures + sizeof(rt_perlin) * desc->n_perlins +
                       sizeof(rt_image) * desc->n_images + desc->n_texel_bytes + sizeof(BNode) * b.bnodes.size() +
                       (b.wide_root != REF_EMPTY ? sizeof(WNode) * b.wnodes.size() : 0) + (sizeof(MSphere) + sizeof(MPrimMeta)) * b.msph.size() + sizeof(MQuad) * b.mquads.size() +
                       sizeof(Instance) * b.insts.size() + sizeof(ChainOp) * b.ops.size() +
                       sizeof(MMedium) * b.media.size();
  return RT_OK;
}

/* ------------------------------------------------------------------ tonemap ------------------------------ */

/* raytracing.c:127-131 with utils.h's clamp: NaN and negatives -> 0 */
__global__ void tonemap_kernel(const float *__restrict__ sum, size_t n, float spp, uint8_t *__restrict__ rgb8) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n)
    return;
  float v = __fsqrt_rn(__fdiv_rn(sum[i], spp));
  v = v > 0.0f ? v : 0.0f;
  v = v < 0.999f ? v : 0.999f;
  rgb8[i] = (uint8_t)(int)(256.0f * v);
}

extern "C" int rt_tonemap_device(const float *radiance_sum_dev, size_t n_values, int spp, uint8_t *rgb8_dev,
                                 void *cuda_stream) {
  if (!radiance_sum_dev || !rgb8_dev || spp <= 0)
    return fail(RT_ERR_INVALID, "rt_tonemap_device: bad argument");
  if (n_values == 0)
    return RT_OK;
  unsigned blocks = (unsigned)((n_values + 255) / 256);
  tonemap_kernel<<<blocks, 256, 0, (cudaStream_t)cuda_stream>>>(radiance_sum_dev, n_values, (float)spp, rgb8_dev);
  CUDA_TRY(cudaGetLastError());
  return RT_OK;
}

/* the same through HOST buffers (upload, kernel, download): for callers that accumulated a float sum themselves,
 * e.g. the progressive / checkpointed passes of rtc_main -- the epilogue still runs on the device */
extern "C" int rt_tonemap(const float *radiance_sum, size_t n_values, int spp, uint8_t *rgb8) {
  if (!radiance_sum || !rgb8 || spp <= 0)
    return fail(RT_ERR_INVALID, "rt_tonemap: bad argument");
  if (rt_device_count() <= 0)
    return fail(RT_ERR_NO_DEVICE, "rt_tonemap: no CUDA device (this library does not compute on the host)");
  if (n_values == 0)
    return RT_OK;
  float *sum_dev = nullptr;
  uint8_t *rgb_dev = nullptr;
  CUDA_TRY(dev_alloc((void **)&sum_dev, n_values * sizeof(float)));
  cudaError_t err = dev_alloc((void **)&rgb_dev, n_values);
  if (err == cudaSuccess)
    err = cudaMemcpy(sum_dev, radiance_sum, n_values * sizeof(float), cudaMemcpyHostToDevice);
  int rc = RT_OK;
  if (err == cudaSuccess)
    rc = rt_tonemap_device(sum_dev, n_values, spp, rgb_dev, nullptr);
  if (err == cudaSuccess && rc == RT_OK)
    err = cudaMemcpy(rgb8, rgb_dev, n_values, cudaMemcpyDeviceToHost);
  dev_free(sum_dev), dev_free(rgb_dev);
  if (err != cudaSuccess)
    return fail(RT_ERR_CUDA, "rt_tonemap: %s", cudaGetErrorString(err));
  return rc;
}

/* ------------------------------------------------------------------ render ------------------------------- */

static int resolve_range(const rt_camera *cam, const rt_opts *opts, int *begin, int *count) {
  if (cam->img_width <= 0 || cam->img_height <= 0 || cam->samples_per_pixel <= 0)
    return fail(RT_ERR_INVALID, "render: empty image or no samples");
  if ((long long)cam->img_width * cam->img_height > (1ll << 26))
    return fail(RT_ERR_INVALID, "render: image larger than 2^26 pixels");
  *begin = opts->sample_begin;
  *count = opts->sample_count > 0 ? opts->sample_count : cam->samples_per_pixel - opts->sample_begin;
  if (*begin < 0 || *count <= 0)
    return fail(RT_ERR_INVALID, "render: empty sample range");
  if (opts->rng_mode > RT_RNG_SAMPLE)
    return fail(RT_ERR_INVALID, "render: unknown rng_mode");
  if (opts->rng_mode == RT_RNG_PIXEL && *begin != 0)
    return fail(RT_ERR_INVALID, "render: RT_RNG_PIXEL streams cannot start at sample %d", *begin);
  return RT_OK;
}

/* enqueue one render into `radiance_dev` on `stream`; records ev[1]..ev[2] around the kernels */
static int enqueue_render(rt_scene *s, const rt_camera *cam, const rt_opts *opts, float *radiance_dev,
                          cudaStream_t stream) {
  int begin = 0, count = 0;
  int rc = resolve_range(cam, opts, &begin, &count);
  if (rc != RT_OK)
    return rc;
  RenderJob job;
  memset(&job, 0, sizeof(job));
  job.cam = *cam;
  job.sample_begin = begin, job.sample_count = count;
  job.rng_mode = opts->rng_mode, job.flags = opts->flags;
  job.radiance = radiance_dev;
  job.counters = s->counters;
  CUDA_TRY(cudaMemsetAsync(s->counters, 0, sizeof(RenderCounters), stream));

  if (opts->kernel == RT_KERNEL_EXACT) {
    if (cam->max_depth > 128)
      return fail(RT_ERR_UNSUPPORTED, "render: the exact kernel unwinds at most 128 levels (max_depth %d)",
                  cam->max_depth);
    int n_threads = 0;
    size_t need = exact_scratch_bytes(*cam, &n_threads);
    rc = ensure(&s->scratch, &s->scratch_bytes, need);
    if (rc != RT_OK)
      return rc;
    CUDA_TRY(cudaEventRecord(s->ev[1], stream));
    CUDA_TRY(exact_render(s->view, job, s->scratch, n_threads, stream));
    CUDA_TRY(cudaEventRecord(s->ev[2], stream));
    s->last_launches = 1;
  } else if (opts->kernel == RT_KERNEL_MEGA) {
    if (!s->mega_ok)
      return fail(RT_ERR_UNSUPPORTED, "render: %s (use RT_KERNEL_EXACT)", s->mega_why.c_str());
    MegaPlan plan;
    const int scene_class = s->view.n_media           ? MEGA_SCENE_MEDIA
                            : s->built_info.n_mquads ? MEGA_SCENE_QUADS
                                                     : MEGA_SCENE_SPHERES;
    CUDA_TRY(mega_plan(*cam, count, opts->rng_mode, s->device, scene_class, (int)s->view.feats | (s->view.wide_root != REF_EMPTY ? MEGA_FEATS_WIDE_TREE : 0), &plan));
    rc = ensure(&s->scratch, &s->scratch_bytes, plan.partial_bytes);
    if (rc != RT_OK)
      return rc;
    CUDA_TRY(cudaEventRecord(s->ev[1], stream));
    CUDA_TRY(mega_render(s->view, job, plan, (float *)s->scratch, stream, &s->last_launches));
    CUDA_TRY(cudaEventRecord(s->ev[2], stream));
  } else if (opts->kernel == RT_KERNEL_WAVE) {
    if (!s->mega_ok)
      return fail(RT_ERR_UNSUPPORTED, "render: %s (use RT_KERNEL_EXACT)", s->mega_why.c_str());
    WavePlan plan;
    CUDA_TRY(wave_plan(*cam, count, opts->rng_mode, s->device, &plan));
    rc = ensure(&s->scratch, &s->scratch_bytes, plan.scratch_bytes);
    if (rc != RT_OK)
      return rc;
    CUDA_TRY(cudaEventRecord(s->ev[1], stream));
    CUDA_TRY(wave_render(s->view, job, plan, s->scratch, stream, &s->last_launches));
    CUDA_TRY(cudaEventRecord(s->ev[2], stream));
  } else {
    return fail(RT_ERR_UNSUPPORTED, "render: kernel %u is not available in this build", opts->kernel);
  }
  return RT_OK;
}

static int collect_stats(rt_scene *s, rt_stats *stats, bool with_total) {
  RenderCounters c;
  CUDA_TRY(cudaMemcpy(&c, s->counters, sizeof(c), cudaMemcpyDeviceToHost));
  rt_stats st;
  memset(&st, 0, sizeof(st));
  st.segments = c.segments, st.light_pdf_evals = c.light_pdf_evals, st.samples = c.samples;
  st.nan_samples = c.nan_samples;
  CUDA_TRY(cudaEventElapsedTime(&st.kernel_ms, s->ev[1], s->ev[2]));
  if (with_total)
    CUDA_TRY(cudaEventElapsedTime(&st.total_ms, s->ev[0], s->ev[3]));
  st.launches = (uint32_t)s->last_launches;
  s->last = st;
  if (stats)
    *stats = st;
  return RT_OK;
}

extern "C" int rt_render_device(rt_scene *scene, const rt_camera *camera, const rt_opts *opts, float *radiance_sum_dev,
                                void *cuda_stream) {
  if (!scene || !camera || !opts || !radiance_sum_dev)
    return fail(RT_ERR_INVALID, "rt_render_device: NULL argument");
  DeviceGuard guard(scene->device);
  return enqueue_render(scene, camera, opts, radiance_sum_dev, (cudaStream_t)cuda_stream);
}

extern "C" int rt_render_device_stats(rt_scene *scene, void *cuda_stream, rt_stats *stats) {
  if (!scene)
    return fail(RT_ERR_INVALID, "rt_render_device_stats: NULL scene");
  DeviceGuard guard(scene->device);
  CUDA_TRY(cudaStreamSynchronize((cudaStream_t)cuda_stream));
  return collect_stats(scene, stats, false);
}

extern "C" int rt_render(rt_scene *scene, const rt_camera *camera, const rt_opts *opts, float *radiance_sum,
                         uint8_t *rgb8, rt_stats *stats) {
  if (!scene || !camera || !opts)
    return fail(RT_ERR_INVALID, "rt_render: NULL argument");
  DeviceGuard guard(scene->device);
  int begin = 0, count = 0;
  int rc = resolve_range(camera, opts, &begin, &count);
  if (rc != RT_OK)
    return rc;
  const size_t n_values = (size_t)camera->img_width * camera->img_height * 3;
  rc = ensure((void **)&scene->radiance, &scene->radiance_bytes, n_values * sizeof(float));
  if (rc == RT_OK)
    rc = ensure((void **)&scene->rgb8, &scene->rgb8_bytes, n_values);
  if (rc != RT_OK)
    return rc;
  cudaStream_t stream = nullptr;
  CUDA_TRY(cudaEventRecord(scene->ev[0], stream));
  rc = enqueue_render(scene, camera, opts, scene->radiance, stream);
  if (rc != RT_OK)
    return rc;
  if (rgb8) {
    rc = rt_tonemap_device(scene->radiance, n_values, count, scene->rgb8, stream);
    if (rc != RT_OK)
      return rc;
    scene->last_launches++;
    CUDA_TRY(cudaMemcpyAsync(rgb8, scene->rgb8, n_values, cudaMemcpyDeviceToHost, stream));
  }
  if (radiance_sum)
    CUDA_TRY(cudaMemcpyAsync(radiance_sum, scene->radiance, n_values * sizeof(float), cudaMemcpyDeviceToHost, stream));
  CUDA_TRY(cudaEventRecord(scene->ev[3], stream));
  CUDA_TRY(cudaStreamSynchronize(stream));
  return collect_stats(scene, stats, true);
}

/* ------------------------------------------------------------------ hit batch ---------------------------- */

extern "C" int rt_hit_batch(rt_scene *scene, uint32_t kernel, size_t n, const float *origins, const float *dirs,
                            float t_min, float t_max, uint64_t *rng_io, float *out_t, uint32_t *out_node, float *out_p,
                            float *out_normal, float *out_uv, uint8_t *out_front, uint32_t *out_material) {
  if (!scene || (n && (!origins || !dirs || !out_t || !out_node)))
    return fail(RT_ERR_INVALID, "rt_hit_batch: NULL argument");
  if (kernel != RT_KERNEL_EXACT && kernel != RT_KERNEL_MEGA && kernel != RT_KERNEL_WAVE)
    return fail(RT_ERR_INVALID, "rt_hit_batch: unknown kernel");
  if (kernel != RT_KERNEL_EXACT && !scene->mega_ok)
    return fail(RT_ERR_UNSUPPORTED, "rt_hit_batch: %s (use RT_KERNEL_EXACT)", scene->mega_why.c_str());
  if (n == 0)
    return RT_OK;
  DeviceGuard guard(scene->device);

  struct Buf {
    void *dev;
    void *host;
    size_t bytes;
    bool in, out;
  };
  std::vector<Buf> bufs;
  auto add = [&](const void *host, size_t bytes, bool in, bool out) -> void * {
    if (!host)
      return nullptr;
    Buf b = {nullptr, const_cast<void *>(host), bytes, in, out};
    if (cudaMalloc(&b.dev, bytes) != cudaSuccess)
      return nullptr;
    bufs.push_back(b);
    return b.dev;
  };
  HitBatch hb;
  memset(&hb, 0, sizeof(hb));
  hb.n = n, hb.t_min = t_min, hb.t_max = t_max;
  hb.origins = (const float *)add(origins, n * 12, true, false);
  hb.dirs = (const float *)add(dirs, n * 12, true, false);
  hb.rng_io = (uint64_t *)add(rng_io, n * 16, true, true);
  hb.out_t = (float *)add(out_t, n * 4, false, true);
  hb.out_node = (uint32_t *)add(out_node, n * 4, false, true);
  hb.out_p = (float *)add(out_p, n * 12, false, true);
  hb.out_normal = (float *)add(out_normal, n * 12, false, true);
  hb.out_uv = (float *)add(out_uv, n * 8, false, true);
  hb.out_front = (uint8_t *)add(out_front, n, false, true);
  hb.out_material = (uint32_t *)add(out_material, n * 4, false, true);

  cudaError_t err = cudaSuccess;
  size_t expected = 4 + (rng_io ? 1 : 0) + (out_p ? 1 : 0) + (out_normal ? 1 : 0) + (out_uv ? 1 : 0) +
                    (out_front ? 1 : 0) + (out_material ? 1 : 0);
  if (bufs.size() != expected)
    err = cudaErrorMemoryAllocation;
  for (size_t k = 0; k < bufs.size() && err == cudaSuccess; k++)
    if (bufs[k].in)
      err = cudaMemcpy(bufs[k].dev, bufs[k].host, bufs[k].bytes, cudaMemcpyHostToDevice);
  if (err == cudaSuccess)
    err = kernel == RT_KERNEL_EXACT ? exact_hit_batch(scene->view, hb, nullptr) : mega_hit_batch(scene->view, hb, nullptr);
  if (err == cudaSuccess)
    err = cudaDeviceSynchronize();
  for (size_t k = 0; k < bufs.size() && err == cudaSuccess; k++)
    if (bufs[k].out)
      err = cudaMemcpy(bufs[k].host, bufs[k].dev, bufs[k].bytes, cudaMemcpyDeviceToHost);
  for (Buf &b : bufs)
    cudaFree(b.dev);
  if (err != cudaSuccess)
    return fail(err == cudaErrorMemoryAllocation ? RT_ERR_NOMEM : RT_ERR_CUDA, "rt_hit_batch: %s",
                cudaGetErrorString(err));
  return RT_OK;
}

/* ------------------------------------------------------------------ single-process multi-GPU ------------- */

__global__ void add_into_kernel(float *__restrict__ acc, const float *__restrict__ other, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n)
    acc[i] += other[i];
}

namespace {

/* NCCL is loaded at run time (libnccl.so.2: the system library, or the one a host program such as PyTorch has
 * already mapped), so that librtc_b200.so has no link-time dependency on it; without it rt_render_multi gathers
 * over peer copies instead. */
struct NcclApi {
  void *handle = nullptr;
  ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*Reduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
  bool ok = false;
};
NcclApi &nccl_api() {
  static NcclApi api;
  static bool tried = false;
  if (tried)
    return api;
  tried = true;
  const char *off = getenv("RT_MULTI_NCCL");
  if (off && off[0] == '0')
    return api;
  /* Which libnccl.so.2: (1) RT_NCCL_LIB, a path given by the host program (the Python mirror points it at the copy
   * PyTorch bundles, so that a later `import torch` in the same process finds the SAME library under that soname and
   * not an older system one -- two NCCL versions cannot share a process); (2) one that is already mapped; (3) the
   * system's.  RTLD_LOCAL: nothing else in the process should bind to it because of us. */
  const char *path = getenv("RT_NCCL_LIB");
  if (path && *path)
    api.handle = dlopen(path, RTLD_NOW | RTLD_LOCAL);
  if (!api.handle)
    api.handle = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL | RTLD_NOLOAD);
  if (!api.handle)
    api.handle = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
  if (!api.handle)
    return api;
#define NCCL_SYM(field, name) api.field = reinterpret_cast<decltype(api.field)>(dlsym(api.handle, name))
  NCCL_SYM(CommInitAll, "ncclCommInitAll");
  NCCL_SYM(CommDestroy, "ncclCommDestroy");
  NCCL_SYM(Reduce, "ncclReduce");
  NCCL_SYM(GroupStart, "ncclGroupStart");
  NCCL_SYM(GroupEnd, "ncclGroupEnd");
  NCCL_SYM(GetErrorString, "ncclGetErrorString");
#undef NCCL_SYM
  api.ok = api.CommInitAll && api.CommDestroy && api.Reduce && api.GroupStart && api.GroupEnd && api.GetErrorString;
  return api;
}

/* what rt_render_multi keeps between calls: one stream and one event per device, the communicator (creating one
 * costs hundreds of milliseconds -- SURVEY.md 8(d) asks for it to be cached), staging for the peer-copy path */
struct MultiContext {
  int n = 0;
  bool nccl = false;
  std::vector<ncclComm_t> comms;
  std::vector<cudaStream_t> streams;
  std::vector<cudaEvent_t> done;
  float *staging = nullptr; /* device 0, peer-copy path only */
  size_t staging_bytes = 0;
};
MultiContext g_multi;
std::mutex g_multi_mutex;

void multi_teardown() {
  MultiContext &m = g_multi;
  for (int g = 0; g < m.n; g++) {
    cudaSetDevice(g);
    if (g < (int)m.comms.size() && m.comms[g])
      nccl_api().CommDestroy(m.comms[g]);
    if (g < (int)m.streams.size() && m.streams[g])
      cudaStreamDestroy(m.streams[g]);
    if (g < (int)m.done.size() && m.done[g])
      cudaEventDestroy(m.done[g]);
  }
  if (m.staging) {
    cudaSetDevice(0);
    dev_free(m.staging);
  }
  m = MultiContext();
}

int multi_setup_fresh(int n_gpus);
int multi_setup(int n_gpus) {
  if (g_multi.n == n_gpus)
    return RT_OK;
  multi_teardown();
  const int rc = multi_setup_fresh(n_gpus);
  if (rc != RT_OK)
    multi_teardown(); /* a half-made context must not look complete to the next call */
  return rc;
}
int multi_setup_fresh(int n_gpus) {
  MultiContext &m = g_multi;
  m.streams.assign(n_gpus, nullptr), m.done.assign(n_gpus, nullptr), m.comms.assign(n_gpus, nullptr);
  m.n = n_gpus;
  for (int g = 0; g < n_gpus; g++) {
    CUDA_TRY(cudaSetDevice(g));
    CUDA_TRY(cudaStreamCreateWithFlags(&m.streams[g], cudaStreamNonBlocking));
    CUDA_TRY(cudaEventCreateWithFlags(&m.done[g], cudaEventDisableTiming));
  }
  if (n_gpus > 1 && nccl_api().ok) { /* (one GPU exchanges nothing: NCCL is not even loaded) */
    NcclApi &api = nccl_api();
    std::vector<int> devs(n_gpus);
    for (int g = 0; g < n_gpus; g++)
      devs[g] = g;
    ncclResult_t r = api.CommInitAll(m.comms.data(), n_gpus, devs.data());
    if (r == ncclSuccess) {
      m.nccl = true;
    } else {
      fprintf(stderr, "rtc_b200: ncclCommInitAll failed (%s); gathering over peer copies\n", api.GetErrorString(r));
      m.comms.assign(n_gpus, nullptr);
    }
  }
  if (!m.nccl)
    for (int g = 1; g < n_gpus; g++) { /* peer copies go over NVLink when access is enabled */
      int can = 0;
      cudaSetDevice(0);
      if (cudaDeviceCanAccessPeer(&can, 0, g) == cudaSuccess && can && cudaDeviceEnablePeerAccess(g, 0) != cudaSuccess)
        cudaGetLastError(); /* already enabled */
    }
  return RT_OK;
}

} // namespace

/* Frees what the library keeps between calls: cached scenes that are not in use, the multi-GPU context. */
extern "C" void rt_shutdown(void) {
  std::vector<rt_scene *> idle;
  {
    std::lock_guard<std::mutex> lock(g_cache_mutex);
    for (size_t i = 0; i < g_cache.size();)
      if (!g_cache[i]->in_use) {
        idle.push_back(g_cache[i]);
        g_cache.erase(g_cache.begin() + (long)i);
      } else {
        i++;
      }
  }
  for (rt_scene *s : idle)
    scene_free(s);
  std::lock_guard<std::mutex> lock(g_multi_mutex);
  int previous = 0;
  cudaGetDevice(&previous);
  multi_teardown();
  cudaSetDevice(previous);
}

extern "C" int rt_render_multi(const rt_scene_desc *desc, const rt_camera *camera, const rt_opts *opts, int n_gpus,
                               float *radiance_sum, uint8_t *rgb8, rt_stats *stats) {
  if (!desc || !camera || !opts || n_gpus < 1)
    return fail(RT_ERR_INVALID, "rt_render_multi: bad argument");
  const int have = rt_device_count();
  if (have <= 0)
    return fail(RT_ERR_NO_DEVICE, "rt_render_multi: no CUDA device");
  if (n_gpus > have)
    return fail(RT_ERR_INVALID, "rt_render_multi: %d GPUs requested, %d present", n_gpus, have);
  int begin = 0, count = 0;
  rt_opts base = *opts;
  base.rng_mode = n_gpus > 1 ? (uint32_t)RT_RNG_SAMPLE : opts->rng_mode;
  int rc = resolve_range(camera, &base, &begin, &count);
  if (rc != RT_OK)
    return rc;
  if (count < n_gpus)
    return fail(RT_ERR_INVALID, "rt_render_multi: fewer samples than GPUs");

  std::lock_guard<std::mutex> lock(g_multi_mutex);
  int previous = 0;
  cudaGetDevice(&previous);
  const size_t n_values = (size_t)camera->img_width * camera->img_height * 3;
  std::vector<rt_scene *> scenes(n_gpus, nullptr);
  /* every exit path: wait for the devices, give the scenes back to the cache, restore the caller's device */
  auto finish = [&](int status) {
    for (int g = 0; g < n_gpus; g++)
      if (scenes[g]) {
        cudaSetDevice(g);
        cudaStreamSynchronize(g_multi.streams[g]);
        rt_scene_release(scenes[g]);
      }
    cudaSetDevice(previous);
    return status;
  };
#define MULTI_TRY(expr)                                                                                              \
  do {                                                                                                               \
    cudaError_t err__ = (expr);                                                                                      \
    if (err__ != cudaSuccess)                                                                                        \
      return finish(fail(RT_ERR_CUDA, "rt_render_multi: %s: %s", #expr, cudaGetErrorString(err__)));                 \
  } while (0)

  rc = multi_setup(n_gpus);
  if (rc != RT_OK) {
    cudaSetDevice(previous);
    return rc;
  }
  MultiContext &m = g_multi;
  for (int g = 0; g < n_gpus; g++) {
    rc = rt_scene_acquire(desc, g, &scenes[g]);
    if (rc == RT_OK) {
      MULTI_TRY(cudaSetDevice(g));
      rc = ensure((void **)&scenes[g]->radiance, &scenes[g]->radiance_bytes, n_values * sizeof(float));
    }
    if (rc != RT_OK)
      return finish(rc);
  }
  MULTI_TRY(cudaSetDevice(0));
  rc = ensure((void **)&scenes[0]->rgb8, &scenes[0]->rgb8_bytes, n_values);
  if (rc == RT_OK && !m.nccl && n_gpus > 1)
    rc = ensure((void **)&m.staging, &m.staging_bytes, n_values * sizeof(float));
  if (rc != RT_OK)
    return finish(rc);
  MULTI_TRY(cudaEventRecord(scenes[0]->ev[0], m.streams[0]));

  /* sample range of GPU g: [begin + g*count/n, begin + (g+1)*count/n) */
  for (int g = 0; g < n_gpus; g++) {
    MULTI_TRY(cudaSetDevice(g));
    rt_opts mine = base;
    mine.sample_begin = begin + (int)((long long)count * g / n_gpus);
    mine.sample_count = begin + (int)((long long)count * (g + 1) / n_gpus) - mine.sample_begin;
    rc = enqueue_render(scenes[g], camera, &mine, scenes[g]->radiance, m.streams[g]);
    if (rc != RT_OK)
      return finish(rc);
    MULTI_TRY(cudaEventRecord(m.done[g], m.streams[g]));
  }
  if (m.nccl && n_gpus > 1) {
    /* the path's one exchange step: sum of the per-device float frames onto device 0 (in place), NCCL over NVLink */
    NcclApi &api = nccl_api();
    ncclResult_t r = api.GroupStart();
    for (int g = 0; g < n_gpus && r == ncclSuccess; g++)
      r = api.Reduce(scenes[g]->radiance, scenes[g]->radiance, n_values, ncclFloat, ncclSum, 0, m.comms[g],
                     m.streams[g]);
    ncclResult_t r_end = api.GroupEnd();
    if (r == ncclSuccess)
      r = r_end;
    if (r != ncclSuccess)
      return finish(fail(RT_ERR_CUDA, "rt_render_multi: ncclReduce: %s", api.GetErrorString(r)));
    MULTI_TRY(cudaSetDevice(0));
  } else {
    /* gather over peer copies and sum on device 0 in GPU order */
    MULTI_TRY(cudaSetDevice(0));
    const unsigned blocks = (unsigned)((n_values + 255) / 256);
    for (int g = 1; g < n_gpus; g++) {
      MULTI_TRY(cudaStreamWaitEvent(m.streams[0], m.done[g], 0));
      MULTI_TRY(cudaMemcpyPeerAsync(m.staging, 0, scenes[g]->radiance, g, n_values * sizeof(float), m.streams[0]));
      add_into_kernel<<<blocks, 256, 0, m.streams[0]>>>(scenes[0]->radiance, m.staging, n_values);
      MULTI_TRY(cudaGetLastError());
    }
  }
  if (rgb8) {
    rc = rt_tonemap_device(scenes[0]->radiance, n_values, count, scenes[0]->rgb8, m.streams[0]);
    if (rc != RT_OK)
      return finish(rc);
    MULTI_TRY(cudaMemcpyAsync(rgb8, scenes[0]->rgb8, n_values, cudaMemcpyDeviceToHost, m.streams[0]));
  }
  if (radiance_sum)
    MULTI_TRY(cudaMemcpyAsync(radiance_sum, scenes[0]->radiance, n_values * sizeof(float), cudaMemcpyDeviceToHost,
                              m.streams[0]));
  MULTI_TRY(cudaEventRecord(scenes[0]->ev[3], m.streams[0]));
  for (int g = 0; g < n_gpus; g++) {
    MULTI_TRY(cudaSetDevice(g));
    MULTI_TRY(cudaStreamSynchronize(m.streams[g]));
  }
  rt_stats total;
  memset(&total, 0, sizeof(total));
  for (int g = 0; g < n_gpus && rc == RT_OK; g++) {
    cudaSetDevice(g);
    rt_stats st;
    rc = collect_stats(scenes[g], &st, g == 0);
    total.segments += st.segments, total.light_pdf_evals += st.light_pdf_evals, total.samples += st.samples;
    total.nan_samples += st.nan_samples;
    total.kernel_ms = st.kernel_ms > total.kernel_ms ? st.kernel_ms : total.kernel_ms;
    total.launches += st.launches;
    if (g == 0)
      total.total_ms = st.total_ms;
  }
  total.launches += (m.nccl ? 0u : (uint32_t)(n_gpus - 1)) + (rgb8 ? 1u : 0u);
  if (stats && rc == RT_OK)
    *stats = total;
  return finish(rc);
#undef MULTI_TRY
}

/* ------------------------------------------------------------------ render without allocation ------------ */

/* Scratch of rt_render_device_noalloc: the render counters (first 256 bytes) and the per-chunk partial sums of the
 * megakernel, in a file-scope array instead of pool memory.  Every GPU that loads the library holds these 32 MiB
 * until the process exits.  NOTE: in host code `g_noalloc_scratch` is a host-side stand-in for the symbol, not a
 * device address -- only cudaGetSymbolAddress gives the pointer a kernel can use (noalloc_scratch below). */
#define NOALLOC_HEAD_BYTES 256u
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_HEAD_BYTES + RT_NOALLOC_SCRATCH_BYTES];

namespace {
std::mutex g_noalloc_mutex;
void *g_noalloc_ptr[64] = {nullptr};

/* device address of the array on `device` (which must be current): looked up once per device */
int noalloc_scratch(int device, unsigned char **out) {
  if (device < 0 || device >= 64)
    return fail(RT_ERR_INVALID, "rt_render_device_noalloc: device %d out of range", device);
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (!g_noalloc_ptr[device]) {
    void *ptr = nullptr;
    cudaError_t err = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
    if (err != cudaSuccess || !ptr)
      return fail(RT_ERR_CUDA, "rt_render_device_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(err));
    g_noalloc_ptr[device] = ptr;
  }
  *out = static_cast<unsigned char *>(g_noalloc_ptr[device]);
  return RT_OK;
}
} // namespace

extern "C" int rt_render_device_noalloc(rt_scene *scene, const rt_camera *camera, const rt_opts *opts,
                                        float *radiance_sum_dev, void *cuda_stream) {
  if (!scene || !camera || !opts || !radiance_sum_dev)
    return fail(RT_ERR_INVALID, "rt_render_device_noalloc: NULL argument");
  if (opts->kernel != RT_KERNEL_MEGA)
    return fail(RT_ERR_UNSUPPORTED, "rt_render_device_noalloc: only RT_KERNEL_MEGA renders without allocating "
                                    "(the exact and wavefront kernels need scratch far beyond the static array)");
  if (!scene->mega_ok)
    return fail(RT_ERR_UNSUPPORTED, "rt_render_device_noalloc: %s", scene->mega_why.c_str());
  int begin = 0, count = 0;
  int rc = resolve_range(camera, opts, &begin, &count);
  if (rc != RT_OK)
    return rc;
  /* the maximum: two partial-sum frames (12 bytes per pixel each) must fit the array */
  const unsigned long long pixels = (unsigned long long)camera->img_width * (unsigned long long)camera->img_height;
  if (pixels > RT_NOALLOC_MAX_PIXELS)
    return fail(RT_ERR_INVALID, "rt_render_device_noalloc: frame %dx%d = %llu pixels exceeds the maximum of %llu pixels "
                                "(static scratch of %u bytes); use rt_render_device",
                camera->img_width, camera->img_height, pixels, (unsigned long long)RT_NOALLOC_MAX_PIXELS,
                (unsigned)RT_NOALLOC_SCRATCH_BYTES);
  DeviceGuard guard(scene->device);
  cudaStream_t stream = (cudaStream_t)cuda_stream;
  unsigned char *scratch = nullptr;
  rc = noalloc_scratch(scene->device, &scratch);
  if (rc != RT_OK)
    return rc;

  MegaPlan plan;
  const int scene_class = scene->view.n_media           ? MEGA_SCENE_MEDIA
                          : scene->built_info.n_mquads ? MEGA_SCENE_QUADS
                                                       : MEGA_SCENE_SPHERES;
  CUDA_TRY(mega_plan(*camera, count, opts->rng_mode, scene->device, scene_class, (int)scene->view.feats | (scene->view.wide_root != REF_EMPTY ? MEGA_FEATS_WIDE_TREE : 0), &plan));
  /* as many work items per pixel as partial-sum frames fit: the trailing items are merged into the last one */
  const size_t frame_bytes = (size_t)pixels * 3 * sizeof(float);
  const int cap = (int)(RT_NOALLOC_SCRATCH_BYTES / frame_bytes);
  if (plan.n_chunks > cap) {
    plan.n_chunks = cap; /* >= 2 by the size check above */
    plan.begin[cap] = count;
  }
  plan.partial_bytes = plan.n_chunks > 1 ? (size_t)plan.n_chunks * frame_bytes : 0;

  RenderJob job;
  memset(&job, 0, sizeof(job));
  job.cam = *camera;
  job.sample_begin = begin, job.sample_count = count;
  job.rng_mode = opts->rng_mode, job.flags = opts->flags;
  job.radiance = radiance_sum_dev;
  job.counters = reinterpret_cast<RenderCounters *>(scratch);
  /* the array keeps the last call's contents: the counters (work-item cursor!) must start at zero; every partial
   * sum a launch reads was written by that launch */
  CUDA_TRY(cudaMemsetAsync(scratch, 0, NOALLOC_HEAD_BYTES, stream));
  int launches = 0;
  CUDA_TRY(mega_render(scene->view, job, plan, reinterpret_cast<float *>(scratch + NOALLOC_HEAD_BYTES), stream,
                       &launches));
  return RT_OK;
}

/* ------------------------------------------------------------------ FP32 roofline probe ------------------ */

__global__ void fp32_peak_kernel(float *out, int iters) {
  float a0 = threadIdx.x * 1e-3f, a1 = a0 + 1.0f, a2 = a0 + 2.0f, a3 = a0 + 3.0f;
  float a4 = a0 + 4.0f, a5 = a0 + 5.0f, a6 = a0 + 6.0f, a7 = a0 + 7.0f;
  const float m = 0.999f, c = 1e-3f;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int k = 0; k < 16; k++) {
      a0 = fmaf(a0, m, c), a1 = fmaf(a1, m, c), a2 = fmaf(a2, m, c), a3 = fmaf(a3, m, c);
      a4 = fmaf(a4, m, c), a5 = fmaf(a5, m, c), a6 = fmaf(a6, m, c), a7 = fmaf(a7, m, c);
    }
  }
  out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
}

extern "C" int rt_fp32_peak(int device, float *tflops) {
  if (!tflops)
    return fail(RT_ERR_INVALID, "rt_fp32_peak: NULL argument");
  if (rt_device_count() <= 0)
    return fail(RT_ERR_NO_DEVICE, "rt_fp32_peak: no CUDA device");
  if (device < 0)
    CUDA_TRY(cudaGetDevice(&device));
  DeviceGuard guard(device);
  int sms = 0;
  CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
  const int blocks = sms * 8, threads = 256, iters = 4096;
  float *out = nullptr;
  CUDA_TRY(cudaMalloc(&out, (size_t)blocks * threads * sizeof(float)));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0), cudaEventCreate(&e1);
  fp32_peak_kernel<<<blocks, threads>>>(out, 64); /* warm-up */
  float best = 0.0f;
  for (int rep = 0; rep < 5; rep++) {
    cudaEventRecord(e0);
    fp32_peak_kernel<<<blocks, threads>>>(out, iters);
    cudaEventRecord(e1);
    cudaError_t err = cudaEventSynchronize(e1);
    if (err != cudaSuccess) {
      cudaFree(out);
      return fail(RT_ERR_CUDA, "rt_fp32_peak: %s", cudaGetErrorString(err));
    }
    float ms = 0.0f;
    cudaEventElapsedTime(&ms, e0, e1);
    double flop = 2.0 * 8 * 16 * (double)iters * blocks * threads;
    float tf = (float)(flop / (ms * 1e-3) * 1e-12);
    best = tf > best ? tf : best;
  }
  cudaEventDestroy(e0), cudaEventDestroy(e1);
  cudaFree(out);
  *tflops = best;
  return RT_OK;
}
