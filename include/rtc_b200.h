/*
 * rtc_b200.h -- C ABI of the sm_100a path-tracing library (librtc_b200.so).
 *
 * Plain C: pointers, sizes and PODs only; no torch, no C++ types.  This is the boundary a host program (C, or
 * anything with an FFI) binds to run the per-pixel render loop of gau-nernst/ray-tracing-c on B200s.  The
 * reference has no FFI of its own; the calls below replace these in-process interfaces of the reference
 * (paths relative to /root/reference):
 *
 *   rt_render / rt_render_device   <- Camera_render()            src/raytracing.c:86-135 (include/raytracing.h:41)
 *                                     Camera_ray_color()         src/raytracing.c:39-84
 *   rt_tonemap_device              <- gamma+quantise epilogue    src/raytracing.c:127-131, include/utils.h:11-17
 *   rt_hit_batch                   <- HittableVTable.hit slot    include/hittable.h:30  (all of src/hittable.c:38-423)
 *   rt_scene_desc (flattened graph)<- World{objects,lights}      include/raytracing.h:9-12 and the struct family in
 *                                     include/hittable.h:35-110, include/material.h:29-36, include/texture.h:5-48
 *   rt_camera                      <- Camera (derived fields)    include/raytracing.h:16-38, src/raytracing.c:13-37
 *
 * Every function returns RT_OK (0) or a negative rt_status; rt_last_error() gives the text (thread-local).
 * There is no CPU fallback: without a CUDA device every compute entry point fails with RT_ERR_NO_DEVICE.
 */
#ifndef RTC_B200_H
#define RTC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_ABI_VERSION 1u
#define RT_NONE 0xFFFFFFFFu

typedef enum rt_status {
  RT_OK = 0,
  RT_ERR_INVALID = -1,     /* bad argument / malformed scene description */
  RT_ERR_NO_DEVICE = -2,   /* no CUDA device (this library never computes on the host) */
  RT_ERR_CUDA = -3,        /* a CUDA runtime call or kernel failed */
  RT_ERR_UNSUPPORTED = -4, /* graph shape outside what the kernels implement */
  RT_ERR_NOMEM = -5,
} rt_status;

/* ---------------------------------------------------------------- flattened scene ------------------------ */

/* One entry per Hittable of the reference graph, numbered in DFS pre-order from world.objects (node 0).
 * Shared sub-objects (a light that is also in `objects`, a medium boundary, BVH n==1 left==right) keep ONE id.
 * The id of a SPHERE / QUAD / MEDIUM node is what rt_hit_batch reports as the hit object id. */
typedef enum rt_node_kind {
  RT_NODE_LIST = 0,      /* HittableList  hittable.h:40-45  a = first slot in children[], b = count          */
  RT_NODE_SPHERE = 1,    /* Sphere        hittable.h:51-56  a = index into spheres[]                          */
  RT_NODE_QUAD = 2,      /* Quad          hittable.h:61-71  a = index into quads[]                            */
  RT_NODE_BVH = 3,       /* BVHNode       hittable.h:77-81  a = left node, b = right node                     */
  RT_NODE_TRANSLATE = 4, /* Translate     hittable.h:86-90  a = child node, f[0..2] = offset                  */
  RT_NODE_ROTATE_Y = 5,  /* RotateY       hittable.h:95-100 a = child node, f[0] = sin, f[1] = cos (host libm) */
  RT_NODE_MEDIUM = 6,    /* ConstantMedium hittable.h:105-110 a = boundary node, b = phase material,
                                                             f[0] = neg_inv_density                           */
} rt_node_kind;

#define RT_NODE_HAS_PDF 1u /* vtable->pdf / ->rand non-NULL in the reference: LIST, SPHERE, QUAD */

typedef struct rt_node {
  uint32_t kind;
  uint32_t a;
  uint32_t b;
  uint32_t flags;
  float bbox[6]; /* Hittable.bbox: xlo,xhi,ylo,yhi,zlo,zhi */
  float f[4];
  uint32_t reserved[2];
} rt_node; /* 64 bytes */

typedef struct rt_sphere {
  float center[3];
  float radius;
  uint32_t material;
  uint32_t node; /* back reference: id of the owning node */
  uint32_t reserved[2];
} rt_sphere; /* 32 bytes */

typedef struct rt_quad {
  float Q[3];
  float D;
  float u[3];
  float area;
  float v[3];
  uint32_t material;
  float normal[3];
  uint32_t node;
  float w[3];
  uint32_t reserved;
} rt_quad; /* 80 bytes */

/* tags equal the reference's MaterialType values (material.h:20-27) */
typedef enum rt_material_tag {
  RT_MAT_SURFACE_NORMAL = 0,
  RT_MAT_LAMBERTIAN = 1,
  RT_MAT_METAL = 2,
  RT_MAT_DIELECTRIC = 3,
  RT_MAT_DIFFUSE_LIGHT = 4,
  RT_MAT_ISOTROPIC = 5,
} rt_material_tag;

typedef struct rt_material {
  uint32_t tag;
  uint32_t texture; /* RT_NONE for DIELECTRIC / SURFACE_NORMAL */
  float param;      /* fuzz (METAL) or eta (DIELECTRIC) */
  uint32_t reserved;
} rt_material; /* 16 bytes */

typedef enum rt_texture_kind {
  RT_TEX_SOLID = 0,   /* color                                   texture.c:8            */
  RT_TEX_CHECKER = 1, /* scale, a = even texture, b = odd texture texture.c:12-22        */
  RT_TEX_IMAGE = 2,   /* a = index into images[]                 texture.c:28-37        */
  RT_TEX_PERLIN = 3,  /* scale, a = index into perlins[], b = depth  texture.c:47-52    */
} rt_texture_kind;

typedef struct rt_texture {
  uint32_t kind;
  uint32_t a;
  uint32_t b;
  float scale;
  float color[3];
  uint32_t reserved;
} rt_texture; /* 32 bytes */

typedef struct rt_perlin {
  float grad[256 * 3]; /* grad_field[i] = grad[3i..3i+2] */
  int32_t perm_x[256];
  int32_t perm_y[256];
  int32_t perm_z[256];
} rt_perlin; /* 6144 bytes */

typedef struct rt_image {
  int32_t width;
  int32_t height;
  uint64_t offset; /* byte offset of the first texel in texels[]; RGB8, row-major, top row first */
} rt_image;

typedef struct rt_scene_desc {
  uint32_t abi_version; /* RT_ABI_VERSION */
  uint32_t n_nodes;
  uint32_t n_children;
  uint32_t n_spheres;
  uint32_t n_quads;
  uint32_t n_materials;
  uint32_t n_textures;
  uint32_t n_perlins;
  uint32_t n_images;
  uint32_t objects_root; /* node id of world.objects (always 0) */
  uint32_t lights_root;  /* node id of world.lights, RT_NONE if that list is empty */
  uint32_t reserved;
  uint64_t n_texel_bytes;
  const rt_node *nodes;
  const uint32_t *children; /* item node ids of every LIST, contiguous per list */
  const rt_sphere *spheres;
  const rt_quad *quads;
  const rt_material *materials;
  const rt_texture *textures;
  const rt_perlin *perlins;
  const rt_image *images;
  const uint8_t *texels;
} rt_scene_desc;

/* ---------------------------------------------------------------- camera / options ----------------------- */

/* The derived fields Camera_init() computes on the host with glibc tanf (raytracing.c:13-37); the device never
 * re-derives them, so camera rays are bit-identical to the reference's. */
typedef struct rt_camera {
  int32_t img_width;
  int32_t img_height;
  int32_t samples_per_pixel;
  int32_t max_depth;
  float background[3];
  float lights_sampling_prob;
  float origin[3]; /* look_from */
  float dof_angle; /* lens sampling iff > 0 (raytracing.c:108) */
  float pixel00[3];
  float reserved0;
  float pixel_delta_u[3];
  float reserved1;
  float pixel_delta_v[3];
  float reserved2;
  float dof_disc_u[3];
  float reserved3;
  float dof_disc_v[3];
  float reserved4;
} rt_camera;

typedef enum rt_kernel {
  RT_KERNEL_EXACT = 0, /* reference traversal order + portable math: bit-identical to oracle (parity anchor)  */
  RT_KERNEL_MEGA = 1,  /* persistent megakernel over the rebuilt SAH BVH (throughput path)                   */
  RT_KERNEL_WAVE = 2,  /* wavefront: generate / extend / shade queues                                         */
} rt_kernel;

typedef enum rt_rng_mode {
  RT_RNG_PIXEL = 0,  /* one PCG32 stream per pixel seeded (17+j, 23+i), shared by its samples: raytracing.c:93-94 */
  RT_RNG_SAMPLE = 1, /* one stream per (pixel, sample): seed ((s<<32)|(17+j), 23+i); equals PIXEL for s = 0      */
} rt_rng_mode;

#define RT_FLAG_DROP_NAN 1u /* skip samples whose radiance is NaN instead of poisoning the pixel (and count them) */

typedef struct rt_opts {
  uint32_t kernel;      /* rt_kernel */
  uint32_t rng_mode;    /* rt_rng_mode */
  int32_t sample_begin; /* first sample index this call renders (multi-GPU sample-range shard) */
  int32_t sample_count; /* how many; 0 = camera.samples_per_pixel - sample_begin. RT_RNG_PIXEL needs the full range */
  uint32_t flags;
  uint32_t reserved[3];
} rt_opts;

typedef struct rt_stats {
  uint64_t segments;        /* closest-hit queries issued by the integrator (calls of objects->hit, raytracing.c:45) */
  uint64_t light_pdf_evals; /* mixture-pdf evaluations (raytracing.c:66-67) */
  uint64_t samples;         /* camera samples traced */
  uint64_t nan_samples;     /* samples whose radiance had a NaN component */
  float kernel_ms;          /* CUDA-event time of the render kernel(s) only */
  float total_ms;           /* CUDA-event time from first H2D to last D2H of the call (rt_render only) */
  uint32_t launches;        /* kernels launched by the call */
  uint32_t reserved;
} rt_stats;

typedef struct rt_scene rt_scene; /* opaque: device copy of one flattened scene on one GPU */

/* ---------------------------------------------------------------- entry points --------------------------- */

const char *rt_last_error(void);
uint32_t rt_abi_version(void);
int rt_device_count(void);

/* Validates `desc`, builds the device representation (reference-order graph + rebuilt SAH BVH groups) and uploads
 * it to CUDA device `device` (-1 = current).  `desc` may be freed after the call. */
int rt_scene_create(const rt_scene_desc *desc, int device, rt_scene **out);
void rt_scene_destroy(rt_scene *scene);

/* The same through a small cache of uploaded scenes, keyed by a 128-bit hash of the CONTENT of `desc` (arrays, counts,
 * roots) and the device: a second call with an equal description returns the resident copy instead of rebuilding the
 * BVH and uploading again.  No pointer of the caller is retained.  Camera_render() uses this pair, so that a host
 * program calling it repeatedly on one World (as the reference's main() would in a loop) pays the scene set-up once.
 * rt_scene_release() hands the scene back (it stays resident, least recently used ones are evicted beyond four per
 * device); rt_scene_destroy() on an acquired scene removes it for good.  rt_shutdown() frees every idle cached scene
 * and the multi-GPU context (streams, NCCL communicator) that rt_render_multi() keeps between calls. */
int rt_scene_acquire(const rt_scene_desc *desc, int device, rt_scene **out);
void rt_scene_release(rt_scene *scene);
void rt_shutdown(void);

/* Host-only: validates `desc` and builds the rebuilt acceleration structure WITHOUT touching a device; reports its
 * shape.  (What rt_scene_create would upload.)  Works on machines without a GPU. */
typedef struct rt_scene_info {
  uint32_t bvh_nodes;     /* binary nodes (each holds both children's boxes) */
  uint32_t bvh_depth;     /* deepest traversal stack the kernels need */
  uint32_t spheres, quads, instances, media, chain_ops;
  uint32_t graph_frames, graph_rays, graph_recs; /* nesting of the reference-order graph (exact kernel limits) */
  int32_t light_nesting;  /* lists under world.lights; -1 = a list nothing can be sampled from */
  uint32_t mega_supported;
  uint64_t device_bytes;  /* bytes rt_scene_create uploads */
  uint32_t pre_prims;     /* oversized primitives kept out of the tree and tested before every walk (0..2) */
  uint32_t wide_nodes;    /* nodes of the 4-wide tree (0 unless built with RT_BVH_WIDE=1) */
} rt_scene_info;
int rt_scene_inspect(const rt_scene_desc *desc, rt_scene_info *info);

/* Camera_render through HOST buffers: uploads nothing but the camera, renders samples
 * [sample_begin, sample_begin+sample_count) of every pixel, and copies back
 *   radiance_sum : float[h][w][3] per-pixel SUM over the rendered samples (may be NULL)
 *   rgb8         : uint8[h][w][3] = tonemap(radiance_sum / sample_count)   (may be NULL)
 * Blocking.  `stats` may be NULL. */
int rt_render(rt_scene *scene, const rt_camera *camera, const rt_opts *opts, float *radiance_sum, uint8_t *rgb8,
              rt_stats *stats);

/* Same render into a DEVICE buffer float[h][w][3] (overwritten, not accumulated), asynchronous on `cuda_stream`
 * (a cudaStream_t, NULL = default stream).  For one-process-per-GPU drivers that reduce with NCCL afterwards.
 * Asynchronous for RT_KERNEL_MEGA and RT_KERNEL_EXACT only: RT_KERNEL_WAVE polls its live-path count from the host
 * between batches of iterations, i.e. it SYNCHRONISES `cuda_stream` (idle on return), and fails with RT_ERR_CUDA
 * (launch timeout) instead of returning an incomplete frame if its iteration bound runs out.  The first call on a
 * scene (or a larger frame) may allocate scratch from the stream-ordered pool; rt_render_device_noalloc never does. */
int rt_render_device(rt_scene *scene, const rt_camera *camera, const rt_opts *opts, float *radiance_sum_dev,
                     void *cuda_stream);
/* rt_render_device WITHOUT ANY ALLOCATION, for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it).  This entry point has no counterpart in the reference.
 * Same render (RT_KERNEL_MEGA only), same arguments, asynchronous on
 * `cuda_stream`; the kernel's scratch -- render counters and per-chunk partial sums -- lives in ONE file-scope device
 * array of RT_NOALLOC_SCRATCH_BYTES (32 MiB, held by every GPU that loads the library until the process exits)
 * instead of pool memory, and the call performs no cudaMalloc / cudaFree / cudaMallocAsync of any kind.
 * Limit: frames of at most RT_NOALLOC_MAX_PIXELS pixels (two partial-sum frames of 12 bytes per pixel must fit);
 * larger frames and the other kernels are refused with RT_ERR_INVALID / RT_ERR_UNSUPPORTED and a message.  Fewer
 * partial frames than rt_render_device would use means a coarser work-item split (the sums differ in the last bits).
 * All calls SHARE the array: calls that overlap on the GPU (two streams, two threads, two scenes on one device) race;
 * serialise them.  The counters are in the array too, so rt_render_device_stats does not see this call. */
#define RT_NOALLOC_SCRATCH_BYTES (32u << 20)
#define RT_NOALLOC_MAX_PIXELS (RT_NOALLOC_SCRATCH_BYTES / (2u * 3u * 4u))
int rt_render_device_noalloc(rt_scene *scene, const rt_camera *camera, const rt_opts *opts, float *radiance_sum_dev,
                             void *cuda_stream);
/* Waits for `cuda_stream` and returns the counters of the last rt_render_device on this scene. */
int rt_render_device_stats(rt_scene *scene, void *cuda_stream, rt_stats *stats);

/* rgb8_dev[i] = (int)(256 * clamp(sqrtf(sum[i] / spp), 0, 0.999)), NaN -> 0   (raytracing.c:127-131). */
int rt_tonemap_device(const float *radiance_sum_dev, size_t n_values, int spp, uint8_t *rgb8_dev, void *cuda_stream);
/* The same epilogue through HOST buffers (upload, kernel, download; blocking): for a caller that summed the float
 * radiance of several sample-range passes itself (rtc_main --passes / --checkpoint). */
int rt_tonemap(const float *radiance_sum, size_t n_values, int spp, uint8_t *rgb8);

/* Closest hit of n rays against world.objects, all HOST arrays, [t_min, t_max] as in the vtable slot.
 *   origins, dirs : float[n][3]
 *   rng_io        : uint64[n][2] {state, inc} or NULL; consumed (and written back) by ConstantMedium nodes
 *   out_t         : float[n]   (+inf on miss)            out_node  : uint32[n] (RT_NONE on miss)
 *   out_p, out_normal : float[n][3] or NULL              out_uv    : float[n][2] or NULL
 *   out_front     : uint8[n] or NULL                     out_material : uint32[n] or NULL
 * kernel = RT_KERNEL_EXACT walks the reference graph; RT_KERNEL_MEGA walks the rebuilt BVH. */
int rt_hit_batch(rt_scene *scene, uint32_t kernel, size_t n, const float *origins, const float *dirs, float t_min,
                 float t_max, uint64_t *rng_io, float *out_t, uint32_t *out_node, float *out_p, float *out_normal,
                 float *out_uv, uint8_t *out_front, uint32_t *out_material);

/* Single-process multi-GPU render through HOST buffers (what Camera_render does under RT_GPUS=N): replicates the
 * scene on devices 0..n_gpus-1 (through the scene cache), gives device g the sample range [g*spp/n, (g+1)*spp/n)
 * (RT_RNG_SAMPLE streams), sums the per-device float frames onto device 0 with ONE ncclReduce(sum, float32, root 0)
 * inside ncclGroupStart/End on a communicator created once with ncclCommInitAll and kept (NCCL is dlopen'ed:
 * libnccl.so.2; without it, or with RT_MULTI_NCCL=0, the frames are gathered over peer copies and added on device 0),
 * and tonemaps there.  One-process-per-GPU drivers use rt_render_device + their own NCCL reduce instead (see
 * ray-tracing-c_b200/dist.py). */
int rt_render_multi(const rt_scene_desc *desc, const rt_camera *camera, const rt_opts *opts, int n_gpus,
                    float *radiance_sum, uint8_t *rgb8, rt_stats *stats);

/* FP32 FMA-chain microbenchmark on `device`: returns achieved TFLOP/s (2 flop per FMA) in *tflops.  Used by
 * bench.py as the measured compute roofline (MEASURED_PEAKS.json only holds HBM and tensor peaks). */
int rt_fp32_peak(int device, float *tflops);

#ifdef __cplusplus
}
#endif
#endif /* RTC_B200_H */
