/*
 * camera.c -- World / Camera set-up and THE drop-in entry point Camera_render().
 *
 *   World_init      /root/reference/src/raytracing.c:8-11
 *   Camera_init     /root/reference/src/raytracing.c:13-37  (kept on the host with glibc tanf; the device only sees
 *                   the derived vectors, so camera rays are bit-identical to the reference's)
 *   Camera_render   /root/reference/src/raytracing.c:86-135 -- here: flatten -> rt_scene_acquire -> rt_render
 *                   (sm_100a kernels) -> rgb8 back in the caller's buffer.  No pixel is computed on the host.
 *
 * Environment knobs read by Camera_render (the signature has no room for options):
 *   RT_KERNEL = mega (default) | exact | wave      RT_RNG = sample (default) | pixel
 *   RT_GPUS   = number of GPUs to shard the sample range over inside this process (default 1)
 */
#include "rtc_host.h"
#include <math.h>
#include <string.h>

void World_init(World *world, size_t max_objects) {
  HittableList_init(&world->objects, max_objects);
  HittableList_init(&world->lights, max_objects);
}

void rtc_camera_defaults(Camera *camera) { /* src/main.c:280-287 */
  memset(camera, 0, sizeof(*camera));
  camera->aspect_ratio = 16.0f / 9.0f;
  camera->img_width = 500;
  camera->samples_per_pixel = 100;
  camera->max_depth = 50;
  camera->vup = vec3(0, 1, 0);
  camera->dof_angle = 0.0f;
  camera->focal_length = 10.0f;
  camera->lights_sampling_prob = 0.5f;
}

void Camera_init(Camera *cam) {
  cam->img_height = (int)((float)cam->img_width / cam->aspect_ratio);

  /* the viewport sits at focal_length; its width uses the INTEGER pixel ratio */
  float vp_h = 2.0f * tanf(cam->vfov * (float)M_PI / 360.0f) * cam->focal_length;
  float vp_w = vp_h * (float)cam->img_width / (float)cam->img_height;

  cam->w = vec3_normalize(vec3_sub_vec3(cam->look_from, cam->look_to));
  cam->u = vec3_cross(cam->vup, cam->w); /* deliberately not normalised, like the reference */
  cam->v = vec3_cross(cam->w, cam->u);

  Vec3 across = vec3_mul_float(cam->u, vp_w);
  Vec3 down = vec3_mul_float(cam->v, -vp_h);
  cam->pixel_delta_u = vec3_div_float(across, (float)cam->img_width);
  cam->pixel_delta_v = vec3_div_float(down, (float)cam->img_height);

  Vec3 corner = vec3_add_vec3(cam->look_from, vec3_mul_float(cam->w, -cam->focal_length));
  corner = vec3_add_vec3(corner, vec3_mul_float(across, -0.5f));
  corner = vec3_add_vec3(corner, vec3_mul_float(down, -0.5f));
  Vec3 first = vec3_add_vec3(corner, vec3_mul_float(cam->pixel_delta_u, 0.5f));
  cam->pixel00_loc = vec3_add_vec3(first, vec3_mul_float(cam->pixel_delta_v, 0.5f));

  float lens_radius = cam->focal_length * tanf(cam->dof_angle * (float)M_PI / 360.0f);
  cam->dof_disc_u = vec3_mul_float(cam->u, lens_radius);
  cam->dof_disc_v = vec3_mul_float(cam->v, lens_radius);
}

int rtc_camera_render_opts(const Camera *camera, const World *world, const rt_opts *opts, int n_gpus, uint8_t *rgb8,
                           float *radiance_sum, rt_stats *stats) {
  rt_scene_desc desc;
  if (rtc_flatten(world, &desc) != 0) {
    fprintf(stderr, "rtc_b200: cannot flatten the scene: %s\n", rtc_flatten_error());
    return RT_ERR_UNSUPPORTED;
  }
  rt_camera cam;
  rtc_camera_pack(camera, &cam);

  int status;
  if (n_gpus <= 1) {
    /* the uploaded scene comes from (and goes back to) the library's cache, keyed by the description's content: a
     * second Camera_render on the same World skips the BVH build and the upload */
    rt_scene *scene = NULL;
    status = rt_scene_acquire(&desc, -1, &scene);
    if (status == RT_OK) {
      status = rt_render(scene, &cam, opts, radiance_sum, rgb8, stats);
      rt_scene_release(scene);
    }
  } else {
    status = rt_render_multi(&desc, &cam, opts, n_gpus, radiance_sum, rgb8, stats);
  }
  rtc_scene_desc_free(&desc);
  return status;
}

static uint32_t env_choice(const char *name, const char *const *words, int n_words, uint32_t fallback) {
  const char *v = getenv(name);
  if (v == NULL || *v == 0)
    return fallback;
  for (int i = 0; i < n_words; i++)
    if (strcmp(v, words[i]) == 0)
      return (uint32_t)i;
  fprintf(stderr, "rtc_b200: ignoring %s=%s\n", name, v);
  return fallback;
}

void Camera_render(const Camera *camera, const World *world, uint8_t *buffer) {
  static const char *const kernels[] = {"exact", "mega", "wave"};
  static const char *const rngs[] = {"pixel", "sample"};
  rt_opts opts;
  memset(&opts, 0, sizeof(opts));
  opts.kernel = env_choice("RT_KERNEL", kernels, 3, RT_KERNEL_MEGA);
  opts.rng_mode = env_choice("RT_RNG", rngs, 2, RT_RNG_SAMPLE);
  int n_gpus = 1;
  const char *g = getenv("RT_GPUS");
  if (g != NULL && atoi(g) > 1)
    n_gpus = atoi(g);
  if (n_gpus > 1)
    opts.rng_mode = RT_RNG_SAMPLE; /* a sample-range shard needs per-sample streams */

  fprintf(stderr, "\rScanlines remaining: %d", camera->img_height); /* the reference's progress line, coarse */
  rt_stats stats;
  int status = rtc_camera_render_opts(camera, world, &opts, n_gpus, buffer, NULL, &stats);
  if (status == RT_ERR_UNSUPPORTED && opts.kernel != RT_KERNEL_EXACT) {
    /* the rebuilt structure has limits the reference does not (255 distinct Translate/RotateY chains, 48 levels of
     * BVH): the reference-order kernel has neither, so the frame is still rendered, only slower */
    fprintf(stderr, "\nrtc_b200: %s -- rendering with the reference-order kernel (RT_KERNEL=exact) instead\n",
            rt_last_error());
    opts.kernel = RT_KERNEL_EXACT;
    status = rtc_camera_render_opts(camera, world, &opts, n_gpus, buffer, NULL, &stats);
  }
  if (status != RT_OK) {
    /* the reference's error convention is assert -> abort; there is no CPU fallback to hide behind */
    fprintf(stderr, "\nrtc_b200: render failed (%d): %s\n", status, rt_last_error());
    abort();
  }
  fprintf(stderr, "\rScanlines remaining: 0");
  fprintf(stderr, "\nDone\n");
  if (getenv("RT_STATS") != NULL)
    fprintf(stderr,
            "{\"segments\": %llu, \"samples\": %llu, \"kernel_ms\": %.3f, \"total_ms\": %.3f, \"mseg_per_s\": %.1f, "
            "\"n_gpus\": %d}\n",
            (unsigned long long)stats.segments, (unsigned long long)stats.samples, stats.kernel_ms, stats.total_ms,
            stats.kernel_ms > 0 ? (double)stats.segments / stats.kernel_ms * 1e-3 : 0.0, n_gpus);
}
