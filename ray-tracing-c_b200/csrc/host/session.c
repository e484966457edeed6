/*
 * session.c -- an opaque handle around one built-in scene (World + Camera), for callers that cannot hold the C
 * structs themselves (Python through ctypes: bench.py, tests).  Everything goes through the public drop-in API:
 * rtc_scene_build -> Camera_init -> rtc_flatten / rtc_camera_render_opts.  Nothing here computes pixels.
 */
#include "rtc_host.h"
#include <string.h>

typedef struct rtc_session {
  World world;
  Camera camera;
  int scene_id;
} rtc_session;

/* width/spp/depth <= 0 and aspect/lights_prob < 0 keep the scene's own value (src/main.c:280-287 + builder). */
rtc_session *rtc_session_open(int scene_id, int width, int spp, int depth, float aspect, float lights_prob) {
  rtc_session *s = calloc(1, sizeof(*s));
  if (s == NULL)
    return NULL;
  rtc_camera_defaults(&s->camera);
  if (width > 0)
    s->camera.img_width = width;
  if (spp > 0)
    s->camera.samples_per_pixel = spp;
  if (rtc_scene_build(scene_id, &s->world, &s->camera) != 0) {
    free(s);
    return NULL;
  }
  if (depth > 0)
    s->camera.max_depth = depth;
  if (aspect > 0)
    s->camera.aspect_ratio = aspect;
  if (lights_prob >= 0)
    s->camera.lights_sampling_prob = lights_prob;
  Camera_init(&s->camera);
  s->scene_id = scene_id;
  return s;
}

void rtc_session_close(rtc_session *s) { free(s); /* the object graph is never freed, like the reference's */ }

int rtc_session_width(const rtc_session *s) { return s->camera.img_width; }
int rtc_session_height(const rtc_session *s) { return s->camera.img_height; }
int rtc_session_spp(const rtc_session *s) { return s->camera.samples_per_pixel; }

void rtc_session_camera(const rtc_session *s, rt_camera *out) { rtc_camera_pack(&s->camera, out); }
/* raw Camera bytes (include/rtc_api.h layout == the reference's, 180 bytes) for cross-checks with ref_tool */
size_t rtc_session_camera_raw(const rtc_session *s, void *out, size_t cap) {
  if (cap >= sizeof(Camera))
    memcpy(out, &s->camera, sizeof(Camera));
  return sizeof(Camera);
}

int rtc_session_flatten(const rtc_session *s, rt_scene_desc *out) { return rtc_flatten(&s->world, out); }

int rtc_session_write_blob(const rtc_session *s, const char *path) {
  rt_scene_desc desc;
  if (rtc_flatten(&s->world, &desc) != 0)
    return -1;
  FILE *f = fopen(path, "wb");
  if (f == NULL) {
    rtc_scene_desc_free(&desc);
    return -1;
  }
  int rc = rtc_scene_desc_write(&desc, f);
  fclose(f);
  rtc_scene_desc_free(&desc);
  return rc;
}

/* The whole of Camera_render with explicit options: flatten -> upload -> kernels -> rgb8 in HOST memory. */
int rtc_session_render(const rtc_session *s, const rt_opts *opts, int n_gpus, uint8_t *rgb8, float *radiance_sum,
                       rt_stats *stats) {
  return rtc_camera_render_opts(&s->camera, &s->world, opts, n_gpus, rgb8, radiance_sum, stats);
}

/* The literal drop-in call (environment-driven options, aborts on failure like the reference's asserts). */
void rtc_session_camera_render(const rtc_session *s, uint8_t *rgb8) { Camera_render(&s->camera, &s->world, rgb8); }
