/*
 * rtc_host.h -- host-side helpers of the backend that the reference has no name for: flattening the Hittable
 * graph into the rt_scene_desc the CUDA library consumes, packing the camera, the built-in scene builders, and
 * (de)serialising a flattened scene.  librtc_host.so exports these next to the drop-in API of rtc_api.h.
 */
#ifndef RTC_HOST_H
#define RTC_HOST_H

#include "rtc_api.h"
#include "rtc_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Walks world->objects then world->lights (the structs of /root/reference/include/hittable.h:35-110,
 * material.h:29-36, texture.h:5-48) and fills `out` with freshly malloc'ed arrays.  Object kinds are recognised by
 * vtable / value-function ADDRESS, learned by constructing one throw-away object per kind through the public
 * constructors that are linked into the program -- so the same code flattens graphs built by this backend's
 * constructors and graphs built by the reference's own hittable.c/material.c/texture.c.
 * Returns 0, or -1 with a message in rtc_flatten_error(). */
int rtc_flatten(const World *world, rt_scene_desc *out);
const char *rtc_flatten_error(void);
void rtc_scene_desc_free(rt_scene_desc *desc);
/* Ids the most recent rtc_flatten() gave to a Hittable* / Material* (RT_NONE if it was not reached). */
uint32_t rtc_flatten_last_node_id(const void *hittable);
uint32_t rtc_flatten_last_material_id(const void *material);

/* One little-endian blob: the rt_scene_desc header (pointers zeroed) followed by the arrays in field order. */
int rtc_scene_desc_write(const rt_scene_desc *desc, FILE *f);
int rtc_scene_desc_read(rt_scene_desc *out, FILE *f);

/* Camera (after Camera_init) -> the POD the kernels take.  include/raytracing.h:16-38. */
void rtc_camera_pack(const Camera *camera, rt_camera *out);

/* The reference's defaults (src/main.c:280-287) and its eight scene builders (src/main.c:9-273), rebuilt.
 * rtc_scene_build returns 0, or -1 for an unknown id (the reference falls back to scene 0; callers decide). */
void rtc_camera_defaults(Camera *camera);
int rtc_scene_build(int scene_id, World *world, Camera *camera);
const char *rtc_scene_title(int scene_id);

/* Image ingest (csrc/host/image_decode.c, image_io.c): PNG, baseline JPEG and binary PPM to RGB8, the bytes
 * stbi_load(..., 3) gives the reference (/root/reference/src/texture.c:38-42).  rtc_load_image returns NULL with
 * *exists = 0 when there is no such file, NULL with *exists = 1 and rtc_image_error() set when it cannot be decoded. */
uint8_t *rtc_load_image(const char *filename, int *width, int *height, int *exists);
const char *rtc_image_error(void);
uint8_t *rtc_load_ppm(const char *filename, int *width, int *height);
void rtc_synthetic_image(int width, int height, uint8_t *rgb);

/* Render through the C ABI with explicit options (Camera_render uses the RT_KERNEL / RT_RNG / RT_GPUS
 * environment variables to fill them).  Returns an rt_status. */
int rtc_camera_render_opts(const Camera *camera, const World *world, const rt_opts *opts, int n_gpus,
                           uint8_t *rgb8, float *radiance_sum, rt_stats *stats);

/* Opaque handle around one built-in scene, for FFI callers that cannot hold World / Camera (session.c).
 * width/spp/depth <= 0 and aspect/lights_prob < 0 keep the scene's own value. */
typedef struct rtc_session rtc_session;
rtc_session *rtc_session_open(int scene_id, int width, int spp, int depth, float aspect, float lights_prob);
void rtc_session_close(rtc_session *s);
int rtc_session_width(const rtc_session *s);
int rtc_session_height(const rtc_session *s);
int rtc_session_spp(const rtc_session *s);
void rtc_session_camera(const rtc_session *s, rt_camera *out);
size_t rtc_session_camera_raw(const rtc_session *s, void *out, size_t cap);
int rtc_session_flatten(const rtc_session *s, rt_scene_desc *out);
int rtc_session_write_blob(const rtc_session *s, const char *path);
int rtc_session_render(const rtc_session *s, const rt_opts *opts, int n_gpus, uint8_t *rgb8, float *radiance_sum,
                       rt_stats *stats);
void rtc_session_camera_render(const rtc_session *s, uint8_t *rgb8);

#ifdef __cplusplus
}
#endif
#endif /* RTC_HOST_H */
