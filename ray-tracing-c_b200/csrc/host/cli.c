/*
 * cli.c -- `rtc_main <scene 0-7> [width spp extra]`: the reference's command line (src/main.c:275-345) on the
 * B200 backend.  Same positional quirk (width needs 3 arguments, spp needs 4), same banners on stderr, same
 * output.tiff; the timer is monotonic with millisecond resolution instead of time().
 *
 * Extras (all optional, anywhere on the line):
 *   --width W  --spp N        the two settings WITHOUT the positional quirk (src/main.c:289-292 reads width only when
 *                             there are three arguments and spp only when there are four)
 *   --depth N  --aspect A  --lights-prob P  --out FILE
 *   --passes K                progressive accumulation: the spp are rendered in K sample-range passes (per-(pixel,
 *                             sample) streams make the ranges additive); after every pass the image so far is written
 *   --checkpoint FILE         ... and the float radiance sum so far is saved (little-endian: magic, width, height,
 *                             spp, samples done, then float[h][w][3])
 *   --resume                  start from --checkpoint FILE if it fits the frame, instead of from sample 0
 *   --stop-after K            end this run after K passes (the checkpoint then holds a partial sum to resume from)
 */
#include "rtc_host.h"
#include <string.h>
#include <time.h>

static double now_s(void) {
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

#define CKPT_MAGIC 0x54435452u /* "RTCT" */
typedef struct {
  uint32_t magic, width, height, spp, done;
} CkptHeader;

static int checkpoint_save(const char *path, const Camera *cam, int done, const float *sum) {
  char tmp[4096];
  snprintf(tmp, sizeof(tmp), "%s.tmp", path);
  FILE *f = fopen(tmp, "wb");
  if (f == NULL)
    return -1;
  CkptHeader h = {CKPT_MAGIC, (uint32_t)cam->img_width, (uint32_t)cam->img_height, (uint32_t)cam->samples_per_pixel,
                  (uint32_t)done};
  size_t n = (size_t)cam->img_width * cam->img_height * 3;
  int ok = fwrite(&h, sizeof(h), 1, f) == 1 && fwrite(sum, sizeof(float), n, f) == n;
  ok = (fclose(f) == 0) && ok;
  return ok && rename(tmp, path) == 0 ? 0 : -1; /* the old checkpoint stays valid until the new one is complete */
}
static int checkpoint_load(const char *path, const Camera *cam, float *sum) {
  FILE *f = fopen(path, "rb");
  if (f == NULL)
    return 0;
  CkptHeader h;
  size_t n = (size_t)cam->img_width * cam->img_height * 3;
  int done = 0;
  if (fread(&h, sizeof(h), 1, f) == 1 && h.magic == CKPT_MAGIC && h.width == (uint32_t)cam->img_width &&
      h.height == (uint32_t)cam->img_height && h.spp == (uint32_t)cam->samples_per_pixel &&
      h.done <= (uint32_t)cam->samples_per_pixel && fread(sum, sizeof(float), n, f) == n)
    done = (int)h.done;
  fclose(f);
  return done;
}

static void save_image(const char *path, const Camera *cam, uint8_t *image) {
  FILE *f = fopen(path, "wb");
  assert((f != NULL) && "Failed to open file");
  write_tiff(f, cam->img_width, cam->img_height, 3, image);
  fclose(f);
}

int main(int argc, char *argv[]) {
  assert(argc > 1);

  /* split "--flag value" pairs from the reference's positional arguments */
  const char *out_path = "output.tiff", *ckpt_path = NULL;
  int depth = -1, width = -1, spp = -1, passes = 1, resume = 0, stop_after = 0;
  float aspect = -1.0f, lights_prob = -1.0f;
  int n_pos = 1;
  for (int i = 1; i < argc; i++) {
    if (strcmp(argv[i], "--resume") == 0) {
      resume = 1;
    } else if (strncmp(argv[i], "--", 2) == 0 && i + 1 < argc) {
      if (strcmp(argv[i], "--depth") == 0)
        depth = atoi(argv[++i]);
      else if (strcmp(argv[i], "--width") == 0)
        width = atoi(argv[++i]);
      else if (strcmp(argv[i], "--spp") == 0)
        spp = atoi(argv[++i]);
      else if (strcmp(argv[i], "--aspect") == 0)
        aspect = (float)atof(argv[++i]);
      else if (strcmp(argv[i], "--out") == 0)
        out_path = argv[++i];
      else if (strcmp(argv[i], "--lights-prob") == 0)
        lights_prob = (float)atof(argv[++i]);
      else if (strcmp(argv[i], "--passes") == 0)
        passes = atoi(argv[++i]);
      else if (strcmp(argv[i], "--checkpoint") == 0)
        ckpt_path = argv[++i];
      else if (strcmp(argv[i], "--stop-after") == 0)
        stop_after = atoi(argv[++i]);
      else
        fprintf(stderr, "rtc_main: unknown option %s\n", argv[i++]);
    } else {
      argv[n_pos++] = argv[i];
    }
  }

  World world;
  memset(&world, 0, sizeof(world));
  Camera camera;
  rtc_camera_defaults(&camera);
  if (n_pos > 3)
    camera.img_width = (int)strtol(argv[2], NULL, 10);
  if (n_pos > 4)
    camera.samples_per_pixel = (int)strtol(argv[3], NULL, 10);
  if (width > 0)
    camera.img_width = width;
  if (spp > 0)
    camera.samples_per_pixel = spp;

  int scene = n_pos > 1 ? (int)strtol(argv[1], NULL, 10) : 0;
  if (rtc_scene_title(scene) == NULL) {
    fprintf(stderr, "Unsupported option. Default to 0\n");
    scene = 0;
  }
  fprintf(stderr, "%s\n", rtc_scene_title(scene));
  rtc_scene_build(scene, &world, &camera);
  if (depth > 0)
    camera.max_depth = depth;
  if (aspect > 0)
    camera.aspect_ratio = aspect; /* e.g. 16/9 at width 3840 for the 4K sweep; scene 7 forces 1.0 otherwise */
  if (lights_prob >= 0)
    camera.lights_sampling_prob = lights_prob;
  Camera_init(&camera);

  const size_t n_values = (size_t)camera.img_width * camera.img_height * 3;
  uint8_t *image = my_malloc(n_values);
  double t0 = now_s();
  if (passes <= 1 && ckpt_path == NULL) {
    Camera_render(&camera, &world, image);
  } else {
    /* progressive / checkpointable accumulation: sample ranges are additive under the per-(pixel, sample) streams */
    if (passes < 1)
      passes = 1;
    if (passes > camera.samples_per_pixel)
      passes = camera.samples_per_pixel;
    float *total = my_malloc(n_values * sizeof(float)), *part = my_malloc(n_values * sizeof(float));
    memset(total, 0, n_values * sizeof(float));
    int done = (resume && ckpt_path != NULL) ? checkpoint_load(ckpt_path, &camera, total) : 0;
    if (done > 0)
      fprintf(stderr, "Resuming at sample %d of %d\n", done, camera.samples_per_pixel);
    else
      memset(total, 0, n_values * sizeof(float));
    const int per_pass = (camera.samples_per_pixel + passes - 1) / passes;
    int passes_run = 0;
    while (done < camera.samples_per_pixel && (stop_after <= 0 || passes_run++ < stop_after)) {
      rt_opts opts;
      memset(&opts, 0, sizeof(opts));
      opts.kernel = RT_KERNEL_MEGA, opts.rng_mode = RT_RNG_SAMPLE;
      opts.sample_begin = done;
      opts.sample_count = camera.samples_per_pixel - done < per_pass ? camera.samples_per_pixel - done : per_pass;
      int status = rtc_camera_render_opts(&camera, &world, &opts, 1, NULL, part, NULL);
      if (status == RT_ERR_UNSUPPORTED) { /* see Camera_render: the reference-order kernel has no structural limits */
        opts.kernel = RT_KERNEL_EXACT;
        status = rtc_camera_render_opts(&camera, &world, &opts, 1, NULL, part, NULL);
      }
      if (status == RT_OK) {
        for (size_t i = 0; i < n_values; i++) /* merging checkpoints is bookkeeping; every sample came from the GPU */
          total[i] += part[i];
        done += opts.sample_count;
        status = rt_tonemap(total, n_values, done, image); /* the epilogue (raytracing.c:127-131) on the device */
      }
      if (status != RT_OK) {
        fprintf(stderr, "\nrtc_b200: render failed (%d): %s\n", status, rt_last_error());
        abort();
      }
      fprintf(stderr, "\rSamples done: %d of %d", done, camera.samples_per_pixel);
      save_image(out_path, &camera, image);
      if (ckpt_path != NULL && checkpoint_save(ckpt_path, &camera, done, total) != 0)
        fprintf(stderr, "\nrtc_main: cannot write the checkpoint %s\n", ckpt_path);
    }
    fprintf(stderr, "\nDone\n");
    free(total), free(part);
  }
  double t1 = now_s();
  fprintf(stderr, "Took %.3f seconds\n", t1 - t0);

  save_image(out_path, &camera, image);
  return 0;
}
