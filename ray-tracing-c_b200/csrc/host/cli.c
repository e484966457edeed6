/*
 * cli.c -- `rtc_main <scene 0-7> [width spp extra]`: the reference's command line (src/main.c:275-345) on the
 * B200 backend.  Same positional quirk (width needs 3 arguments, spp needs 4), same banners on stderr, same
 * output.tiff; the timer is monotonic with millisecond resolution instead of time().
 *
 * Extras (all optional, after the positional arguments):  --depth N  --aspect A  --out FILE  --lights-prob P
 */
#include "rtc_host.h"
#include <string.h>
#include <time.h>

static double now_s(void) {
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

int main(int argc, char *argv[]) {
  assert(argc > 1);

  /* split "--flag value" pairs from the reference's positional arguments */
  const char *out_path = "output.tiff";
  int depth = -1;
  float aspect = -1.0f, lights_prob = -1.0f;
  int n_pos = 1;
  for (int i = 1; i < argc; i++) {
    if (strncmp(argv[i], "--", 2) == 0 && i + 1 < argc) {
      if (strcmp(argv[i], "--depth") == 0)
        depth = atoi(argv[++i]);
      else if (strcmp(argv[i], "--aspect") == 0)
        aspect = (float)atof(argv[++i]);
      else if (strcmp(argv[i], "--out") == 0)
        out_path = argv[++i];
      else if (strcmp(argv[i], "--lights-prob") == 0)
        lights_prob = (float)atof(argv[++i]);
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

  int scene = (int)strtol(argv[1], NULL, 10);
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

  uint8_t *image = my_malloc((size_t)camera.img_width * camera.img_height * 3);
  double t0 = now_s();
  Camera_render(&camera, &world, image);
  double t1 = now_s();
  fprintf(stderr, "Took %.3f seconds\n", t1 - t0);

  FILE *f = fopen(out_path, "wb");
  assert((f != NULL) && "Failed to open file");
  write_tiff(f, camera.img_width, camera.img_height, 3, image);
  fclose(f);
  return 0;
}
