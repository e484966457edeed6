/*
 * ref_tool.c -- TEST INFRASTRUCTURE.  A thin driver around the UNMODIFIED reference, compiled from the sources
 * where they lie under /root/reference (see oracle/Makefile; output goes to oracle/_ref/, never into history).
 * It exists to pin the oracle restatement (oracle/rtc_oracle.c) and the host layer (constructors, BVH build,
 * flattener) against the real thing, and to generate the golden fixtures in tests/golden/.
 *
 * It pulls the reference's scene builders in by including its main.c with `main` renamed, links the reference's
 * hittable.c / material.c / texture.c / raytracing.c / vec3.c / pcg32.c / utils.c as they are, and links THIS
 * repo's flatten.c so that the very same flattener runs over a graph built by the reference's constructors.
 *
 *   ref_tool render  <scene> <width> <spp> <depth> <aspect|0> <out.rgb8> <out.camera>
 *                                                                           Camera_render of the reference; prints
 *                                                                           `render_seconds=<s>` (monotonic clock around
 *                                                                           Camera_render only: the scope of the
 *                                                                           reference's own timer, src/main.c:334-338)
 *   ref_tool flatten <scene> <out.blob>                                     rt_scene_desc of the reference's graph
 *   ref_tool hits    <scene> <rays.bin> <out.bin>                           world.objects.hit() on a ray batch
 *   ref_tool tiff    0 <width> <height> <channels> <out.tiff>               the reference's write_tiff (src/tiff.c:15-49)
 *                                                                           on the byte pattern (7 i + 3) mod 251
 *   ref_tool image   0 <file> <out.rgb8>                                    the reference's Image_init (src/texture.c:38-42:
 *                                                                           stbi_load, 3 channels) -> "w h\n" + RGB8 bytes
 *
 * Scenes 3 and 7 load "earthmap.jpg" from the working directory; the tool first writes the synthetic image there
 * as a binary PPM under that name (stb_image sniffs content, not the extension).
 */
#define main reference_main_unused
#include "main.c" /* /root/reference/src/main.c via -I/root/reference/src */
#undef main

#include "rtc_b200.h"
#include <math.h>
#include <string.h>
#include <time.h>

/* from this repo's flatten.c / image_io.c (declared by hand: rtc_host.h would redeclare the reference's types) */
int rtc_flatten(const World *world, rt_scene_desc *out);
const char *rtc_flatten_error(void);
int rtc_scene_desc_write(const rt_scene_desc *desc, FILE *f);
uint32_t rtc_flatten_last_material_id(const void *material);
void rtc_synthetic_image(int width, int height, uint8_t *rgb);

_Static_assert(sizeof(Hittable) == 32 && sizeof(Sphere) == 56 && sizeof(Quad) == 112 && sizeof(BVHNode) == 48 &&
                   sizeof(HitRecord) == 48 && sizeof(Material) == 24 && sizeof(Perlin) == 6160 &&
                   sizeof(Camera) == 180,
               "reference struct layout changed");

static void write_earthmap(void) {
  enum { W = 1024, H = 512 };
  uint8_t *rgb = malloc((size_t)W * H * 3);
  rtc_synthetic_image(W, H, rgb);
  FILE *f = fopen("earthmap.jpg", "wb");
  assert(f != NULL);
  fprintf(f, "P6\n%d %d\n255\n", W, H);
  fwrite(rgb, 1, (size_t)W * H * 3, f);
  fclose(f);
  free(rgb);
}

static void build(int scene, World *world, Camera *camera) {
  camera->aspect_ratio = 16.0f / 9.0f;
  camera->img_width = 500;
  camera->samples_per_pixel = 100;
  camera->max_depth = 50;
  camera->vup = vec3(0, 1, 0);
  camera->dof_angle = 0.0f;
  camera->focal_length = 10.0f;
  camera->lights_sampling_prob = 0.5f;
  if (scene == 3 || scene == 7)
    write_earthmap();
  switch (scene) {
  case 0: scene_metal_and_lambertian(world, camera); break;
  case 1: scene_book1_final(world, camera); break;
  case 2: scene_checker(world, camera); break;
  case 3: scene_earth(world, camera); break;
  case 4: scene_perlin(world, camera); break;
  case 5: scene_simple_light(world, camera); break;
  case 6: scene_cornell_box(world, camera); break;
  case 7: scene_book2_final(world, camera, true); break;
  default: fprintf(stderr, "ref_tool: scene must be 0..7\n"); exit(2);
  }
}

/* one ray of a `hits` batch: 48 bytes in, 64 bytes out */
typedef struct RayIn {
  float o[3], d[3], t_min, t_max;
  uint64_t rng_state, rng_inc;
} RayIn;
typedef struct HitOut {
  float t, p[3], normal[3], u, v;
  uint32_t hit, front_face, material;
  uint64_t rng_state;
  uint32_t pad[2];
} HitOut;
_Static_assert(sizeof(RayIn) == 48 && sizeof(HitOut) == 64, "batch record sizes");

int main(int argc, char **argv) {
  if (argc < 3) {
    fprintf(stderr, "usage: ref_tool render|flatten|hits <scene> ...\n");
    return 2;
  }
  const char *cmd = argv[1];
  if (strcmp(cmd, "tiff") == 0 && argc == 7) {
    int w = atoi(argv[3]), h = atoi(argv[4]), c = atoi(argv[5]);
    size_t n = (size_t)w * h * c;
    uint8_t *px = malloc(n);
    for (size_t i = 0; i < n; i++)
      px[i] = (uint8_t)((7 * i + 3) % 251);
    FILE *f = fopen(argv[6], "wb");
    assert(f != NULL);
    int rc = write_tiff(f, w, h, c, px);
    fclose(f);
    return rc;
  }
  if (strcmp(cmd, "image") == 0 && argc == 5) {
    Image im;
    Image_init(&im, argv[3]); /* asserts when stb cannot decode the file, like the reference */
    FILE *f = fopen(argv[4], "wb");
    assert(f != NULL);
    fprintf(f, "%d %d\n", im.width, im.height);
    fwrite(im.buffer, 1, (size_t)im.width * im.height * 3, f);
    fclose(f);
    return 0;
  }
  int scene = atoi(argv[2]);
  World world = {0};
  Camera camera;
  memset(&camera, 0, sizeof(camera));
  build(scene, &world, &camera);

  if (strcmp(cmd, "render") == 0 && argc == 9) {
    camera.img_width = atoi(argv[3]);
    camera.samples_per_pixel = atoi(argv[4]);
    camera.max_depth = atoi(argv[5]);
    if (atof(argv[6]) > 0)
      camera.aspect_ratio = (float)atof(argv[6]);
    Camera_init(&camera);
    size_t n = (size_t)camera.img_width * camera.img_height * 3;
    uint8_t *image = malloc(n);
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    Camera_render(&camera, &world, image);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    printf("render_seconds=%.6f\n", (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec));
    FILE *f = fopen(argv[7], "wb");
    assert(f != NULL);
    fwrite(image, 1, n, f);
    fclose(f);
    f = fopen(argv[8], "w"); /* sidecar: the derived camera, for cross-checking Camera_init */
    assert(f != NULL);
    fwrite(&camera, sizeof(camera), 1, f);
    fclose(f);
    return 0;
  }

  rt_scene_desc desc;
  if (rtc_flatten(&world, &desc) != 0) {
    fprintf(stderr, "ref_tool: flatten failed: %s\n", rtc_flatten_error());
    return 1;
  }

  if (strcmp(cmd, "flatten") == 0 && argc == 4) {
    FILE *f = fopen(argv[3], "wb");
    assert(f != NULL);
    int rc = rtc_scene_desc_write(&desc, f);
    fclose(f);
    return rc ? 1 : 0;
  }

  if (strcmp(cmd, "hits") == 0 && argc == 5) {
    FILE *fin = fopen(argv[3], "rb");
    assert(fin != NULL);
    fseek(fin, 0, SEEK_END);
    size_t n = (size_t)ftell(fin) / sizeof(RayIn);
    fseek(fin, 0, SEEK_SET);
    RayIn *rays = malloc(n * sizeof(RayIn));
    HitOut *out = calloc(n, sizeof(HitOut));
    if (fread(rays, sizeof(RayIn), n, fin) != n)
      return 1;
    fclose(fin);

    const Hittable *objects = &world.objects.hittable;
    for (size_t i = 0; i < n; i++) {
      Ray ray = {{{rays[i].o[0], rays[i].o[1], rays[i].o[2]}}, {{rays[i].d[0], rays[i].d[1], rays[i].d[2]}}};
      PCG32 rng = {rays[i].rng_state, rays[i].rng_inc};
      HitRecord rec;
      memset(&rec, 0, sizeof(rec)); /* fields a hit does not write (medium: normal,u,v,front) read as 0 or stale */
      bool hit = objects->vtable->hit(objects, &ray, rays[i].t_min, rays[i].t_max, &rec, &rng);
      out[i].hit = hit;
      out[i].rng_state = rng.state;
      if (hit) {
        out[i].t = rec.t;
        memcpy(out[i].p, rec.p.values, 12);
        memcpy(out[i].normal, rec.normal.values, 12);
        out[i].u = rec.u;
        out[i].v = rec.v;
        out[i].front_face = rec.front_face;
        out[i].material = rtc_flatten_last_material_id(rec.material);
      } else {
        out[i].t = INFINITY;
        out[i].material = RT_NONE;
      }
    }
    FILE *fo = fopen(argv[4], "wb");
    assert(fo != NULL);
    fwrite(out, sizeof(HitOut), n, fo);
    fclose(fo);
    return 0;
  }

  fprintf(stderr, "ref_tool: bad arguments\n");
  return 2;
}
