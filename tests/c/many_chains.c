/* Test program (tests/test_gpu_boundary.py): a World of N Translate(RotateY(Sphere)) items built through the public
 * constructors of include/*.h -- more distinct transform chains than the rebuilt structure can tag (255) -- rendered
 * with the literal drop-in call.  Camera_render must fall back to the reference-order kernel, not abort. */
#include "hittable.h"
#include "material.h"
#include "raytracing.h"
#include "texture.h"
#include <stdio.h>
#include <stdlib.h>

int main(int argc, char **argv) {
  int n = argc > 1 ? atoi(argv[1]) : 300;
  World world = {0};
  World_init(&world, (size_t)n + 4);
  Material *grey = Lambertian_new(Solid_new(vec3(0.6f, 0.6f, 0.6f)));
  Material *red = Lambertian_new(Solid_new(vec3(0.8f, 0.2f, 0.2f)));
  HittableList_append(&world.objects, Sphere_new(vec3(0, -1000, 0), 1000, grey));
  for (int k = 0; k < n; k++) {
    Hittable *s = Sphere_new(vec3(0, 0.4f, 0), 0.4f, red);
    Hittable *r = RotateY_new(s, (float)(k % 90));
    HittableList_append(&world.objects, Translate_new(r, vec3((float)(k % 20) - 10.0f, 0, (float)(k / 20) - 8.0f)));
  }
  Camera camera = {0};
  camera.aspect_ratio = 2.0f;
  camera.img_width = 64;
  camera.samples_per_pixel = 4;
  camera.max_depth = 8;
  camera.background = vec3(0.7f, 0.8f, 1.0f);
  camera.vfov = 40;
  camera.look_from = vec3(0, 6, 18);
  camera.look_to = vec3(0, 0, 0);
  camera.vup = vec3(0, 1, 0);
  camera.dof_angle = 0;
  camera.focal_length = 10;
  camera.lights_sampling_prob = 0;
  Camera_init(&camera);
  size_t bytes = (size_t)camera.img_width * camera.img_height * 3;
  uint8_t *image = malloc(bytes);
  Camera_render(&camera, &world, image);
  unsigned long sum = 0;
  for (size_t i = 0; i < bytes; i++)
    sum += image[i];
  printf("%d %d %lu\n", camera.img_width, camera.img_height, sum);
  return 0;
}
