/*
 * scenes.c -- the eight built-in scenes, rebuilt against the drop-in API (rtc_api.h).
 *
 * Same objects, in the same order, from the same PCG32 draws as /root/reference/src/main.c:9-273, so that the
 * flattened scene is bit-identical to the one the reference's builders produce under gcc (tests compare the two
 * blobs).  Where the reference leaves the order of RNG draws to the compiler (arguments of one call, main.c:58,62)
 * the gcc/x86-64 order -- right-most argument first -- is written out explicitly here.
 */
#include "rtc_host.h"
#include <math.h>

static Texture *rgb(float r, float g, float b) { return Solid_new(vec3(r, g, b)); }

static void add(World *world, Hittable *h) { HittableList_append(&world->objects, h); }
static void add_light(World *world, Hittable *h) {
  HittableList_append(&world->objects, h);
  HittableList_append(&world->lights, h);
}

/* Replace the flat object list by one BVH over it (main.c:75-78). */
static void wrap_objects_in_bvh(World *world, PCG32 *rng) {
  Hittable *tree = BVHNode_new(&world->objects, rng);
  free(world->objects.items);
  HittableList_init(&world->objects, 1);
  HittableList_append(&world->objects, tree);
}

static void look(Camera *cam, float vfov, Vec3 from, Vec3 to, Vec3 background) {
  cam->vfov = vfov;
  cam->look_from = from;
  cam->look_to = to;
  cam->background = background;
}

static const Vec3 SKY = {{0.7f, 0.8f, 1.0f}};

/* 0 -- Book 1: two lambertian and two metal spheres in a plain list (main.c:9-30) */
static void scene_metal_and_lambertian(World *world, Camera *cam) {
  World_init(world, 4);
  add(world, Sphere_new(vec3(0, -100.5f, -1), 100, Lambertian_new(rgb(0.8f, 0.8f, 0.0f))));
  add(world, Sphere_new(vec3(0, 0, -1), 0.5f, Lambertian_new(rgb(0.7f, 0.3f, 0.3f))));
  add(world, Sphere_new(vec3(-1, 0, -1), 0.5f, Metal_new(rgb(0.8f, 0.8f, 0.8f), 0.3f)));
  add(world, Sphere_new(vec3(1, 0, -1), 0.5f, Metal_new(rgb(0.8f, 0.6f, 0.2f), 1.0f)));
  look(cam, 90.0f, VEC3_ZERO, vec3(0, 0, -1), SKY);
}

/* 1 -- Book 1 final: ground + 3 big + up to 484 small spheres, BVH, thin lens (main.c:32-85) */
static void scene_book1_final(World *world, Camera *cam) {
  World_init(world, 4 + 22 * 22);
  add(world, Sphere_new(vec3(0, -1000, -1), 1000, Lambertian_new(rgb(0.5f, 0.5f, 0.5f))));
  add(world, Sphere_new(vec3(0, 1, 0), 1, Dielectric_new(1.5f)));
  add(world, Sphere_new(vec3(-4, 1, 0), 1, Lambertian_new(rgb(0.4f, 0.2f, 0.1f))));
  add(world, Sphere_new(vec3(4, 1, 0), 1, Metal_new(rgb(0.7f, 0.6f, 0.5f), 0)));

  PCG32 rng;
  pcg32_seed(&rng, 19, 29);
  const Vec3 keep_clear = vec3(4, 0.2f, 0);
  const float r = 0.2f;

  for (int a = -11; a < 11; a++)
    for (int b = -11; b < 11; b++) {
      float pick = pcg32_f32(&rng);
      float jitter_z = pcg32_f32(&rng); /* z before x: right-to-left argument evaluation */
      float jitter_x = pcg32_f32(&rng);
      Vec3 c = vec3((float)a + 0.9f * jitter_x, r, (float)b + 0.9f * jitter_z);
      if (!(vec3_length(vec3_sub_vec3(c, keep_clear)) > 0.9f))
        continue;

      Material *m;
      if (pick < 0.8f) {
        Vec3 second = vec3_rand(&rng); /* right operand of the product is drawn first */
        Vec3 first = vec3_rand(&rng);
        m = Lambertian_new(Solid_new(vec3_mul_vec3(first, second)));
      } else if (pick < 0.95f) {
        Vec3 tint = vec3_rand_between(&rng, 0.5f, 1);
        Texture *t = Solid_new(tint);
        m = Metal_new(t, pcg32_f32(&rng) * 0.5f);
      } else {
        m = Dielectric_new(1.5f);
      }
      add(world, Sphere_new(c, r, m));
    }

  wrap_objects_in_bvh(world, &rng);
  look(cam, 20.0f, vec3(13, 2, 3), VEC3_ZERO, SKY);
  cam->dof_angle = 0.6f;
}

/* 2 -- Book 2: uv-checker spheres (main.c:87-99) */
static void scene_checker(World *world, Camera *cam) {
  World_init(world, 2);
  Material *m = Lambertian_new(Checker_new(0.01f, rgb(0.2f, 0.3f, 0.1f), rgb(0.9f, 0.9f, 0.9f)));
  add(world, Sphere_new(vec3(0, -10, 0), 10, m));
  add(world, Sphere_new(vec3(0, 10, 0), 10, m));
  look(cam, 20.0f, vec3(13, 2, 3), VEC3_ZERO, SKY);
}

/* 3 -- Book 2: image-textured globe (main.c:101-111) */
static void scene_earth(World *world, Camera *cam) {
  World_init(world, 1);
  add(world, Sphere_new(VEC3_ZERO, 2, Lambertian_new(Image_new("earthmap.jpg"))));
  look(cam, 20.0f, vec3(13, 2, 3), VEC3_ZERO, SKY);
}

/* 4 -- Book 2: marble Perlin spheres (main.c:113-129) */
static void scene_perlin(World *world, Camera *cam) {
  World_init(world, 2);
  PCG32 rng;
  pcg32_seed(&rng, 19, 29);
  Material *m = Lambertian_new(Perlin_new(4.0f, 7, &rng));
  add(world, Sphere_new(vec3(0, -1000, 0), 1000, m));
  add(world, Sphere_new(vec3(0, 2, 0), 2, m));
  look(cam, 20.0f, vec3(13, 2, 3), VEC3_ZERO, SKY);
  cam->dof_angle = 0.0f;
  cam->focal_length = 10.0f;
}

/* 5 -- Book 2: quad + sphere emitters, both importance-sampled (main.c:131-155) */
static void scene_simple_light(World *world, Camera *cam) {
  World_init(world, 4);
  PCG32 rng;
  pcg32_seed(&rng, 19, 29);
  Material *marble = Lambertian_new(Perlin_new(4.0f, 7, &rng));
  Material *glow = DiffuseLight_new(rgb(4, 4, 4));
  add(world, Sphere_new(vec3(0, -1000, 0), 1000, marble));
  add(world, Sphere_new(vec3(0, 2, 0), 2, marble));
  add_light(world, Quad_new(vec3(3, 1, -2), vec3(2, 0, 0), vec3(0, 2, 0), glow));
  add_light(world, Sphere_new(vec3(0, 7, 0), 2, glow));
  look(cam, 20.0f, vec3(26, 3, 6), vec3(0, 2, 0), VEC3_ZERO);
}

/* 6 -- Cornell box with the Book-3 mixture pdf (main.c:157-190) */
static void scene_cornell_box(World *world, Camera *cam) {
  World_init(world, 6 + 2);
  Material *red = Lambertian_new(rgb(0.65f, 0.05f, 0.05f));
  Material *white = Lambertian_new(rgb(0.73f, 0.73f, 0.73f));
  Material *green = Lambertian_new(rgb(0.12f, 0.45f, 0.15f));
  Material *lamp = DiffuseLight_new(rgb(15, 15, 15));

  add(world, Quad_new(vec3(555, 0, 0), vec3(0, 555, 0), vec3(0, 0, 555), green));
  add(world, Quad_new(vec3(0, 0, 0), vec3(0, 555, 0), vec3(0, 0, 555), red));
  add(world, Quad_new(vec3(0, 0, 0), vec3(555, 0, 0), vec3(0, 0, 555), white));
  add(world, Quad_new(vec3(555, 555, 555), vec3(-555, 0, 0), vec3(0, 0, -555), white));
  add(world, Quad_new(vec3(0, 0, 555), vec3(555, 0, 0), vec3(0, 555, 0), white));
  add_light(world, Quad_new(vec3(343, 554, 332), vec3(-130, 0, 0), vec3(0, 0, -105), lamp));

  Hittable *tall = Box_new(vec3(0, 0, 0), vec3(165, 330, 165), white);
  add(world, Translate_new(RotateY_new(tall, 15), vec3(265, 0, 295)));
  Hittable *cube = Box_new(vec3(0, 0, 0), vec3(165, 165, 165), white);
  add(world, Translate_new(RotateY_new(cube, -18), vec3(130, 0, 65)));

  cam->aspect_ratio = 1.0f;
  look(cam, 40.0f, vec3(278, 278, -800), vec3(278, 278, 0), VEC3_ZERO);
}

/* 7 -- Book 2 final: box floor BVH, light, glass, metal, two media, image, perlin, rotated sphere cloud
 * (main.c:192-273; the reference always passes enable_bvh = true) */
static void scene_book2_final(World *world, Camera *cam) {
  World_init(world, 11);
  PCG32 rng;
  pcg32_seed(&rng, 19, 29);

  const int side = 20;
  const float cell = 100.0f;
  Material *ground = Lambertian_new(rgb(0.48f, 0.83f, 0.53f));
  HittableList *floor_boxes = (HittableList *)HittableList_new((size_t)side * side);
  for (int i = 0; i < side; i++)
    for (int j = 0; j < side; j++) {
      Vec3 lo = vec3(-1000.0f + i * cell, 0.0f, -1000.0f + j * cell);
      float top = pcg32_f32_between(&rng, 1, 101);
      Vec3 hi = vec3(-1000.0f + (i + 1) * cell, top, -1000.0f + (j + 1) * cell);
      HittableList_append(floor_boxes, Box_new(lo, hi, ground));
    }
  Hittable *floor_tree = BVHNode_new(floor_boxes, &rng);
  free(floor_boxes->items);
  free(floor_boxes);
  add(world, floor_tree);

  add_light(world, Quad_new(vec3(123, 554, 147), vec3(300, 0, 0), vec3(0, 0, 265), DiffuseLight_new(rgb(7, 7, 7))));

  add(world, Sphere_new(vec3(400, 400, 200), 50, Lambertian_new(rgb(0.7f, 0.3f, 0.1f))));
  Material *glass = Dielectric_new(1.5f);
  add(world, Sphere_new(vec3(260, 150, 45), 50, glass));
  add(world, Sphere_new(vec3(0, 150, 145), 50, Metal_new(rgb(0.8f, 0.8f, 0.9f), 1.0f)));

  /* glass ball filled with blue smoke: the boundary is a list item AND the medium's boundary */
  Hittable *shell = Sphere_new(vec3(360, 150, 145), 70, glass);
  add(world, shell);
  add(world, ConstantMedium_new(shell, 0.2f, rgb(0.2f, 0.4f, 0.9f)));
  /* thin mist over everything; its boundary is not itself visible */
  add(world, ConstantMedium_new(Sphere_new(vec3(0, 0, 0), 5000, glass), 0.0001f, rgb(1, 1, 1)));

  add(world, Sphere_new(vec3(400, 200, 400), 100, Lambertian_new(Image_new("earthmap.jpg"))));
  add(world, Sphere_new(vec3(220, 280, 300), 80, Lambertian_new(Perlin_new(0.1f, 7, &rng))));

  const int n_cloud = 1000;
  Material *white = Lambertian_new(rgb(0.73f, 0.73f, 0.73f));
  HittableList *cloud = (HittableList *)HittableList_new(n_cloud);
  for (int i = 0; i < n_cloud; i++)
    HittableList_append(cloud, Sphere_new(vec3_rand_between(&rng, 0, 165), 10, white));
  Hittable *cloud_tree = BVHNode_new(cloud, &rng);
  free(cloud->items);
  free(cloud);
  add(world, Translate_new(RotateY_new(cloud_tree, 15.0f), vec3(-100, 270, 395)));

  cam->aspect_ratio = 1.0f;
  look(cam, 40.0f, vec3(478, 278, -600), vec3(278, 278, 0), VEC3_ZERO);
}

static const struct {
  const char *title; /* the reference's stderr banner, main.c:298-326 */
  void (*build)(World *, Camera *);
} SCENES[] = {
    {"Book 1: Metal and Lambertian", scene_metal_and_lambertian},
    {"Book 1: Final scene", scene_book1_final},
    {"Book 2: Checker", scene_checker},
    {"Book 2: Earth", scene_earth},
    {"Book 2: Perlin noise", scene_perlin},
    {"Book 2: Simple light", scene_simple_light},
    {"Book 2: Cornell box", scene_cornell_box},
    {"Book 2: Final scene", scene_book2_final},
};
enum { N_SCENES = sizeof(SCENES) / sizeof(SCENES[0]) };

const char *rtc_scene_title(int id) { return (id >= 0 && id < N_SCENES) ? SCENES[id].title : NULL; }

int rtc_scene_build(int id, World *world, Camera *cam) {
  if (id < 0 || id >= N_SCENES)
    return -1;
  SCENES[id].build(world, cam);
  return 0;
}
