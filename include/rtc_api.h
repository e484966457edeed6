/*
 * rtc_api.h -- the host-side object model of the B200 path-tracing backend.
 *
 * This single header declares, with the SAME names, field order and x86-64 layout, the public C surface of
 * gau-nernst/ray-tracing-c that sits above the render hot path, so that a scene builder written against the
 * reference (its src/main.c) compiles and links against this backend unchanged.  The thin headers
 * vec3.h / pcg32.h / hittable.h / material.h / texture.h / raytracing.h / tiff.h / utils.h next to this file
 * only include it.
 *
 * Reference interface each block replaces (paths are /root/reference/...):
 *   PCG32                      include/pcg32.h:7-16
 *   Vec3 + generic arithmetic  include/vec3.h:8-61
 *   Ray / AABB / Hittable*     include/hittable.h:12-113
 *   HitRecord / Material / ONB include/material.h:10-66
 *   Texture family             include/texture.h:5-51
 *   World / Camera             include/raytracing.h:9-41
 *   write_tiff                 include/tiff.h:7
 *   my_malloc, min/max/clamp   include/utils.h:8-31
 *
 * What is different behind the names: there is no CPU integrator.  The vtable slots of a Hittable and the
 * `value` slot of a Texture are identity tags (they point at functions that abort with a message), the
 * constructors only record geometry, and Camera_render() flattens the graph and runs sm_100a CUDA kernels
 * through the C ABI in rtc_b200.h.
 */
#ifndef RTC_API_H
#define RTC_API_H

#include <assert.h>
#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ utils (include/utils.h:8-31) ---------- */

#ifndef M_PI
#define M_PI 3.14159265358979323846264338327950288
#endif
#ifndef M_1_PI
#define M_1_PI 0.318309886183790671537767526745028724
#endif

#ifndef __cplusplus
#ifndef min
#define min(a, b) ((a) < (b) ? (a) : (b))
#endif
#ifndef max
#define max(a, b) ((a) > (b) ? (a) : (b))
#endif
#ifndef clamp
#define clamp(x, lo, hi) min(max(x, lo), hi)
#endif
#endif

void *my_malloc(size_t size); /* aborts (assert) on NULL, like the reference */

/* ------------------------------------------------------------------ PCG32 (include/pcg32.h) -------------- */

typedef struct PCG32 {
  uint64_t state;
  uint64_t inc;
} PCG32;

void pcg32_seed(PCG32 *rng, uint64_t initstate, uint64_t initseq);
uint32_t pcg32_u32(PCG32 *rng);
void pcg32_advance(PCG32 *rng, uint64_t delta); /* extension: skip `delta` draws in O(log delta) */
uint32_t pcg32_u32_between(PCG32 *rng, uint32_t lo, uint32_t hi); /* lo + u32 % (hi-lo), biased on purpose */
float pcg32_f32(PCG32 *rng);                                      /* top 24 bits / 2^24, in [0,1) */
float pcg32_f32_between(PCG32 *rng, float lo, float hi);

/* ------------------------------------------------------------------ Vec3 (include/vec3.h) ---------------- */

typedef union Vec3 {
  struct {
    float x;
    float y;
    float z;
  };
  float values[3];
} Vec3;

extern const Vec3 VEC3_ZERO;

Vec3 vec3(float x, float y, float z);
Vec3 *Vec3_new(float x, float y, float z);
Vec3 vec3_neg(Vec3 a);
Vec3 vec3_inv(Vec3 a);

Vec3 vec3_add_vec3(Vec3 a, Vec3 b);
Vec3 vec3_mul_vec3(Vec3 a, Vec3 b);
Vec3 vec3_sub_vec3(Vec3 a, Vec3 b); /* a + (-b) */
Vec3 vec3_div_vec3(Vec3 a, Vec3 b); /* a * (1/b): reciprocal first, as the reference does */

Vec3 vec3_add_float(Vec3 a, float s);
Vec3 vec3_mul_float(Vec3 a, float s);
Vec3 vec3_sub_float(Vec3 a, float s);
Vec3 vec3_div_float(Vec3 a, float s); /* a * (1/s) */

Vec3 vec3_lerp(Vec3 a, Vec3 b, float w);
Vec3 vec3_min(Vec3 a, Vec3 b);
Vec3 vec3_max(Vec3 a, Vec3 b);

float vec3_length2(Vec3 a);
float vec3_length(Vec3 a);
float vec3_dot(Vec3 a, Vec3 b);
Vec3 vec3_cross(Vec3 a, Vec3 b);
Vec3 vec3_normalize(Vec3 a);
bool vec3_near_zero(Vec3 a);

Vec3 vec3_rand(PCG32 *rng);                                /* draw order z, y, x (gcc evaluates right-to-left) */
Vec3 vec3_rand_between(PCG32 *rng, float lo, float hi);    /* same order */
Vec3 vec3_rand_unit_vector(PCG32 *rng);                    /* cube rejection, then normalise */
Vec3 vec3_rand_hemisphere(Vec3 normal, PCG32 *rng);

#ifndef __cplusplus
/* Type-generic, variadic (2..4 operands, folded left) spellings used by the reference's scene code. */
#define RTC_V3_PICK(op, rhs) _Generic((rhs), Vec3: vec3_##op##_vec3, default: vec3_##op##_float)
#define RTC_V3_2(op, a, b) RTC_V3_PICK(op, b)(a, b)
#define RTC_V3_3(op, a, b, c) RTC_V3_2(op, RTC_V3_2(op, a, b), c)
#define RTC_V3_4(op, a, b, c, d) RTC_V3_2(op, RTC_V3_3(op, a, b, c), d)
#define RTC_V3_NTH(_1, _2, _3, _4, _5, name, ...) name
#define RTC_V3_FOLD(op, ...) RTC_V3_NTH(op, __VA_ARGS__, RTC_V3_4, RTC_V3_3, RTC_V3_2, )(op, __VA_ARGS__)
#define vec3_add(...) RTC_V3_FOLD(add, __VA_ARGS__)
#define vec3_mul(...) RTC_V3_FOLD(mul, __VA_ARGS__)
#define vec3_sub(a, b) RTC_V3_2(sub, a, b)
#define vec3_div(a, b) RTC_V3_2(div, a, b)
#endif

/* ------------------------------------------------------------------ geometry (include/hittable.h) -------- */

typedef struct Texture Texture;
typedef struct Material Material;
typedef struct HitRecord HitRecord;

typedef struct Ray {
  Vec3 origin;
  Vec3 direction; /* not normalised anywhere on the path */
} Ray;

Vec3 ray_at(const Ray *ray, float t);

typedef union AABB {
  struct {
    float x[2];
    float y[2];
    float z[2];
  };
  float values[3][2]; /* [axis][lo,hi] */
} AABB;

typedef struct Hittable Hittable;
typedef struct HittableVTable {
  bool (*hit)(const Hittable *self, const Ray *ray, float t_min, float t_max, HitRecord *rec, PCG32 *rng);
  float (*pdf)(const Hittable *self, const Ray *ray, PCG32 *rng);
  Vec3 (*rand)(const Hittable *self, Vec3 origin, PCG32 *rng);
} HittableVTable;

struct Hittable {
  HittableVTable *vtable; /* kind tag: compared by address by the flattener */
  AABB bbox;
};

typedef struct HittableList {
  Hittable hittable;
  size_t max_size;
  size_t size;
  Hittable **items;
} HittableList;

void HittableList_init(HittableList *self, size_t max_size);
Hittable *HittableList_new(size_t max_size);
void HittableList_append(HittableList *self, Hittable *item);

typedef struct Sphere {
  Hittable hittable;
  Vec3 center;
  float radius;
  Material *material;
} Sphere;

void Sphere_init(Sphere *self, Vec3 center, float radius, Material *mat);
Hittable *Sphere_new(Vec3 center, float radius, Material *mat);

typedef struct Quad {
  Hittable hittable;
  Vec3 Q;
  Vec3 u;
  Vec3 v;
  Vec3 normal; /* unit */
  float D;     /* normal . Q */
  Vec3 w;      /* (u x v) / |u x v|^2 */
  Material *material;
  float area;
} Quad;

void Quad_init(Quad *self, Vec3 Q, Vec3 u, Vec3 v, Material *mat);
Hittable *Quad_new(Vec3 Q, Vec3 u, Vec3 v, Material *mat);
Hittable *Box_new(Vec3 a, Vec3 b, Material *mat); /* HittableList of 6 quads */

typedef struct BVHNode {
  Hittable hittable;
  Hittable *left;
  Hittable *right;
} BVHNode;

void BVHNode_init(BVHNode *self, const HittableList *list, PCG32 *rng);
Hittable *BVHNode_new(const HittableList *list, PCG32 *rng);

typedef struct Translate {
  Hittable hittable;
  Hittable *object;
  Vec3 offset;
} Translate;

void Translate_init(Translate *self, Hittable *object, Vec3 offset);
Hittable *Translate_new(Hittable *object, Vec3 offset);

typedef struct RotateY {
  Hittable hittable;
  Hittable *object;
  float sin_theta;
  float cos_theta;
} RotateY;

void RotateY_init(RotateY *self, Hittable *object, float angle_degrees);
Hittable *RotateY_new(Hittable *object, float angle_degrees);

typedef struct ConstantMedium {
  Hittable hittable;
  Hittable *boundary;
  float neg_inv_density;
  Material *phase_fn;
} ConstantMedium;

void ConstantMedium_init(ConstantMedium *self, Hittable *boundary, float density, Texture *albedo);
Hittable *ConstantMedium_new(Hittable *boundary, float density, Texture *albedo);

/* ------------------------------------------------------------------ shading (include/material.h) --------- */

struct HitRecord {
  Vec3 p;
  Vec3 normal;
  Material *material;
  float t;
  float u;
  float v;
  bool front_face;
};

typedef enum MaterialType {
  SURFACE_NORMAL,
  LAMBERTIAN,
  METAL,
  DIELECTRIC,
  DIFFUSE_LIGHT,
  ISOTROPIC,
} MaterialType;

struct Material {
  MaterialType tag;
  Texture *albedo;
  union {
    float fuzz; /* METAL */
    float eta;  /* DIELECTRIC */
  };
};

void SurfaceNormal_init(Material *self);
Material *SurfaceNormal_new(void);
void Lambertian_init(Material *self, Texture *albedo);
Material *Lambertian_new(Texture *albedo);
void Metal_init(Material *self, Texture *albedo, float fuzz);
Material *Metal_new(Texture *albedo, float fuzz);
void Dielectric_init(Material *self, float eta);
Material *Dielectric_new(float eta);
void DiffuseLight_init(Material *self, Texture *albedo);
Material *DiffuseLight_new(Texture *albedo);
void Isotropic_init(Material *self, Texture *albedo);
Material *Isotropic_new(Texture *albedo);

typedef struct ONB {
  Vec3 u;
  Vec3 v;
  Vec3 w;
} ONB;
Vec3 ONB_local(const ONB *self, Vec3 a);
void ONB_from_w(ONB *self, Vec3 w);

/* ------------------------------------------------------------------ textures (include/texture.h) --------- */

struct Texture {
  Vec3 (*value)(const Texture *self, float u, float v, Vec3 p); /* kind tag, compared by address */
};

typedef struct Solid {
  Texture texture;
  Vec3 color;
} Solid;
void Solid_init(Solid *self, Vec3 color);
Texture *Solid_new(Vec3 color);

typedef struct Checker {
  Texture texture;
  float scale;
  Texture *even;
  Texture *odd;
} Checker;
void Checker_init(Checker *self, float scale, Texture *even, Texture *odd);
Texture *Checker_new(float scale, Texture *even, Texture *odd);

typedef struct Image {
  Texture texture;
  int width;
  int height;
  uint8_t *buffer; /* RGB8, row-major, top row first */
} Image;
void Image_init(Image *self, char *filename);
Texture *Image_new(char *filename);

#define N_PERLIN 256
typedef struct Perlin {
  Texture texture;
  float scale;
  int depth;
  Vec3 grad_field[N_PERLIN];
  int perm_x[N_PERLIN];
  int perm_y[N_PERLIN];
  int perm_z[N_PERLIN];
} Perlin;
void Perlin_init(Perlin *self, float scale, int depth, PCG32 *rng);
Texture *Perlin_new(float scale, int depth, PCG32 *rng);

/* ------------------------------------------------------------------ camera (include/raytracing.h) -------- */

typedef struct World {
  HittableList objects;
  HittableList lights; /* aliases of objects that are also in `objects` */
} World;

void World_init(World *world, size_t max_objects);

typedef struct Camera {
  float aspect_ratio;
  int img_width;
  int img_height;
  int samples_per_pixel;
  int max_depth;
  Vec3 background;
  float vfov;
  Vec3 look_from;
  Vec3 look_to;
  Vec3 vup;
  float dof_angle;
  float focal_length;
  Vec3 pixel00_loc;
  Vec3 pixel_delta_u;
  Vec3 pixel_delta_v;
  Vec3 u;
  Vec3 v;
  Vec3 w;
  Vec3 dof_disc_u;
  Vec3 dof_disc_v;
  float lights_sampling_prob;
} Camera;

void Camera_init(Camera *camera);
/* THE drop-in boundary (src/raytracing.c:86).  Fills img_height*img_width*3 bytes, RGB, top row first. */
void Camera_render(const Camera *camera, const World *world, uint8_t *buffer);

/* ------------------------------------------------------------------ output (include/tiff.h) -------------- */

int write_tiff(FILE *f, int width, int height, int n_channels, uint8_t *buffer);

#ifdef __cplusplus
}
#endif
#endif /* RTC_API_H */
