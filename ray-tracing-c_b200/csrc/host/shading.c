/*
 * shading.c -- host constructors for materials and textures (rtc_api.h).  Evaluation (scatter, emit, texture
 * lookup, Perlin noise) happens on the GPU; here we only record parameters and build the tables the device reads.
 *
 *   material constructors   /root/reference/src/material.c:19-21, 42-43, 59-60, 87-91, 99-101
 *   ONB                     /root/reference/src/material.c:144-152
 *   Solid / Checker         /root/reference/src/texture.c:8-26
 *   Image                   /root/reference/src/texture.c:38-43   (stb_image there; here a binary-PPM reader, and a
 *                           procedural stand-in when the asset is absent -- the repo ships no earthmap.jpg)
 *   Perlin tables           /root/reference/src/texture.c:53-74   (256 unit gradients, then 3 Fisher-Yates shuffles)
 */
#include "rtc_api.h"
#include "rtc_host.h"
#include <math.h>
#include <string.h>

/* ---- materials ---- */

static Material *material_make(MaterialType tag, Texture *albedo, float param) {
  Material *m = my_malloc(sizeof(Material));
  memset(m, 0, sizeof(*m));
  m->tag = tag;
  m->albedo = albedo;
  m->fuzz = param; /* union with eta */
  return m;
}

void SurfaceNormal_init(Material *self) { self->tag = SURFACE_NORMAL; }
Material *SurfaceNormal_new(void) { return material_make(SURFACE_NORMAL, NULL, 0.0f); }

void Lambertian_init(Material *self, Texture *albedo) {
  self->tag = LAMBERTIAN;
  self->albedo = albedo;
  self->fuzz = 0.0f;
}
Material *Lambertian_new(Texture *albedo) { return material_make(LAMBERTIAN, albedo, 0.0f); }

void Metal_init(Material *self, Texture *albedo, float fuzz) {
  self->tag = METAL;
  self->albedo = albedo;
  self->fuzz = fuzz;
}
Material *Metal_new(Texture *albedo, float fuzz) { return material_make(METAL, albedo, fuzz); }

void Dielectric_init(Material *self, float eta) {
  self->tag = DIELECTRIC;
  self->albedo = NULL;
  self->eta = eta;
}
Material *Dielectric_new(float eta) { return material_make(DIELECTRIC, NULL, eta); }

void DiffuseLight_init(Material *self, Texture *albedo) {
  self->tag = DIFFUSE_LIGHT;
  self->albedo = albedo;
  self->fuzz = 0.0f;
}
Material *DiffuseLight_new(Texture *albedo) { return material_make(DIFFUSE_LIGHT, albedo, 0.0f); }

void Isotropic_init(Material *self, Texture *albedo) {
  self->tag = ISOTROPIC;
  self->albedo = albedo;
  self->fuzz = 0.0f;
}
Material *Isotropic_new(Texture *albedo) { return material_make(ISOTROPIC, albedo, 0.0f); }

/* ---- orthonormal basis (pure math, kept for API completeness) ---- */

void ONB_from_w(ONB *self, Vec3 w) {
  self->w = vec3_normalize(w);
  Vec3 helper = fabsf(self->w.x) > 0.9f ? vec3(0, 1, 0) : vec3(1, 0, 0);
  self->v = vec3_normalize(vec3_cross(self->w, helper));
  self->u = vec3_cross(self->w, self->v);
}
Vec3 ONB_local(const ONB *self, Vec3 a) {
  Vec3 r = vec3_add_vec3(vec3_mul_float(self->u, a.x), vec3_mul_float(self->v, a.y));
  return vec3_add_vec3(r, vec3_mul_float(self->w, a.z));
}

/* ---- texture identity tags ---- */

static void no_host_texture(const char *what) {
  fprintf(stderr, "rtc_b200: %s was called on the host; textures are evaluated on the GPU.\n", what);
  abort();
}
static Vec3 Solid_value_tag(const Texture *t, float u, float v, Vec3 p) {
  (void)t, (void)u, (void)v;
  no_host_texture("Solid.value");
  return p;
}
static Vec3 Checker_value_tag(const Texture *t, float u, float v, Vec3 p) {
  (void)t, (void)u, (void)v;
  no_host_texture("Checker.value");
  return p;
}
static Vec3 Image_value_tag(const Texture *t, float u, float v, Vec3 p) {
  (void)t, (void)u, (void)v;
  no_host_texture("Image.value");
  return p;
}
static Vec3 Perlin_value_tag(const Texture *t, float u, float v, Vec3 p) {
  (void)t, (void)u, (void)v;
  no_host_texture("Perlin.value");
  return p;
}

void Solid_init(Solid *self, Vec3 color) {
  self->texture.value = Solid_value_tag;
  self->color = color;
}
Texture *Solid_new(Vec3 color) {
  Solid *s = my_malloc(sizeof(Solid));
  Solid_init(s, color);
  return &s->texture;
}

void Checker_init(Checker *self, float scale, Texture *even, Texture *odd) {
  self->texture.value = Checker_value_tag;
  self->scale = scale;
  self->even = even;
  self->odd = odd;
}
Texture *Checker_new(float scale, Texture *even, Texture *odd) {
  Checker *c = my_malloc(sizeof(Checker));
  Checker_init(c, scale, even, odd);
  return &c->texture;
}

/* ---- image texture ---- */

void Image_init(Image *self, char *filename) {
  /* texture.c:38-42: stbi_load(filename, &w, &h, NULL, 3) and an assert.  Here: PNG / baseline JPEG / binary PPM
   * decoded by image_decode.c (same bytes as stb); a file that EXISTS but cannot be decoded stops the program like the
   * reference's assert does; only when there is no such file does the synthetic procedural image stand in (the asset
   * is not part of the reference repository; BASELINE.json: "synthetic procedural image if the asset is absent"). */
  self->texture.value = Image_value_tag;
  int exists = 0;
  self->buffer = rtc_load_image(filename, &self->width, &self->height, &exists);
  if (self->buffer == NULL && exists) {
    fprintf(stderr, "rtc_b200: cannot decode image '%s': %s\n", filename, rtc_image_error());
    assert((self->buffer != NULL) && "Failed to load image");
    abort(); /* (also under NDEBUG) */
  }
  if (self->buffer == NULL) {
    fprintf(stderr, "rtc_b200: image '%s' is absent; using the synthetic 1024x512 image\n", filename);
    self->width = 1024;
    self->height = 512;
    self->buffer = my_malloc((size_t)self->width * self->height * 3);
    rtc_synthetic_image(self->width, self->height, self->buffer);
  }
}
Texture *Image_new(char *filename) {
  Image *im = my_malloc(sizeof(Image));
  Image_init(im, filename);
  return &im->texture;
}

/* ---- Perlin tables ---- */

static void shuffle_identity(int perm[N_PERLIN], PCG32 *rng) {
  for (int i = 0; i < N_PERLIN; i++)
    perm[i] = i;
  for (int i = N_PERLIN - 1; i > 0; i--) {
    uint32_t j = pcg32_u32_between(rng, 0, (uint32_t)i + 1);
    int t = perm[i];
    perm[i] = perm[j];
    perm[j] = t;
  }
}

void Perlin_init(Perlin *self, float scale, int depth, PCG32 *rng) {
  self->texture.value = Perlin_value_tag;
  self->scale = scale;
  self->depth = depth;
  for (int i = 0; i < N_PERLIN; i++)
    self->grad_field[i] = vec3_rand_unit_vector(rng);
  shuffle_identity(self->perm_x, rng);
  shuffle_identity(self->perm_y, rng);
  shuffle_identity(self->perm_z, rng);
}
Texture *Perlin_new(float scale, int depth, PCG32 *rng) {
  Perlin *p = my_malloc(sizeof(Perlin));
  Perlin_init(p, scale, depth, rng);
  return &p->texture;
}
