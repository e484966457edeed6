/*
 * geometry.c -- host constructors of the Hittable family (rtc_api.h), i.e. everything a scene builder calls
 * before Camera_render().  Nothing here intersects rays: the vtable slots are identity tags whose functions abort
 * (the render loop and the hit tests live in csrc/device/ and are reached through rtc_b200.h).
 *
 * The bounding boxes, quad plane constants and the BVH topology are computed with the reference's exact
 * arithmetic and draw order so that the flattened scene is bit-identical to the one the reference would build:
 *   AABB helpers        /root/reference/src/hittable.c:9-37
 *   HittableList        /root/reference/src/hittable.c:109-118
 *   Sphere / Quad / Box /root/reference/src/hittable.c:180-184, 230-264
 *   BVH build           /root/reference/src/hittable.c:57-72, 279-323  (axis drawn at EVERY node, libc qsort, n/2 split)
 *   Translate / RotateY /root/reference/src/hittable.c:336-345, 369-390
 *   ConstantMedium      /root/reference/src/hittable.c:425-434
 */
#include "rtc_api.h"
#include <math.h>

/* ---- identity-tag vtables ---- */

static void no_host_path(const char *what) {
  fprintf(stderr,
          "rtc_b200: %s was called on the host. This backend has no CPU integrator; rays are traced on the GPU "
          "through Camera_render() / rt_hit_batch() (include/rtc_b200.h).\n",
          what);
  abort();
}

#define RTC_TAG_FUNCS(Kind)                                                                                            \
  static bool Kind##_hit_tag(const Hittable *s, const Ray *r, float a, float b, HitRecord *rec, PCG32 *g) {            \
    (void)s, (void)r, (void)a, (void)b, (void)rec, (void)g;                                                            \
    no_host_path(#Kind ".hit");                                                                                        \
    return false;                                                                                                      \
  }                                                                                                                    \
  static float Kind##_pdf_tag(const Hittable *s, const Ray *r, PCG32 *g) {                                             \
    (void)s, (void)r, (void)g;                                                                                         \
    no_host_path(#Kind ".pdf");                                                                                        \
    return 0.0f;                                                                                                       \
  }                                                                                                                    \
  static Vec3 Kind##_rand_tag(const Hittable *s, Vec3 o, PCG32 *g) {                                                   \
    (void)s, (void)g;                                                                                                  \
    no_host_path(#Kind ".rand");                                                                                       \
    return o;                                                                                                          \
  }

RTC_TAG_FUNCS(HittableList)
RTC_TAG_FUNCS(Sphere)
RTC_TAG_FUNCS(Quad)

#define RTC_TAG_HIT_ONLY(Kind)                                                                                         \
  static bool Kind##_hit_tag(const Hittable *s, const Ray *r, float a, float b, HitRecord *rec, PCG32 *g) {            \
    (void)s, (void)r, (void)a, (void)b, (void)rec, (void)g;                                                            \
    no_host_path(#Kind ".hit");                                                                                        \
    return false;                                                                                                      \
  }

RTC_TAG_HIT_ONLY(BVHNode)
RTC_TAG_HIT_ONLY(Translate)
RTC_TAG_HIT_ONLY(RotateY)
RTC_TAG_HIT_ONLY(ConstantMedium)

/* pdf/rand are NULL exactly where the reference leaves them NULL (hittable.c:278,335,368,424) */
static HittableVTable LIST_KIND = {HittableList_hit_tag, HittableList_pdf_tag, HittableList_rand_tag};
static HittableVTable SPHERE_KIND = {Sphere_hit_tag, Sphere_pdf_tag, Sphere_rand_tag};
static HittableVTable QUAD_KIND = {Quad_hit_tag, Quad_pdf_tag, Quad_rand_tag};
static HittableVTable BVH_KIND = {BVHNode_hit_tag, NULL, NULL};
static HittableVTable TRANSLATE_KIND = {Translate_hit_tag, NULL, NULL};
static HittableVTable ROTATE_Y_KIND = {RotateY_hit_tag, NULL, NULL};
static HittableVTable MEDIUM_KIND = {ConstantMedium_hit_tag, NULL, NULL};

/* ---- boxes ---- */

static AABB box_empty(void) {
  AABB b;
  for (int a = 0; a < 3; a++) {
    b.values[a][0] = INFINITY;
    b.values[a][1] = -INFINITY;
  }
  return b;
}
static AABB box_of_points(Vec3 p, Vec3 q) {
  AABB b;
  for (int a = 0; a < 3; a++) {
    b.values[a][0] = fminf(p.values[a], q.values[a]);
    b.values[a][1] = fmaxf(p.values[a], q.values[a]);
  }
  return b;
}
static AABB box_union(AABB p, AABB q) {
  AABB b;
  for (int a = 0; a < 3; a++) {
    b.values[a][0] = fminf(p.values[a][0], q.values[a][0]);
    b.values[a][1] = fmaxf(p.values[a][1], q.values[a][1]);
  }
  return b;
}
/* flat quads get +-1e-4 on their thin axis (hittable.c:24-37) */
static AABB box_thicken(AABB in) {
  const float delta = 1e-4f;
  AABB b = in;
  for (int a = 0; a < 3; a++)
    if (in.values[a][1] - in.values[a][0] < delta) {
      b.values[a][0] = in.values[a][0] - delta;
      b.values[a][1] = in.values[a][1] + delta;
    }
  return b;
}

/* ---- HittableList ---- */

void HittableList_init(HittableList *self, size_t max_size) {
  self->hittable.vtable = &LIST_KIND;
  self->hittable.bbox = box_empty();
  self->max_size = max_size;
  self->size = 0;
  self->items = my_malloc(sizeof(Hittable *) * max_size);
}
Hittable *HittableList_new(size_t max_size) {
  HittableList *l = my_malloc(sizeof(HittableList));
  HittableList_init(l, max_size);
  return &l->hittable;
}
void HittableList_append(HittableList *self, Hittable *item) {
  assert((self->size < self->max_size) && "List is full");
  self->items[self->size] = item;
  self->size += 1;
  self->hittable.bbox = box_union(self->hittable.bbox, item->bbox);
}

/* ---- Sphere ---- */

void Sphere_init(Sphere *self, Vec3 center, float radius, Material *mat) {
  self->hittable.vtable = &SPHERE_KIND;
  self->hittable.bbox = box_of_points(vec3_sub_float(center, radius), vec3_add_float(center, radius));
  self->center = center;
  self->radius = radius;
  self->material = mat;
}
Hittable *Sphere_new(Vec3 center, float radius, Material *mat) {
  Sphere *s = my_malloc(sizeof(Sphere));
  Sphere_init(s, center, radius, mat);
  return &s->hittable;
}

/* ---- Quad / Box ---- */

void Quad_init(Quad *self, Vec3 Q, Vec3 u, Vec3 v, Material *mat) {
  Vec3 far_corner = vec3_add_vec3(vec3_add_vec3(Q, u), v);
  Vec3 n = vec3_cross(u, v);
  self->hittable.vtable = &QUAD_KIND;
  self->hittable.bbox = box_thicken(box_of_points(Q, far_corner));
  self->Q = Q;
  self->u = u;
  self->v = v;
  self->normal = vec3_normalize(n);
  self->D = vec3_dot(self->normal, Q);
  self->w = vec3_div_float(n, vec3_length2(n));
  self->material = mat;
  self->area = vec3_length(n);
}
Hittable *Quad_new(Vec3 Q, Vec3 u, Vec3 v, Material *mat) {
  Quad *q = my_malloc(sizeof(Quad));
  Quad_init(q, Q, u, v, mat);
  return &q->hittable;
}

Hittable *Box_new(Vec3 a, Vec3 b, Material *mat) {
  Vec3 lo = vec3_min(a, b), hi = vec3_max(a, b);
  Vec3 ex = vec3(hi.x - lo.x, 0, 0), ey = vec3(0, hi.y - lo.y, 0), ez = vec3(0, 0, hi.z - lo.z);
  /* face order = traversal order = tie-break order of the reference: front, right, back, left, top, bottom */
  struct {
    Vec3 corner, e0, e1;
  } face[6] = {
      {vec3(lo.x, lo.y, hi.z), ex, ey},           {vec3(hi.x, lo.y, hi.z), vec3_neg(ez), ey},
      {vec3(hi.x, lo.y, lo.z), vec3_neg(ex), ey}, {vec3(lo.x, lo.y, lo.z), ez, ey},
      {vec3(lo.x, hi.y, hi.z), ex, vec3_neg(ez)}, {vec3(lo.x, lo.y, lo.z), ex, ez},
  };
  HittableList *sides = (HittableList *)HittableList_new(6);
  for (int i = 0; i < 6; i++)
    HittableList_append(sides, Quad_new(face[i].corner, face[i].e0, face[i].e1, mat));
  return &sides->hittable;
}

/* ---- BVH (host build; the GPU consumes the flattened tree, and separately rebuilds a SAH tree) ---- */

static int g_sort_axis; /* qsort has no context argument; the build is single-threaded like the reference's */

static int order_on_axis(const Hittable *p, const Hittable *q, int axis) {
  float lo_p = p->bbox.values[axis][0], lo_q = q->bbox.values[axis][0];
  return lo_p < lo_q ? -1 : (lo_p > lo_q ? 1 : 0);
}
static int qsort_on_axis(const void *p, const void *q) {
  return order_on_axis(*(Hittable *const *)p, *(Hittable *const *)q, g_sort_axis);
}

static void bvh_build(BVHNode *node, Hittable **src, size_t n, PCG32 *rng) {
  node->hittable.vtable = &BVH_KIND;

  Hittable **work = my_malloc(sizeof(Hittable *) * n); /* private copy: the sort must not disturb the caller */
  for (size_t i = 0; i < n; i++)
    work[i] = src[i];

  int axis = (int)pcg32_u32_between(rng, 0, 3); /* consumed even when n is 1 or 2 */

  if (n == 1) {
    node->left = node->right = work[0];
  } else if (n == 2) {
    bool keep = order_on_axis(work[0], work[1], axis) < 0; /* ties put work[1] on the left */
    node->left = keep ? work[0] : work[1];
    node->right = keep ? work[1] : work[0];
  } else {
    g_sort_axis = axis;
    qsort(work, n, sizeof(work[0]), qsort_on_axis);
    size_t half = n / 2;
    BVHNode *lo = my_malloc(sizeof(BVHNode));
    bvh_build(lo, work, half, rng);
    BVHNode *hi = my_malloc(sizeof(BVHNode));
    bvh_build(hi, work + half, n - half, rng);
    node->left = &lo->hittable;
    node->right = &hi->hittable;
  }
  free(work);
  node->hittable.bbox = box_union(node->left->bbox, node->right->bbox);
}

void BVHNode_init(BVHNode *self, const HittableList *list, PCG32 *rng) {
  bvh_build(self, list->items, list->size, rng);
}
Hittable *BVHNode_new(const HittableList *list, PCG32 *rng) {
  BVHNode *n = my_malloc(sizeof(BVHNode));
  BVHNode_init(n, list, rng);
  return &n->hittable;
}

/* ---- instances ---- */

void Translate_init(Translate *self, Hittable *object, Vec3 offset) {
  AABB b = object->bbox;
  for (int a = 0; a < 3; a++) {
    b.values[a][0] += offset.values[a];
    b.values[a][1] += offset.values[a];
  }
  self->hittable.vtable = &TRANSLATE_KIND;
  self->hittable.bbox = b;
  self->object = object;
  self->offset = offset;
}
Hittable *Translate_new(Hittable *object, Vec3 offset) {
  Translate *t = my_malloc(sizeof(Translate));
  Translate_init(t, object, offset);
  return &t->hittable;
}

void RotateY_init(RotateY *self, Hittable *object, float angle_degrees) {
  float radians = angle_degrees * (float)M_PI / 180.f;
  float s = sinf(radians), c = cosf(radians);
  AABB in = object->bbox;
  Vec3 lo = vec3(INFINITY, INFINITY, INFINITY), hi = vec3(-INFINITY, -INFINITY, -INFINITY);
  for (int i = 0; i < 2; i++)
    for (int j = 0; j < 2; j++)
      for (int k = 0; k < 2; k++) {
        float x = (float)i * in.x[1] + (float)(1 - i) * in.x[0];
        float y = (float)j * in.y[1] + (float)(1 - j) * in.y[0];
        float z = (float)k * in.z[1] + (float)(1 - k) * in.z[0];
        Vec3 corner = vec3(c * x + s * z, y, -s * x + c * z); /* object -> world = inverse of the ray rotation */
        lo = vec3_min(lo, corner);
        hi = vec3_max(hi, corner);
      }
  self->hittable.vtable = &ROTATE_Y_KIND;
  self->hittable.bbox = box_of_points(lo, hi);
  self->object = object;
  self->sin_theta = s;
  self->cos_theta = c;
}
Hittable *RotateY_new(Hittable *object, float angle_degrees) {
  RotateY *r = my_malloc(sizeof(RotateY));
  RotateY_init(r, object, angle_degrees);
  return &r->hittable;
}

/* ---- participating medium ---- */

void ConstantMedium_init(ConstantMedium *self, Hittable *boundary, float density, Texture *albedo) {
  self->hittable.vtable = &MEDIUM_KIND;
  self->hittable.bbox = boundary->bbox;
  self->boundary = boundary;
  self->neg_inv_density = -1.0f / density;
  self->phase_fn = Isotropic_new(albedo);
}
Hittable *ConstantMedium_new(Hittable *boundary, float density, Texture *albedo) {
  ConstantMedium *m = my_malloc(sizeof(ConstantMedium));
  ConstantMedium_init(m, boundary, density, albedo);
  return &m->hittable;
}
