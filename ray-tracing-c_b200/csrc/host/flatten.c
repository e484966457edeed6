/*
 * flatten.c -- Hittable graph -> rt_scene_desc (the host half of the drop-in boundary).
 *
 * The reference's scene is a pointer graph of heap objects whose dynamic type is only observable through the
 * address of a `static` vtable (hittable.c:108,179,229,278,335,368,424) or texture value function
 * (texture.c:8,12,28,47).  We learn those addresses once by building one probe object per kind with the public
 * constructors that are linked into the running program, then walk world.objects / world.lights depth-first and
 * emit index-linked arrays the GPU can read with no pointer chasing on the host side.
 *
 * Node ids are DFS pre-order from world.objects (id 0).  A pointer that is reached twice (lights alias objects,
 * a medium's boundary is also a list item, BVH n==1 has left==right) keeps its first id.
 */
#include "rtc_host.h"
#include <math.h>
#include <string.h>
#include <unistd.h>

static char g_error[256];
const char *rtc_flatten_error(void) { return g_error; }

/* ------------------------------------------------------------------ kind discovery ----------------------- */

typedef struct KindTable {
  bool ready;
  const HittableVTable *list, *sphere, *quad, *bvh, *translate, *rotate_y, *medium;
  Vec3 (*solid)(const Texture *, float, float, Vec3);
  Vec3 (*checker)(const Texture *, float, float, Vec3);
  Vec3 (*image)(const Texture *, float, float, Vec3);
  Vec3 (*perlin)(const Texture *, float, float, Vec3);
} KindTable;

static KindTable g_kinds;

static void learn_kinds(void) {
  if (g_kinds.ready)
    return;
  PCG32 rng;
  pcg32_seed(&rng, 1, 1);

  Texture *solid = Solid_new(vec3(0, 0, 0));
  Material *mat = Lambertian_new(solid);
  Hittable *sphere = Sphere_new(vec3(0, 0, 0), 1.0f, mat);
  Hittable *quad = Quad_new(vec3(0, 0, 0), vec3(1, 0, 0), vec3(0, 1, 0), mat);
  HittableList *list = (HittableList *)HittableList_new(1);
  HittableList_append(list, sphere);

  g_kinds.list = list->hittable.vtable;
  g_kinds.sphere = sphere->vtable;
  g_kinds.quad = quad->vtable;
  g_kinds.bvh = BVHNode_new(list, &rng)->vtable;
  g_kinds.translate = Translate_new(sphere, vec3(0, 0, 0))->vtable;
  g_kinds.rotate_y = RotateY_new(sphere, 0.0f)->vtable;
  g_kinds.medium = ConstantMedium_new(sphere, 1.0f, solid)->vtable;

  g_kinds.solid = solid->value;
  g_kinds.checker = Checker_new(1.0f, solid, solid)->value;
  g_kinds.perlin = Perlin_new(1.0f, 1, &rng)->value;

  /* An Image can only be constructed from a file: write a 1x1 binary PPM, load it, remove it. */
  char path[] = "/tmp/rtc_probe_XXXXXX";
  int fd = mkstemp(path);
  if (fd >= 0) {
    static const char ppm[] = "P6\n1 1\n255\n\x10\x20\x30";
    if (write(fd, ppm, sizeof(ppm) - 1) == (ssize_t)(sizeof(ppm) - 1))
      g_kinds.image = Image_new(path)->value;
    close(fd);
    unlink(path);
  }
  g_kinds.ready = true; /* the probe objects are leaked on purpose: a few hundred bytes, once per process */
}

/* ------------------------------------------------------------------ growable arrays + pointer map -------- */

typedef struct Vec {
  char *data;
  size_t len, cap, item;
} Vec;

static void *vec_push(Vec *v) {
  if (v->len == v->cap) {
    v->cap = v->cap ? v->cap * 2 : 64;
    v->data = realloc(v->data, v->cap * v->item);
    assert(v->data != NULL);
  }
  void *slot = v->data + v->len * v->item;
  memset(slot, 0, v->item);
  v->len++;
  return slot;
}
#define VEC_AT(v, T, i) (((T *)(v).data)[i])

typedef struct PtrMap { /* open addressing, pointer -> u32 */
  const void **keys;
  uint32_t *vals;
  size_t cap, used;
} PtrMap;

static size_t ptr_hash(const void *p) {
  uint64_t x = (uint64_t)(uintptr_t)p;
  x ^= x >> 33;
  x *= 0xff51afd7ed558ccdULL;
  x ^= x >> 33;
  return (size_t)x;
}
static void ptrmap_put(PtrMap *m, const void *key, uint32_t val);
static void ptrmap_grow(PtrMap *m) {
  PtrMap bigger = {0};
  bigger.cap = m->cap ? m->cap * 2 : 256;
  bigger.keys = calloc(bigger.cap, sizeof(void *));
  bigger.vals = calloc(bigger.cap, sizeof(uint32_t));
  for (size_t i = 0; i < m->cap; i++)
    if (m->keys[i] != NULL)
      ptrmap_put(&bigger, m->keys[i], m->vals[i]);
  free(m->keys);
  free(m->vals);
  *m = bigger;
}
static void ptrmap_put(PtrMap *m, const void *key, uint32_t val) {
  if ((m->used + 1) * 2 > m->cap)
    ptrmap_grow(m);
  size_t i = ptr_hash(key) & (m->cap - 1);
  while (m->keys[i] != NULL && m->keys[i] != key)
    i = (i + 1) & (m->cap - 1);
  if (m->keys[i] == NULL)
    m->used++;
  m->keys[i] = key;
  m->vals[i] = val;
}
static uint32_t ptrmap_get(const PtrMap *m, const void *key) {
  if (m->cap == 0)
    return RT_NONE;
  size_t i = ptr_hash(key) & (m->cap - 1);
  while (m->keys[i] != NULL) {
    if (m->keys[i] == key)
      return m->vals[i];
    i = (i + 1) & (m->cap - 1);
  }
  return RT_NONE;
}
static void ptrmap_free(PtrMap *m) {
  free(m->keys);
  free(m->vals);
  memset(m, 0, sizeof(*m));
}

/* ------------------------------------------------------------------ the walk ----------------------------- */

typedef struct Flattener {
  Vec nodes, children, spheres, quads, materials, textures, perlins, images, texels;
  PtrMap node_of, material_of, texture_of;
  bool failed;
} Flattener;

static void fail(Flattener *fl, const char *msg) {
  if (!fl->failed)
    snprintf(g_error, sizeof(g_error), "%s", msg);
  fl->failed = true;
}

static uint32_t emit_texture(Flattener *fl, const Texture *tex) {
  if (tex == NULL)
    return RT_NONE;
  uint32_t id = ptrmap_get(&fl->texture_of, tex);
  if (id != RT_NONE)
    return id;
  id = (uint32_t)fl->textures.len;
  vec_push(&fl->textures);
  ptrmap_put(&fl->texture_of, tex, id);

  rt_texture t = {0};
  if (tex->value == g_kinds.solid) {
    const Solid *s = (const Solid *)tex;
    t.kind = RT_TEX_SOLID;
    memcpy(t.color, s->color.values, sizeof(t.color));
  } else if (tex->value == g_kinds.checker) {
    const Checker *c = (const Checker *)tex;
    t.kind = RT_TEX_CHECKER;
    t.scale = c->scale;
    t.a = emit_texture(fl, c->even);
    t.b = emit_texture(fl, c->odd);
  } else if (tex->value == g_kinds.perlin) {
    const Perlin *p = (const Perlin *)tex;
    t.kind = RT_TEX_PERLIN;
    t.scale = p->scale;
    t.b = (uint32_t)p->depth;
    t.a = (uint32_t)fl->perlins.len;
    rt_perlin *dst = vec_push(&fl->perlins);
    for (int i = 0; i < N_PERLIN; i++) {
      memcpy(&dst->grad[3 * i], p->grad_field[i].values, 3 * sizeof(float));
      dst->perm_x[i] = p->perm_x[i];
      dst->perm_y[i] = p->perm_y[i];
      dst->perm_z[i] = p->perm_z[i];
    }
  } else if (g_kinds.image != NULL && tex->value == g_kinds.image) {
    const Image *im = (const Image *)tex;
    t.kind = RT_TEX_IMAGE;
    t.a = (uint32_t)fl->images.len;
    rt_image *dst = vec_push(&fl->images);
    dst->width = im->width;
    dst->height = im->height;
    dst->offset = fl->texels.len;
    size_t n = (size_t)im->width * im->height * 3;
    size_t padded = (n + 15) & ~(size_t)15; /* keep every image 16-byte aligned in the blob */
    if (fl->texels.len + padded > fl->texels.cap) {
      fl->texels.cap = (fl->texels.len + padded) * 2;
      fl->texels.data = realloc(fl->texels.data, fl->texels.cap);
      assert(fl->texels.data != NULL);
    }
    memcpy(fl->texels.data + fl->texels.len, im->buffer, n);
    memset(fl->texels.data + fl->texels.len + n, 0, padded - n);
    fl->texels.len += padded;
  } else {
    fail(fl, "texture with an unknown value function (user-defined textures cannot run on the device)");
  }
  VEC_AT(fl->textures, rt_texture, id) = t;
  return id;
}

static uint32_t emit_material(Flattener *fl, const Material *mat) {
  if (mat == NULL) {
    fail(fl, "primitive with a NULL material");
    return RT_NONE;
  }
  uint32_t id = ptrmap_get(&fl->material_of, mat);
  if (id != RT_NONE)
    return id;
  id = (uint32_t)fl->materials.len;
  vec_push(&fl->materials);
  ptrmap_put(&fl->material_of, mat, id);

  rt_material m = {0};
  m.tag = (uint32_t)mat->tag;
  m.texture = RT_NONE;
  switch (mat->tag) {
  case LAMBERTIAN:
  case DIFFUSE_LIGHT:
  case ISOTROPIC:
    m.texture = emit_texture(fl, mat->albedo);
    break;
  case METAL:
    m.texture = emit_texture(fl, mat->albedo);
    m.param = mat->fuzz;
    break;
  case DIELECTRIC:
    m.param = mat->eta;
    break;
  case SURFACE_NORMAL:
    break;
  default:
    fail(fl, "material with an unknown tag");
  }
  if (m.texture == RT_NONE && (mat->tag == LAMBERTIAN || mat->tag == DIFFUSE_LIGHT || mat->tag == ISOTROPIC ||
                               mat->tag == METAL))
    fail(fl, "material that needs an albedo texture has none");
  VEC_AT(fl->materials, rt_material, id) = m;
  return id;
}

static void copy_bbox(rt_node *n, const Hittable *h) {
  for (int a = 0; a < 3; a++) {
    n->bbox[2 * a] = h->bbox.values[a][0];
    n->bbox[2 * a + 1] = h->bbox.values[a][1];
  }
}

static uint32_t emit_node(Flattener *fl, const Hittable *h, int depth) {
  if (h == NULL) {
    fail(fl, "NULL Hittable in the graph");
    return RT_NONE;
  }
  if (depth > 256) {
    fail(fl, "Hittable graph deeper than 256 levels (cycle?)");
    return RT_NONE;
  }
  uint32_t id = ptrmap_get(&fl->node_of, h);
  if (id != RT_NONE)
    return id;
  id = (uint32_t)fl->nodes.len;
  vec_push(&fl->nodes);
  ptrmap_put(&fl->node_of, h, id);

  /* `n` is filled locally and stored at the end: recursive calls may reallocate fl->nodes. */
  rt_node n = {0};
  copy_bbox(&n, h);
  const HittableVTable *vt = h->vtable;

  if (vt == g_kinds.list) {
    const HittableList *l = (const HittableList *)h;
    n.kind = RT_NODE_LIST;
    n.flags = RT_NODE_HAS_PDF;
    n.a = (uint32_t)fl->children.len;
    n.b = (uint32_t)l->size;
    for (size_t i = 0; i < l->size; i++)
      vec_push(&fl->children); /* reserve this list's slots before descending */
    for (size_t i = 0; i < l->size; i++) {
      uint32_t child = emit_node(fl, l->items[i], depth + 1);
      VEC_AT(fl->children, uint32_t, n.a + i) = child;
    }
  } else if (vt == g_kinds.sphere) {
    const Sphere *s = (const Sphere *)h;
    n.kind = RT_NODE_SPHERE;
    n.flags = RT_NODE_HAS_PDF;
    n.a = (uint32_t)fl->spheres.len;
    vec_push(&fl->spheres);
    rt_sphere out = {0};
    memcpy(out.center, s->center.values, sizeof(out.center));
    out.radius = s->radius;
    out.material = emit_material(fl, s->material);
    out.node = id;
    VEC_AT(fl->spheres, rt_sphere, n.a) = out;
  } else if (vt == g_kinds.quad) {
    const Quad *q = (const Quad *)h;
    n.kind = RT_NODE_QUAD;
    n.flags = RT_NODE_HAS_PDF;
    n.a = (uint32_t)fl->quads.len;
    vec_push(&fl->quads);
    rt_quad out = {0};
    memcpy(out.Q, q->Q.values, 12);
    memcpy(out.u, q->u.values, 12);
    memcpy(out.v, q->v.values, 12);
    memcpy(out.normal, q->normal.values, 12);
    memcpy(out.w, q->w.values, 12);
    out.D = q->D;
    out.area = q->area;
    out.material = emit_material(fl, q->material);
    out.node = id;
    VEC_AT(fl->quads, rt_quad, n.a) = out;
  } else if (vt == g_kinds.bvh) {
    const BVHNode *b = (const BVHNode *)h;
    n.kind = RT_NODE_BVH;
    n.a = emit_node(fl, b->left, depth + 1);
    n.b = emit_node(fl, b->right, depth + 1);
  } else if (vt == g_kinds.translate) {
    const Translate *t = (const Translate *)h;
    n.kind = RT_NODE_TRANSLATE;
    memcpy(n.f, t->offset.values, 12);
    n.a = emit_node(fl, t->object, depth + 1);
  } else if (vt == g_kinds.rotate_y) {
    const RotateY *r = (const RotateY *)h;
    n.kind = RT_NODE_ROTATE_Y;
    n.f[0] = r->sin_theta;
    n.f[1] = r->cos_theta;
    n.a = emit_node(fl, r->object, depth + 1);
  } else if (vt == g_kinds.medium) {
    const ConstantMedium *m = (const ConstantMedium *)h;
    n.kind = RT_NODE_MEDIUM;
    n.f[0] = m->neg_inv_density;
    n.b = emit_material(fl, m->phase_fn);
    n.a = emit_node(fl, m->boundary, depth + 1);
  } else {
    fail(fl, "Hittable with an unknown vtable (user-defined hittables cannot run on the device)");
  }
  VEC_AT(fl->nodes, rt_node, id) = n;
  return id;
}

static PtrMap g_last_nodes, g_last_materials;
uint32_t rtc_flatten_last_node_id(const void *hittable) { return ptrmap_get(&g_last_nodes, hittable); }
uint32_t rtc_flatten_last_material_id(const void *material) { return ptrmap_get(&g_last_materials, material); }

static void *vec_steal(Vec *v) {
  void *p = v->data;
  memset(v, 0, sizeof(*v));
  return p;
}

int rtc_flatten(const World *world, rt_scene_desc *out) {
  learn_kinds();
  g_error[0] = 0;
  memset(out, 0, sizeof(*out));

  Flattener fl = {0};
  fl.nodes.item = sizeof(rt_node);
  fl.children.item = sizeof(uint32_t);
  fl.spheres.item = sizeof(rt_sphere);
  fl.quads.item = sizeof(rt_quad);
  fl.materials.item = sizeof(rt_material);
  fl.textures.item = sizeof(rt_texture);
  fl.perlins.item = sizeof(rt_perlin);
  fl.images.item = sizeof(rt_image);
  fl.texels.item = 1;

  uint32_t objects_root = emit_node(&fl, &world->objects.hittable, 0);
  uint32_t lights_root = RT_NONE;
  if (world->lights.size > 0)
    lights_root = emit_node(&fl, &world->lights.hittable, 0);

  /* keep the pointer -> id maps of the most recent walk for rtc_flatten_last_*_id() (test/oracle tooling) */
  ptrmap_free(&g_last_nodes);
  ptrmap_free(&g_last_materials);
  g_last_nodes = fl.node_of;
  g_last_materials = fl.material_of;
  ptrmap_free(&fl.texture_of);

  out->abi_version = RT_ABI_VERSION;
  out->n_nodes = (uint32_t)fl.nodes.len;
  out->n_children = (uint32_t)fl.children.len;
  out->n_spheres = (uint32_t)fl.spheres.len;
  out->n_quads = (uint32_t)fl.quads.len;
  out->n_materials = (uint32_t)fl.materials.len;
  out->n_textures = (uint32_t)fl.textures.len;
  out->n_perlins = (uint32_t)fl.perlins.len;
  out->n_images = (uint32_t)fl.images.len;
  out->n_texel_bytes = fl.texels.len;
  out->objects_root = objects_root;
  out->lights_root = lights_root;
  out->nodes = vec_steal(&fl.nodes);
  out->children = vec_steal(&fl.children);
  out->spheres = vec_steal(&fl.spheres);
  out->quads = vec_steal(&fl.quads);
  out->materials = vec_steal(&fl.materials);
  out->textures = vec_steal(&fl.textures);
  out->perlins = vec_steal(&fl.perlins);
  out->images = vec_steal(&fl.images);
  out->texels = vec_steal(&fl.texels);

  if (fl.failed) {
    rtc_scene_desc_free(out);
    return -1;
  }
  return 0;
}

void rtc_scene_desc_free(rt_scene_desc *d) {
  free((void *)d->nodes);
  free((void *)d->children);
  free((void *)d->spheres);
  free((void *)d->quads);
  free((void *)d->materials);
  free((void *)d->textures);
  free((void *)d->perlins);
  free((void *)d->images);
  free((void *)d->texels);
  memset(d, 0, sizeof(*d));
}

/* ------------------------------------------------------------------ (de)serialisation -------------------- */

static const char BLOB_MAGIC[8] = {'R', 'T', 'C', 'S', 'C', 'N', '0', '1'};

typedef struct BlobPart {
  const void *data;
  size_t bytes;
} BlobPart;

static void blob_parts(const rt_scene_desc *d, BlobPart part[9]) {
  part[0] = (BlobPart){d->nodes, sizeof(rt_node) * d->n_nodes};
  part[1] = (BlobPart){d->children, sizeof(uint32_t) * d->n_children};
  part[2] = (BlobPart){d->spheres, sizeof(rt_sphere) * d->n_spheres};
  part[3] = (BlobPart){d->quads, sizeof(rt_quad) * d->n_quads};
  part[4] = (BlobPart){d->materials, sizeof(rt_material) * d->n_materials};
  part[5] = (BlobPart){d->textures, sizeof(rt_texture) * d->n_textures};
  part[6] = (BlobPart){d->perlins, sizeof(rt_perlin) * d->n_perlins};
  part[7] = (BlobPart){d->images, sizeof(rt_image) * d->n_images};
  part[8] = (BlobPart){d->texels, d->n_texel_bytes};
}

int rtc_scene_desc_write(const rt_scene_desc *desc, FILE *f) {
  rt_scene_desc header = *desc;
  header.nodes = NULL, header.children = NULL, header.spheres = NULL, header.quads = NULL;
  header.materials = NULL, header.textures = NULL, header.perlins = NULL, header.images = NULL;
  header.texels = NULL;
  if (fwrite(BLOB_MAGIC, 1, 8, f) != 8 || fwrite(&header, sizeof(header), 1, f) != 1)
    return -1;
  BlobPart part[9];
  blob_parts(desc, part);
  for (int i = 0; i < 9; i++)
    if (part[i].bytes && fwrite(part[i].data, 1, part[i].bytes, f) != part[i].bytes)
      return -1;
  return 0;
}

int rtc_scene_desc_read(rt_scene_desc *out, FILE *f) {
  char magic[8];
  if (fread(magic, 1, 8, f) != 8 || memcmp(magic, BLOB_MAGIC, 8) != 0)
    return -1;
  if (fread(out, sizeof(*out), 1, f) != 1 || out->abi_version != RT_ABI_VERSION)
    return -1;
  BlobPart part[9];
  blob_parts(out, part); /* sizes only; the pointers in a stored header are NULL */
  void *loaded[9] = {0};
  for (int i = 0; i < 9; i++) {
    loaded[i] = malloc(part[i].bytes ? part[i].bytes : 1);
    if (loaded[i] == NULL || (part[i].bytes && fread(loaded[i], 1, part[i].bytes, f) != part[i].bytes)) {
      for (int k = 0; k <= i; k++)
        free(loaded[k]);
      memset(out, 0, sizeof(*out));
      return -1;
    }
  }
  out->nodes = loaded[0], out->children = loaded[1], out->spheres = loaded[2], out->quads = loaded[3];
  out->materials = loaded[4], out->textures = loaded[5], out->perlins = loaded[6], out->images = loaded[7];
  out->texels = loaded[8];
  return 0;
}

/* ------------------------------------------------------------------ camera ------------------------------- */

void rtc_camera_pack(const Camera *c, rt_camera *out) {
  memset(out, 0, sizeof(*out));
  out->img_width = c->img_width;
  out->img_height = c->img_height;
  out->samples_per_pixel = c->samples_per_pixel;
  out->max_depth = c->max_depth;
  memcpy(out->background, c->background.values, 12);
  out->lights_sampling_prob = c->lights_sampling_prob;
  memcpy(out->origin, c->look_from.values, 12);
  out->dof_angle = c->dof_angle;
  memcpy(out->pixel00, c->pixel00_loc.values, 12);
  memcpy(out->pixel_delta_u, c->pixel_delta_u.values, 12);
  memcpy(out->pixel_delta_v, c->pixel_delta_v.values, 12);
  memcpy(out->dof_disc_u, c->dof_disc_u.values, 12);
  memcpy(out->dof_disc_v, c->dof_disc_v.values, 12);
}
