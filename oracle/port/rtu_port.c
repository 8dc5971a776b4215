/* oracle/port/rtu_port.c — TEST INFRASTRUCTURE ONLY (CPU oracle; never linked into the product).
 *
 * Plain-C restatement of the simx RT-unit path ("layout B"): quantized-BVH4 traversal with restart trail and
 * 5-entry short stack (sim/simx/rt_traversal.cpp:26-213, ShortStack sim/common/types.h:1808-1841), its intersection
 * primitives (:231-339), the RTUnit any-hit=accept driver (sim/simx/rt_unit.cpp:98-116,199-202 with
 * tests/regression/raytracing/shaders/anyhit.cpp:28-35), the device kernel's ray generation
 * (tests/regression/raytracing/kernel.cpp:13-39) and the closest-hit / miss shaders (shaders/closest.cpp:11-130,
 * shaders/miss.cpp:7-14, rtx_shading.h:5-18,55-67, common.h:149-162).
 *
 * Parity pin: bit-identical hit records and frames to oracle/_ref/libref_rtu.so (the reference's rt_traversal.cpp
 * compiled verbatim + its own headers' arithmetic) — tests/test_oracle_rtu.py — and to tests/golden/rtu_*.npz.
 * The reference ships no golden vectors for this path.  Compiled -O2 -ffp-contract=off.
 *
 * Everything operates on a flat memory image: Vortex address A is mem[A - base] (what dcache_read sees).
 */
#include <pthread.h>
#include <stdlib.h>
#include "rc_math.h"

#define RTU_WIDTH 4            /* RT_BVH_WIDTH */
#define RTU_TRAIL 32           /* MAX_TRAIL_LEVEL, rt_traversal.h:8 */
#define RTU_STACK 5            /* TRAVERSAL_STACK_CAPACITY, rt_traversal.h:9 */
#define RTU_NODE_BYTES 52      /* sizeof(BVHNode), rt_traversal.h:19-33 */
#define RTU_BLAS_STRIDE 160    /* rt_traversal.cpp:112 */
#define RTU_TRI_BYTES 36       /* sizeof(Triangle), rt_traversal.h:40-42 */

typedef struct { float dist, bx, by, bz; uint32_t blasIdx, triIdx; } rtu_hit;
typedef struct { float ox, oy, oz, dx, dy, dz; } rtu_ray;
typedef struct {
  const uint8_t *mem; uint64_t base, size;
  uint32_t tlas, blas, qbvh, tri;           /* DCRs 0x006..0x009 (rt_traversal.cpp:35-38) */
} rtu_image;
typedef struct { uint64_t calls, nodes, blas, tris, restarts, wraps; } rtu_work;

#pragma pack(push, 1)
typedef struct { uint8_t meta; uint8_t q[6]; } rtu_child;
typedef struct {
  float px, py, pz; int8_t ex, ey, ez; uint8_t imask;
  uint32_t leftFirst, leafData;
  rtu_child children[RTU_WIDTH];
} rtu_node;
#pragma pack(pop)

/* The reference computes child addresses from `base_ptr`, which BVHTraverser::pop does NOT reset when the popped entry is a
 * TLAS node (rt_traversal.cpp:179-213,92): after an instance was visited, the children of a popped TLAS *inner* node are
 * fetched relative to that instance's BVH — geometry behind a TLAS of more than one inner level is silently missed.
 * 0 (default) reproduces this; 1 addresses the children of TLAS nodes from the TLAS base (what the author evidently meant),
 * used only to check the CUDA fast order, which has no such defect. */
static int g_fix_tlas_base = 0;
void port_rtu_set_fix_tlas_base(int on) { g_fix_tlas_base = on; }

static void img_read(const rtu_image *im, void *dst, uint64_t addr, uint32_t n) {   /* RTUnit::dcache_read */
  if (addr < im->base || addr + n > im->base + im->size) { memset(dst, 0, n); return; }
  memcpy(dst, im->mem + (addr - im->base), n);
}

/* ---- ShortStack<TraversalStackEntry,5> (types.h:1808-1841): a ring that overwrites its oldest entry ---- */
typedef struct { uint32_t node_ptr; int last; } rtu_entry;
typedef struct { uint32_t head, bottom, count; rtu_entry e[RTU_STACK]; } rtu_stack;
static void st_push(rtu_stack *s, rtu_entry v) {
  if (s->count == RTU_STACK) s->bottom = (s->bottom + 1) % RTU_STACK; else s->count++;
  s->e[s->head] = v;
  s->head = (s->head + 1) % RTU_STACK;
}
static rtu_entry st_pop(rtu_stack *s) {
  rtu_entry z = {0, 0};
  if (s->count == 0) return z;
  s->head = (s->head == 0) ? (RTU_STACK - 1) : (s->head - 1);
  s->count--;
  return s->e[s->head];
}

/* ---- rt_traversal.cpp:318-339 ---- */
static float ray_box(const rtu_ray *r, float min_x, float min_y, float min_z, float max_x, float max_y, float max_z) {
  float idir_x = 1.0f / r->dx, idir_y = 1.0f / r->dy, idir_z = 1.0f / r->dz;
  float tx1 = (min_x - r->ox) * idir_x, tx2 = (max_x - r->ox) * idir_x;
  float tmin = std_minf(tx1, tx2), tmax = std_maxf(tx1, tx2);
  float ty1 = (min_y - r->oy) * idir_y, ty2 = (max_y - r->oy) * idir_y;
  tmin = std_maxf(tmin, std_minf(ty1, ty2));
  tmax = std_minf(tmax, std_maxf(ty1, ty2));
  float tz1 = (min_z - r->oz) * idir_z, tz2 = (max_z - r->oz) * idir_z;
  tmin = std_maxf(tmin, std_minf(tz1, tz2));
  tmax = std_minf(tmax, std_maxf(tz1, tz2));
  return (tmax < tmin || tmax <= 0) ? RC_LARGE_FLOAT : tmin;
}

/* ---- rt_traversal.cpp:263-316 ---- */
static float ray_tri(const rtu_ray *r, const float *t, float *bx, float *by, float *bz) {
  float e1x = t[3] - t[0], e1y = t[4] - t[1], e1z = t[5] - t[2];
  float e2x = t[6] - t[0], e2y = t[7] - t[1], e2z = t[8] - t[2];
  float hx = r->dy * e2z - r->dz * e2y, hy = r->dz * e2x - r->dx * e2z, hz = r->dx * e2y - r->dy * e2x;
  float a = e1x * hx + e1y * hy + e1z * hz;
  if (fabs(a) < RC_EPSILON) return RC_LARGE_FLOAT;
  float f = 1 / a;
  float sx = r->ox - t[0], sy = r->oy - t[1], sz = r->oz - t[2];
  float w1 = f * (sx * hx + sy * hy + sz * hz);
  if (w1 < 0 || w1 > 1) return RC_LARGE_FLOAT;
  float qx = sy * e1z - sz * e1y, qy = sz * e1x - sx * e1z, qz = sx * e1y - sy * e1x;
  float w2 = f * (r->dx * qx + r->dy * qy + r->dz * qz);
  if (w2 < 0 || w1 + w2 > 1) return RC_LARGE_FLOAT;
  float tf = f * (e2x * qx + e2y * qy + e2z * qz);
  if (tf <= RC_EPSILON) return RC_LARGE_FLOAT;
  *bx = w1, *by = w2, *bz = 1 - w1 - w2;
  return tf;
}

/* ---- rt_traversal.cpp:231-261: 3x4 transform, direction not re-normalised ---- */
static rtu_ray ray_xform(const rtu_ray *r, const float *m) {
  rtu_ray o;
  o.ox = m[0] * r->ox + m[1] * r->oy + m[2] * r->oz + m[3];
  o.oy = m[4] * r->ox + m[5] * r->oy + m[6] * r->oz + m[7];
  o.oz = m[8] * r->ox + m[9] * r->oy + m[10] * r->oz + m[11];
  o.dx = m[0] * r->dx + m[1] * r->dy + m[2] * r->dz;
  o.dy = m[4] * r->dx + m[5] * r->dy + m[6] * r->dz;
  o.dz = m[8] * r->dx + m[9] * r->dy + m[10] * r->dz;
  return o;
}

static int is_top(const rtu_node *n) { return n->imask == 1; }                          /* :219-221 */
static int is_leaf(const rtu_node *n) {                                                  /* :223-225 */
  return (is_top(n) && n->leafData != 0xFFFFFFFFu) || (!is_top(n) && n->leafData != 0);
}

/* ---- BVHTraverser::pop (:179-213) ---- */
static int trav_pop(const rtu_image *im, uint32_t *base_ptr, uint32_t *node_ptr, uint32_t *level, uint32_t *trail,
                    rtu_stack *stack, rtu_work *w) {
  int parent = -1;
  for (int i = (int)*level - 1; i >= 0; i--)                                             /* findNextParentLevel :170-177 */
    if (trail[i] != RTU_WIDTH) { parent = i; break; }
  if (parent < 0) return 1;
  trail[parent]++;
  for (int i = parent + 1; i < RTU_TRAIL; i++) trail[i] = 0;
  if (stack->count == 0) {
    *base_ptr = im->tlas, *node_ptr = im->tlas, *level = 0;                              /* restart from the root */
    if (w) w->restarts++;
  } else {
    rtu_entry e = st_pop(stack);
    *node_ptr = e.node_ptr;
    if (e.last) trail[parent] = RTU_WIDTH;
    *level = (uint32_t)parent + 1;
  }
  return 0;
}

/* ---- BVHTraverser::traverse (:26-168).  Returns 1 when traversal completed, 0 when a candidate is pending. ---- */
static int traverse(const rtu_image *im, const rtu_ray *ray, rtu_hit *hit, float *pending, uint32_t *trail, rtu_stack *stack,
                    rtu_work *w) {
  uint32_t level = 0, base_ptr = im->tlas, node_ptr = im->tlas, blasIdx = 0;
  rtu_ray cur = *ray;
  rtu_node node;
  int done = 0;
  if (w) w->calls++;
  while (!done) {
    img_read(im, &node, node_ptr, RTU_NODE_BYTES);
    if (w) w->nodes++;
    if (!is_leaf(&node)) {
      float dist[RTU_WIDTH]; uint32_t idx[RTU_WIDTH]; int n = 0;
      for (int i = 0; i < RTU_WIDTH; i++) {
        if (node.children[i].meta == 0) continue;
        const uint8_t *q = node.children[i].q;
        float min_x = node.px + ldexpf((float)q[0], node.ex), min_y = node.py + ldexpf((float)q[1], node.ey),
              min_z = node.pz + ldexpf((float)q[2], node.ez);
        float max_x = node.px + ldexpf((float)q[3], node.ex), max_y = node.py + ldexpf((float)q[4], node.ey),
              max_z = node.pz + ldexpf((float)q[5], node.ez);
        float d = ray_box(is_top(&node) ? ray : &cur, min_x, min_y, min_z, max_x, max_y, max_z);
        if (d < hit->dist) {
          /* std::sort with "a.dist > b.dist" on <= 4 elements is libstdc++'s insertion sort: stable, farthest first */
          int p = n;
          while (p > 0 && d > dist[p - 1]) { dist[p] = dist[p - 1]; idx[p] = idx[p - 1]; p--; }
          dist[p] = d, idx[p] = (uint32_t)i;
          n++;
        }
      }
      uint32_t k = trail[level];
      /* k == WIDTH: keep only the farthest.  With n == 0 the reference's "size()-1" wraps and its loop spins over an
         empty vector (a no-op); the outcome is the same as dropping nothing. */
      uint32_t drop = (k == RTU_WIDTH) ? (n > 0 ? (uint32_t)n - 1 : 0) : k;
      if (w && k == RTU_WIDTH && n == 0) w->wraps++;   /* where the compiled reference spins for 2^32 iterations */
      while (drop-- > 0 && n > 0) n--;                                                   /* pop_back = drop the nearest */
      if (n == 0) {
        done = trav_pop(im, &base_ptr, &node_ptr, &level, trail, stack, w);
      } else {
        uint32_t closest = idx[n - 1];
        n--;
        if (g_fix_tlas_base && is_top(&node)) base_ptr = im->tlas;
        node_ptr = base_ptr + (node.leftFirst + closest) * RTU_NODE_BYTES;
        if (n == 0) trail[level] = RTU_WIDTH;
        else
          for (int i = 0; i < n; i++) {
            rtu_entry e = {base_ptr + (node.leftFirst + idx[i]) * RTU_NODE_BYTES, i == 0};
            st_push(stack, e);
          }
        level++;
      }
    } else if (is_top(&node)) {
      blasIdx = node.leafData;
      struct { uint32_t bvh_offset; float inv[12]; } b;                                  /* BLASNode, rt_traversal.h:35-38 */
      img_read(im, &b, im->blas + blasIdx * RTU_BLAS_STRIDE, sizeof(b));
      if (w) w->blas++;
      cur = ray_xform(ray, b.inv);
      base_ptr = im->qbvh + b.bvh_offset * RTU_NODE_BYTES;
      node_ptr = base_ptr;
    } else {
      uint32_t triCount = node.leafData, leftFirst = node.leftFirst;
      for (uint32_t i = 0; i < triCount; ++i) {
        uint32_t triIdx = leftFirst + i;
        float t[9];
        img_read(im, t, im->tri + triIdx * RTU_TRI_BYTES, RTU_TRI_BYTES);
        if (w) w->tris++;
        float bx, by, bz;
        float d = ray_tri(&cur, t, &bx, &by, &bz);
        if (d < hit->dist) {
          *pending = d;
          hit->bx = bx, hit->by = by, hit->bz = bz, hit->blasIdx = blasIdx, hit->triIdx = triIdx;
          while (stack->count) st_pop(stack);
          return 0;
        }
      }
      done = trav_pop(im, &base_ptr, &node_ptr, &level, trail, stack, w);
    }
  }
  return 1;
}

/* one ray through the RT unit with the shipped any-hit shader (always ACCEPT) */
static void trace_max(const rtu_image *im, const rtu_ray *ray, rtu_hit *hit, float tmax, rtu_work *w) {
  uint32_t trail[RTU_TRAIL];
  rtu_stack stack;
  memset(trail, 0, sizeof(trail));
  memset(&stack, 0, sizeof(stack));
  hit->dist = tmax, hit->bx = hit->by = hit->bz = 0.0f, hit->blasIdx = hit->triIdx = 0;
  float pending = 0.0f;
  while (!traverse(im, ray, hit, &pending, trail, &stack, w)) hit->dist = pending;      /* commit(ACCEPT), rt_unit.cpp:199-202 */
}
static void trace_one(const rtu_image *im, const rtu_ray *ray, rtu_hit *hit, rtu_work *w) { trace_max(im, ray, hit, RC_LARGE_FLOAT, w); }

/* ================================================================ device program: raygen + shaders ================ */
#pragma pack(push, 1)
typedef struct {                       /* raytracing/common.h:164-195 — 216 B */
  uint32_t dst_width, dst_height; uint64_t dst_addr;
  uint64_t tri_addr, triEx_addr, triIdx_addr, mat_addr, tex_addr, bvh_addr, qBvh_addr, blas_addr, tlas_addr;
  uint32_t tlas_root;
  v3 camera_pos, camera_forward, camera_right, camera_up; v2 viewplane;
  uint32_t samples_per_pixel, max_depth;
  v3 light_pos, light_color, ambient_color, background_color;
  uint32_t _pad; uint64_t sbt_addr;
} rtu_karg;
typedef struct {                       /* raytracing/common.h:86-99 — 160 B */
  uint32_t bvh_offset; float invTransform[16]; float transform[16]; uint32_t _p0;
  uint64_t mat_offset; uint32_t tex_width, tex_height; float reflectivity; uint32_t _p1;
} rtu_blas;
typedef struct { v3 N0, N1, N2; v2 uv0, uv1, uv2; uint32_t texId; } rtu_triex;          /* common.h:39-43 — 64 B */
typedef struct {                       /* common.h:20-36 — 88 B */
  v3 ambient, diffuse, specular, emissive; float shininess, ior, dissolve, reflectivity;
  int32_t diffuse_tex_id, illum; uint32_t tex_width, tex_height; uint64_t tex_offset;
} rtu_material;
#pragma pack(pop)
_Static_assert(sizeof(rtu_karg) == 216 && sizeof(rtu_blas) == 160 && sizeof(rtu_triex) == 64 && sizeof(rtu_material) == 88 &&
               sizeof(rtu_node) == 52, "layout B");

/* kernel.cpp:28-39 — the camera the shipped kernel hard-codes (camera 0).  u,v: double arithmetic narrowed to float. */
static rtu_ray gen_fixed(uint32_t x, uint32_t y, const rtu_karg *a) {
  v3 pos = V3(0.0f, 100.0f, 0.0f), front = V3(1.0f, 0.0f, 0.0f);
  float FOV = 1.0f;
  float u = (float)((x * 2.0 - a->dst_width) / a->dst_height);
  float v = (float)((y * 2.0 - a->dst_height) / a->dst_height);
  v3 right = vcross(front, V3(0.0f, 1.0f, 0.0f));
  v3 up = vcross(right, front);
  v3 dir = vnormalize(vadd(vadd(sscale(u, right), sscale(v, up)), sscale(FOV, front)));
  rtu_ray r = {pos.x, pos.y, pos.z, dir.x, dir.y, dir.z};
  return r;
}
/* kernel.cpp:13-26 — kernel_arg_t pinhole camera (camera 1) */
static rtu_ray gen_pinhole(uint32_t x, uint32_t y, const rtu_karg *a) {
  float x_ndc = (float)((x + 0.5f) / a->dst_width - 0.5);
  float y_ndc = (float)((y + 0.5f) / a->dst_height - 0.5);
  float x_vp = x_ndc * a->viewplane.x, y_vp = y_ndc * a->viewplane.y;
  v3 pt_cam = vadd(vadd(sscale(x_vp, a->camera_right), sscale(y_vp, a->camera_up)), a->camera_forward);
  v3 pt_w = vadd(pt_cam, a->camera_pos);
  v3 d = vnormalize(vsub(pt_w, a->camera_pos));
  rtu_ray r = {a->camera_pos.x, a->camera_pos.y, a->camera_pos.z, d.x, d.y, d.z};
  return r;
}

static v3 rgb8_to_f(uint32_t c) {                                                         /* common.h:155-162 */
  float s = 1 / 256.0f;
  int r = (c >> 16) & 255, g = (c >> 8) & 255, b = c & 255;
  return V3(r * s, g * s, b * s);
}
static uint32_t f_to_rgb8(v3 c) {                                                          /* common.h:148-153 */
  int r = (int)(std_minf(c.x, 1.f) * 255), g = (int)(std_minf(c.y, 1.f) * 255), b = (int)(std_minf(c.z, 1.f) * 255);
  return (uint32_t)((r << 16) + (g << 8) + b);
}

/* closest.cpp:11-130 (+ miss.cpp:12); `bounce` = ray_payload_t::bounce */
static v3 shade(const rtu_image *im, const rtu_karg *a, const rtu_ray *ray, uint32_t bounce, rtu_work *w, uint64_t *shaded) {
  rtu_hit h;
  trace_one(im, ray, &h, w);
  if (h.dist == RC_LARGE_FLOAT) return a->background_color;
  if (shaded) ++*shaded;
  rtu_blas blas; rtu_triex ex; rtu_material mat;
  img_read(im, &blas, a->blas_addr + (uint64_t)h.blasIdx * sizeof(rtu_blas), sizeof(blas));
  img_read(im, &ex, a->triEx_addr + (uint64_t)h.triIdx * sizeof(rtu_triex), sizeof(ex));
  img_read(im, &mat, a->mat_addr + (uint64_t)ex.texId * sizeof(rtu_material), sizeof(mat));
  v3 org = V3(ray->ox, ray->oy, ray->oz), dir = V3(ray->dx, ray->dy, ray->dz);
  v3 radiance = V3(0, 0, 0);
  float throughput = 1.0f;
  v3 I = vadd(org, vscale(dir, h.dist));
  v3 N = vadd(vadd(vscale(ex.N1, h.bx), vscale(ex.N2, h.by)), vscale(ex.N0, h.bz));
  m4 inv; memcpy(inv.c, blas.invTransform, 64);
  m4 invT = m4_transposed3(&inv);
  N = vnormalize(xform_vec(N, &invT));
  v2 uv;
  uv.x = ex.uv1.x * h.bx + ex.uv2.x * h.by + ex.uv0.x * h.bz;
  uv.y = ex.uv1.y * h.bx + ex.uv2.y * h.by + ex.uv0.y * h.bz;
  v3 texColor;
  if (mat.diffuse_tex_id >= 0) {                                                          /* texSample, rtx_shading.h:5-18 */
    uint32_t iu = (uint32_t)(uv.x * mat.tex_width), iv = (uint32_t)(uv.y * mat.tex_height);
    iu %= mat.tex_width, iv %= mat.tex_height;
    uint32_t texel;
    img_read(im, &texel, a->tex_addr + mat.tex_offset + 4ull * (iu + iv * mat.tex_width), 4);
    texColor = rgb8_to_f(texel);
  } else {
    texColor = mat.diffuse;
  }
  v3 L = vsub(a->light_pos, I);                                                           /* diffuseLighting, rtx_shading.h:55-67 */
  float dist = vlength(L);
  L = vscale(L, 1.0f / dist);
  float att = 1.0f / (1.0f + dist * 0.1f);
  float NdotL = std_maxf(0.0f, vdot(N, L));
  v3 diffuse = vmul(texColor, vadd(a->ambient_color, vscale(sscale(att, a->light_color), NdotL)));
  float reflectivity = blas.reflectivity;
  radiance = vadd(radiance, vscale(sscale(throughput, diffuse), 1 - reflectivity));
  throughput *= reflectivity;
  if (reflectivity > 0.0f && bounce + 1 < a->max_depth) {
    v3 R = vnormalize(vsub(dir, vscale(sscale(2.0f, N), vdot(N, dir))));
    v3 so = vadd(I, vscale(R, 0.001f));
    rtu_ray sec = {so.x, so.y, so.z, R.x, R.y, R.z};
    v3 c = shade(im, a, &sec, bounce + 1, w, shaded);                                     /* nested traceRay (closest.cpp:104-117) */
    radiance = vadd(radiance, vscale(c, throughput));
  } else {
    radiance = vadd(radiance, vscale(a->background_color, throughput));
  }
  return radiance;
}

/* ---------------------------------------------------------------- extended modes (no reference counterpart) ---------
 * TraceExt for the RT-unit layout: BASELINE configs 1-3 ask for shadow rays and diffuse bounces, which the reference does not
 * have on either path.  The definition is the one of the raycast layout (oracle/port/rc_shade.c:rc_trace_ext, kept expression
 * for expression; the CUDA side is rt_rtu.cu:shade_ext_b), with this layout's own pieces: the RT unit's traversal
 * (traverse(), i.e. sim/simx/rt_traversal.cpp restated — the only intersection primitive) and the closest-hit shader's
 * surface evaluation (closest.cpp:61-90: normal, uv, material / texture colour).  mode bit0 = shadow rays, bit1 = diffuse. */
typedef struct { v3 I, N, texColor; float reflectivity; } rtu_surface;
static rtu_surface surface_b(const rtu_image *im, const rtu_karg *a, const rtu_ray *ray, const rtu_hit *h) {
  rtu_blas blas; rtu_triex ex; rtu_material mat;
  img_read(im, &blas, a->blas_addr + (uint64_t)h->blasIdx * sizeof(rtu_blas), sizeof(blas));
  img_read(im, &ex, a->triEx_addr + (uint64_t)h->triIdx * sizeof(rtu_triex), sizeof(ex));
  img_read(im, &mat, a->mat_addr + (uint64_t)ex.texId * sizeof(rtu_material), sizeof(mat));
  v3 org = V3(ray->ox, ray->oy, ray->oz), dir = V3(ray->dx, ray->dy, ray->dz);
  rtu_surface s;
  s.I = vadd(org, vscale(dir, h->dist));
  v3 N = vadd(vadd(vscale(ex.N1, h->bx), vscale(ex.N2, h->by)), vscale(ex.N0, h->bz));
  m4 inv; memcpy(inv.c, blas.invTransform, 64);
  m4 invT = m4_transposed3(&inv);
  s.N = vnormalize(xform_vec(N, &invT));
  v2 uv;
  uv.x = ex.uv1.x * h->bx + ex.uv2.x * h->by + ex.uv0.x * h->bz;
  uv.y = ex.uv1.y * h->bx + ex.uv2.y * h->by + ex.uv0.y * h->bz;
  if (mat.diffuse_tex_id >= 0) {
    uint32_t iu = (uint32_t)(uv.x * mat.tex_width), iv = (uint32_t)(uv.y * mat.tex_height);
    iu %= mat.tex_width, iv %= mat.tex_height;
    uint32_t texel;
    img_read(im, &texel, a->tex_addr + mat.tex_offset + 4ull * (iu + iv * mat.tex_width), 4);
    s.texColor = rgb8_to_f(texel);
  } else {
    s.texColor = mat.diffuse;
  }
  s.reflectivity = blas.reflectivity;
  return s;
}
static uint32_t wang_hash_b(uint32_t s) {            /* raytracing/common.h WangHash == raycast/common.h:85-91 */
  s = (s ^ 61) ^ (s >> 16);
  s *= 9;
  s = s ^ (s >> 4);
  s *= 0x27d4eb2d;
  s = s ^ (s >> 15);
  return s;
}
static uint32_t random_int_b(uint32_t *s) {          /* xorshift32 (common.h:93-99) */
  *s ^= *s << 13;
  *s ^= *s >> 17;
  *s ^= *s << 5;
  return *s;
}
static float random_float_b(uint32_t *s) { return random_int_b(s) * 2.3283064365387e-10f; }

static v3 trace_ext_b(const rtu_image *im, const rtu_karg *a, const rtu_ray *ray, uint32_t mode, uint32_t *seed, rtu_work *w,
                      uint64_t *shaded) {
  rtu_ray cur = *ray;
  v3 radiance = V3(0, 0, 0);
  v3 thr = V3(1, 1, 1);
  for (uint32_t bounce = 0; bounce < a->max_depth; ++bounce) {
    rtu_hit hit;
    trace_one(im, &cur, &hit, w);
    if (hit.dist == RC_LARGE_FLOAT) {
      radiance = vadd(radiance, vmul(a->background_color, thr));
      break;
    }
    if (shaded) ++*shaded;
    rtu_surface s = surface_b(im, a, &cur, &hit);
    v3 dir = V3(cur.dx, cur.dy, cur.dz);
    v3 Ns = (vdot(s.N, dir) > 0.0f) ? vneg(s.N) : s.N;
    v3 O = vadd(s.I, vscale(Ns, 0.001f));
    v3 L = vsub(a->light_pos, s.I);
    float dist = vlength(L);
    L = vscale(L, 1.0f / dist);
    float att = 1.0f / (1.0f + dist * 0.1f);
    float NdotL = std_maxf(0.0f, vdot(s.N, L));
    if ((mode & 1u) && NdotL > 0.0f) {
      v3 Lv = vsub(a->light_pos, O);
      float d2 = vlength(Lv);
      v3 D = vscale(Lv, 1.0f / d2);
      rtu_ray sr = {O.x, O.y, O.z, D.x, D.y, D.z};
      rtu_hit sh;
      trace_max(im, &sr, &sh, d2, w);                 /* occluded iff something is hit before the light */
      if (sh.dist < d2) NdotL = 0.0f;
    }
    v3 diffuse = vmul(s.texColor, vadd(a->ambient_color, vscale(sscale(att, a->light_color), NdotL)));
    if (mode & 2u) {
      radiance = vadd(radiance, vmul(thr, diffuse));
      thr = vmul(thr, s.texColor);
      if (bounce + 1 < a->max_depth) {
        float px, py, pz, l2;
        do {
          px = random_float_b(seed) * 2.0f - 1.0f;
          py = random_float_b(seed) * 2.0f - 1.0f;
          pz = random_float_b(seed) * 2.0f - 1.0f;
          l2 = px * px + py * py + pz * pz;
        } while (l2 > 1.0f || l2 < 1e-6f);
        float il = 1.0f / sqrtf(l2);
        v3 D = vadd(Ns, V3(px * il, py * il, pz * il));
        float dl2 = vdot(D, D);
        if (dl2 < 1e-12f) D = Ns;
        else D = vscale(D, 1.0f / sqrtf(dl2));
        cur.ox = O.x, cur.oy = O.y, cur.oz = O.z, cur.dx = D.x, cur.dy = D.y, cur.dz = D.z;
        continue;
      }
      break;
    }
    radiance = vadd(radiance, vscale(vmul(thr, diffuse), 1 - s.reflectivity));
    thr = vscale(thr, s.reflectivity);
    if (s.reflectivity > 0.0f && bounce + 1 < a->max_depth) {
      v3 R = vnormalize(vsub(dir, vscale(sscale(2.0f, s.N), vdot(s.N, dir))));
      v3 so = vadd(s.I, vscale(R, 0.001f));
      cur.ox = so.x, cur.oy = so.y, cur.oz = so.z, cur.dx = R.x, cur.dy = R.y, cur.dz = R.z;
      continue;
    }
    radiance = vadd(radiance, vmul(thr, a->background_color));
    break;
  }
  return radiance;
}

/* ================================================================ C API (same shapes as oracle/ref_harness) ======== */
typedef struct {
  rtu_image im; const rtu_karg *arg; uint32_t camera; uint32_t *out;
  const float *rays; uint64_t n; rtu_hit *hits; uint8_t *flags;
  uint32_t next; pthread_mutex_t mu; rtu_work work; uint64_t shaded, total_rays; int count;
  uint32_t mode;   /* 0 = the shipped shaders; bit0 shadow rays, bit1 diffuse bounces (trace_ext_b) */
} rtu_job;

static void *trace_worker(void *p) {
  rtu_job *j = (rtu_job *)p;
  rtu_work w; memset(&w, 0, sizeof(w));
  for (;;) {
    pthread_mutex_lock(&j->mu);
    uint64_t b = (uint64_t)j->next * 1024; j->next++;
    pthread_mutex_unlock(&j->mu);
    if (b >= j->n) break;
    uint64_t e = b + 1024 < j->n ? b + 1024 : j->n;
    for (uint64_t i = b; i < e; ++i) {
      rtu_ray r = {j->rays[6 * i], j->rays[6 * i + 1], j->rays[6 * i + 2], j->rays[6 * i + 3], j->rays[6 * i + 4], j->rays[6 * i + 5]};
      uint64_t wraps0 = w.wraps;
      trace_one(&j->im, &r, &j->hits[i], (j->count || j->flags) ? &w : NULL);
      if (j->flags) j->flags[i] = (uint8_t)(w.wraps != wraps0);
    }
  }
  pthread_mutex_lock(&j->mu);
  j->work.calls += w.calls, j->work.nodes += w.nodes, j->work.blas += w.blas, j->work.tris += w.tris, j->work.restarts += w.restarts;
  j->work.wraps += w.wraps;
  pthread_mutex_unlock(&j->mu);
  return NULL;
}

static void *render_worker(void *p) {
  rtu_job *j = (rtu_job *)p;
  rtu_work w; memset(&w, 0, sizeof(w));
  uint64_t shaded = 0;
  const rtu_karg *a = j->arg;
  for (;;) {
    pthread_mutex_lock(&j->mu);
    uint32_t y = j->next++;
    pthread_mutex_unlock(&j->mu);
    if (y >= a->dst_height) break;
    for (uint32_t x = 0; x < a->dst_width; ++x) {
      rtu_ray r = j->camera ? gen_pinhole(x, y, a) : gen_fixed(x, y, a);
      v3 c;
      if (j->mode == 0) {
        c = shade(&j->im, a, &r, 0, j->count ? &w : NULL, &shaded);    /* kernel.cpp:66-81: samples overwrite one payload */
      } else {
        uint32_t seed = wang_hash_b((x + y * a->dst_width) * a->samples_per_pixel + 1);
        if (seed == 0) seed = 1;
        c = trace_ext_b(&j->im, a, &r, j->mode, &seed, j->count ? &w : NULL, &shaded);
      }
      j->out[x + y * a->dst_width] = f_to_rgb8(c);
    }
  }
  pthread_mutex_lock(&j->mu);
  j->work.calls += w.calls, j->work.nodes += w.nodes, j->work.blas += w.blas, j->work.tris += w.tris, j->work.restarts += w.restarts;
  j->shaded += shaded;
  pthread_mutex_unlock(&j->mu);
  return NULL;
}

static void run_job(rtu_job *j, void *(*fn)(void *), uint32_t threads) {
  pthread_mutex_init(&j->mu, NULL);
  if (threads <= 1) fn(j);
  else {
    pthread_t *t = (pthread_t *)malloc(sizeof(pthread_t) * threads);
    for (uint32_t i = 0; i < threads; ++i) pthread_create(&t[i], NULL, fn, j);
    for (uint32_t i = 0; i < threads; ++i) pthread_join(t[i], NULL);
    free(t);
  }
  pthread_mutex_destroy(&j->mu);
}

/* work (optional, 6 x u64): traverse calls, nodes read, BLAS entered, triangles tested, restarts, shaded hits (render) or
   size()-1 wrap events (trace_rays) */
void port_rtu_trace_rays(const uint8_t *mem, uint64_t base, uint64_t size, uint32_t tlas, uint32_t blas, uint32_t qbvh, uint32_t tri,
                         const float *rays, uint64_t n, void *hits, uint32_t threads, uint64_t *work, uint8_t *wrap_flags) {
  rtu_job j; memset(&j, 0, sizeof(j));
  j.flags = wrap_flags;   /* optional, one byte per ray: 1 = this ray reaches the reference's size()-1 wrap (rt_traversal.cpp:81) */
  j.im.mem = mem, j.im.base = base, j.im.size = size, j.im.tlas = tlas, j.im.blas = blas, j.im.qbvh = qbvh, j.im.tri = tri;
  j.rays = rays, j.n = n, j.hits = (rtu_hit *)hits, j.count = work != NULL;
  run_job(&j, trace_worker, threads);
  if (work) work[0] = j.work.calls, work[1] = j.work.nodes, work[2] = j.work.blas, work[3] = j.work.tris, work[4] = j.work.restarts, work[5] = j.work.wraps;
}

void port_rtuh_render(const uint8_t *mem, uint64_t base, uint64_t size, const void *karg, const uint32_t *dcr, uint32_t camera,
                      uint32_t *out, uint32_t threads, uint64_t *work) {
  rtu_job j; memset(&j, 0, sizeof(j));
  j.im.mem = mem, j.im.base = base, j.im.size = size, j.im.tlas = dcr[0], j.im.blas = dcr[1], j.im.qbvh = dcr[2], j.im.tri = dcr[3];
  j.arg = (const rtu_karg *)karg, j.camera = camera, j.out = out, j.count = work != NULL;
  run_job(&j, render_worker, threads);
  if (work) work[0] = j.work.calls, work[1] = j.work.nodes, work[2] = j.work.blas, work[3] = j.work.tris, work[4] = j.work.restarts, work[5] = j.shaded;
}

/* the same frame with the extended modes (mode bit0 shadow rays, bit1 diffuse bounces) */
void port_rtuh_render_ext(const uint8_t *mem, uint64_t base, uint64_t size, const void *karg, const uint32_t *dcr, uint32_t camera,
                          uint32_t mode, uint32_t *out, uint32_t threads, uint64_t *work) {
  rtu_job j; memset(&j, 0, sizeof(j));
  j.im.mem = mem, j.im.base = base, j.im.size = size, j.im.tlas = dcr[0], j.im.blas = dcr[1], j.im.qbvh = dcr[2], j.im.tri = dcr[3];
  j.arg = (const rtu_karg *)karg, j.camera = camera, j.out = out, j.count = work != NULL, j.mode = mode;
  run_job(&j, render_worker, threads);
  if (work) work[0] = j.work.calls, work[1] = j.work.nodes, work[2] = j.work.blas, work[3] = j.work.tris, work[4] = j.work.restarts, work[5] = j.shaded;
}

void port_rtuh_primary_rays(const void *karg, uint32_t camera, float *rays) {
  const rtu_karg *a = (const rtu_karg *)karg;
  for (uint32_t y = 0; y < a->dst_height; ++y)
    for (uint32_t x = 0; x < a->dst_width; ++x) {
      rtu_ray r = camera ? gen_pinhole(x, y, a) : gen_fixed(x, y, a);
      memcpy(rays + 6 * ((uint64_t)y * a->dst_width + x), &r, 24);
    }
}
