/* oracle/port/rc_api.c — TEST INFRASTRUCTURE ONLY (CPU oracle; never linked into the product).
 *
 * C entry points of the oracle port, with the same surface as the reference harness (oracle/ref_harness/ref_raycast.cpp)
 * so oracle/pyoracle.py binds either with one prefix switch:
 *   scene glue  = Scene::init (tests/regression/raycast/scene.cpp:17-82) around rc_bvh_build / rc_tlas_build
 *   port_setup  = Tracer::setup (tracer.cpp:152-196) + Scene::computeFramingCamera (scene.cpp:138-172)
 *   port_render = Tracer::render (tracer.cpp:249-263), optionally row-parallel over pthreads (no arithmetic change)
 */
#include <pthread.h>
#include <stdlib.h>
#include <unistd.h>
#include "rc_scene.h"

void rc_tlas_build(rc_scene *sc);

void *port_scene_create(uint32_t n_meshes, const uint32_t *tri_counts, const float *tris, const float *triex, const uint32_t *tex_w,
                        const uint32_t *tex_h, const uint32_t *texels, const float *reflectivity, const float *transforms) {
  rc_scene *sc = (rc_scene *)calloc(1, sizeof(rc_scene));
  uint64_t num_tris = 0, tex_px = 0;
  for (uint32_t i = 0; i < n_meshes; ++i) { num_tris += tri_counts[i]; tex_px += (uint64_t)tex_w[i] * tex_h[i]; }
  sc->n_meshes = n_meshes;
  sc->n_tris = num_tris;
  sc->tri_counts = (uint32_t *)malloc(sizeof(uint32_t) * (n_meshes ? n_meshes : 1));
  sc->bvh_counts = (uint32_t *)calloc(n_meshes ? n_meshes : 1, sizeof(uint32_t));
  sc->tri = (rc_tri *)malloc(sizeof(rc_tri) * (num_tris ? num_tris : 1));
  sc->triEx = (rc_tri_ex *)malloc(sizeof(rc_tri_ex) * (num_tris ? num_tris : 1));
  sc->triIdx = (uint32_t *)malloc(4 * (num_tris ? num_tris : 1));
  sc->centroids = (v3 *)malloc(sizeof(v3) * (num_tris ? num_tris : 1));
  sc->bvh = (rc_bvh_node *)calloc(num_tris ? num_tris * 2 : 1, sizeof(rc_bvh_node));   /* std::vector::resize zero-fills */
  sc->blas = (rc_blas_node *)calloc(n_meshes ? n_meshes : 1, sizeof(rc_blas_node));
  sc->tlas = (rc_tlas_node *)calloc(n_meshes ? 2 * n_meshes - 1 : 1, sizeof(rc_tlas_node));
  sc->tlas_indices = (uint32_t *)calloc(n_meshes ? n_meshes : 1, sizeof(uint32_t));
  sc->tex_bytes = tex_px * 4;
  sc->tex = (uint8_t *)malloc(sc->tex_bytes ? sc->tex_bytes : 1);
  memcpy(sc->tri, tris, num_tris * sizeof(rc_tri));
  memcpy(sc->triEx, triex, num_tris * sizeof(rc_tri_ex));
  memcpy(sc->tex, texels, sc->tex_bytes);
  uint32_t bvh_offset = 0, tri_offset = 0;
  uint64_t tex_offset = 0;
  for (uint32_t i = 0; i < n_meshes; ++i) {
    sc->tri_counts[i] = tri_counts[i];
    for (uint32_t j = 0; j < tri_counts[i]; ++j) {                  /* scene.cpp:48-53 */
      const rc_tri *t = &sc->tri[tri_offset + j];
      sc->triIdx[tri_offset + j] = tri_offset + j;
      sc->centroids[tri_offset + j] = vdivs(vadd(vadd(t->v0, t->v1), t->v2), 3);
    }
    rc_blas_node *b = &sc->blas[i];                                 /* scene.cpp:55-63 */
    b->bvh_offset = bvh_offset;
    b->tex_offset = tex_offset;
    b->transform = m4_identity();
    b->invTransform = m4_identity();
    b->tex_width = tex_w[i];
    b->tex_height = tex_h[i];
    b->reflectivity = reflectivity[i];
    /* BVH over the GLOBAL triangle/centroid arrays, writing this mesh's slice of nodes and indices (scene.cpp:66) */
    sc->bvh_counts[i] = rc_bvh_build(sc->tri, sc->centroids, tri_counts[i], sc->bvh + bvh_offset, sc->triIdx + tri_offset);
    bvh_offset += sc->bvh_counts[i];
    tri_offset += tri_counts[i];
    tex_offset += (uint64_t)tex_w[i] * tex_h[i] * 4;
  }
  if (transforms) {
    for (uint32_t i = 0; i < n_meshes; ++i) {                       /* blas_node_t::applyTransform, common.h:56-59 */
      m4 T;
      memcpy(T.c, transforms + 16 * i, 64);
      sc->blas[i].transform = m4_mul(&T, &sc->blas[i].transform);
      sc->blas[i].invTransform = m4_inverted(&sc->blas[i].transform);
    }
  }
  return sc;
}

void port_scene_apply_transform(void *h, const float *m16) {        /* Scene::applyTransform, scene.cpp:216-220 */
  rc_scene *sc = (rc_scene *)h;
  m4 T;
  memcpy(T.c, m16, 64);
  for (uint32_t i = 0; i < sc->n_meshes; ++i) {
    sc->blas[i].transform = m4_mul(&T, &sc->blas[i].transform);
    sc->blas[i].invTransform = m4_inverted(&sc->blas[i].transform);
  }
}

void port_scene_build(void *h) { rc_tlas_build((rc_scene *)h); }

void port_scene_view(void *h, rc_scene_view *v) {
  rc_scene *sc = (rc_scene *)h;
  v->tri = sc->tri;       v->tri_bytes = sc->n_tris * sizeof(rc_tri);
  v->triEx = sc->triEx;   v->triEx_bytes = sc->n_tris * sizeof(rc_tri_ex);
  v->triIdx = sc->triIdx; v->triIdx_bytes = sc->n_tris * 4;
  v->bvh = sc->bvh;       v->bvh_bytes = sc->n_tris * 2 * sizeof(rc_bvh_node);
  v->blas = sc->blas;     v->blas_bytes = (uint64_t)sc->n_meshes * sizeof(rc_blas_node);
  v->tlas = sc->tlas;     v->tlas_bytes = (uint64_t)(sc->n_meshes ? 2 * sc->n_meshes - 1 : 0) * sizeof(rc_tlas_node);
  v->tex = sc->tex;       v->tex_bytes = sc->tex_bytes;
  v->tlas_root = sc->tlas_root;
  uint32_t used = 0;
  for (uint32_t i = 0; i < sc->n_meshes; ++i) used += sc->bvh_counts[i];
  v->bvh_nodes_used = used;
}

void port_scene_free(void *h) {
  rc_scene *sc = (rc_scene *)h;
  if (!sc) return;
  free(sc->tri_counts); free(sc->bvh_counts); free(sc->tri); free(sc->triEx); free(sc->triIdx); free(sc->centroids);
  free(sc->bvh); free(sc->blas); free(sc->tlas); free(sc->tlas_indices); free(sc->tex);
  free(sc);
}

/* Tracer::setup — tracer.cpp:152-196 (camera_vfov in DEGREES: framing uses vfov*DEG2RAD, the viewplane uses the raw
 * value in tan(), exactly as the reference does) */
void port_setup(void *h, void *arg_out, uint32_t width, uint32_t height, uint32_t spp, uint32_t max_depth, float camera_vfov, float zoom,
                int rotate_scene, const float *light_pos, const float *light_color, const float *ambient, const float *background) {
  rc_scene *sc = (rc_scene *)h;
  rc_kernel_arg a;
  memset(&a, 0, sizeof(a));
  a.dst_width = width; a.dst_height = height;
  a.samples_per_pixel = spp; a.max_depth = max_depth;
  if (rotate_scene) {
    m4 rx = m4_rotate_x(-RC_PI / 4), ry = m4_rotate_y(RC_PI / 4);
    m4 T = m4_mul(&rx, &ry);
    port_scene_apply_transform(sc, T.c);
  }
  rc_tlas_build(sc);
  a.tlas_root = sc->tlas_root;
  {                                                              /* Scene::computeFramingCamera, scene.cpp:138-172 */
    float vfov = camera_vfov * RC_DEG2RAD;
    v3 bmin = v3s(+RC_LARGE_FLOAT), bmax = v3s(-RC_LARGE_FLOAT);
    uint32_t n_nodes = sc->n_meshes ? 2 * sc->n_meshes - 1 : 0;
    for (uint32_t i = 0; i < n_nodes; ++i) {
      const rc_tlas_node *n = &sc->tlas[i];
      bmin.x = std_minf(bmin.x, n->aabbMin.x); bmin.y = std_minf(bmin.y, n->aabbMin.y); bmin.z = std_minf(bmin.z, n->aabbMin.z);
      bmax.x = std_maxf(bmax.x, n->aabbMax.x); bmax.y = std_maxf(bmax.y, n->aabbMax.y); bmax.z = std_maxf(bmax.z, n->aabbMax.z);
    }
    v3 center = vscale(vadd(bmin, bmax), 0.5f);
    float radius = vlength(vsub(bmax, center));
    v3 target = center, cup = V3(0.0f, 1.0f, 0.0f);
    float distance = radius / tanf(vfov);
    distance *= zoom;
    v3 pos = vsub(center, vscale(V3(0.0f, 0.0f, 1.0f), distance));
    v3 forward = vnormalize(vsub(target, pos));                  /* tracer.cpp:168-175 */
    v3 right = vnormalize(vcross(forward, cup));
    v3 up = vcross(right, forward);
    a.camera_pos = pos; a.camera_forward = forward; a.camera_right = right; a.camera_up = up;
  }
  {
    float aspect_ratio = (float)width / height;                  /* tracer.cpp:179-183 */
    float viewport_height = 2.0f * tanf(camera_vfov * 0.5f);
    float viewport_width = viewport_height * aspect_ratio;
    a.viewplane.x = viewport_width; a.viewplane.y = viewport_height;
  }
  a.light_pos = V3(light_pos[0], light_pos[1], light_pos[2]);
  a.light_color = V3(light_color[0], light_color[1], light_color[2]);
  a.ambient_color = V3(ambient[0], ambient[1], ambient[2]);
  a.background_color = V3(background[0], background[1], background[2]);
  a.tri_addr = (uint64_t)(uintptr_t)sc->tri;                     /* host pointers, as the use_cpu_ path (tracer.cpp:102-109) */
  a.triEx_addr = (uint64_t)(uintptr_t)sc->triEx;
  a.triIdx_addr = (uint64_t)(uintptr_t)sc->triIdx;
  a.bvh_addr = (uint64_t)(uintptr_t)sc->bvh;
  a.tlas_addr = (uint64_t)(uintptr_t)sc->tlas;
  a.blas_addr = (uint64_t)(uintptr_t)sc->blas;
  a.tex_addr = (uint64_t)(uintptr_t)sc->tex;
  memcpy(arg_out, &a, sizeof(a));
}
