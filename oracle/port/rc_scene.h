/* oracle/port/rc_scene.h — TEST INFRASTRUCTURE ONLY. */
#ifndef RC_SCENE_H
#define RC_SCENE_H
#include "rc_layout.h"

uint32_t rc_bvh_build(const rc_tri *tris, const v3 *centroids, uint32_t triCount, rc_bvh_node *nodes, uint32_t *triIdx);

typedef struct {
  uint32_t n_meshes;
  uint64_t n_tris;
  uint32_t *tri_counts;
  rc_tri *tri;
  rc_tri_ex *triEx;
  uint32_t *triIdx;
  v3 *centroids;
  rc_bvh_node *bvh;        /* 2 * n_tris nodes allocated, as scene.cpp:32 */
  uint32_t *bvh_counts;    /* nodes used per mesh */
  rc_blas_node *blas;
  rc_tlas_node *tlas;      /* 2 * n_meshes - 1 */
  uint32_t *tlas_indices;
  uint32_t tlas_root;
  uint8_t *tex;
  uint64_t tex_bytes;
} rc_scene;

/* the same view struct as oracle/ref_harness/ref_raycast.cpp:ref_scene_view_t */
typedef struct {
  const void *tri;    uint64_t tri_bytes;
  const void *triEx;  uint64_t triEx_bytes;
  const void *triIdx; uint64_t triIdx_bytes;
  const void *bvh;    uint64_t bvh_bytes;
  const void *blas;   uint64_t blas_bytes;
  const void *tlas;   uint64_t tlas_bytes;
  const void *tex;    uint64_t tex_bytes;
  uint32_t tlas_root;
  uint32_t bvh_nodes_used;
} rc_scene_view;
#endif
