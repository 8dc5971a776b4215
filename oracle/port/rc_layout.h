/* oracle/port/rc_layout.h — TEST INFRASTRUCTURE ONLY.
 * Byte layouts of the reference's raycast buffers (tests/regression/raycast/common.h:19-147, geometry.h:1401-1405),
 * declared for the C oracle and size-checked against SURVEY.md 8(a). */
#ifndef RC_LAYOUT_H
#define RC_LAYOUT_H
#include <stddef.h>
#include "rc_math.h"

typedef struct { v3 v0, v1, v2; } rc_tri;                                   /* tri_t, 36 B */
typedef struct { v3 N0, N1, N2; v2 uv0, uv1, uv2; } rc_tri_ex;              /* tri_ex_t, 60 B */
typedef struct { float dist; v3 bcoords; uint32_t blasIdx, triIdx; } rc_hit; /* ray_hit_t, 24 B */
typedef struct { v3 aabbMin; uint32_t leftFirst; v3 aabbMax; uint32_t triCount; } rc_bvh_node;  /* bvh_node_t, 32 B */
typedef struct {                                                            /* blas_node_t, 160 B */
  m4 transform, invTransform;
  uint32_t bvh_offset, _pad0;
  uint64_t tex_offset;
  uint32_t tex_width, tex_height;
  float reflectivity;
  uint32_t _pad1;
} rc_blas_node;
typedef struct { v3 aabbMin; uint32_t leftRight; v3 aabbMax; uint32_t blasIdx; } rc_tlas_node;   /* tlas_node_t, 32 B */
typedef struct {                                                            /* kernel_arg_t, 192 B */
  uint32_t dst_width, dst_height;
  uint64_t dst_addr, tri_addr, triEx_addr, triIdx_addr, tex_addr, bvh_addr, blas_addr, tlas_addr;
  uint32_t tlas_root;
  v3 camera_pos, camera_forward, camera_right, camera_up;
  v2 viewplane;
  uint32_t samples_per_pixel, max_depth;
  v3 light_pos, light_color, ambient_color, background_color;
  uint32_t _tail;
} rc_kernel_arg;
typedef struct { v3 orig, dir; } rc_ray;

_Static_assert(sizeof(rc_tri) == 36 && sizeof(rc_tri_ex) == 60 && sizeof(rc_hit) == 24, "tri layouts");
_Static_assert(sizeof(rc_bvh_node) == 32 && sizeof(rc_tlas_node) == 32 && sizeof(rc_blas_node) == 160, "node layouts");
_Static_assert(offsetof(rc_blas_node, bvh_offset) == 128 && offsetof(rc_blas_node, tex_offset) == 136 &&
               offsetof(rc_blas_node, reflectivity) == 152, "blas layout");
_Static_assert(sizeof(rc_kernel_arg) == 192 && offsetof(rc_kernel_arg, tlas_root) == 72 && offsetof(rc_kernel_arg, viewplane) == 124 &&
               offsetof(rc_kernel_arg, light_pos) == 140 && offsetof(rc_kernel_arg, background_color) == 176, "kernel_arg_t");

/* work counters (SURVEY 8d "algorithmic work per ray"); NULL = do not count */
typedef struct { uint64_t rays, box_tests, tri_tests, blas_entered, shaded_hits; } rc_counters;

/* rc_render.c */
rc_ray rc_generate_ray(uint32_t x, uint32_t y, const rc_kernel_arg *arg);
void rc_tlas_intersect(const rc_ray *ray, const rc_kernel_arg *arg, rc_hit *hit, rc_counters *cnt);
v3 rc_trace(const rc_ray *ray, const rc_kernel_arg *arg, rc_counters *cnt);
v3 rc_trace_ext(const rc_ray *ray, const rc_kernel_arg *arg, uint32_t mode, uint32_t *seed, rc_counters *cnt);
uint32_t rc_rgb32f_to_rgb8(v3 c);
uint32_t rc_wang_hash(uint32_t s);
#endif
