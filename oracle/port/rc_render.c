/* oracle/port/rc_render.c — TEST INFRASTRUCTURE ONLY (CPU oracle; never linked into the product).
 *
 * Plain-C restatement of the reference's per-ray hot path (tests/regression/raycast/render.h and the two
 * ray_t::intersect overloads of geometry.h).  Same operand order, same comparisons, same traversal order, so the
 * results are bit-identical to the reference's own build (checked against oracle/_ref by tests/test_oracle.py).
 */
#include "rc_layout.h"

/* ray_t::intersect(tri) — geometry.h:1416-1445 (Moller-Trumbore, two-sided) */
static inline int ray_tri(const rc_ray *r, const rc_tri *tri, float *dist, v3 *bc) {
  v3 edge1 = vsub(tri->v1, tri->v0);
  v3 edge2 = vsub(tri->v2, tri->v0);
  v3 h = vcross(r->dir, edge2);
  float a = vdot(edge1, h);
  if (fabs(a) < RC_EPSILON) return 0;
  float f = 1 / a;
  v3 s = vsub(r->orig, tri->v0);
  float w1 = f * vdot(s, h);
  if (w1 < 0 || w1 > 1) return 0;
  v3 q = vcross(s, edge1);
  float w2 = f * vdot(r->dir, q);
  if (w2 < 0 || w1 + w2 > 1) return 0;
  float t = f * vdot(edge2, q);
  if (t <= RC_EPSILON) return 0;
  *dist = t;
  bc->x = w1;
  bc->y = w2;
  bc->z = 1 - w1 - w2;
  return 1;
}

/* ray_t::intersect(aabbMin, aabbMax) — geometry.h:1447-1468 (slab test; reciprocal recomputed per call) */
static inline float ray_aabb(const rc_ray *r, v3 mn, v3 mx) {
  float idir_x = 1.0f / r->dir.x;
  float idir_y = 1.0f / r->dir.y;
  float idir_z = 1.0f / r->dir.z;
  float tx1 = (mn.x - r->orig.x) * idir_x;
  float tx2 = (mx.x - r->orig.x) * idir_x;
  float tmin = std_minf(tx1, tx2);
  float tmax = std_maxf(tx1, tx2);
  float ty1 = (mn.y - r->orig.y) * idir_y;
  float ty2 = (mx.y - r->orig.y) * idir_y;
  tmin = std_maxf(tmin, std_minf(ty1, ty2)), tmax = std_minf(tmax, std_maxf(ty1, ty2));
  float tz1 = (mn.z - r->orig.z) * idir_z;
  float tz2 = (mx.z - r->orig.z) * idir_z;
  tmin = std_maxf(tmin, std_minf(tz1, tz2)), tmax = std_minf(tmax, std_maxf(tz1, tz2));
  if (tmax < tmin || tmax <= 0) return RC_LARGE_FLOAT;
  return tmin;
}

/* BVHIntersect — render.h:74-126.  Note the order: when both children are hit the NEARER one is pushed first
 * (swap at :113-115 when dLeft < dRight), so the FARTHER child is popped and visited first. */
static void bvh_intersect(const rc_ray *ray, uint32_t blasIdx, const rc_bvh_node *bvh, const uint32_t *triIdxBuf, const rc_tri *triBuf,
                          rc_hit *hit, rc_counters *cnt) {
  uint32_t stack[RC_STACK];
  uint32_t sp = 0;
  stack[sp++] = 0;
  while (sp != 0) {
    uint32_t nodeIdx = stack[--sp];
    const rc_bvh_node *node = &bvh[nodeIdx];
    if (node->triCount != 0) {
      for (uint32_t i = 0; i < node->triCount; ++i) {
        uint32_t triIdx = triIdxBuf[node->leftFirst + i];
        float dist;
        v3 bc;
        if (cnt) cnt->tri_tests++;
        if (ray_tri(ray, &triBuf[triIdx], &dist, &bc) && dist < hit->dist) {
          hit->dist = dist;
          hit->bcoords = bc;
          hit->blasIdx = blasIdx;
          hit->triIdx = triIdx;
        }
      }
    } else {
      uint32_t left = node->leftFirst;
      uint32_t right = left + 1;
      float dLeft = ray_aabb(ray, bvh[left].aabbMin, bvh[left].aabbMax);
      float dRight = ray_aabb(ray, bvh[right].aabbMin, bvh[right].aabbMax);
      if (cnt) cnt->box_tests += 2;
      int hitLeft = (dLeft != RC_LARGE_FLOAT) && (dLeft < hit->dist);
      int hitRight = (dRight != RC_LARGE_FLOAT) && (dRight < hit->dist);
      if (hitLeft && hitRight) {
        if (dLeft < dRight) { uint32_t t = left; left = right; right = t; }
        stack[sp++] = right;
        stack[sp++] = left;
      } else if (hitLeft) {
        stack[sp++] = left;
      } else if (hitRight) {
        stack[sp++] = right;
      }
    }
  }
}

/* BLASIntersect — render.h:128-141; ray_t::transform geometry.h:1411-1414 (direction first, not renormalised) */
static void blas_intersect(const rc_ray *ray, uint32_t blasIdx, const rc_blas_node *blasBuf, const rc_bvh_node *bvhBuf,
                           const uint32_t *triIdxBuf, const rc_tri *triBuf, rc_hit *hit, rc_counters *cnt) {
  const rc_blas_node *blas = &blasBuf[blasIdx];
  rc_ray backup = *ray;
  backup.dir = xform_vec(backup.dir, &blas->invTransform);
  backup.orig = xform_pos(backup.orig, &blas->invTransform);
  if (cnt) cnt->blas_entered++;
  bvh_intersect(&backup, blasIdx, bvhBuf + blas->bvh_offset, triIdxBuf, triBuf, hit, cnt);
}

/* TLASIntersect — render.h:143-188 (near child visited first here: swap when dLeft > dRight) */
void rc_tlas_intersect(const rc_ray *ray, const rc_kernel_arg *arg, rc_hit *hit, rc_counters *cnt) {
  const rc_tlas_node *tlas = (const rc_tlas_node *)(uintptr_t)arg->tlas_addr;
  const rc_blas_node *blasBuf = (const rc_blas_node *)(uintptr_t)arg->blas_addr;
  const rc_bvh_node *bvhBuf = (const rc_bvh_node *)(uintptr_t)arg->bvh_addr;
  const uint32_t *triIdxBuf = (const uint32_t *)(uintptr_t)arg->triIdx_addr;
  const rc_tri *triBuf = (const rc_tri *)(uintptr_t)arg->tri_addr;
  uint32_t stack[RC_STACK];
  uint32_t sp = 0;
  stack[sp++] = arg->tlas_root;
  if (cnt) cnt->rays++;
  while (sp != 0) {
    uint32_t nodeIdx = stack[--sp];
    const rc_tlas_node *node = &tlas[nodeIdx];
    if (node->leftRight == 0) {
      blas_intersect(ray, node->blasIdx, blasBuf, bvhBuf, triIdxBuf, triBuf, hit, cnt);
    } else {
      uint32_t left = node->leftRight & 0xFFFF;
      uint32_t right = node->leftRight >> 16;
      float dLeft = ray_aabb(ray, tlas[left].aabbMin, tlas[left].aabbMax);
      float dRight = ray_aabb(ray, tlas[right].aabbMin, tlas[right].aabbMax);
      if (cnt) cnt->box_tests += 2;
      int hitLeft = (dLeft != RC_LARGE_FLOAT) && (dLeft < hit->dist);
      int hitRight = (dRight != RC_LARGE_FLOAT) && (dRight < hit->dist);
      if (hitLeft && hitRight) {
        if (dLeft > dRight) { uint32_t t = left; left = right; right = t; }
        stack[sp++] = right;
        stack[sp++] = left;
      } else if (hitLeft) {
        stack[sp++] = left;
      } else if (hitRight) {
        stack[sp++] = right;
      }
    }
  }
}

/* GenerateRay — render.h:190-208.  "- 0.5" is a double literal: the subtraction is done in double, then narrowed. */
rc_ray rc_generate_ray(uint32_t x, uint32_t y, const rc_kernel_arg *arg) {
  float x_ndc = (x + 0.5f) / arg->dst_width - 0.5;
  float y_ndc = (y + 0.5f) / arg->dst_height - 0.5;
  float x_vp = x_ndc * arg->viewplane.x;
  float y_vp = y_ndc * arg->viewplane.y;
  v3 pt_cam = vadd(vadd(sscale(x_vp, arg->camera_right), sscale(y_vp, arg->camera_up)), arg->camera_forward);
  v3 pt_w = vadd(pt_cam, arg->camera_pos);
  v3 dir = vnormalize(vsub(pt_w, arg->camera_pos));
  rc_ray r = {arg->camera_pos, dir};
  return r;
}
