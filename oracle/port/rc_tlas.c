/* oracle/port/rc_tlas.c — TEST INFRASTRUCTURE ONLY (CPU oracle; never linked into the product).
 * TLAS::build / buildRecursive / partition restated in C (tests/regression/raycast/bvh.cpp:181-346). */
#include <stdlib.h>
#include "rc_scene.h"

#define BINS 8

static inline float centroid_of(const rc_tlas_node *n, int axis) { return (v3get(n->aabbMin, axis) + v3get(n->aabbMax, axis)) / 2; }

static uint32_t tlas_partition(rc_scene *sc, int start, int end, int axis, float splitPos) {   /* bvh.cpp:308-346 */
  int left = start, right = end;
  uint32_t *idx = sc->tlas_indices;
  while (left <= right) {
    while (left <= end) {
      if (centroid_of(&sc->tlas[idx[left]], axis) < splitPos) left++;
      else break;
    }
    while (right >= start) {
      if (centroid_of(&sc->tlas[idx[right]], axis) >= splitPos) right--;
      else break;
    }
    if (left < right) {
      uint32_t t = idx[left];
      idx[left] = idx[right];
      idx[right] = t;
      left++;
      right--;
    }
  }
  if (right < start) return (uint32_t)end;
  if (left > end) return (uint32_t)start;
  return (uint32_t)right;
}

static int cmp_float(const void *a, const void *b) {
  float x = *(const float *)a, y = *(const float *)b;
  return (x > y) - (x < y);
}

static uint32_t tlas_build_rec(rc_scene *sc, uint32_t start, uint32_t end, uint32_t *nextInternal) {   /* bvh.cpp:214-306 */
  uint32_t *idx = sc->tlas_indices;
  if (start == end) return idx[start];
  v3 aabbMin = sc->tlas[idx[start]].aabbMin, aabbMax = sc->tlas[idx[start]].aabbMax;
  for (uint32_t i = start + 1; i <= end; ++i) {
    aabbMin = vmin(aabbMin, sc->tlas[idx[i]].aabbMin);
    aabbMax = vmax(aabbMax, sc->tlas[idx[i]].aabbMax);
  }
  int splitAxis = 0;
  float splitPos = 0.0f;
  float bestCost = INFINITY;
  v3 extent = vsub(aabbMax, aabbMin);
  for (uint32_t axis = 0; axis < 3; ++axis) {
    if (v3get(extent, axis) <= 0) continue;
    float binWidth = v3get(extent, axis) / BINS;
    for (uint32_t i = 1; i < BINS; ++i) {
      float candidatePos = v3get(aabbMin, axis) + i * binWidth;
      uint32_t leftTris = 0, rightTris = 0;
      v3 leftMin = v3s(RC_LARGE_FLOAT), leftMax = v3s(-RC_LARGE_FLOAT);
      v3 rightMin = v3s(RC_LARGE_FLOAT), rightMax = v3s(-RC_LARGE_FLOAT);
      for (uint32_t j = start; j <= end; ++j) {
        const rc_tlas_node *node = &sc->tlas[idx[j]];
        if (centroid_of(node, (int)axis) < candidatePos) {
          leftMin = vmin(leftMin, node->aabbMin);
          leftMax = vmax(leftMax, node->aabbMax);
          leftTris += sc->tri_counts[node->blasIdx];
        } else {
          rightMin = vmin(rightMin, node->aabbMin);
          rightMax = vmax(rightMax, node->aabbMax);
          rightTris += sc->tri_counts[node->blasIdx];
        }
      }
      if (leftTris == 0 || rightTris == 0) continue;
      float leftArea = surface_area(leftMin, leftMax);
      float rightArea = surface_area(rightMin, rightMax);
      float cost = leftArea * leftTris + rightArea * rightTris;
      if (cost < bestCost) {
        bestCost = cost;
        splitAxis = (int)axis;
        splitPos = candidatePos;
      }
    }
  }
  if (bestCost == INFINITY) {   /* median fallback */
    splitAxis = (extent.x > extent.y) ? ((extent.x > extent.z) ? 0 : 2) : ((extent.y > extent.z) ? 1 : 2);
    uint32_t n = end - start + 1;
    float *cs = (float *)malloc(n * sizeof(float));
    for (uint32_t i = start; i <= end; ++i) {
      const rc_tlas_node *node = &sc->tlas[idx[i]];
      cs[i - start] = (v3get(node->aabbMin, splitAxis) + v3get(node->aabbMax, splitAxis)) * 0.5f;
    }
    qsort(cs, n, sizeof(float), cmp_float);
    splitPos = cs[n / 2];
    free(cs);
  }
  uint32_t mid = tlas_partition(sc, (int)start, (int)end, splitAxis, splitPos);
  if (mid == start || mid == end) mid = (start + end) / 2;
  uint32_t leftChild = tlas_build_rec(sc, start, mid, nextInternal);
  uint32_t rightChild = tlas_build_rec(sc, mid + 1, end, nextInternal);
  uint32_t nodeIndex = (*nextInternal)++;
  rc_tlas_node *node = &sc->tlas[nodeIndex];
  node->leftRight = (rightChild << 16) | leftChild;   /* common.h:72-74 */
  node->aabbMin = aabbMin;
  node->aabbMax = aabbMax;
  return nodeIndex;
}

void rc_tlas_build(rc_scene *sc) {   /* bvh.cpp:181-212 */
  if (sc->n_meshes == 0) return;
  uint32_t bvh_offset = 0;
  for (uint32_t i = 0; i < sc->n_meshes; ++i) {
    const rc_bvh_node *root = &sc->bvh[sc->blas[i].bvh_offset];
    (void)bvh_offset;
    v3 mn = root->aabbMin, mx = root->aabbMax;
    v3 bmin = v3s(RC_LARGE_FLOAT), bmax = v3s(-RC_LARGE_FLOAT);
    for (int c = 0; c < 8; ++c) {
      v3 pos = V3(c & 1 ? mx.x : mn.x, c & 2 ? mx.y : mn.y, c & 4 ? mx.z : mn.z);
      v3 p = xform_pos(pos, &sc->blas[i].transform);
      bmin = vmin(bmin, p);
      bmax = vmax(bmax, p);
    }
    sc->tlas[i].aabbMin = bmin;
    sc->tlas[i].aabbMax = bmax;
    sc->tlas[i].blasIdx = i;
    sc->tlas[i].leftRight = 0;
    sc->tlas_indices[i] = i;
  }
  uint32_t next = sc->n_meshes;
  sc->tlas_root = tlas_build_rec(sc, 0, sc->n_meshes - 1, &next);
}
