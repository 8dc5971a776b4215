/* oracle/port/rc_bvh.c — TEST INFRASTRUCTURE ONLY (CPU oracle; never linked into the product).
 *
 * Plain-C restatement of the reference's host-side acceleration-structure builders, so that the oracle can produce the
 * exact buffers a Vortex raycast host uploads: binned-SAH BVH2 per mesh (tests/regression/raycast/bvh.cpp:23-161),
 * SAH TLAS over instances (bvh.cpp:165-346), mat4 inverse (geometry.h:1149-1192).
 */
#include <stdlib.h>
#include "rc_scene.h"

#define BINS 8 /* bvh.cpp:6 */

/* mat4_t::inverted — geometry.h:1149-1192 (MESA cofactor expansion).  Each entry is the left-to-right sum of six signed
 * triple products (c[i]*c[j])*c[k]; negation and "x - y == x + (-y)" are exact, so a signed table reproduces the rounding. */
static const signed char kCof[16][6][4] = {
    {{+1, 5, 10, 15}, {-1, 5, 11, 14}, {-1, 9, 6, 15}, {+1, 9, 7, 14}, {+1, 13, 6, 11}, {-1, 13, 7, 10}},
    {{-1, 1, 10, 15}, {+1, 1, 11, 14}, {+1, 9, 2, 15}, {-1, 9, 3, 14}, {-1, 13, 2, 11}, {+1, 13, 3, 10}},
    {{+1, 1, 6, 15}, {-1, 1, 7, 14}, {-1, 5, 2, 15}, {+1, 5, 3, 14}, {+1, 13, 2, 7}, {-1, 13, 3, 6}},
    {{-1, 1, 6, 11}, {+1, 1, 7, 10}, {+1, 5, 2, 11}, {-1, 5, 3, 10}, {-1, 9, 2, 7}, {+1, 9, 3, 6}},
    {{-1, 4, 10, 15}, {+1, 4, 11, 14}, {+1, 8, 6, 15}, {-1, 8, 7, 14}, {-1, 12, 6, 11}, {+1, 12, 7, 10}},
    {{+1, 0, 10, 15}, {-1, 0, 11, 14}, {-1, 8, 2, 15}, {+1, 8, 3, 14}, {+1, 12, 2, 11}, {-1, 12, 3, 10}},
    {{-1, 0, 6, 15}, {+1, 0, 7, 14}, {+1, 4, 2, 15}, {-1, 4, 3, 14}, {-1, 12, 2, 7}, {+1, 12, 3, 6}},
    {{+1, 0, 6, 11}, {-1, 0, 7, 10}, {-1, 4, 2, 11}, {+1, 4, 3, 10}, {+1, 8, 2, 7}, {-1, 8, 3, 6}},
    {{+1, 4, 9, 15}, {-1, 4, 11, 13}, {-1, 8, 5, 15}, {+1, 8, 7, 13}, {+1, 12, 5, 11}, {-1, 12, 7, 9}},
    {{-1, 0, 9, 15}, {+1, 0, 11, 13}, {+1, 8, 1, 15}, {-1, 8, 3, 13}, {-1, 12, 1, 11}, {+1, 12, 3, 9}},
    {{+1, 0, 5, 15}, {-1, 0, 7, 13}, {-1, 4, 1, 15}, {+1, 4, 3, 13}, {+1, 12, 1, 7}, {-1, 12, 3, 5}},
    {{-1, 0, 5, 11}, {+1, 0, 7, 9}, {+1, 4, 1, 11}, {-1, 4, 3, 9}, {-1, 8, 1, 7}, {+1, 8, 3, 5}},
    {{-1, 4, 9, 14}, {+1, 4, 10, 13}, {+1, 8, 5, 14}, {-1, 8, 6, 13}, {-1, 12, 5, 10}, {+1, 12, 6, 9}},
    {{+1, 0, 9, 14}, {-1, 0, 10, 13}, {-1, 8, 1, 14}, {+1, 8, 2, 13}, {+1, 12, 1, 10}, {-1, 12, 2, 9}},
    {{-1, 0, 5, 14}, {+1, 0, 6, 13}, {+1, 4, 1, 14}, {-1, 4, 2, 13}, {-1, 12, 1, 6}, {+1, 12, 2, 5}},
    {{+1, 0, 5, 10}, {-1, 0, 6, 9}, {-1, 4, 1, 10}, {+1, 4, 2, 9}, {+1, 8, 1, 6}, {-1, 8, 2, 5}},
};

m4 m4_inverted(const m4 *m) {
  const float *c = m->c;
  float inv[16];
  for (int e = 0; e < 16; ++e) {
    float acc = 0.0f;
    for (int t = 0; t < 6; ++t) {
      const signed char *q = kCof[e][t];
      float first = (q[0] < 0) ? -c[q[1]] : c[q[1]];
      float term = first * c[q[2]] * c[q[3]];
      acc = (t == 0) ? term : acc + term;
    }
    inv[e] = acc;
  }
  const float det = c[0] * inv[0] + c[1] * inv[4] + c[2] * inv[8] + c[3] * inv[12];
  m4 r = m4_identity();
  if (det != 0) {
    const float invdet = 1.0f / det;
    for (int i = 0; i < 16; i++) r.c[i] = inv[i] * invdet;
  }
  return r;
}

/* ---- AABB (bvh.h:9-31) ---- */
typedef struct { v3 bmin, bmax; } aabb_t;
static inline aabb_t aabb_empty(void) { aabb_t b = {{RC_LARGE_FLOAT, RC_LARGE_FLOAT, RC_LARGE_FLOAT}, {-RC_LARGE_FLOAT, -RC_LARGE_FLOAT, -RC_LARGE_FLOAT}}; return b; }
static inline void aabb_grow(aabb_t *b, v3 p) { b->bmin = vmin(b->bmin, p); b->bmax = vmax(b->bmax, p); }
static inline void aabb_grow_box(aabb_t *b, const aabb_t *o) {
  if (o->bmin.x != RC_LARGE_FLOAT) { aabb_grow(b, o->bmin); aabb_grow(b, o->bmax); }
}
static inline float aabb_area(const aabb_t *b) {
  v3 e = vsub(b->bmax, b->bmin);
  return e.x * e.y + e.y * e.z + e.z * e.x;
}

/* ---- BVH (bvh.cpp:10-161) ---- */
typedef struct {
  const rc_tri *tris;
  const v3 *centroids;
  uint32_t *triIdx;       /* this mesh's slice of the index buffer */
  rc_bvh_node *nodes;     /* this mesh's slice of the node buffer */
  uint32_t nodeCount;
} bvh_builder;

static void update_bounds(bvh_builder *B, rc_bvh_node *node, v3 *cmin, v3 *cmax) {     /* bvh.cpp:142-161 */
  node->aabbMin = v3s(RC_LARGE_FLOAT);
  node->aabbMax = v3s(-RC_LARGE_FLOAT);
  v3 mn = v3s(RC_LARGE_FLOAT), mx = v3s(-RC_LARGE_FLOAT);
  for (uint32_t first = node->leftFirst, i = 0; i < node->triCount; i++) {
    uint32_t t = B->triIdx[first + i];
    const rc_tri *tri = &B->tris[t];
    node->aabbMin = vmin(node->aabbMin, tri->v0);
    node->aabbMin = vmin(node->aabbMin, tri->v1);
    node->aabbMin = vmin(node->aabbMin, tri->v2);
    node->aabbMax = vmax(node->aabbMax, tri->v0);
    node->aabbMax = vmax(node->aabbMax, tri->v1);
    node->aabbMax = vmax(node->aabbMax, tri->v2);
    mn = vmin(mn, B->centroids[t]);
    mx = vmax(mx, B->centroids[t]);
  }
  *cmin = mn;
  *cmax = mx;
}

static float find_best_split(bvh_builder *B, const rc_bvh_node *node, v3 cmin, v3 cmax, uint32_t *axis, uint32_t *splitPos) {  /* bvh.cpp:89-140 */
  float bestCost = RC_LARGE_FLOAT;
  for (uint32_t a = 0; a < 3; a++) {
    float boundsMin = v3get(cmin, a), boundsMax = v3get(cmax, a);
    if (boundsMin == boundsMax) continue;
    float scale = BINS / (boundsMax - boundsMin);
    float leftCountArea[BINS - 1], rightCountArea[BINS - 1];
    int leftSum = 0, rightSum = 0;
    aabb_t binBounds[BINS];
    int binCount[BINS];
    for (int i = 0; i < BINS; ++i) { binBounds[i] = aabb_empty(); binCount[i] = 0; }
    for (uint32_t i = 0; i < node->triCount; i++) {
      uint32_t t = B->triIdx[node->leftFirst + i];
      const rc_tri *tri = &B->tris[t];
      int binIdx = std_mini(BINS - 1, (int)((v3get(B->centroids[t], a) - boundsMin) * scale));
      binCount[binIdx]++;
      aabb_grow(&binBounds[binIdx], tri->v0);
      aabb_grow(&binBounds[binIdx], tri->v1);
      aabb_grow(&binBounds[binIdx], tri->v2);
    }
    aabb_t leftBox = aabb_empty(), rightBox = aabb_empty();
    for (int i = 0; i < BINS - 1; i++) {
      leftSum += binCount[i];
      aabb_grow_box(&leftBox, &binBounds[i]);
      leftCountArea[i] = leftSum * aabb_area(&leftBox);
      rightSum += binCount[BINS - 1 - i];
      aabb_grow_box(&rightBox, &binBounds[BINS - 1 - i]);
      rightCountArea[BINS - 2 - i] = rightSum * aabb_area(&rightBox);
    }
    for (int i = 0; i < BINS - 1; i++) {
      const float planeCost = leftCountArea[i] + rightCountArea[i];
      if (planeCost < bestCost) {
        *axis = a;
        *splitPos = (uint32_t)(i + 1);
        bestCost = planeCost;
      }
    }
  }
  return bestCost;
}

static uint32_t partition_tris(bvh_builder *B, const rc_bvh_node *node, uint32_t axis, uint32_t splitPos, v3 cmin, v3 cmax) {  /* bvh.cpp:68-87 */
  float scale = BINS / (v3get(cmax, axis) - v3get(cmin, axis));
  uint32_t *triPtr = B->triIdx + node->leftFirst;
  uint32_t i = 0;
  uint32_t j = node->triCount - 1;
  while (i <= j) {
    uint32_t t = triPtr[i];
    uint32_t bin = (uint32_t)std_maxi(0, std_mini((int)((v3get(B->centroids[t], axis) - v3get(cmin, axis)) * scale), BINS - 1));
    if (bin < splitPos) {
      i++;
    } else {
      uint32_t tmp = triPtr[i];
      triPtr[i] = triPtr[j];
      triPtr[j] = tmp;
      j--;
    }
  }
  return i;
}

static void init_node(bvh_builder *B, uint32_t nodeIdx, uint32_t first, uint32_t count);

static void subdivide(bvh_builder *B, uint32_t nodeIdx, v3 cmin, v3 cmax) {            /* bvh.cpp:40-66 */
  rc_bvh_node *node = &B->nodes[nodeIdx];
  uint32_t axis = 0, splitPos = 0;
  float splitCost = find_best_split(B, node, cmin, cmax, &axis, &splitPos);
  float nosplitCost = surface_area(node->aabbMin, node->aabbMax) * node->triCount;     /* common.h:41-43 */
  if (splitCost >= nosplitCost) return;
  uint32_t leftCount = partition_tris(B, node, axis, splitPos, cmin, cmax);
  if (leftCount == 0 || leftCount == node->triCount) return;
  uint32_t leftChildIdx = B->nodeCount++;
  uint32_t rightChildIdx = B->nodeCount++;
  init_node(B, leftChildIdx, node->leftFirst, leftCount);
  init_node(B, rightChildIdx, node->leftFirst + leftCount, node->triCount - leftCount);
  node->leftFirst = leftChildIdx;
  node->triCount = 0;
}

static void init_node(bvh_builder *B, uint32_t nodeIdx, uint32_t first, uint32_t count) {   /* bvh.cpp:29-38 */
  rc_bvh_node *node = &B->nodes[nodeIdx];
  node->leftFirst = first;
  node->triCount = count;
  v3 cmin, cmax;
  update_bounds(B, node, &cmin, &cmax);
  if (count > 1) subdivide(B, nodeIdx, cmin, cmax);
}

uint32_t rc_bvh_build(const rc_tri *tris, const v3 *centroids, uint32_t triCount, rc_bvh_node *nodes, uint32_t *triIdx) {   /* bvh.cpp:10-27 */
  bvh_builder B = {tris, centroids, triIdx, nodes, 0};
  uint32_t root = B.nodeCount++;
  init_node(&B, root, 0, triCount);
  return B.nodeCount;
}
