// bvh2_builder.h — host-side builders of the layout-A acceleration structures (product code, CPU, no CUDA).
//
// Produces byte-for-byte the buffers the reference host builds (tests/regression/raycast/bvh.cpp:23-161 binned-SAH BVH2,
// :181-346 SAH TLAS), but organised for many-core hosts: the BVH is grown as a task tree (independent subtrees are built
// by a thread pool, each node's binning pass keeps the reference's sequential order so that every float result and every
// triangle permutation is identical), then flattened in one pass that hands out node indices in the reference's order
// (children pair first, then the whole left subtree, then the right one).
#pragma once
#include <cstdint>
#include <vector>

namespace vxh {

struct Vec3 { float x, y, z; };

struct Tri { Vec3 v0, v1, v2; };                        // tri_t      36 B
struct TriEx { float n[9]; float uv[6]; };              // tri_ex_t   60 B
struct BvhNode { Vec3 lo; uint32_t leftFirst; Vec3 hi; uint32_t triCount; };   // bvh_node_t 32 B
struct TlasNode { Vec3 lo; uint32_t leftRight; Vec3 hi; uint32_t blasIdx; };   // tlas_node_t 32 B
struct BlasNode {                                        // blas_node_t 160 B
  float transform[16], invTransform[16];
  uint32_t bvh_offset, pad0;
  uint64_t tex_offset;
  uint32_t tex_width, tex_height;
  float reflectivity;
  uint32_t pad1;
};
static_assert(sizeof(Tri) == 36 && sizeof(TriEx) == 60 && sizeof(BvhNode) == 32 && sizeof(TlasNode) == 32 && sizeof(BlasNode) == 160,
              "layout A records");

// Builds the BVH2 of triangles idx[0..count) (GLOBAL indices into tris/centroids; permuted in place) into nodes[0..].
// Returns the number of nodes used (<= 2*count-1).  threads <= 1: single-threaded.
uint32_t build_bvh2(const Tri *tris, const Vec3 *centroids, uint32_t count, uint32_t *idx, BvhNode *nodes, int threads);

// TLAS over `n` instances: leaves 0..n-1 (world boxes of the instance roots), inner nodes n..2n-2; returns the root index.
uint32_t build_tlas(const BvhNode *bvh, const BlasNode *blas, const uint32_t *tri_counts, uint32_t n, TlasNode *tlas);

void mat4_mul(const float *a, const float *b, float *out);      // geometry.h:1220-1231
void mat4_inverse(const float *m, float *out);                  // geometry.h:1149-1192

}  // namespace vxh
