// rt_build.h — GPU BVH construction (layout A BVH2 and layout B quantized BVH4), see rt_build.cu.
#pragma once
#include <cuda_runtime.h>
#include "rt_layouts.h"

namespace vxb {
// Builds the BVH2 of triangles tri[0..n) into nodes[0..2n-1) and triIdx[0..n) (values tri_first + i, leaf order).
// Synchronous on `stream`.  Returns 0 or a negative cudaError_t; *launches = kernels launched (the CUB sort counted once).
// Leaves: subtrees of at most max_leaf triangles become one leaf where the surface-area heuristic prices the leaf no higher than
// the split (c_trav = cost of one node visit relative to one triangle test); max_leaf = 1 keeps one-triangle leaves.
int build_lbvh(const TriA *tri, uint32_t n, BvhNodeA *nodes, uint32_t *triIdx, uint32_t tri_first, cudaStream_t stream, int *launches,
               float c_trav = 1.0f, uint32_t max_leaf = 4);
// Layout B (second half): the same radix tree collapsed to a BVH4 and quantized into the RT unit's 52-byte nodes
// (raytracing/bvh.cpp:163-258), one-triangle leaves.  tri[0..n) and triEx[0..n) (may be null) are PERMUTED IN PLACE into leaf
// order, because layout B addresses triangles without an index array; leaf.leftFirst = tri_first + position.  Node indices are
// local to this mesh (its blas_node_t.bvh_offset is added by the RT unit).  qbvh_capacity in nodes (2n always suffices).
// root_box6 (host memory, may be null) receives the mesh's bounds for the TLAS.
int build_qbvh4(TriA *tri, TriExB *triEx, uint32_t n, uint32_t tri_first, QNodeB *qbvh, uint64_t qbvh_capacity, cudaStream_t stream,
                uint32_t *nodes_used, float *root_box6, int *launches);
// build_lbvh without any allocation: the scratch arrays come from a static __device__ array of kBuildStaticScratchBytes (per GPU,
// held from library load to exit), which limits it to kBuildStaticMaxTris triangles.  Same kernels, same result as build_lbvh.
// Calls on one GPU share the array: they must not overlap on the GPU.
constexpr uint32_t kBuildStaticMaxTris = 65536;
constexpr size_t kBuildStaticScratchBytes = 8u << 20;   // 104 B per triangle in 11 arrays + the sort's histograms
int build_lbvh_static(const TriA *tri, uint32_t n, BvhNodeA *nodes, uint32_t *triIdx, uint32_t tri_first, cudaStream_t stream, int *launches,
                      float c_trav = 1.0f, uint32_t max_leaf = 4);
// frees the current device's builder scratch block (kept between builds)
void build_release_scratch();
}  // namespace vxb
