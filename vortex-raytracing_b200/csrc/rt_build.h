// rt_build.h — GPU BVH construction (layout A), see rt_build.cu.
#pragma once
#include <cuda_runtime.h>
#include "rt_layouts.h"

namespace vxb {
// Builds the BVH2 of triangles tri[0..n) into nodes[0..2n-1) and triIdx[0..n) (values tri_first + i, leaf order).
// Synchronous on `stream`.  Returns 0 or a negative cudaError_t; *launches = kernels launched (the CUB sort counted once).
int build_lbvh(const TriA *tri, uint32_t n, BvhNodeA *nodes, uint32_t *triIdx, uint32_t tri_first, cudaStream_t stream, int *launches);
}  // namespace vxb
