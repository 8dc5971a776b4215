// rt_build.cu — GPU construction of a layout-A BVH2: replaces the recursive host builder
// (tests/regression/raycast/bvh.cpp:23-161, O(N log N) on one core; 5.4 s for 10 M triangles even with this repo's
// task-parallel host builder) for hosts that want the acceleration structure built where the triangles already are.
//
// Output is the reference's own byte layout, consumable unchanged by the reference CPU path, simx, the CPU oracle and the
// b200 kernels: bvh_node_t[2N-1] (raycast/common.h:32-44: a node's two children adjacent at leftFirst / leftFirst+1, leaf iff
// triCount != 0, leaf.leftFirst = first slot of the instance's triIdx range) and triIdx[N] (global triangle indices in leaf
// order, scene.cpp:51).  Boxes are the exact min/max of the vertex coordinates (what updateNodeBounds computes, bvh.cpp:
// 145-160), so any traversal finds the same closest hit as on the host-built tree; only the tree topology (and with it the
// order in which equal-t triangles are met) differs.
//
// Algorithm: linear BVH — 63-bit Morton codes of the triangle centroids (21 bits per axis, duplicates disambiguated by
// index), one CUB radix sort, Karras' parallel radix-tree construction (one thread per internal node), bottom-up box fit
// with one atomic counter per internal node, then a relabelling pass that lays the children of internal node i at slots
// 1+2i and 2+2i so the reference's "children are adjacent" addressing holds.  One-triangle leaves.
#include <cub/device/device_radix_sort.cuh>
#include <cfloat>
#include <cstdint>
#include "rt_layouts.h"
#include "rt_build.h"

namespace vxb {
namespace {

struct Box3 {
  float lo[3], hi[3];
};

__device__ __forceinline__ void atomic_min_f(float *addr, float v) {   // order-preserving int view of IEEE floats
  if (v >= 0) atomicMin(reinterpret_cast<int *>(addr), __float_as_int(v));
  else atomicMax(reinterpret_cast<unsigned *>(addr), __float_as_uint(v));
}
__device__ __forceinline__ void atomic_max_f(float *addr, float v) {
  if (v >= 0) atomicMax(reinterpret_cast<int *>(addr), __float_as_int(v));
  else atomicMin(reinterpret_cast<unsigned *>(addr), __float_as_uint(v));
}

__global__ void k_init_bounds(float *b) {
  if (threadIdx.x < 3) b[threadIdx.x] = FLT_MAX;
  else if (threadIdx.x < 6) b[threadIdx.x] = -FLT_MAX;
}

// leaf boxes (exact vertex min/max) + centroid bounds of the mesh
__global__ void k_leaf_boxes(const float *__restrict__ tri, uint32_t n, Box3 *__restrict__ leaf_box, float *__restrict__ cb) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  float c[3] = {0, 0, 0};
  bool live = i < n;
  if (live) {
    const float *t = tri + (size_t)i * 9;
    Box3 b;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      float v0 = t[a], v1 = t[3 + a], v2 = t[6 + a];
      b.lo[a] = fminf(fminf(v0, v1), v2);
      b.hi[a] = fmaxf(fmaxf(v0, v1), v2);
      c[a] = (v0 + v1 + v2) * (1.0f / 3.0f);
    }
    leaf_box[i] = b;
  }
  // warp-reduce the centroid bounds, one atomic per warp and axis
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    float lo = live ? c[a] : FLT_MAX, hi = live ? c[a] : -FLT_MAX;
    for (int o = 16; o > 0; o >>= 1) {
      lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
      hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if ((threadIdx.x & 31) == 0 && lo <= hi) {
      atomic_min_f(cb + a, lo);
      atomic_max_f(cb + 3 + a, hi);
    }
  }
}

__device__ __forceinline__ uint64_t spread21(uint32_t v) {   // 21 bits -> every third bit
  uint64_t x = v & 0x1FFFFFu;
  x = (x | x << 32) & 0x1F00000000FFFFull;
  x = (x | x << 16) & 0x1F0000FF0000FFull;
  x = (x | x << 8) & 0x100F00F00F00F00Full;
  x = (x | x << 4) & 0x10C30C30C30C30C3ull;
  x = (x | x << 2) & 0x1249249249249249ull;
  return x;
}

__global__ void k_morton(const float *__restrict__ tri, uint32_t n, const float *__restrict__ cb, uint64_t *__restrict__ keys,
                         uint32_t *__restrict__ vals) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float *t = tri + (size_t)i * 9;
  uint32_t q[3];
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    float c = (t[a] + t[3 + a] + t[6 + a]) * (1.0f / 3.0f);
    float ext = cb[3 + a] - cb[a];
    float u = ext > 0 ? (c - cb[a]) / ext : 0.0f;
    u = fminf(fmaxf(u, 0.0f), 1.0f);
    q[a] = min((uint32_t)(u * 2097151.0f), 2097151u);
  }
  keys[i] = spread21(q[0]) << 2 | spread21(q[1]) << 1 | spread21(q[2]);
  vals[i] = i;
}

// common-prefix length of sorted keys i and j (Karras 2012); equal keys fall back to the index bits
__device__ __forceinline__ int delta(const uint64_t *__restrict__ k, int n, int i, int j) {
  if (j < 0 || j >= n) return -1;
  uint64_t a = k[i], b = k[j];
  if (a != b) return __clzll((long long)(a ^ b));
  return 64 + __clz(i ^ j);
}

// children encoding: bit 31 set = leaf (sorted position), else internal node index
constexpr uint32_t kLeafBit = 0x80000000u;

__global__ void k_radix_tree(const uint64_t *__restrict__ keys, int n, uint2 *__restrict__ children, uint32_t *__restrict__ parent_int,
                             uint32_t *__restrict__ parent_leaf) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n - 1) return;
  int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
  int dmin = delta(keys, n, i, i - d);
  int lmax = 2;
  while (delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
  int l = 0;
  for (int t = lmax / 2; t >= 1; t /= 2)
    if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
  int j = i + l * d;
  int dnode = delta(keys, n, i, j);
  int s = 0, t = l;
  do {
    t = (t + 1) / 2;
    if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
  } while (t > 1);
  int gamma = i + s * d + min(d, 0);
  int lo = min(i, j), hi = max(i, j);
  uint32_t left = (lo == gamma) ? (kLeafBit | (uint32_t)gamma) : (uint32_t)gamma;
  uint32_t right = (hi == gamma + 1) ? (kLeafBit | (uint32_t)(gamma + 1)) : (uint32_t)(gamma + 1);
  children[i] = make_uint2(left, right);
  if (left & kLeafBit) parent_leaf[left & ~kLeafBit] = i; else parent_int[left] = i;
  if (right & kLeafBit) parent_leaf[right & ~kLeafBit] = i; else parent_int[right] = i;
  if (i == 0) parent_int[0] = 0xFFFFFFFFu;
}

// bottom-up: the second thread to reach an internal node merges its children's boxes and climbs on
__global__ void k_fit(const uint2 *__restrict__ children, const uint32_t *__restrict__ parent_int, const uint32_t *__restrict__ parent_leaf,
                      const uint32_t *__restrict__ vals, const Box3 *__restrict__ leaf_box, Box3 *__restrict__ int_box,
                      uint32_t *__restrict__ arrived, uint32_t n) {
  uint32_t l = blockIdx.x * blockDim.x + threadIdx.x;
  if (l >= n) return;
  uint32_t cur = parent_leaf[l];
  while (cur != 0xFFFFFFFFu) {
    __threadfence();
    if (atomicAdd(&arrived[cur], 1u) == 0) return;
    uint2 c = children[cur];
    const volatile Box3 *a = (c.x & kLeafBit) ? &leaf_box[vals[c.x & ~kLeafBit]] : &int_box[c.x];
    const volatile Box3 *b = (c.y & kLeafBit) ? &leaf_box[vals[c.y & ~kLeafBit]] : &int_box[c.y];
    Box3 m;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      m.lo[k] = fminf(a->lo[k], b->lo[k]);
      m.hi[k] = fmaxf(a->hi[k], b->hi[k]);
    }
    int_box[cur] = m;
    cur = parent_int[cur];
  }
}

__device__ __forceinline__ void write_node(BvhNodeA *out, const Box3 &b, uint32_t leftFirst, uint32_t triCount) {
  float4 *p = reinterpret_cast<float4 *>(out);
  p[0] = make_float4(b.lo[0], b.lo[1], b.lo[2], __uint_as_float(leftFirst));
  p[1] = make_float4(b.hi[0], b.hi[1], b.hi[2], __uint_as_float(triCount));
}

// relabel into the reference layout: root at 0, the children of internal node i at 1+2i and 2+2i
__global__ void k_emit(const uint2 *__restrict__ children, const uint32_t *__restrict__ vals, const Box3 *__restrict__ leaf_box,
                       const Box3 *__restrict__ int_box, uint32_t n, BvhNodeA *__restrict__ nodes, uint32_t *__restrict__ triIdx,
                       uint32_t tri_first) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) triIdx[i] = tri_first + vals[i];
  if (n == 1) {
    if (i == 0) write_node(nodes, leaf_box[vals[0]], 0u, 1u);
    return;
  }
  if (i >= n - 1) return;
  if (i == 0) write_node(nodes, int_box[0], 1u, 0u);
  uint2 c = children[i];
  uint32_t cc[2] = {c.x, c.y};
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    BvhNodeA *dst = nodes + 1 + 2 * (size_t)i + k;
    if (cc[k] & kLeafBit) {
      uint32_t pos = cc[k] & ~kLeafBit;
      write_node(dst, leaf_box[vals[pos]], pos, 1u);          // leaf: triIdx slot `pos` (instance-local), one triangle
    } else {
      write_node(dst, int_box[cc[k]], 1u + 2u * cc[k], 0u);   // inner: its own children pair
    }
  }
}

template <class T>
cudaError_t alloc(T **p, size_t count) { return cudaMalloc(reinterpret_cast<void **>(p), count * sizeof(T)); }

}  // namespace

int build_lbvh(const TriA *tri, uint32_t n, BvhNodeA *nodes, uint32_t *triIdx, uint32_t tri_first, cudaStream_t stream, int *launches) {
  if (n == 0) return -(int)cudaErrorInvalidValue;
  const float *tf = reinterpret_cast<const float *>(tri);
  float *cb = nullptr;
  Box3 *leaf_box = nullptr, *int_box = nullptr;
  uint64_t *keys = nullptr, *keys2 = nullptr;
  uint32_t *vals = nullptr, *vals2 = nullptr, *parent_int = nullptr, *parent_leaf = nullptr, *arrived = nullptr;
  uint2 *children = nullptr;
  void *tmp = nullptr;
  size_t tmp_bytes = 0;
  cudaError_t e = cudaSuccess;
  int nl = 0;
#define VXB_TRY(x)                     \
  do {                                 \
    if (e == cudaSuccess) e = (x);     \
  } while (0)
  VXB_TRY(alloc(&cb, 6));
  VXB_TRY(alloc(&leaf_box, n));
  VXB_TRY(alloc(&int_box, n));
  VXB_TRY(alloc(&keys, n));
  VXB_TRY(alloc(&keys2, n));
  VXB_TRY(alloc(&vals, n));
  VXB_TRY(alloc(&vals2, n));
  VXB_TRY(alloc(&parent_int, n));
  VXB_TRY(alloc(&parent_leaf, n));
  VXB_TRY(alloc(&arrived, n));
  VXB_TRY(alloc(&children, n));
  cub::DoubleBuffer<uint64_t> dk(keys, keys2);
  cub::DoubleBuffer<uint32_t> dv(vals, vals2);
  VXB_TRY(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, dk, dv, (int)n, 0, 63, stream));
  VXB_TRY(cudaMalloc(&tmp, tmp_bytes ? tmp_bytes : 16));
  if (e == cudaSuccess) {
    const unsigned blocks = (n + 255) / 256;
    k_init_bounds<<<1, 32, 0, stream>>>(cb);
    k_leaf_boxes<<<blocks, 256, 0, stream>>>(tf, n, leaf_box, cb);
    k_morton<<<blocks, 256, 0, stream>>>(tf, n, cb, keys, vals);
    nl += 3;
    VXB_TRY(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, dk, dv, (int)n, 0, 63, stream));
    nl += 1;   // counted as one library step (CUB launches several kernels)
    VXB_TRY(cudaMemsetAsync(arrived, 0, (size_t)n * 4, stream));
    if (n > 1) {
      k_radix_tree<<<(n - 1 + 255) / 256, 256, 0, stream>>>(dk.Current(), (int)n, children, parent_int, parent_leaf);
      k_fit<<<blocks, 256, 0, stream>>>(children, parent_int, parent_leaf, dv.Current(), leaf_box, int_box, arrived, n);
      nl += 2;
    }
    k_emit<<<blocks, 256, 0, stream>>>(children, dv.Current(), leaf_box, int_box, n, nodes, triIdx, tri_first);
    nl += 1;
    VXB_TRY(cudaGetLastError());
    VXB_TRY(cudaStreamSynchronize(stream));
  }
#undef VXB_TRY
  cudaFree(cb), cudaFree(leaf_box), cudaFree(int_box), cudaFree(keys), cudaFree(keys2), cudaFree(vals), cudaFree(vals2);
  cudaFree(parent_int), cudaFree(parent_leaf), cudaFree(arrived), cudaFree(children), cudaFree(tmp);
  if (launches) *launches = nl;
  return e == cudaSuccess ? 0 : -(int)e;
}

}  // namespace vxb
