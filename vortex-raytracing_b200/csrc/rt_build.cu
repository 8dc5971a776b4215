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
// 1+2i and 2+2i so the reference's "children are adjacent" addressing holds.
// Leaves (layout A): the bottom-up pass that fits the boxes also prices every subtree with the surface-area heuristic —
// cost(node) = min(n * A * C_tri, C_trav * A + cost(left) + cost(right)), the criterion of the reference's subdivide()
// (raycast/bvh.cpp:40-48: split only if the children cost less than the unsplit node) with a traversal term — and a subtree of
// at most max_leaf triangles whose leaf cost is the smaller one becomes ONE leaf (its triangles are a contiguous range of the
// Morton order, i.e. of triIdx, exactly the reference's leaf format).  The two triangles of a quad share their box, so they
// end up in one leaf as in the reference's trees; nodes below a collapsed node stay in the buffer as unreachable fillers.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
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
  // one grid step for all three axes (the largest extent): a thin axis — the height of a terrain — then only differs in its low
  // bits, so the radix tree does not cut along it before the other two axes have been cut to the same scale.  Normalising every
  // axis by its own extent made every third split a useless one across the terrain's thickness (measured: 4K frame 1.52x the
  // SAH tree's time before, profiles/r02_sweep_gpu_build.txt)
  const float ext = fmaxf(fmaxf(cb[3] - cb[0], cb[4] - cb[1]), cb[5] - cb[2]);
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    float c = (t[a] + t[3 + a] + t[6 + a]) * (1.0f / 3.0f);
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
constexpr uint32_t kCollapsed = 0x80000000u;   // flag in SahInfo::count: this internal node is emitted as one leaf
struct SahInfo {                                // per internal node, written by the thread that fits its box
  uint32_t count;                               // triangles below (| kCollapsed)
  uint32_t first;                               // their first position in Morton order
  float cost;                                   // SAH cost of the subtree as it will be emitted
};
struct SahParams {
  float c_trav, c_tri;
  uint32_t max_leaf;                            // 1 = one-triangle leaves (no pricing)
};
__device__ __forceinline__ float half_area(float lx, float ly, float lz, float hx, float hy, float hz) {
  const float ex = hx - lx, ey = hy - ly, ez = hz - lz;
  return ex * ey + ey * ez + ez * ex;
}

__global__ void k_fit(const uint2 *__restrict__ children, const uint32_t *__restrict__ parent_int, const uint32_t *__restrict__ parent_leaf,
                      const uint32_t *__restrict__ vals, const Box3 *__restrict__ leaf_box, Box3 *__restrict__ int_box,
                      uint32_t *__restrict__ arrived, uint32_t n, SahInfo *__restrict__ sah, SahParams sp) {
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
    if (sah) {
      const volatile SahInfo *sa = (c.x & kLeafBit) ? nullptr : &sah[c.x], *sb = (c.y & kLeafBit) ? nullptr : &sah[c.y];
      const uint32_t na = sa ? (sa->count & ~kCollapsed) : 1u, nb = sb ? (sb->count & ~kCollapsed) : 1u;
      const uint32_t fa = sa ? sa->first : (c.x & ~kLeafBit), fb = sb ? sb->first : (c.y & ~kLeafBit);
      const float ca = sa ? sa->cost : sp.c_tri * half_area(a->lo[0], a->lo[1], a->lo[2], a->hi[0], a->hi[1], a->hi[2]);
      const float cb = sb ? sb->cost : sp.c_tri * half_area(b->lo[0], b->lo[1], b->lo[2], b->hi[0], b->hi[1], b->hi[2]);
      const float A = half_area(m.lo[0], m.lo[1], m.lo[2], m.hi[0], m.hi[1], m.hi[2]);
      const uint32_t nn = na + nb;
      const float leaf_cost = sp.c_tri * (float)nn * A, split_cost = sp.c_trav * A + ca + cb;
      const bool col = nn <= sp.max_leaf && leaf_cost <= split_cost;
      sah[cur].count = nn | (col ? kCollapsed : 0u);
      sah[cur].first = fa < fb ? fa : fb;
      sah[cur].cost = col ? leaf_cost : split_cost;
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
                       uint32_t tri_first, const SahInfo *__restrict__ sah, const uint32_t *__restrict__ parent_int) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) triIdx[i] = tri_first + vals[i];
  if (n == 1) {
    if (i == 0) write_node(nodes, leaf_box[vals[0]], 0u, 1u);
    return;
  }
  if (i >= n - 1) return;
  // below a collapsed node (or being one) this node's children are unreachable: they are written as one-triangle fillers
  bool below = sah && (sah[i].count & kCollapsed);
  if (sah)
    for (uint32_t p = parent_int[i]; !below && p != 0xFFFFFFFFu; p = parent_int[p]) below = (sah[p].count & kCollapsed) != 0;
  if (i == 0) {
    if (below) write_node(nodes, int_box[0], sah[0].first, sah[0].count & ~kCollapsed);   // the whole mesh is one leaf
    else write_node(nodes, int_box[0], 1u, 0u);
  }
  uint2 c = children[i];
  uint32_t cc[2] = {c.x, c.y};
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    BvhNodeA *dst = nodes + 1 + 2 * (size_t)i + k;
    if (cc[k] & kLeafBit) {
      uint32_t pos = cc[k] & ~kLeafBit;
      write_node(dst, leaf_box[vals[pos]], pos, 1u);          // leaf: triIdx slot `pos` (instance-local), one triangle
    } else if (below) {
      write_node(dst, int_box[cc[k]], sah[cc[k]].first, 1u);  // filler (a leaf, so later passes skip it)
    } else if (sah && (sah[cc[k]].count & kCollapsed)) {
      write_node(dst, int_box[cc[k]], sah[cc[k]].first, sah[cc[k]].count & ~kCollapsed);   // a priced subtree as one leaf
    } else {
      write_node(dst, int_box[cc[k]], 1u + 2u * cc[k], 0u);   // inner: its own children pair
    }
  }
}

// Scratch memory of the builders: one grow-only block per CUDA device, carved up by a bump allocator and kept between builds —
// cudaMalloc / cudaFree of a dozen arrays cost more than the 1 M-triangle build itself (38 ms vs 11 ms per frame of
// bench.py's dynamic scene).  Builds of one device are serialised by the runtime's device mutex.
struct Scratch {
  uint8_t *base = nullptr;
  size_t cap = 0, used = 0;
};
Scratch g_scratch[64];
Scratch &scratch() {
  int d = 0;
  cudaGetDevice(&d);
  return g_scratch[d & 63];
}
cudaError_t scratch_reserve(size_t bytes) {   // start of a build: everything handed out before is forgotten
  Scratch &s = scratch();
  s.used = 0;
  if (s.cap >= bytes) return cudaSuccess;
  if (s.base) cudaFree(s.base);
  s.base = nullptr, s.cap = 0;
  const size_t want = bytes + bytes / 8 + (1 << 20);
  cudaError_t e = cudaMalloc(reinterpret_cast<void **>(&s.base), want);
  if (e == cudaSuccess) s.cap = want;
  return e;
}
template <class T>
cudaError_t alloc(T **p, size_t count) {
  Scratch &s = scratch();
  const size_t bytes = (count * sizeof(T) + 255) & ~size_t(255);
  if (s.used + bytes > s.cap) return cudaErrorMemoryAllocation;
  *p = reinterpret_cast<T *>(s.base + s.used);
  s.used += bytes;
  return cudaSuccess;
}

}  // namespace

namespace {
// the radix tree of a mesh, on the device: Karras arrays + fitted boxes
struct Lbvh {
  float *cb = nullptr;
  Box3 *leaf_box = nullptr, *int_box = nullptr;
  uint64_t *keys = nullptr, *keys2 = nullptr;
  uint32_t *vals = nullptr, *vals2 = nullptr, *parent_int = nullptr, *parent_leaf = nullptr, *arrived = nullptr;
  uint2 *children = nullptr;
  SahInfo *sah = nullptr;            // layout A with max_leaf > 1
  void *tmp = nullptr;
  const uint32_t *order = nullptr;   // sorted position -> triangle (points into vals / vals2)
  int launches = 0;
  void release() {
    // scratch memory: nothing to free (see Scratch)
  }
};

#define VXB_TRY(x)                     \
  do {                                 \
    if (e == cudaSuccess) e = (x);     \
  } while (0)

cudaError_t lbvh_tree(const float *tf, uint32_t n, cudaStream_t stream, Lbvh &t, SahParams sp = SahParams{0.f, 1.f, 1u}) {
  size_t tmp_bytes = 0;
  cudaError_t e = cudaSuccess;
  {   // everything this build (and, for layout B, its emission passes) takes from the scratch block
    size_t sort_bytes = 0;
    cub::DoubleBuffer<uint64_t> k0(nullptr, nullptr);
    cub::DoubleBuffer<uint32_t> v0(nullptr, nullptr);
    cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, k0, v0, (int)n, 0, 63, stream);
    VXB_TRY(scratch_reserve((size_t)n * 256 + sort_bytes + (1 << 16)));
  }
  VXB_TRY(alloc(&t.cb, 6));
  VXB_TRY(alloc(&t.leaf_box, n));
  VXB_TRY(alloc(&t.int_box, n));
  VXB_TRY(alloc(&t.keys, n));
  VXB_TRY(alloc(&t.keys2, n));
  VXB_TRY(alloc(&t.vals, n));
  VXB_TRY(alloc(&t.vals2, n));
  VXB_TRY(alloc(&t.parent_int, n));
  VXB_TRY(alloc(&t.parent_leaf, n));
  VXB_TRY(alloc(&t.arrived, n));
  VXB_TRY(alloc(&t.children, n));
  if (sp.max_leaf > 1) VXB_TRY(alloc(&t.sah, n));
  cub::DoubleBuffer<uint64_t> dk(t.keys, t.keys2);
  cub::DoubleBuffer<uint32_t> dv(t.vals, t.vals2);
  VXB_TRY(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, dk, dv, (int)n, 0, 63, stream));
  VXB_TRY(alloc(reinterpret_cast<uint8_t **>(&t.tmp), tmp_bytes ? tmp_bytes : 16));
  if (e != cudaSuccess) return e;
  const unsigned blocks = (n + 255) / 256;
  k_init_bounds<<<1, 32, 0, stream>>>(t.cb);
  k_leaf_boxes<<<blocks, 256, 0, stream>>>(tf, n, t.leaf_box, t.cb);
  k_morton<<<blocks, 256, 0, stream>>>(tf, n, t.cb, t.keys, t.vals);
  t.launches += 3;
  VXB_TRY(cub::DeviceRadixSort::SortPairs(t.tmp, tmp_bytes, dk, dv, (int)n, 0, 63, stream));
  t.launches += 1;   // counted as one library step (CUB launches several kernels)
  VXB_TRY(cudaMemsetAsync(t.arrived, 0, (size_t)n * 4, stream));
  t.order = dv.Current();
  if (n > 1 && e == cudaSuccess) {
    k_radix_tree<<<(n - 1 + 255) / 256, 256, 0, stream>>>(dk.Current(), (int)n, t.children, t.parent_int, t.parent_leaf);
    k_fit<<<blocks, 256, 0, stream>>>(t.children, t.parent_int, t.parent_leaf, t.order, t.leaf_box, t.int_box, t.arrived, n, t.sah, sp);
    t.launches += 2;
  }
  VXB_TRY(cudaGetLastError());
  return e;
}

// ------------------------------------------------------------------------------------------------ layout B emission
// BVH2 -> BVH4: the internal nodes at even depth survive, each adopting its grandchildren (a leaf child stays one slot);
// every triangle is a one-triangle leaf node.  Children of a surviving node are consecutive, as the RT unit addresses them
// (calcNodePtr(base, leftFirst + k), sim/simx/rt_traversal.cpp:69-75).
__global__ void k_b_parity(const uint32_t *__restrict__ parent_int, uint32_t n_int, uint8_t *__restrict__ even) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_int) return;
  uint32_t d = 0;
  for (uint32_t c = i; parent_int[c] != 0xFFFFFFFFu; c = parent_int[c]) ++d;
  even[i] = (d & 1u) ? 0 : 1;
}
__global__ void k_b_count(const uint2 *__restrict__ children, const uint8_t *__restrict__ even, uint32_t n_int, uint32_t *__restrict__ nchild) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_int) return;
  uint32_t c = 0;
  if (even[i]) {
    uint2 ch = children[i];
    c = ((ch.x & kLeafBit) ? 1u : 2u) + ((ch.y & kLeafBit) ? 1u : 2u);
  }
  nchild[i] = c;
}
// the (up to four) children of surviving node i, in slot order
__device__ __forceinline__ uint32_t b_children(const uint2 *__restrict__ children, uint32_t i, uint32_t out[4]) {
  uint2 ch = children[i];
  uint32_t cc[2] = {ch.x, ch.y}, k = 0;
#pragma unroll
  for (int s = 0; s < 2; ++s) {
    if (cc[s] & kLeafBit) {
      out[k++] = cc[s];
    } else {
      uint2 g = children[cc[s]];
      out[k++] = g.x, out[k++] = g.y;
    }
  }
  return k;
}
__global__ void k_b_assign(const uint2 *__restrict__ children, const uint8_t *__restrict__ even, const uint32_t *__restrict__ first,
                           uint32_t n_int, uint32_t *__restrict__ idx_int, uint32_t *__restrict__ idx_leaf) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_int || !even[i]) return;
  if (i == 0) idx_int[0] = 0;
  uint32_t kids[4];
  uint32_t nk = b_children(children, i, kids), base = 1u + first[i];
  for (uint32_t k = 0; k < nk; ++k) {
    if (kids[k] & kLeafBit) idx_leaf[kids[k] & ~kLeafBit] = base + k;
    else idx_int[kids[k]] = base + k;
  }
}

// quantization of one node (raytracing/bvh.cpp:200-258): origin = the node's own min corner, one power-of-two step per
// axis, child corners floor/ceil'ed to 8 bits — and then widened until the box the RT unit decodes
// (origin + ldexp(q, e), rt_traversal.cpp:61-67) really contains the child, as the host builder does.
__device__ __forceinline__ int b_exponent(float extent) {
  if (!(extent > 0.0f)) return 0;
  float e = ceilf(log2f(extent / 255.0f));
  return (int)fminf(fmaxf(e, -126.0f), 127.0f);
}
__device__ __forceinline__ bool b_quantize_axis(float origin, int e, float lo, float hi, uint8_t &qlo, uint8_t &qhi) {
  float scale = ldexpf(1.0f, e);
  float a = floorf((lo - origin) / scale), b = ceilf((hi - origin) / scale);
  if (!(a >= 0.0f)) a = 0.0f;
  if (!(b >= 0.0f)) b = 0.0f;
  while (a > 0.0f && origin + ldexpf(a, e) > lo) a -= 1.0f;
  while (b <= 255.0f && origin + ldexpf(b, e) < hi) b += 1.0f;
  if (b > 255.0f || a > 255.0f) return false;
  qlo = (uint8_t)a, qhi = (uint8_t)b;
  return true;
}
__device__ void b_write_node(uint8_t *dst, const Box3 &self, const Box3 *kid, uint32_t nk, uint32_t leftFirst, uint32_t leafIdx) {
  int e[3];
  uint8_t q[4][6];
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    e[a] = b_exponent(self.hi[a] - self.lo[a]);
    for (;;) {
      bool ok = true;
      for (uint32_t k = 0; k < nk && ok; ++k) ok = b_quantize_axis(self.lo[a], e[a], kid[k].lo[a], kid[k].hi[a], q[k][a], q[k][3 + a]);
      if (ok || e[a] == 127) break;
      e[a]++;
    }
  }
  uint32_t w[13];   // 52 bytes, the record is only 4-byte aligned
  w[0] = __float_as_uint(self.lo[0]), w[1] = __float_as_uint(self.lo[1]), w[2] = __float_as_uint(self.lo[2]);
  w[3] = (uint32_t)(uint8_t)(int8_t)e[0] | (uint32_t)(uint8_t)(int8_t)e[1] << 8 | (uint32_t)(uint8_t)(int8_t)e[2] << 16;   // imask 0 = BLAS node
  w[4] = leftFirst, w[5] = leafIdx;
  uint8_t cb[28];
  for (uint32_t k = 0; k < 4; ++k) {
    cb[7 * k] = k < nk ? 1 : 0;
    for (int c = 0; c < 6; ++c) cb[7 * k + 1 + c] = k < nk ? q[k][c] : 0;
  }
  for (int i = 0; i < 7; ++i) w[6 + i] = (uint32_t)cb[4 * i] | (uint32_t)cb[4 * i + 1] << 8 | (uint32_t)cb[4 * i + 2] << 16 | (uint32_t)cb[4 * i + 3] << 24;
  uint32_t *o = reinterpret_cast<uint32_t *>(dst);
  for (int i = 0; i < 13; ++i) o[i] = w[i];
}
__global__ void k_b_emit_inner(const uint2 *__restrict__ children, const uint8_t *__restrict__ even, const uint32_t *__restrict__ first,
                               const uint32_t *__restrict__ idx_int, const uint32_t *__restrict__ order, const Box3 *__restrict__ leaf_box,
                               const Box3 *__restrict__ int_box, uint32_t n_int, uint8_t *__restrict__ qbvh) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_int || !even[i]) return;
  uint32_t kids[4];
  uint32_t nk = b_children(children, i, kids);
  Box3 kb[4];
  for (uint32_t k = 0; k < nk; ++k) kb[k] = (kids[k] & kLeafBit) ? leaf_box[order[kids[k] & ~kLeafBit]] : int_box[kids[k]];
  b_write_node(qbvh + (size_t)idx_int[i] * sizeof(QNodeB), int_box[i], kb, nk, 1u + first[i], 0u);
}
// one-triangle leaves + the triangles and their attributes moved into leaf order (layout B has no index indirection)
__global__ void k_b_emit_leaf(const uint32_t *__restrict__ idx_leaf, const uint32_t *__restrict__ order, const Box3 *__restrict__ leaf_box,
                              uint32_t n, uint32_t tri_first, uint8_t *__restrict__ qbvh, const float *__restrict__ tri_in,
                              float *__restrict__ tri_out, const uint4 *__restrict__ ex_in, uint4 *__restrict__ ex_out) {
  uint32_t pos = blockIdx.x * blockDim.x + threadIdx.x;
  if (pos >= n) return;
  uint32_t src = order[pos];
  uint32_t node = n == 1 ? 0u : idx_leaf[pos];
  b_write_node(qbvh + (size_t)node * sizeof(QNodeB), leaf_box[src], nullptr, 0, tri_first + pos, 1u);
#pragma unroll
  for (int k = 0; k < 9; ++k) tri_out[(size_t)pos * 9 + k] = tri_in[(size_t)src * 9 + k];
  if (ex_in) {
#pragma unroll
    for (int k = 0; k < 4; ++k) ex_out[(size_t)pos * 4 + k] = ex_in[(size_t)src * 4 + k];
  }
}
__global__ void k_b_root_box(const Box3 *__restrict__ leaf_box, const Box3 *__restrict__ int_box, const uint32_t *__restrict__ order,
                             uint32_t n, float *__restrict__ out6) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const Box3 &b = n == 1 ? leaf_box[order[0]] : int_box[0];
  for (int a = 0; a < 3; ++a) out6[a] = b.lo[a], out6[3 + a] = b.hi[a];
}

}  // namespace

int build_lbvh(const TriA *tri, uint32_t n, BvhNodeA *nodes, uint32_t *triIdx, uint32_t tri_first, cudaStream_t stream, int *launches,
               float c_trav, uint32_t max_leaf) {
  if (n == 0) return -(int)cudaErrorInvalidValue;
  if (max_leaf < 1) max_leaf = 1;
  if (max_leaf > 14) max_leaf = 14;   // the kernels' leaf cursors carry counts below 15
  Lbvh t;
  cudaError_t e = lbvh_tree(reinterpret_cast<const float *>(tri), n, stream, t, SahParams{c_trav, 1.0f, max_leaf});
  if (e == cudaSuccess) {
    k_emit<<<(n + 255) / 256, 256, 0, stream>>>(t.children, t.order, t.leaf_box, t.int_box, n, nodes, triIdx, tri_first, t.sah, t.parent_int);
    t.launches += 1;
    VXB_TRY(cudaGetLastError());
    VXB_TRY(cudaStreamSynchronize(stream));
  }
  t.release();
  if (launches) *launches = t.launches;
  return e == cudaSuccess ? 0 : -(int)e;
}

int build_qbvh4(TriA *tri, TriExB *triEx, uint32_t n, uint32_t tri_first, QNodeB *qbvh, uint64_t qbvh_capacity, cudaStream_t stream,
                uint32_t *nodes_used, float *root_box6, int *launches) {
  if (n == 0) return -(int)cudaErrorInvalidValue;
  Lbvh t;
  const uint32_t n_int = n - 1;
  uint8_t *even = nullptr;
  uint32_t *nchild = nullptr, *first = nullptr, *idx_int = nullptr, *idx_leaf = nullptr;
  float *tri_tmp = nullptr, *box_dev = nullptr;
  uint4 *ex_tmp = nullptr;
  void *scan_tmp = nullptr;
  size_t scan_bytes = 0;
  uint32_t total = 1;
  cudaError_t e = lbvh_tree(reinterpret_cast<const float *>(tri), n, stream, t);
  VXB_TRY(alloc(&even, n));
  VXB_TRY(alloc(&nchild, n));
  VXB_TRY(alloc(&first, n));
  VXB_TRY(alloc(&idx_int, n));
  VXB_TRY(alloc(&idx_leaf, n));
  VXB_TRY(alloc(&tri_tmp, (size_t)n * 9));
  VXB_TRY(alloc(&box_dev, 6));
  if (triEx) VXB_TRY(alloc(&ex_tmp, (size_t)n * 4));
  VXB_TRY(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, nchild, first, (int)(n_int ? n_int : 1), stream));
  VXB_TRY(alloc(reinterpret_cast<uint8_t **>(&scan_tmp), scan_bytes ? scan_bytes : 16));
  if (e == cudaSuccess && n_int) {
    const unsigned ib = (n_int + 255) / 256;
    k_b_parity<<<ib, 256, 0, stream>>>(t.parent_int, n_int, even);
    k_b_count<<<ib, 256, 0, stream>>>(t.children, even, n_int, nchild);
    VXB_TRY(cub::DeviceScan::ExclusiveSum(scan_tmp, scan_bytes, nchild, first, (int)n_int, stream));
    t.launches += 3;
    uint32_t last[2] = {0, 0};
    VXB_TRY(cudaMemcpyAsync(&last[0], first + n_int - 1, 4, cudaMemcpyDeviceToHost, stream));
    VXB_TRY(cudaMemcpyAsync(&last[1], nchild + n_int - 1, 4, cudaMemcpyDeviceToHost, stream));
    VXB_TRY(cudaStreamSynchronize(stream));
    total = 1 + last[0] + last[1];
  }
  if (e == cudaSuccess && total > qbvh_capacity) e = cudaErrorInvalidValue;   // caller's node buffer is too small
  if (e == cudaSuccess) {
    VXB_TRY(cudaMemcpyAsync(tri_tmp, tri, (size_t)n * 36, cudaMemcpyDeviceToDevice, stream));
    if (triEx) VXB_TRY(cudaMemcpyAsync(ex_tmp, triEx, (size_t)n * 64, cudaMemcpyDeviceToDevice, stream));
    if (n_int) {
      const unsigned ib = (n_int + 255) / 256;
      k_b_assign<<<ib, 256, 0, stream>>>(t.children, even, first, n_int, idx_int, idx_leaf);
      k_b_emit_inner<<<ib, 256, 0, stream>>>(t.children, even, first, idx_int, t.order, t.leaf_box, t.int_box, n_int,
                                             reinterpret_cast<uint8_t *>(qbvh));
      t.launches += 2;
    }
    k_b_emit_leaf<<<(n + 255) / 256, 256, 0, stream>>>(idx_leaf, t.order, t.leaf_box, n, tri_first, reinterpret_cast<uint8_t *>(qbvh),
                                                       tri_tmp, reinterpret_cast<float *>(tri), ex_tmp, reinterpret_cast<uint4 *>(triEx));
    k_b_root_box<<<1, 32, 0, stream>>>(t.leaf_box, t.int_box, t.order, n, box_dev);
    t.launches += 2;
    VXB_TRY(cudaGetLastError());
    if (root_box6) VXB_TRY(cudaMemcpyAsync(root_box6, box_dev, 24, cudaMemcpyDeviceToHost, stream));
    VXB_TRY(cudaStreamSynchronize(stream));
  }
  t.release();
  // scratch memory: nothing to free (see Scratch)
  if (nodes_used) *nodes_used = total;
  if (launches) *launches = t.launches;
  return e == cudaSuccess ? 0 : -(int)e;
}
#undef VXB_TRY

void build_release_scratch() {
  Scratch &s = scratch();
  if (s.base) cudaFree(s.base);
  s = Scratch{};
}

}  // namespace vxb

// ------------------------------------------------------------------------------------------------ build without allocation
// vx_b200_build_bvh_static: the layout-A build for callers that must not allocate (CUDA graph capture, serving loops, harnesses
// that forbid cudaMalloc).  Its scratch arrays are carved from this array, which every GPU that loads the library holds until
// exit; all calls on one GPU share it (calls that overlap on the GPU would race) and it keeps the previous call's contents.
static __device__ __align__(256) uint8_t g_build_static_scratch[vxb::kBuildStaticScratchBytes];

namespace vxb {
namespace {
uint8_t *g_static_base[64];   // per CUDA device: where the array lives THERE (the symbol itself is only a host-side handle)
struct Bump {
  uint8_t *base;
  size_t cap, used;
  template <class T>
  cudaError_t take(T **p, size_t count) {
    const size_t bytes = (count * sizeof(T) + 255) & ~size_t(255);
    if (used + bytes > cap) return cudaErrorMemoryAllocation;
    *p = reinterpret_cast<T *>(base + used);
    used += bytes;
    return cudaSuccess;
  }
};
}  // namespace

int build_lbvh_static(const TriA *tri, uint32_t n, BvhNodeA *nodes, uint32_t *triIdx, uint32_t tri_first, cudaStream_t stream, int *launches,
                      float c_trav, uint32_t max_leaf) {
  if (n == 0 || n > kBuildStaticMaxTris) return -(int)cudaErrorInvalidValue;
  if (max_leaf < 1) max_leaf = 1;
  if (max_leaf > 14) max_leaf = 14;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return -(int)e;
  uint8_t *&base = g_static_base[dev & 63];
  if (!base) {   // once per device, with that device current
    uint8_t *ptr = nullptr;
    e = cudaGetSymbolAddress(reinterpret_cast<void **>(&ptr), g_build_static_scratch);
    if (e != cudaSuccess) return -(int)e;
    base = ptr;
  }
  Bump b{base, kBuildStaticScratchBytes, 0};
  const SahParams sp{c_trav, 1.0f, max_leaf};
  const float *tf = reinterpret_cast<const float *>(tri);
  float *cb = nullptr;
  Box3 *leaf_box = nullptr, *int_box = nullptr;
  uint64_t *keys = nullptr, *keys2 = nullptr;
  uint32_t *vals = nullptr, *vals2 = nullptr, *parent_int = nullptr, *parent_leaf = nullptr, *arrived = nullptr;
  uint2 *children = nullptr;
  SahInfo *sah = nullptr;
  uint8_t *tmp = nullptr;
#define VXB_TRY_S(x)                 \
  do {                               \
    if (e == cudaSuccess) e = (x);   \
  } while (0)
  VXB_TRY_S(b.take(&cb, 6));
  VXB_TRY_S(b.take(&leaf_box, n));
  VXB_TRY_S(b.take(&int_box, n));
  VXB_TRY_S(b.take(&keys, n));
  VXB_TRY_S(b.take(&keys2, n));
  VXB_TRY_S(b.take(&vals, n));
  VXB_TRY_S(b.take(&vals2, n));
  VXB_TRY_S(b.take(&parent_int, n));
  VXB_TRY_S(b.take(&parent_leaf, n));
  VXB_TRY_S(b.take(&arrived, n));
  VXB_TRY_S(b.take(&children, n));
  if (sp.max_leaf > 1) VXB_TRY_S(b.take(&sah, n));
  cub::DoubleBuffer<uint64_t> dk(keys, keys2);
  cub::DoubleBuffer<uint32_t> dv(vals, vals2);
  size_t tmp_bytes = 0;
  VXB_TRY_S(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, dk, dv, (int)n, 0, 63, stream));
  VXB_TRY_S(b.take(&tmp, tmp_bytes ? tmp_bytes : 16));
  if (e != cudaSuccess) return -(int)e;
  int nl = 0;
  const unsigned blocks = (n + 255) / 256;
  k_init_bounds<<<1, 32, 0, stream>>>(cb);                       // the array keeps the last call's contents:
  VXB_TRY_S(cudaMemsetAsync(arrived, 0, (size_t)n * 4, stream));   // bounds and arrival flags start from scratch
  k_leaf_boxes<<<blocks, 256, 0, stream>>>(tf, n, leaf_box, cb);
  k_morton<<<blocks, 256, 0, stream>>>(tf, n, cb, keys, vals);
  nl += 3;
  VXB_TRY_S(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, dk, dv, (int)n, 0, 63, stream));
  nl += 1;
  const uint32_t *order = dv.Current();
  if (n > 1 && e == cudaSuccess) {
    k_radix_tree<<<(n - 1 + 255) / 256, 256, 0, stream>>>(dk.Current(), (int)n, children, parent_int, parent_leaf);
    k_fit<<<blocks, 256, 0, stream>>>(children, parent_int, parent_leaf, order, leaf_box, int_box, arrived, n, sah, sp);
    nl += 2;
  }
  if (e == cudaSuccess) {
    k_emit<<<blocks, 256, 0, stream>>>(children, order, leaf_box, int_box, n, nodes, triIdx, tri_first, sah, parent_int);
    nl += 1;
  }
  VXB_TRY_S(cudaGetLastError());
  VXB_TRY_S(cudaStreamSynchronize(stream));
#undef VXB_TRY_S
  if (launches) *launches = nl;
  return e == cudaSuccess ? 0 : -(int)e;
}

}  // namespace vxb
