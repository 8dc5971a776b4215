// rt_rtu.cu — sm_100a kernels for the layout-B hot path: the simx RT unit's quantized-BVH4 traversal + the
// raytracing regression kernel's ray generation and closest-hit / miss shading.
//
// Reference behaviour being reproduced:
//   BVHTraverser::traverse / pop            sim/simx/rt_traversal.cpp:26-213  (trail + 5-entry short stack + restart)
//   ray_transform / ray_tri / ray_box       sim/simx/rt_traversal.cpp:231-339
//   RTUnit any-hit = ACCEPT driver          sim/simx/rt_unit.cpp:98-116,199-202; tests/regression/raytracing/shaders/anyhit.cpp:28-35
//   GenerateRay (both cameras)              tests/regression/raytracing/kernel.cpp:13-39
//   closest-hit / miss shaders              tests/regression/raytracing/shaders/closest.cpp:11-130, miss.cpp:7-14
//   texSample / diffuseLighting / RGB pack  tests/regression/raytracing/rtx_shading.h:5-18,55-67; common.h:148-162
//
// B200 design (not a translation of the above):
//   * the 52-byte, 4-byte-aligned quantized nodes are decoded ONCE per scene upload into 128-byte SoA records (four
//     child boxes as floats + the node's own header): a visit is seven 128-bit loads of one cache line instead of
//     thirteen 32-bit loads + 24 int->float conversions + 24 ldexp;  origin + ldexp(q, e) is evaluated with the same
//     operations, so the boxes are bit-identical to what the RT unit computes per visit;
//   * triangles become 48-byte edge-form records (same single subtractions the RT unit performs per test);
//   * fast order (default): one full stack per ray (no restarts, no re-visits), children visited nearest first, popped
//     entries re-culled against the current closest hit.  It evaluates a subset of the reference's box/triangle tests
//     with identical results, so it returns the same closest hit except when two triangles have EQUAL t (first-visited
//     wins in both, the visiting order differs);
//   * strict order: the reference's algorithm visit for visit (restart trail, ring stack that forgets its oldest entry,
//     re-traversal after every accepted candidate) — used to prove arithmetic parity: hit records AND work counters
//     equal the oracle's exactly;
//   * persistent warps pull 8x4 screen tiles from an atomic counter; FP32 only, -fmad=false.
//   * tried and rejected (measured on B200, 1080p / 100 K triangles): Aila-Laine "while-while" control flow (all lanes descend
//     inner nodes until every lane holds a leaf) — 4.6 vs 6.9 Grays/s: with 4-wide nodes a ray visits only ~10 nodes, so
//     waiting for the slowest lane's descent costs more than the leaf-phase convergence saves.
#include <cstdio>
#include "rt_device.cuh"
#include "rt_rtu.h"

namespace vxb {
namespace {
using namespace dev;

#ifndef VXB_RTU_MIN_BLOCKS
#define VXB_RTU_MIN_BLOCKS 4   // 64 registers: +3-6 % once the schedule feedback removed the tail (no gain before it)
#endif
constexpr uint32_t kFlagTlas = 1u, kFlagLeaf = 2u;
constexpr uint32_t kStackB = 96;          // fast-order stack entries (3 pushes per level at most)
constexpr uint32_t kTlasBit = 0x80000000u;
constexpr uint32_t kMaxBounce = 16;       // nested traceRay depth kept for the unwind (closest.cpp:104-117)

// ---------------------------------------------------------------- scene re-flatten
__global__ void k_rtu_decode_nodes(const uint8_t *__restrict__ raw, GpuNodeB *__restrict__ out, uint64_t n) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t *w = reinterpret_cast<const uint32_t *>(raw + i * sizeof(QNodeB));   // 52 B = 13 words, 4-byte aligned
  uint32_t r[13];
#pragma unroll
  for (int k = 0; k < 13; ++k) r[k] = __ldg(w + k);
  float px = __uint_as_float(r[0]), py = __uint_as_float(r[1]), pz = __uint_as_float(r[2]);
  int ex = (int8_t)(r[3] & 0xFF), ey = (int8_t)((r[3] >> 8) & 0xFF), ez = (int8_t)((r[3] >> 16) & 0xFF);
  uint32_t imask = r[3] >> 24;
  uint32_t leftFirst = r[4], leafData = r[5];
  const uint8_t *cb = reinterpret_cast<const uint8_t *>(r + 6);   // 4 x {meta, q[6]}
  GpuNodeB g;
  uint32_t valid = 0;
  const float qnan = __uint_as_float(0x7FC00000u);
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const uint8_t *c = cb + 7 * k;
    if (c[0] != 0) {
      valid |= 1u << k;
      // rt_traversal.cpp:61-67: node.p + std::ldexp(float(q), e)
      g.lox[k] = px + ldexpf((float)c[1], ex), g.loy[k] = py + ldexpf((float)c[2], ey), g.loz[k] = pz + ldexpf((float)c[3], ez);
      g.hix[k] = px + ldexpf((float)c[4], ex), g.hiy[k] = py + ldexpf((float)c[5], ey), g.hiz[k] = pz + ldexpf((float)c[6], ez);
    } else {
      g.lox[k] = g.loy[k] = g.loz[k] = g.hix[k] = g.hiy[k] = g.hiz[k] = qnan;
    }
  }
  bool top = imask == 1;                                                        // isTopLevel, :219-221
  bool leaf = (top && leafData != 0xFFFFFFFFu) || (!top && leafData != 0);      // isLeaf, :223-225
  g.leftFirst = leftFirst, g.leafData = leafData;
  g.flags = (top ? kFlagTlas : 0u) | (leaf ? kFlagLeaf : 0u) | (valid << 4);
  g._pad[0] = g._pad[1] = g._pad[2] = g._pad[3] = g._pad[4] = 0;
  out[i] = g;
}

__global__ void k_rtu_gather_tris(const float *__restrict__ tri, GpuTri *__restrict__ gtri, uint64_t n) {
  uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n) return;
  const float *t = tri + p * 9;
  GpuTri g;
  float v0x = t[0], v0y = t[1], v0z = t[2];
  g.v0x = v0x, g.v0y = v0y, g.v0z = v0z;
  g.e1x = t[3] - v0x, g.e1y = t[4] - v0y, g.e1z = t[5] - v0z;   // edge1 (rt_traversal.cpp:271-273)
  g.e2x = t[6] - v0x, g.e2y = t[7] - v0y, g.e2z = t[8] - v0z;   // edge2 (:275-277)
  g.triIdx = (uint32_t)p, g._p1 = g._p2 = 0;
  gtri[p] = g;
}

// ---------------------------------------------------------------- per-ray helpers
struct ObjRay {
  f3 o, d, id;
  bool nan_safe;
};
__device__ __forceinline__ void set_inv(ObjRay &r) {
  r.id = mk3(1.0f / r.d.x, 1.0f / r.d.y, 1.0f / r.d.z);   // ray_box_intersect recomputes this per call (:323-325): same IEEE divide
  r.nan_safe = isinf(r.id.x) || isinf(r.id.y) || isinf(r.id.z) || isnan(r.id.x) || isnan(r.id.y) || isnan(r.id.z);
}
// ray_transform (:231-261): 3x4, no "+ m3*0" on the direction, direction not re-normalised
__device__ __forceinline__ ObjRay to_object(const Ray &w, const float *__restrict__ m) {
  float M[12];
#pragma unroll
  for (int i = 0; i < 12; ++i) M[i] = __ldg(m + i);
  ObjRay r;
  r.o = mk3(M[0] * w.o.x + M[1] * w.o.y + M[2] * w.o.z + M[3], M[4] * w.o.x + M[5] * w.o.y + M[6] * w.o.z + M[7],
            M[8] * w.o.x + M[9] * w.o.y + M[10] * w.o.z + M[11]);
  r.d = mk3(M[0] * w.d.x + M[1] * w.d.y + M[2] * w.d.z, M[4] * w.d.x + M[5] * w.d.y + M[6] * w.d.z,
            M[8] * w.d.x + M[9] * w.d.y + M[10] * w.d.z);
  set_inv(r);
  return r;
}

struct NodeBoxes {
  float4 lox, loy, loz, hix, hiy, hiz;
};
__device__ __forceinline__ NodeBoxes load_boxes(const GpuNodeB *n) {
  const uint8_t *p = reinterpret_cast<const uint8_t *>(n);
  return NodeBoxes{ldg4(p), ldg4(p + 16), ldg4(p + 32), ldg4(p + 48), ldg4(p + 64), ldg4(p + 80)};
}
__device__ __forceinline__ uint4 load_info(const GpuNodeB *n) {
  return __ldg(reinterpret_cast<const uint4 *>(reinterpret_cast<const uint8_t *>(n) + 96));
}
// header + boxes of one node issued back to back (one 128-byte line, one round trip), before anything looks at the header
__device__ __forceinline__ void load_node(const GpuNodeB *n, uint4 &info, NodeBoxes &b) {
  const uint8_t *p = reinterpret_cast<const uint8_t *>(n);
  float4 h = ldg4_pinned(p + 96);
  b = NodeBoxes{ldg4_pinned(p), ldg4_pinned(p + 16), ldg4_pinned(p + 32), ldg4_pinned(p + 48), ldg4_pinned(p + 64), ldg4_pinned(p + 80)};
  info = make_uint4(__float_as_uint(h.x), __float_as_uint(h.y), __float_as_uint(h.z), __float_as_uint(h.w));
}
__device__ __forceinline__ float comp(const float4 &v, int k) { return k == 0 ? v.x : (k == 1 ? v.y : (k == 2 ? v.z : v.w)); }

// ray_box_intersect (:318-339) of child k; returns the reference's value (tmin, or LARGE_FLOAT on a miss; NaN stays NaN)
__device__ __forceinline__ float child_box(const NodeBoxes &b, int k, const ObjRay &r) {
  float4 lo = make_float4(comp(b.lox, k), comp(b.loy, k), comp(b.loz, k), 0.f);
  float4 hi = make_float4(comp(b.hix, k), comp(b.hiy, k), comp(b.hiz, k), 0.f);
  return slab(lo, hi, r.o, r.id, r.nan_safe);
}

// leaf of a BLAS: ray_tri_intersect over [first, first+count) (:126-160).  FIRST_ONLY = the reference's behaviour of
// returning at the first candidate (strict order); otherwise every triangle is tested and the closest kept.
template <bool STATS, bool FIRST_ONLY>
__device__ __forceinline__ bool leaf_tris(const RtuScene &sc, uint32_t first, uint32_t count, const ObjRay &r, uint32_t blasIdx, Hit &hit,
                                          float &pending, Local &lc) {
  Ray rr{r.o, r.d};
  for (uint32_t i = 0; i < count; ++i) {
    uint32_t triIdx = first + i;
    if (triIdx >= sc.n_tri) break;
    const uint8_t *tp = reinterpret_cast<const uint8_t *>(sc.gtri + triIdx);
    float4 q0 = ldg4(tp), q1 = ldg4(tp + 16), q2 = ldg4(tp + 32);
    float t, w1, w2;
    if (STATS) lc.tri++;
    if (moller_trumbore(rr, mk3(q0.x, q0.y, q0.z), mk3(q1.x, q1.y, q1.z), mk3(q2.x, q2.y, q2.z), t, w1, w2) && t < hit.dist) {
      hit.bx = w1, hit.by = w2, hit.bz = 1 - w1 - w2;
      hit.blas = blasIdx, hit.tri = triIdx;
      if (FIRST_ONLY) {
        pending = t;
        return true;
      }
      hit.dist = t;
    }
  }
  return false;
}

struct Instance {
  ObjRay ray;
  uint32_t base, blasIdx;
};
__device__ __forceinline__ bool enter_instance(const RtuScene &sc, const Ray &wray, uint32_t blasIdx, Instance &in) {
  if (blasIdx >= sc.n_blas) return false;
  const BlasNodeB *b = sc.blas + blasIdx;               // stride 160 (rt_traversal.cpp:112)
  in.ray = to_object(wray, b->invTransform);
  in.base = __ldg(&b->bvh_offset);
  in.blasIdx = blasIdx;
  return in.base < sc.n_bvh;
}

// ---------------------------------------------------------------- fast order
template <bool STATS>
__device__ __forceinline__ void trace_fast(const RtuScene &sc, const Ray &wray, Hit &hit, Local &lc, uint32_t *stk_e, float *stk_d) {
  lc.rays++;
  ObjRay world;
  world.o = wray.o, world.d = wray.d;
  set_inv(world);
  Instance in;
  in.base = 0, in.blasIdx = 0, in.ray = world;
  uint32_t sp = 0;
  uint32_t cur = kTlasBit | 0u;                       // TLAS node 0 (calcNodePtr(tlas_ptr, 0), :41)
  float unused;
  for (;;) {
    simt_step<STATS>(lc);
    const bool top = (cur & kTlasBit) != 0;
    const uint32_t idx = cur & ~kTlasBit;
    bool ok = top ? (idx < sc.n_tlas) : (idx < sc.n_bvh);
    if (ok) {
      const GpuNodeB *n = (top ? sc.gtlas : sc.gbvh) + idx;
      uint4 info;
      NodeBoxes b;
      load_node(n, info, b);
      if (STATS) lc.box++;
      if (!(info.z & kFlagLeaf)) {
        const ObjRay &r = top ? world : in.ray;
        const uint32_t child0 = (top ? 0u : in.base) + info.x;
        float d[4];
        uint32_t e[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          float dk = child_box(b, k, r);
          bool h = ((info.z >> (4 + k)) & 1u) && (dk < hit.dist);
          d[k] = h ? dk : __int_as_float(0x7F800000);
          e[k] = (top ? kTlasBit : 0u) | (child0 + k);
        }
        // sort by (distance ascending, child index descending) — the reference's visiting order: nearest first and, among
        // equal distances, the LATER child first (its stable far->near sort is consumed from the back, :76-90), so that
        // equal-t hits resolve to the same triangle; 5-comparator network
#define VXB_CSWAP(i, j)                       \
  if (d[j] < d[i] || (d[j] == d[i] && e[j] > e[i])) { \
    float td = d[i]; d[i] = d[j]; d[j] = td;  \
    uint32_t te = e[i]; e[i] = e[j]; e[j] = te; \
  }
        VXB_CSWAP(0, 1) VXB_CSWAP(2, 3) VXB_CSWAP(0, 2) VXB_CSWAP(1, 3) VXB_CSWAP(1, 2)
#undef VXB_CSWAP
        if (d[0] < __int_as_float(0x7F800000)) {
          if (sp + 3 <= kStackB) {
            if (d[3] < __int_as_float(0x7F800000)) { stk_e[sp] = e[3]; stk_d[sp] = d[3]; ++sp; }
            if (d[2] < __int_as_float(0x7F800000)) { stk_e[sp] = e[2]; stk_d[sp] = d[2]; ++sp; }
            if (d[1] < __int_as_float(0x7F800000)) { stk_e[sp] = e[1]; stk_d[sp] = d[1]; ++sp; }
          } else {
            lc.err |= 1u;   // stack overflow: vx_ready_wait reports the run as failed
          }
          cur = e[0];
          continue;
        }
      } else if (top) {
        if (enter_instance(sc, wray, info.y, in)) {
          if (STATS) lc.blas++;
          cur = in.base;                               // node 0 of the instance's BVH
          continue;
        }
      } else {
        leaf_tris<STATS, false>(sc, info.x, info.y, in.ray, in.blasIdx, hit, unused, lc);
      }
    }
    bool found = false;
    while (sp != 0) {
      --sp;
      if (stk_d[sp] < hit.dist) {
        cur = stk_e[sp];
        found = true;
        break;
      }
    }
    if (!found) break;
  }
}

// ---------------------------------------------------------------- strict order: the reference algorithm visit for visit
struct Ring {            // ShortStack<TraversalStackEntry, 5> (sim/common/types.h:1808-1841)
  uint32_t e[5];         // bit 31 TLAS space, bit 30 "last", low bits node index
  uint32_t head, count;
};
__device__ __forceinline__ void ring_push(Ring &s, uint32_t v) {
  if (s.count < 5) s.count++;        // full: the oldest entry is overwritten
  s.e[s.head] = v;
  s.head = (s.head + 1) % 5;
}
__device__ __forceinline__ uint32_t ring_pop(Ring &s) {
  s.head = (s.head == 0) ? 4 : (s.head - 1);
  s.count--;
  return s.e[s.head];
}
constexpr uint32_t kLastBit = 0x40000000u;
constexpr uint32_t kTrailLevels = 32;   // MAX_TRAIL_LEVEL

// BVHTraverser::pop (:179-213).  Returns true when traversal is complete.
__device__ __forceinline__ bool strict_pop(uint32_t &cur, uint32_t &base, uint32_t &level, uint8_t *trail, Ring &stack) {
  int parent = -1;
  for (int i = (int)level - 1; i >= 0; i--)
    if (trail[i] != 4) {
      parent = i;
      break;
    }
  if (parent < 0) return true;
  trail[parent]++;
  for (uint32_t i = parent + 1; i < kTrailLevels; i++) trail[i] = 0;
  if (stack.count == 0) {
    cur = kTlasBit | 0u;   // restart from the root (base_ptr = node_ptr = tlas_ptr)
    base = kTlasBit | 0u;
    level = 0;
  } else {
    uint32_t e = ring_pop(stack);
    cur = e & ~kLastBit;
    if (e & kLastBit) trail[parent] = 4;
    level = (uint32_t)parent + 1;
  }
  return false;
}

// BVHTraverser::traverse (:26-168): true = completed, false = candidate pending (hit.b*, ids, `pending` set)
template <bool STATS>
__device__ __forceinline__ bool strict_traverse(const RtuScene &sc, const Ray &wray, const ObjRay &world, Hit &hit, float &pending,
                                                uint8_t *trail, Ring &stack, Local &lc) {
  uint32_t level = 0;
  uint32_t cur = kTlasBit | 0u;
  // `base` is the reference's base_ptr as (buffer, node index): child k of a node lives at base + leftFirst + k.  pop()
  // does NOT restore it when the popped entry is a TLAS node (:179-213), so after an instance was visited the children
  // of a popped TLAS inner node are fetched from that instance's BVH.  Reproduced here address for address; the node's own
  // imask (not the buffer it was fetched from) decides how it is interpreted, as in the reference.
  uint32_t base = kTlasBit | 0u;
  Instance in;
  in.base = 0, in.blasIdx = 0, in.ray = world;
  bool done = false;
  while (!done) {
    simt_step<STATS>(lc);
    const bool in_tlas_buf = (cur & kTlasBit) != 0;
    const uint32_t idx = cur & ~kTlasBit;
    if (!(in_tlas_buf ? (idx < sc.n_tlas) : (idx < sc.n_bvh))) {   // hostile index: treat as an empty inner node
      done = strict_pop(cur, base, level, trail, stack);
      continue;
    }
    const GpuNodeB *n = (in_tlas_buf ? sc.gtlas : sc.gbvh) + idx;
    uint4 info = load_info(n);
    const bool top = (info.z & kFlagTlas) != 0;
    if (STATS) lc.box++;
    if (!(info.z & kFlagLeaf)) {
      NodeBoxes b = load_boxes(n);
      const ObjRay &r = top ? world : in.ray;
      float d[4];
      uint32_t ci[4];
      int cnt = 0;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (!((info.z >> (4 + k)) & 1u)) continue;
        float dk = child_box(b, k, r);
        if (dk < hit.dist) {
          // std::sort(a.dist > b.dist) on <= 4 elements == stable insertion sort, farthest first (:76-78)
          int p = cnt;
          while (p > 0 && dk > d[p - 1]) {
            d[p] = d[p - 1], ci[p] = ci[p - 1];
            p--;
          }
          d[p] = dk, ci[p] = (uint32_t)k;
          cnt++;
        }
      }
      uint32_t k = level < kTrailLevels ? trail[level] : 0;
      uint32_t drop = (k == 4) ? (cnt > 0 ? (uint32_t)cnt - 1 : 0) : k;   // :80-86 (the size()-1 wrap on an empty list drops nothing)
      while (drop-- > 0 && cnt > 0) cnt--;
      if (cnt == 0) {
        done = strict_pop(cur, base, level, trail, stack);
      } else {
        const uint32_t space = base & kTlasBit;
        const uint32_t child0 = (base & ~kTlasBit) + info.x;
        uint32_t closest = ci[cnt - 1];
        cnt--;
        cur = space | (child0 + closest);
        if (cnt == 0) {
          if (level < kTrailLevels) trail[level] = 4;
        } else {
          for (int i = 0; i < cnt; ++i) ring_push(stack, space | (child0 + ci[i]) | (i == 0 ? kLastBit : 0u));
        }
        level++;
      }
    } else if (top) {
      if (enter_instance(sc, wray, info.y, in)) {
        if (STATS) lc.blas++;
        cur = in.base;
        base = in.base;
      } else {
        done = strict_pop(cur, base, level, trail, stack);
      }
    } else {
      if (leaf_tris<STATS, true>(sc, info.x, info.y, in.ray, in.blasIdx, hit, pending, lc)) {
        stack.count = 0, stack.head = 0;   // "clear stack for now to ensure correctness" (:148-152)
        return false;
      }
      done = strict_pop(cur, base, level, trail, stack);
    }
  }
  return true;
}

template <bool STATS>
__device__ __forceinline__ void trace_strict(const RtuScene &sc, const Ray &wray, Hit &hit, Local &lc) {
  lc.rays++;
  ObjRay world;
  world.o = wray.o, world.d = wray.d;
  set_inv(world);
  uint8_t trail[kTrailLevels];
#pragma unroll
  for (uint32_t i = 0; i < kTrailLevels; ++i) trail[i] = 0;
  Ring stack;
  stack.head = 0, stack.count = 0;
  float pending = 0.0f;
  // RTUnit::traverse + commit(ACCEPT) (rt_unit.cpp:98-116,199-202): the shipped any-hit shader accepts every candidate
  while (!strict_traverse<STATS>(sc, wray, world, hit, pending, trail, stack, lc)) hit.dist = pending;
}

template <bool STRICT, bool STATS>
__device__ __forceinline__ void trace_b(const RtuScene &sc, const Ray &wray, Hit &hit, Local &lc, uint32_t *stk_e, float *stk_d) {
  if (STRICT) trace_strict<STATS>(sc, wray, hit, lc);
  else trace_fast<STATS>(sc, wray, hit, lc, stk_e, stk_d);
}

// ---------------------------------------------------------------- ray generation (kernel.cpp:13-39)
__device__ __forceinline__ Ray generate_ray_b(uint32_t x, uint32_t y, const KernelArgB &a, uint32_t camera) {
  if (camera == 0) {
    // the camera the shipped kernel hard-codes; u and v are DOUBLE expressions narrowed to float (kernel.cpp:32-33)
    f3 pos = mk3(0.0f, 100.0f, 0.0f), front = mk3(1.0f, 0.0f, 0.0f);
    float FOV = 1.0f;
    float u = (float)(((double)x * 2.0 - (double)a.dst_width) / (double)a.dst_height);
    float v = (float)(((double)y * 2.0 - (double)a.dst_height) / (double)a.dst_height);
    f3 right = cross3(front, mk3(0.0f, 1.0f, 0.0f));
    f3 up = cross3(right, front);
    f3 dir = normalize3(u * right + v * up + FOV * front);
    return Ray{pos, dir};
  }
  float x_ndc = (x + 0.5f) / (float)a.dst_width - 0.5f;    // "- 0.5" in double then narrowed == one exact float subtraction
  float y_ndc = (y + 0.5f) / (float)a.dst_height - 0.5f;
  float x_vp = x_ndc * a.viewplane.x, y_vp = y_ndc * a.viewplane.y;
  f3 right = mk3(a.camera_right.x, a.camera_right.y, a.camera_right.z);
  f3 up = mk3(a.camera_up.x, a.camera_up.y, a.camera_up.z);
  f3 fwd = mk3(a.camera_forward.x, a.camera_forward.y, a.camera_forward.z);
  f3 pos = mk3(a.camera_pos.x, a.camera_pos.y, a.camera_pos.z);
  f3 pt_cam = x_vp * right + y_vp * up + fwd;
  f3 pt_w = pt_cam + pos;
  return Ray{pos, normalize3(pt_w - pos)};
}

// ---------------------------------------------------------------- closest-hit / miss shading (closest.cpp:11-130, miss.cpp:12)
struct SurfaceB {
  f3 I, N, texColor;
  float reflectivity;
};
__device__ __forceinline__ SurfaceB surface_b(const RtuScene &sc, const Ray &ray, const Hit &hit) {
  const BlasNodeB *blas = sc.blas + hit.blas;
  const float *tx = reinterpret_cast<const float *>(sc.triEx + hit.tri);   // N0 N1 N2 uv0 uv1 uv2 texId (16 words)
  f3 N0 = mk3(__ldg(tx + 0), __ldg(tx + 1), __ldg(tx + 2));
  f3 N1 = mk3(__ldg(tx + 3), __ldg(tx + 4), __ldg(tx + 5));
  f3 N2 = mk3(__ldg(tx + 6), __ldg(tx + 7), __ldg(tx + 8));
  float u0 = __ldg(tx + 9), v0 = __ldg(tx + 10), u1 = __ldg(tx + 11), v1 = __ldg(tx + 12), u2 = __ldg(tx + 13), v2 = __ldg(tx + 14);
  uint32_t texId = __float_as_uint(__ldg(tx + 15));
  SurfaceB s;
  s.I = ray.o + ray.d * hit.dist;
  f3 N = N1 * hit.bx + N2 * hit.by + N0 * hit.bz;                          // closest.cpp:66
  const float *inv = blas->invTransform;                                   // transposed(): 3x3 only, cells 3,7,11 = 0
  float T[12] = {__ldg(inv + 0), __ldg(inv + 4), __ldg(inv + 8), 0.0f, __ldg(inv + 1), __ldg(inv + 5), __ldg(inv + 9), 0.0f,
                 __ldg(inv + 2), __ldg(inv + 6), __ldg(inv + 10), 0.0f};
  s.N = normalize3(xform_vec(N, T));
  float uvx = u1 * hit.bx + u2 * hit.by + u0 * hit.bz;
  float uvy = v1 * hit.bx + v2 * hit.by + v0 * hit.bz;
  s.texColor = mk3(0.f, 0.f, 0.f);
  if (sc.mat && texId < sc.n_mat) {
    const MaterialB *m = sc.mat + texId;
    if (__ldg(&m->diffuse_tex_id) >= 0) {                                   // texSample (rtx_shading.h:5-18)
      uint32_t tw = __ldg(&m->tex_width), th = __ldg(&m->tex_height);
      if (tw && th && sc.tex) {
        uint32_t iu = f2u_x86(uvx * tw), iv = f2u_x86(uvy * th);
        iu %= tw, iv %= th;
        uint64_t off = __ldg(&m->tex_offset) + 4ull * (iu + iv * tw);
        if (off + 4 <= sc.tex_bytes) s.texColor = rgb8_to_rgb32f(__ldg(reinterpret_cast<const uint32_t *>(sc.tex + off)));
      }
    } else {
      s.texColor = mk3(__ldg(&m->diffuse.x), __ldg(&m->diffuse.y), __ldg(&m->diffuse.z));
    }
  }
  s.reflectivity = __ldg(&blas->reflectivity);
  return s;
}

// The shaders recurse (closest-hit traces the mirror ray and adds secPayload.color * throughput, closest.cpp:104-119);
// evaluated here as a forward pass that records (radiance_k, throughput_k) per level and an unwind in the same order of
// floating-point operations.
template <bool STRICT, bool STATS>
__device__ __forceinline__ f3 shade_b(const RtuScene &sc, const KernelArgB &a, Ray cur, Local &lc, uint32_t *stk_e, float *stk_d) {
  const f3 light_pos = mk3(a.light_pos.x, a.light_pos.y, a.light_pos.z);
  const f3 light_color = mk3(a.light_color.x, a.light_color.y, a.light_color.z);
  const f3 ambient = mk3(a.ambient_color.x, a.ambient_color.y, a.ambient_color.z);
  const f3 background = mk3(a.background_color.x, a.background_color.y, a.background_color.z);
  f3 rad[kMaxBounce];
  float thr[kMaxBounce];
  uint32_t depth = 0;
  f3 color;
  for (;;) {
    Hit hit{kLargeFloat, 0, 0, 0, 0, 0};
    trace_b<STRICT, STATS>(sc, cur, hit, lc, stk_e, stk_d);
    if (hit.dist == kLargeFloat) {
      color = background;                                                   // miss.cpp:12
      break;
    }
    if (STATS) lc.shaded++;
    SurfaceB s = surface_b(sc, cur, hit);
    f3 L = light_pos - s.I;                                                 // diffuseLighting (rtx_shading.h:55-67)
    float dist = length3(L);
    L = L * (1.0f / dist);
    float att = 1.0f / (1.0f + dist * 0.1f);
    float NdotL = std_max(0.0f, dot3(s.N, L));
    f3 diffuse = s.texColor * (ambient + att * light_color * NdotL);
    f3 radiance = mk3(0, 0, 0);
    float throughput = 1.0f;
    radiance = radiance + throughput * diffuse * (1 - s.reflectivity);
    throughput *= s.reflectivity;
    if (s.reflectivity > 0.0f && depth + 1 < a.max_depth && depth + 1 < kMaxBounce) {
      f3 R = normalize3(cur.d - 2.0f * s.N * dot3(s.N, cur.d));
      rad[depth] = radiance, thr[depth] = throughput;
      ++depth;
      cur.o = s.I + R * 0.001f;
      cur.d = R;
      continue;
    }
    color = radiance + background * throughput;
    break;
  }
  while (depth > 0) {
    --depth;
    color = rad[depth] + color * thr[depth];
  }
  return color;
}

// Extended modes (no reference counterpart; BASELINE configs 1-3 ask for shadow rays and diffuse bounces): the definition of
// the raycast layout (rt_raycast.cu:trace_ext == oracle/port/rc_shade.c:rc_trace_ext) with this layout's traversal and surface
// evaluation, expression for expression the CPU checker's oracle/port/rtu_port.c:trace_ext_b.  mode bit0 shadow, bit1 diffuse.
template <bool STRICT, bool STATS>
__device__ __forceinline__ f3 shade_ext_b(const RtuScene &sc, const KernelArgB &a, Ray cur, uint32_t mode, uint32_t &seed, Local &lc,
                                          uint32_t *stk_e, float *stk_d) {
  const f3 light_pos = mk3(a.light_pos.x, a.light_pos.y, a.light_pos.z);
  const f3 light_color = mk3(a.light_color.x, a.light_color.y, a.light_color.z);
  const f3 ambient = mk3(a.ambient_color.x, a.ambient_color.y, a.ambient_color.z);
  const f3 background = mk3(a.background_color.x, a.background_color.y, a.background_color.z);
  f3 radiance = mk3(0, 0, 0);
  f3 thr = mk3(1, 1, 1);
  for (uint32_t bounce = 0; bounce < a.max_depth; ++bounce) {
    Hit hit{kLargeFloat, 0, 0, 0, 0, 0};
    trace_b<STRICT, STATS>(sc, cur, hit, lc, stk_e, stk_d);
    if (hit.dist == kLargeFloat) {
      radiance = radiance + background * thr;
      break;
    }
    if (STATS) lc.shaded++;
    SurfaceB s = surface_b(sc, cur, hit);
    f3 Ns = (dot3(s.N, cur.d) > 0.0f) ? neg(s.N) : s.N;
    f3 O = s.I + Ns * 0.001f;
    f3 L = light_pos - s.I;
    float dist = length3(L);
    L = L * (1.0f / dist);
    float att = 1.0f / (1.0f + dist * 0.1f);
    float NdotL = std_max(0.0f, dot3(s.N, L));
    if ((mode & 1u) && NdotL > 0.0f) {
      f3 Lv = light_pos - O;
      float d2 = length3(Lv);
      f3 D = Lv * (1.0f / d2);
      Hit sh{d2, 0, 0, 0, 0, 0};                       // occluded iff something is hit before the light
      trace_b<STRICT, STATS>(sc, Ray{O, D}, sh, lc, stk_e, stk_d);
      if (sh.dist < d2) NdotL = 0.0f;
    }
    f3 diffuse = s.texColor * (ambient + att * light_color * NdotL);
    if (mode & 2u) {
      radiance = radiance + thr * diffuse;
      thr = thr * s.texColor;
      if (bounce + 1 < a.max_depth) {
        float px, py, pz, l2;
        do {
          px = random_float(seed) * 2.0f - 1.0f;
          py = random_float(seed) * 2.0f - 1.0f;
          pz = random_float(seed) * 2.0f - 1.0f;
          l2 = px * px + py * py + pz * pz;
        } while (l2 > 1.0f || l2 < 1e-6f);
        float il = 1.0f / sqrtf(l2);
        f3 D = Ns + mk3(px * il, py * il, pz * il);
        float dl2 = dot3(D, D);
        if (dl2 < 1e-12f) D = Ns;
        else D = D * (1.0f / sqrtf(dl2));
        cur.o = O;
        cur.d = D;
        continue;
      }
      break;
    }
    radiance = radiance + thr * diffuse * (1 - s.reflectivity);
    thr = thr * s.reflectivity;
    if (s.reflectivity > 0.0f && bounce + 1 < a.max_depth) {
      f3 R = normalize3(cur.d - 2.0f * s.N * dot3(s.N, cur.d));
      cur.o = s.I + R * 0.001f;
      cur.d = R;
      continue;
    }
    radiance = radiance + thr * background;
    break;
  }
  return radiance;
}

// ---------------------------------------------------------------- the pixel loop (kernel.cpp:41-133)
template <bool STRICT, bool STATS, bool EXT = false>
__global__ void __launch_bounds__(kBlockThreads, (STRICT || EXT) ? 2 : VXB_RTU_MIN_BLOCKS) k_rtu_render(const __grid_constant__ RtuRender p) {
  uint32_t stk_e[STRICT ? 1 : kStackB];
  float stk_d[STRICT ? 1 : kStackB];
  const KernelArgB &a = p.arg;
  const uint32_t W = a.dst_width, H = a.dst_height;
  const uint32_t lane = threadIdx.x & 31;
  const uint32_t tcols = (W + kTileW - 1) / kTileW;
  const uint32_t band_rows = p.part.band_rows, world = p.part.world, rank = p.part.rank;
  const uint32_t nbands_total = (H + band_rows - 1) / band_rows;
  const uint32_t nbands = nbands_total > rank ? (nbands_total - rank + world - 1) / world : 0;
  const uint32_t tiles_per_band = (band_rows / kTileH) * tcols;
  const uint32_t ntiles = nbands * tiles_per_band;
  Local lc{};
  for (;;) {
    uint32_t tile = 0;
    if (lane == 0) tile = atomicAdd(p.tile_counter, 1u);
    tile = __shfl_sync(0xffffffffu, tile, 0);
    if (tile >= ntiles) break;
    uint32_t grp;
    tile = sched_tile(p.sched, tile, ntiles, grp);
    const uint32_t t0 = (uint32_t)clock();
    uint32_t band_local = tile / tiles_per_band, r = tile % tiles_per_band;
    uint32_t trow = r / tcols, tcol = r % tcols;
    uint32_t x = tcol * kTileW + (lane & (kTileW - 1));
    uint32_t y = (band_local * world + rank) * band_rows + trow * kTileH + (lane / kTileW);
    if (x < W && y < H) {
      // kernel.cpp:66-81: every sample traces the same ray into the same payload, so one evaluation is the result
      Ray ray = generate_ray_b(x, y, a, p.camera);
      f3 color;
      if (EXT) {
        uint32_t seed = wang_hash((y * W + x) * a.samples_per_pixel + 1);
        if (seed == 0) seed = 1;
        color = shade_ext_b<STRICT, STATS>(p.scene, a, ray, p.mode, seed, lc, stk_e, stk_d);
      } else {
        color = shade_b<STRICT, STATS>(p.scene, a, ray, lc, stk_e, stk_d);
      }
      p.out[y * W + x] = rgb32f_to_rgb8(color);
    }
    __syncwarp();
    if (lane == 0 && grp != 0xFFFFFFFFu && p.sched.cost) atomicMax(p.sched.cost + grp, (uint32_t)clock() - t0);   // longest tile of the group
  }
  flush_counters(p.counters, lc, STATS);
}

template <bool STRICT, bool STATS>
__global__ void __launch_bounds__(kBlockThreads) k_rtu_trace(const __grid_constant__ RtuTrace p) {
  uint32_t stk_e[STRICT ? 1 : kStackB];
  float stk_d[STRICT ? 1 : kStackB];
  Local lc{};
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  const uint64_t n_round = (p.count + 31) / 32 * 32;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += stride) {
    if (i < p.count) {
      Ray ray;
      if (p.rays) {
        const float *q = p.rays + 6 * i;
        ray.o = mk3(__ldg(q), __ldg(q + 1), __ldg(q + 2));
        ray.d = mk3(__ldg(q + 3), __ldg(q + 4), __ldg(q + 5));
      } else {
        ray = generate_ray_b((uint32_t)(i % p.arg.dst_width), (uint32_t)(i / p.arg.dst_width), p.arg, p.camera);
      }
      Hit hit{kLargeFloat, 0, 0, 0, 0, 0};
      trace_b<STRICT, STATS>(p.scene, ray, hit, lc, stk_e, stk_d);
      p.hits[i] = HitA{hit.dist, hit.bx, hit.by, hit.bz, hit.blas, hit.tri};
    }
  }
  __syncwarp();
  flush_counters(p.counters, lc, STATS);
}

#include "vx_raytrace_b200.cuh"
#include "rt_rtu_programs.cuh"

template <class K, class P>
int launch_persistent(K kern, const P &p, int num_sms, cudaStream_t stream, uint64_t want_blocks) {
  int per_sm = 0;
  cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kBlockThreads, 0);
  if (e != cudaSuccess) return -(int)e;
  if (per_sm < 1) per_sm = 1;
  uint64_t blocks = (uint64_t)num_sms * per_sm;
  if (want_blocks && want_blocks < blocks) blocks = want_blocks;
  kern<<<(unsigned)blocks, kBlockThreads, 0, stream>>>(p);
  e = cudaGetLastError();
  return e == cudaSuccess ? 1 : -(int)e;
}

}  // namespace

cudaError_t launch_rtu_decode_nodes(const QNodeB *nodes, GpuNodeB *out, uint64_t n, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  k_rtu_decode_nodes<<<(unsigned)((n + 127) / 128), 128, 0, stream>>>(reinterpret_cast<const uint8_t *>(nodes), out, n);
  return cudaGetLastError();
}
cudaError_t launch_rtu_gather_tris(const TriA *tri, GpuTri *gtri, uint64_t n, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  k_rtu_gather_tris<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(reinterpret_cast<const float *>(tri), gtri, n);
  return cudaGetLastError();
}

int launch_rtu_render(const RtuRender &p, int num_sms, cudaStream_t stream) {
  if (p.mode & 3u) {   // shadow rays / diffuse bounces
    if (p.strict_order) return p.stats ? launch_persistent(k_rtu_render<true, true, true>, p, num_sms, stream, 0)
                                       : launch_persistent(k_rtu_render<true, false, true>, p, num_sms, stream, 0);
    return p.stats ? launch_persistent(k_rtu_render<false, true, true>, p, num_sms, stream, 0)
                   : launch_persistent(k_rtu_render<false, false, true>, p, num_sms, stream, 0);
  }
  if (p.strict_order) return p.stats ? launch_persistent(k_rtu_render<true, true>, p, num_sms, stream, 0)
                                     : launch_persistent(k_rtu_render<true, false>, p, num_sms, stream, 0);
  return p.stats ? launch_persistent(k_rtu_render<false, true>, p, num_sms, stream, 0)
                 : launch_persistent(k_rtu_render<false, false>, p, num_sms, stream, 0);
}

namespace {
template <class P>
int launch_program_t(const RtuRender &p, int num_sms, cudaStream_t stream) {
  if (p.strict_order) return p.stats ? launch_persistent(k_rtu_program<P, true, true>, p, num_sms, stream, 0)
                                     : launch_persistent(k_rtu_program<P, true, false>, p, num_sms, stream, 0);
  return p.stats ? launch_persistent(k_rtu_program<P, false, true>, p, num_sms, stream, 0)
                 : launch_persistent(k_rtu_program<P, false, false>, p, num_sms, stream, 0);
}
}  // namespace

// shader programs over the traceRay/getWork/getAttr/commit shim (vx_raytrace_b200.cuh); `program` indexes rt_rtu_programs.cuh
int launch_rtu_program(const RtuRender &p, uint32_t program, int num_sms, cudaStream_t stream) {
  // the closest-hit shader recurses through the SBT for mirror rays: give the device stack room for kMaxBounce frames
  size_t cur = 0;   // per device, so queried at every launch (a host call, microseconds)
  if (cudaDeviceGetLimit(&cur, cudaLimitStackSize) != cudaSuccess || cur < 12288) {
    cudaError_t e = cudaDeviceSetLimit(cudaLimitStackSize, 12288);
    if (e != cudaSuccess) return -(int)e;
  }
  switch (program) {
    case 0: return launch_program_t<ShaderSet<false>>(p, num_sms, stream);
    case 1: return launch_program_t<ShaderSet<true>>(p, num_sms, stream);
    default: return -(int)cudaErrorInvalidValue;
  }
}

int launch_rtu_trace(const RtuTrace &p, int num_sms, cudaStream_t stream) {
  if (p.count == 0) return 0;
  uint64_t want = (p.count + kBlockThreads - 1) / kBlockThreads;
  if (p.strict_order) return p.stats ? launch_persistent(k_rtu_trace<true, true>, p, num_sms, stream, want)
                                     : launch_persistent(k_rtu_trace<true, false>, p, num_sms, stream, want);
  return p.stats ? launch_persistent(k_rtu_trace<false, true>, p, num_sms, stream, want)
                 : launch_persistent(k_rtu_trace<false, false>, p, num_sms, stream, want);
}

}  // namespace vxb
