// rt_machine_b.cu — the layout-B (simx RT unit, quantized BVH4) frame and hit-record kernels as the persistent, warp-scheduled
// ray state machine of rt_machine.cu (fast order only; the strict order — the reference's restart trail, visit for visit —
// stays in rt_rtu.cu's one-ray-per-lane kernel, as do frames with more than kMachineBounce nested mirror rays).
//
// Reference behaviour being reproduced (same as rt_rtu.cu):
//   BVHTraverser::traverse            sim/simx/rt_traversal.cpp:26-168     ray_transform / ray_tri / ray_box  :231-339
//   GenerateRay (both cameras)        tests/regression/raytracing/kernel.cpp:13-39
//   closest-hit / miss shaders        tests/regression/raytracing/shaders/closest.cpp:61-127, miss.cpp:12
//
// What is different from rt_rtu.cu's fast order:
//   * every lane owns a ray SLOT; a warp runs the operation most of its lanes wait for — GEN (pull pixels / rays), TRAV (all
//     lanes that hold an inner BLAS node visit it together; leaves, instance entries and TLAS nodes are postponed until fewer
//     than `quorum_inner` lanes hold an inner node), POST (shade, queue the mirror ray in the same slot or write the pixel);
//   * a node's record carries one CURSOR per child (k_rtu_link, run once per upload): an inner child's node index, a small
//     leaf's (count, first triangle), an instance's blasIdx — so a leaf or an instance node is never fetched, and whether a
//     lane holds an inner node is known without a load;
//   * the newest VXB_RTU_RING stack entries live in a shared-memory ring (rt_machine.cu's measurement for incoherent rays).
// Arithmetic is unchanged: same decoded boxes, same slab and Moller-Trumbore expressions, same shading expressions and the
// same (distance ascending, child index descending) visiting order as rt_rtu.cu's trace_fast, so the hit records and frames
// are the ones that kernel produces (tests/test_gpu_rtu.py compares both with the CPU port and the compiled reference).
#include "rt_device.cuh"
#include "rt_rtu.h"

namespace vxb {
namespace {
using namespace dev;

#ifndef VXB_RTU_RING
#define VXB_RTU_RING 0
#endif
#ifndef VXB_RTU_TRACE_RING
#define VXB_RTU_TRACE_RING 8
#endif
#ifndef VXB_RTUM_MIN_BLOCKS
#define VXB_RTUM_MIN_BLOCKS 4
#endif

constexpr uint32_t kTlasBit = 0x80000000u;
constexpr uint32_t kKindShift = 27;
constexpr uint32_t kPay = (1u << kKindShift) - 1;
constexpr uint32_t kKindInstance = 14u, kKindRead = 15u;   // 0 = inner node, 1..13 = BLAS leaf with that many triangles
constexpr uint32_t kFlagLeaf = 2u;
constexpr uint32_t kStackM = 96;
constexpr uint32_t kMachineBounce = 4;                      // nested mirror rays kept per slot (rad/thr columns)

constexpr uint32_t S_EMPTY = 0, S_TRAV = 1, S_DONE = 2, S_MASK = 3;
constexpr uint32_t F_NAN = 8, F_INBLAS = 16;

enum : uint32_t {
  C_PIX, C_WOX, C_WOY, C_WOZ, C_WDX, C_WDY, C_WDZ, C_T0, C_GRP, C_DEPTH,
  C_RAD0,                                                   // kMachineBounce x {rad.x, rad.y, rad.z, thr}
  C_END = C_RAD0 + 4 * kMachineBounce,
  C_TRACE_END = C_T0                                        // hit-record mode keeps only the pixel / ray index and the world ray
};

struct Lane {
  f3 o, d, id;
  float dist, bx, by;
  uint32_t hblas, htri;
  uint32_t cur, sp, st, blas;
  uint32_t steps, base;
};

struct Smem {
  float *base;
  uint32_t tid;
  __device__ __forceinline__ float &f(uint32_t col) const { return base[col * kBlockThreads + tid]; }
  __device__ __forceinline__ uint32_t &u(uint32_t col) const { return reinterpret_cast<uint32_t *>(base)[col * kBlockThreads + tid]; }
  __device__ __forceinline__ f3 get3(uint32_t col) const { return mk3(f(col), f(col + 1), f(col + 2)); }
  __device__ __forceinline__ void put3(uint32_t col, f3 v) const { f(col) = v.x, f(col + 1) = v.y, f(col + 2) = v.z; }
};

struct Stk {
  uint32_t *e;
  float *d;
  uint2 *ring;
};
template <int RING>
__device__ __forceinline__ void stk_push(Lane &L, const Stk &s, uint32_t e, float d) {
  if (RING > 0) {
    constexpr uint32_t K = RING > 0 ? RING : 1;
    if (L.sp - L.base == K) {
      const uint2 o = s.ring[(L.base % K) * kBlockThreads];
      s.e[L.base] = o.x, s.d[L.base] = __uint_as_float(o.y);
      ++L.base;
    }
    s.ring[(L.sp % K) * kBlockThreads] = make_uint2(e, __float_as_uint(d));
  } else {
    s.e[L.sp] = e, s.d[L.sp] = d;
  }
  ++L.sp;
}
template <int RING>
__device__ __forceinline__ void pop(Lane &L, const Stk &s) {
  for (;;) {
    if (L.sp == 0) { L.st = (L.st & ~S_MASK) | S_DONE; return; }
    --L.sp;
    uint32_t e;
    float d;
    if (RING > 0) {
      constexpr uint32_t K = RING > 0 ? RING : 1;
      if (L.sp < L.base) {
        L.base = L.sp;
        e = s.e[L.sp], d = s.d[L.sp];
      } else {
        const uint2 o = s.ring[(L.sp % K) * kBlockThreads];
        e = o.x, d = __uint_as_float(o.y);
      }
    } else {
      e = s.e[L.sp], d = s.d[L.sp];
    }
    if (d < L.dist) { L.cur = e; return; }
  }
}

__device__ __forceinline__ void set_recip(Lane &L) {
  L.id = mk3(1.0f / L.d.x, 1.0f / L.d.y, 1.0f / L.d.z);   // ray_box_intersect recomputes this per call (:323-325): same IEEE divide
  bool bad = isinf(L.id.x) || isinf(L.id.y) || isinf(L.id.z) || isnan(L.id.x) || isnan(L.id.y) || isnan(L.id.z);
  L.st = bad ? (L.st | F_NAN) : (L.st & ~F_NAN);
}

__device__ __forceinline__ void start_ray(Lane &L, const Smem &sm, f3 o, f3 d, Local &lc) {
  sm.put3(C_WOX, o);
  sm.put3(C_WDX, d);
  L.o = o, L.d = d;
  L.st = S_TRAV;
  set_recip(L);
  L.dist = kLargeFloat, L.bx = 0.f, L.by = 0.f, L.hblas = 0, L.htri = 0;
  L.cur = kTlasBit | (kKindRead << kKindShift) | 0u;       // TLAS node 0 (calcNodePtr(tlas_ptr, 0), :41)
  L.sp = 0, L.base = 0, L.steps = 0;
  lc.rays++;
}

template <bool STATS>
__device__ __forceinline__ void lat_note(uint32_t *lat, uint32_t row, uint32_t st, bool active, bool exhausted) {
  if (STATS) {
    const uint32_t s = st & S_MASK;
    const uint32_t col = active ? 3u : (s == S_TRAV ? 2u : (s == S_DONE ? 0u : (exhausted ? 4u : 1u)));
    lat[row * 5u + col]++;
  }
}

// one visit of a node record: four slab tests, nearest child next, the others pushed farthest first.  Order = (distance
// ascending, child index descending), the reference's own (rt_rtu.cu trace_fast; rt_traversal.cpp:76-90).
template <bool STATS, int RING>
__device__ __forceinline__ void visit4(Lane &L, const GpuNodeB *n, bool careful, const Stk &stk, Local &lc) {
  const uint8_t *w = reinterpret_cast<const uint8_t *>(n);
  float4 lox = ldg4(w), loy = ldg4(w + 16), loz = ldg4(w + 32), hix = ldg4(w + 48), hiy = ldg4(w + 64), hiz = ldg4(w + 80);
  uint4 ds = __ldg(reinterpret_cast<const uint4 *>(w + 112));
  if (STATS) lc.box++;
  float d[4];
  uint32_t e[4] = {ds.x, ds.y, ds.z, ds.w};
  uint32_t ki[4] = {0u, 1u, 2u, 3u};
  const float kInf = __int_as_float(0x7F800000);
  {
    float t;
    bool h;
    h = slab_hit(make_float4(lox.x, loy.x, loz.x, 0.f), make_float4(hix.x, hiy.x, hiz.x, 0.f), L.o, L.id, careful, L.dist, t);
    d[0] = (h && e[0] != 0xFFFFFFFFu) ? t : kInf;
    h = slab_hit(make_float4(lox.y, loy.y, loz.y, 0.f), make_float4(hix.y, hiy.y, hiz.y, 0.f), L.o, L.id, careful, L.dist, t);
    d[1] = (h && e[1] != 0xFFFFFFFFu) ? t : kInf;
    h = slab_hit(make_float4(lox.z, loy.z, loz.z, 0.f), make_float4(hix.z, hiy.z, hiz.z, 0.f), L.o, L.id, careful, L.dist, t);
    d[2] = (h && e[2] != 0xFFFFFFFFu) ? t : kInf;
    h = slab_hit(make_float4(lox.w, loy.w, loz.w, 0.f), make_float4(hix.w, hiy.w, hiz.w, 0.f), L.o, L.id, careful, L.dist, t);
    d[3] = (h && e[3] != 0xFFFFFFFFu) ? t : kInf;
  }
#define VXB_CSWAP(i, j)                                        \
  if (d[j] < d[i] || (d[j] == d[i] && ki[j] > ki[i])) {        \
    float td = d[i]; d[i] = d[j]; d[j] = td;                   \
    uint32_t te = e[i]; e[i] = e[j]; e[j] = te;                \
    uint32_t tk = ki[i]; ki[i] = ki[j]; ki[j] = tk;            \
  }
  VXB_CSWAP(0, 1) VXB_CSWAP(2, 3) VXB_CSWAP(0, 2) VXB_CSWAP(1, 3) VXB_CSWAP(1, 2)
#undef VXB_CSWAP
  if (d[0] < kInf) {
    if (L.sp + 3 <= kStackM) {
      if (d[3] < kInf) stk_push<RING>(L, stk, e[3], d[3]);
      if (d[2] < kInf) stk_push<RING>(L, stk, e[2], d[2]);
      if (d[1] < kInf) stk_push<RING>(L, stk, e[1], d[1]);
    } else {
      lc.err |= 1u;
    }
    L.cur = e[0];
  } else {
    pop<RING>(L, stk);
  }
}

// everything that is not an inner BLAS node: TLAS nodes, instance entry, nodes that must be read first, leaves
template <bool STATS, int RING>
__device__ __forceinline__ void step_other(Lane &L, const Smem &sm, const RtuScene &sc, const Stk &stk, Local &lc) {
  const uint32_t kind = (L.cur >> kKindShift) & 15u, pay = L.cur & kPay;
  if (L.cur & kTlasBit) {
    if (L.st & F_INBLAS) {   // back from an instance: the active ray becomes the world-space ray again
      L.o = sm.get3(C_WOX), L.d = sm.get3(C_WDX);
      L.st &= ~F_INBLAS;
      set_recip(L);
    }
    if (kind == kKindInstance) {
      // TLAS leaf (:110-122): ray_transform with the first 12 floats of invTransform (stride 160), jump to qBvh + bvh_offset
      if (STATS) lc.box++;                          // the instance's TLAS node counts as a visited node, as in rt_rtu.cu
      if (pay >= sc.n_blas) { pop<RING>(L, stk); return; }
      const BlasNodeB *b = sc.blas + pay;
      float M[12];
#pragma unroll
      for (int i = 0; i < 12; ++i) M[i] = __ldg(b->invTransform + i);
      const f3 o = L.o, d = L.d;
      L.o = mk3(M[0] * o.x + M[1] * o.y + M[2] * o.z + M[3], M[4] * o.x + M[5] * o.y + M[6] * o.z + M[7],
                M[8] * o.x + M[9] * o.y + M[10] * o.z + M[11]);
      L.d = mk3(M[0] * d.x + M[1] * d.y + M[2] * d.z, M[4] * d.x + M[5] * d.y + M[6] * d.z, M[8] * d.x + M[9] * d.y + M[10] * d.z);
      const uint32_t base = __ldg(&b->bvh_offset);
      if (base >= sc.n_bvh || base > kPay) {       // hostile offset: the instance is skipped, the world ray comes back
        L.o = o, L.d = d;
        pop<RING>(L, stk);
        return;
      }
      L.st |= F_INBLAS;
      set_recip(L);
      L.blas = pay;
      L.cur = (kKindRead << kKindShift) | base;    // node 0 of the instance's BVH
      if (STATS) lc.blas++;
      return;
    }
    if (pay >= sc.n_tlas) { pop<RING>(L, stk); return; }
    const GpuNodeB *n = sc.gtlas + pay;
    if (kind == kKindRead) {
      uint4 info = __ldg(reinterpret_cast<const uint4 *>(reinterpret_cast<const uint8_t *>(n) + 96));
      if (info.z & kFlagLeaf) {
        L.cur = kTlasBit | (kKindInstance << kKindShift) | (info.y & kPay);
        if (info.y > kPay) pop<RING>(L, stk);
        return;
      }
    }
    visit4<STATS, RING>(L, n, (L.st & F_NAN) != 0, stk, lc);   // a TLAS inner node
    return;
  }
  uint32_t first = pay, count = kind;
  if (kind == kKindRead) {
    if (pay >= sc.n_bvh) { pop<RING>(L, stk); return; }
    uint4 info = __ldg(reinterpret_cast<const uint4 *>(reinterpret_cast<const uint8_t *>(sc.gbvh + pay) + 96));
    if (!(info.z & kFlagLeaf)) {
      L.cur = pay;                                  // an inner node after all: visited by the inner loop
      return;
    }
    first = info.x, count = info.y;
  }
  if (STATS) lc.box++;                              // the leaf node
  Ray r{L.o, L.d};
  for (uint32_t i = 0; i < count; ++i) {           // BLAS leaf (:126-160), every triangle tested, the closest kept
    const uint32_t triIdx = first + i;
    if (triIdx >= sc.n_tri) break;
    const uint8_t *q = reinterpret_cast<const uint8_t *>(sc.gtri + triIdx);
    float4 q0 = ldg4(q), q1 = ldg4(q + 16), q2 = ldg4(q + 32);
    float t, w1, w2;
    if (STATS) lc.tri++;
    if (moller_trumbore(r, mk3(q0.x, q0.y, q0.z), mk3(q1.x, q1.y, q1.z), mk3(q2.x, q2.y, q2.z), t, w1, w2) && t < L.dist) {
      L.dist = t, L.bx = w1, L.by = w2;
      L.hblas = L.blas, L.htri = triIdx;
    }
  }
  pop<RING>(L, stk);
}

__device__ __forceinline__ bool is_inner(const Lane &L) { return (L.st & S_MASK) == S_TRAV && (L.cur >> kKindShift) == 0u; }   // BLAS space, kind 0

template <bool STATS, bool COST, int RING>
__device__ __forceinline__ void traverse(Lane &L, const Smem &sm, const RtuScene &sc, uint32_t quorum, uint32_t inner_quorum, const Stk &stk,
                                         Local &lc, uint32_t *lat, bool exhausted) {
  for (;;) {
    const bool careful = __any_sync(0xffffffffu, (L.st & (S_MASK | F_NAN)) == (S_TRAV | F_NAN));
    for (;;) {
      bool inner = is_inner(L);
      unsigned m = __ballot_sync(0xffffffffu, inner);
      if ((uint32_t)__popc(m) < inner_quorum) break;
      lat_note<STATS>(lat, 2, L.st, inner, exhausted);
      if (inner) {
        simt_step<STATS>(lc);
        if (STATS) L.steps++;
        if (COST) sm.u(C_T0) += 1u;
        visit4<STATS, RING>(L, sc.gbvh + (L.cur & kPay), careful, stk, lc);
      }
    }
    const bool other = (L.st & S_MASK) == S_TRAV && !is_inner(L);
    lat_note<STATS>(lat, 2, L.st, other, exhausted);
    if (other) {
      simt_step<STATS>(lc);
      if (STATS) L.steps++;
      if (COST) sm.u(C_T0) += 1u;
      step_other<STATS, RING>(L, sm, sc, stk, lc);
    }
    unsigned t = __ballot_sync(0xffffffffu, (L.st & S_MASK) == S_TRAV);
    if ((uint32_t)__popc(t) < quorum) return;
  }
}

// ---- GenerateRay, both cameras (kernel.cpp:13-39) ----
__device__ __forceinline__ Ray generate_ray_b(uint32_t x, uint32_t y, const KernelArgB &a, uint32_t camera) {
  if (camera == 0) {   // the camera the shipped kernel hard-codes; u and v are DOUBLE expressions narrowed to float (:32-33)
    f3 pos = mk3(0.0f, 100.0f, 0.0f), front = mk3(1.0f, 0.0f, 0.0f);
    float FOV = 1.0f;
    float u = (float)(((double)x * 2.0 - (double)a.dst_width) / (double)a.dst_height);
    float v = (float)(((double)y * 2.0 - (double)a.dst_height) / (double)a.dst_height);
    f3 right = cross3(front, mk3(0.0f, 1.0f, 0.0f));
    f3 up = cross3(right, front);
    return Ray{pos, normalize3(u * right + v * up + FOV * front)};
  }
  float x_ndc = (x + 0.5f) / (float)a.dst_width - 0.5f;
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

struct SurfaceB {
  f3 I, N, texColor;
  float reflectivity;
};
__device__ __forceinline__ SurfaceB surface_b(const RtuScene &sc, const Ray &ray, float dist, float bx, float by, uint32_t hblas, uint32_t htri) {
  const float bz = 1 - bx - by;
  const BlasNodeB *blas = sc.blas + hblas;
  const float *tx = reinterpret_cast<const float *>(sc.triEx + htri);   // N0 N1 N2 uv0 uv1 uv2 texId (16 words)
  f3 N0 = mk3(__ldg(tx + 0), __ldg(tx + 1), __ldg(tx + 2));
  f3 N1 = mk3(__ldg(tx + 3), __ldg(tx + 4), __ldg(tx + 5));
  f3 N2 = mk3(__ldg(tx + 6), __ldg(tx + 7), __ldg(tx + 8));
  float u0 = __ldg(tx + 9), v0 = __ldg(tx + 10), u1 = __ldg(tx + 11), v1 = __ldg(tx + 12), u2 = __ldg(tx + 13), v2 = __ldg(tx + 14);
  uint32_t texId = __float_as_uint(__ldg(tx + 15));
  SurfaceB s;
  s.I = ray.o + ray.d * dist;
  f3 N = N1 * bx + N2 * by + N0 * bz;                                    // closest.cpp:66
  const float *inv = blas->invTransform;                                 // transposed(): 3x3 only, cells 3,7,11 = 0
  float T[12] = {__ldg(inv + 0), __ldg(inv + 4), __ldg(inv + 8), 0.0f, __ldg(inv + 1), __ldg(inv + 5), __ldg(inv + 9), 0.0f,
                 __ldg(inv + 2), __ldg(inv + 6), __ldg(inv + 10), 0.0f};
  s.N = normalize3(xform_vec(N, T));
  float uvx = u1 * bx + u2 * by + u0 * bz;
  float uvy = v1 * bx + v2 * by + v0 * bz;
  s.texColor = mk3(0.f, 0.f, 0.f);
  if (sc.mat && texId < sc.n_mat) {
    const MaterialB *m = sc.mat + texId;
    if (__ldg(&m->diffuse_tex_id) >= 0) {                                 // texSample (rtx_shading.h:5-18)
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

constexpr uint32_t kModeFrame = 0, kModeTrace = 4;

// POST: closest-hit / miss shaders.  The shaders recurse (closest-hit traces the mirror ray and adds secPayload.color *
// throughput, closest.cpp:104-119): the forward pass records (radiance_k, throughput_k) per level in the slot's columns, the
// unwind evaluates them in the shaders' own order of floating-point operations.
template <uint32_t MODE, bool STATS, bool COST>
__device__ __forceinline__ void post(Lane &L, const Smem &sm, const RtuMachine &p, Local &lc) {
  if (MODE == kModeTrace) {
    const bool hit = L.dist != kLargeFloat;
    p.hits[sm.u(C_PIX)] = HitA{L.dist, hit ? L.bx : 0.f, hit ? L.by : 0.f, hit ? 1 - L.bx - L.by : 0.f, L.hblas, L.htri};
    L.st = S_EMPTY;
    return;
  }
  const KernelArgB &a = p.arg;
  const f3 background = mk3(a.background_color.x, a.background_color.y, a.background_color.z);
  uint32_t depth = sm.u(C_DEPTH);
  f3 color;
  if (L.dist == kLargeFloat) {
    color = background;                                                   // miss.cpp:12
  } else {
    if (STATS) lc.shaded++;
    const f3 light_pos = mk3(a.light_pos.x, a.light_pos.y, a.light_pos.z);
    const f3 light_color = mk3(a.light_color.x, a.light_color.y, a.light_color.z);
    const f3 ambient = mk3(a.ambient_color.x, a.ambient_color.y, a.ambient_color.z);
    Ray wr{sm.get3(C_WOX), sm.get3(C_WDX)};
    SurfaceB s = surface_b(p.scene, wr, L.dist, L.bx, L.by, L.hblas, L.htri);
    f3 Lv = light_pos - s.I;                                              // diffuseLighting (rtx_shading.h:55-67)
    float dist = length3(Lv);
    Lv = Lv * (1.0f / dist);
    float att = 1.0f / (1.0f + dist * 0.1f);
    float NdotL = std_max(0.0f, dot3(s.N, Lv));
    f3 diffuse = s.texColor * (ambient + att * light_color * NdotL);
    f3 radiance = mk3(0, 0, 0);
    float throughput = 1.0f;
    radiance = radiance + throughput * diffuse * (1 - s.reflectivity);
    throughput *= s.reflectivity;
    if (s.reflectivity > 0.0f && depth + 1 < a.max_depth && depth + 1 < kMachineBounce + 1) {
      f3 R = normalize3(wr.d - 2.0f * s.N * dot3(s.N, wr.d));
      sm.put3(C_RAD0 + 4 * depth, radiance);
      sm.f(C_RAD0 + 4 * depth + 3) = throughput;
      sm.u(C_DEPTH) = depth + 1;
      start_ray(L, sm, s.I + R * 0.001f, R, lc);
      return;
    }
    color = radiance + background * throughput;
  }
  while (depth > 0) {
    --depth;
    color = sm.get3(C_RAD0 + 4 * depth) + color * sm.f(C_RAD0 + 4 * depth + 3);
  }
  p.out[sm.u(C_PIX)] = rgb32f_to_rgb8(color);
  if (COST && sm.u(C_GRP) != 0xFFFFFFFFu) atomicMax(p.sched.cost + sm.u(C_GRP), sm.u(C_T0));
  L.st = S_EMPTY;
}

template <uint32_t MODE, bool STATS, bool COST>
__global__ void __launch_bounds__(kBlockThreads, VXB_RTUM_MIN_BLOCKS) k_rtu_machine(const __grid_constant__ RtuMachine p) {
  constexpr uint32_t kCols = (MODE == kModeTrace) ? C_TRACE_END : C_END;
  constexpr int RING = (MODE == kModeTrace) ? VXB_RTU_TRACE_RING : VXB_RTU_RING;
  __shared__ float s_cols[kCols * kBlockThreads];
  extern __shared__ uint2 s_ring[];
  uint32_t stk_e[kStackM];
  float stk_d[kStackM];
  const Stk stk{stk_e, stk_d, RING > 0 ? s_ring + threadIdx.x : nullptr};
  const Smem sm{s_cols, threadIdx.x};
  const KernelArgB &a = p.arg;
  const uint32_t lane = threadIdx.x & 31;
  const uint32_t W = a.dst_width, H = a.dst_height;
  const uint32_t tcols = (W + kTileW - 1) / kTileW;
  const uint32_t band_rows = p.part.band_rows, world = p.part.world, rank = p.part.rank;
  const uint32_t nbands_total = (H + band_rows - 1) / band_rows;
  const uint32_t nbands = nbands_total > rank ? (nbands_total - rank + world - 1) / world : 0;
  const uint32_t tiles_per_band = (band_rows / kTileH) * tcols;
  uint32_t total = nbands * tiles_per_band * 32u;
  if (MODE != kModeTrace && p.tile_end) total = (min(p.tile_end, nbands * tiles_per_band) - p.tile_first) * 32u;
  if (MODE == kModeTrace) total = (uint32_t)((p.count + 31u) / 32u * 32u);

  Lane L;
  L.st = S_EMPTY, L.sp = 0, L.base = 0, L.cur = 0, L.dist = kLargeFloat, L.bx = L.by = 0.f, L.hblas = L.htri = 0, L.blas = 0, L.steps = 0;
  L.o = L.d = L.id = mk3(0, 0, 0);
  Local lc{};
  uint32_t lat[STATS ? 15 : 1];
  if (STATS)
    for (int k = 0; k < 15; ++k) lat[k] = 0;
  uint32_t chunk_next = 0, chunk_end = 0;
  bool exhausted = (total == 0);

  for (;;) {
    unsigned trav = __ballot_sync(0xffffffffu, (L.st & S_MASK) == S_TRAV);
    unsigned done = __ballot_sync(0xffffffffu, (L.st & S_MASK) == S_DONE);
    const uint32_t quorum = exhausted ? 1u : p.quorum_refill;
    if (trav && ((uint32_t)__popc(trav) >= quorum || (!done && exhausted))) {
      traverse<STATS, COST, RING>(L, sm, p.scene, quorum, exhausted ? 1u : p.quorum_inner, stk, lc, lat, exhausted);
      continue;
    }
    if (done) {
      lat_note<STATS>(lat, 1, L.st, (L.st & S_MASK) == S_DONE, exhausted);
      if (STATS && (L.st & S_MASK) == S_DONE) atomicAdd(&p.counters->steps_hist[L.steps ? min(15, 32 - __clz(L.steps)) : 0], 1ull);
      if ((L.st & S_MASK) == S_DONE) post<MODE, STATS, COST>(L, sm, p, lc);
      continue;
    }
    if (exhausted) {
      if (!trav) break;
      continue;
    }
    unsigned need = __ballot_sync(0xffffffffu, (L.st & S_MASK) == S_EMPTY);
    while (need) {
      if (chunk_next == chunk_end) {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(p.tile_counter, p.chunk);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= total) {
          exhausted = true;
          break;
        }
        chunk_next = base;
        chunk_end = min(base + p.chunk, total);
      }
      uint32_t avail = chunk_end - chunk_next;
      bool mine = (need >> lane) & 1u;
      uint32_t my_rank = __popc(need & ((1u << lane) - 1u));
      bool take = mine && my_rank < avail;
      if (take) {
        uint32_t item = chunk_next + my_rank;
        if (MODE == kModeTrace) {
          if (item < p.count) {
            sm.u(C_PIX) = item;
            Ray r;
            if (p.rays) {
              const float *q = p.rays + 6ull * item;
              r.o = mk3(__ldg(q), __ldg(q + 1), __ldg(q + 2));
              r.d = mk3(__ldg(q + 3), __ldg(q + 4), __ldg(q + 5));
            } else {
              r = generate_ray_b(item % W, item / W, a, p.camera);
            }
            start_ray(L, sm, r.o, r.d, lc);
          }
        } else {
          uint32_t grp, in_tile = item & 31u;
          uint32_t tile = sched_tile(p.sched, (item >> 5) + p.tile_first, total >> 5, grp);
          uint32_t band_local = tile / tiles_per_band, r = tile % tiles_per_band;
          uint32_t trow = r / tcols, tcol = r % tcols;
          uint32_t x = tcol * kTileW + (in_tile & (kTileW - 1));
          uint32_t y = (band_local * world + rank) * band_rows + trow * kTileH + (in_tile / kTileW);
          if (x < W && y < H) {
            sm.u(C_PIX) = y * W + x;
            sm.u(C_DEPTH) = 0u;
            if (COST) sm.u(C_GRP) = (in_tile & 7u) ? 0xFFFFFFFFu : grp, sm.u(C_T0) = 0u;
            // kernel.cpp:66-81: every sample traces the same ray into the same payload, so one evaluation is the result
            Ray ray = generate_ray_b(x, y, a, p.camera);
            start_ray(L, sm, ray.o, ray.d, lc);
          }
        }
      }
      lat_note<STATS>(lat, 0, L.st, take, false);
      unsigned took = __ballot_sync(0xffffffffu, take);
      chunk_next += (uint32_t)__popc(took);
      need &= ~took;
    }
  }
  __syncwarp();
  flush_counters(p.counters, lc, STATS);
  if (STATS)
    for (int k = 0; k < 15; ++k) {
      const unsigned v = __reduce_add_sync(0xffffffffu, lat[k]);
      if (lane == 0 && v) atomicAdd(&p.counters->latency[k], (unsigned long long)v);
    }
}

// ---- frames, one ray per lane, over the same cursor-carrying records ----
// Coherent primary rays leave the state machine nothing to schedule (measured: 7 % slower than one ray per lane), but the child
// cursors pay on their own: a leaf or an instance node is never fetched (2 of the ~10 node reads of a primary ray), and the
// per-lane loop below is the machine's own step functions without the warp scheduling.  Persistent warps pull 8x4 tiles like
// rt_rtu.cu's k_rtu_render; the tile's clock count feeds the schedule feedback the same way.
template <bool STATS>
__global__ void __launch_bounds__(kBlockThreads, VXB_RTUM_MIN_BLOCKS) k_rtu_direct(const __grid_constant__ RtuMachine p) {
  __shared__ float s_cols[C_END * kBlockThreads];
  uint32_t stk_e[kStackM];
  float stk_d[kStackM];
  const Stk stk{stk_e, stk_d, nullptr};
  const Smem sm{s_cols, threadIdx.x};
  const KernelArgB &a = p.arg;
  const uint32_t W = a.dst_width, H = a.dst_height;
  const uint32_t lane = threadIdx.x & 31;
  const uint32_t tcols = (W + kTileW - 1) / kTileW;
  const uint32_t band_rows = p.part.band_rows, world = p.part.world, rank = p.part.rank;
  const uint32_t nbands_total = (H + band_rows - 1) / band_rows;
  const uint32_t nbands = nbands_total > rank ? (nbands_total - rank + world - 1) / world : 0;
  const uint32_t tiles_per_band = (band_rows / kTileH) * tcols;
  const uint32_t ntiles = nbands * tiles_per_band;
  Lane L;
  L.st = S_EMPTY, L.sp = 0, L.base = 0, L.cur = 0, L.dist = kLargeFloat, L.bx = L.by = 0.f, L.hblas = L.htri = 0, L.blas = 0, L.steps = 0;
  L.o = L.d = L.id = mk3(0, 0, 0);
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
      sm.u(C_PIX) = y * W + x;
      sm.u(C_DEPTH) = 0u;
      Ray ray = generate_ray_b(x, y, a, p.camera);
      start_ray(L, sm, ray.o, ray.d, lc);
      do {   // one ray after the other of this pixel's path: the primary ray, then the nested mirror rays POST queues
        while ((L.st & S_MASK) == S_TRAV) {
          simt_step<STATS>(lc);
          if (is_inner(L)) visit4<STATS, 0>(L, p.scene.gbvh + (L.cur & kPay), (L.st & F_NAN) != 0, stk, lc);
          else step_other<STATS, 0>(L, sm, p.scene, stk, lc);
        }
        post<kModeFrame, STATS, false>(L, sm, p, lc);
      } while ((L.st & S_MASK) != S_EMPTY);
    }
    __syncwarp();
    if (lane == 0 && grp != 0xFFFFFFFFu && p.sched.cost) atomicMax(p.sched.cost + grp, (uint32_t)clock() - t0);   // longest tile of the group
  }
  __syncwarp();
  flush_counters(p.counters, lc, STATS);
}

// ---- child cursors of the decoded records (once per upload) ----
__global__ void k_rtu_link(GpuNodeB *__restrict__ nodes, uint64_t n_tlas, uint64_t n_bvh, const uint32_t *__restrict__ bases, uint32_t n_bases) {
  uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= n_tlas + n_bvh) return;
  GpuNodeB *me = nodes + g;
  const uint32_t flags = me->flags;
  uint32_t desc[4] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu};
  if (!(flags & kFlagLeaf)) {
    const bool top = g < n_tlas;
    uint64_t base = 0;
    if (!top && n_bases) {   // the instance this node belongs to: last entry of `bases` that is <= its index
      const uint64_t i = g - n_tlas;
      uint32_t lo = 0, hi = n_bases;
      while (hi - lo > 1) {
        uint32_t mid = (lo + hi) / 2;
        if (bases[mid] <= i) lo = mid;
        else hi = mid;
      }
      base = bases[lo];
    }
    const uint64_t limit = top ? n_tlas : n_bvh;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (!((flags >> (4 + k)) & 1u)) continue;
      const uint64_t c = base + (uint64_t)me->leftFirst + k;   // child k lives at base + leftFirst + k (:96-99)
      if (c >= limit || c > kPay) continue;                    // hostile index: the slot stays unused
      const GpuNodeB *ch = nodes + (top ? 0 : n_tlas) + c;
      const uint32_t cf = ch->flags;
      uint32_t d;
      if (!(cf & kFlagLeaf)) d = (top ? kTlasBit : 0u) | (uint32_t)c;
      else if (top) d = ch->leafData <= kPay ? (kTlasBit | (kKindInstance << kKindShift) | ch->leafData) : 0xFFFFFFFFu;
      else if (ch->leafData < kKindInstance && ch->leftFirst <= kPay) d = (ch->leafData << kKindShift) | ch->leftFirst;
      else d = (kKindRead << kKindShift) | (uint32_t)c;
      desc[k] = d;
    }
  }
  me->_pad[1] = desc[0], me->_pad[2] = desc[1], me->_pad[3] = desc[2], me->_pad[4] = desc[3];
}

template <uint32_t MODE, bool STATS>
int launch_bm(const RtuMachine &p, int num_sms, cudaStream_t stream) {
  auto kern = (MODE != kModeTrace && p.sched.cost) ? k_rtu_machine<MODE, STATS, MODE != kModeTrace> : k_rtu_machine<MODE, STATS, false>;
  int per_sm = 0;
  const size_t dyn = (size_t)((MODE == kModeTrace) ? VXB_RTU_TRACE_RING : VXB_RTU_RING) * kBlockThreads * sizeof(uint2);
  cudaError_t e = cudaSuccess;
  if (dyn) e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
  if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kBlockThreads, dyn);
  if (e != cudaSuccess) return -(int)e;
  if (per_sm < 1) per_sm = 1;
  kern<<<num_sms * per_sm, kBlockThreads, dyn, stream>>>(p);
  e = cudaGetLastError();
  return e == cudaSuccess ? 1 : -(int)e;
}

}  // namespace

cudaError_t launch_rtu_link(GpuNodeB *nodes, uint64_t n_tlas, uint64_t n_bvh, const uint32_t *bases, uint32_t n_bases, cudaStream_t stream) {
  const uint64_t n = n_tlas + n_bvh;
  if (n == 0) return cudaSuccess;
  k_rtu_link<<<(unsigned)((n + 127) / 128), 128, 0, stream>>>(nodes, n_tlas, n_bvh, bases, n_bases);
  return cudaGetLastError();
}

bool rtu_machine_supports(const KernelArgB &a) { return a.max_depth <= kMachineBounce + 1; }

int launch_rtu_direct(const RtuMachine &p, int num_sms, cudaStream_t stream) {
  auto kern = p.stats ? k_rtu_direct<true> : k_rtu_direct<false>;
  int per_sm = 0;
  cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kBlockThreads, 0);
  if (e != cudaSuccess) return -(int)e;
  if (per_sm < 1) per_sm = 1;
  kern<<<num_sms * per_sm, kBlockThreads, 0, stream>>>(p);
  e = cudaGetLastError();
  return e == cudaSuccess ? 1 : -(int)e;
}

int launch_rtu_machine(const RtuMachine &p, int num_sms, cudaStream_t stream) {
  if (p.hits) {
    if (p.count == 0) return 0;
    return p.stats ? launch_bm<kModeTrace, true>(p, num_sms, stream) : launch_bm<kModeTrace, false>(p, num_sms, stream);
  }
  return p.stats ? launch_bm<kModeFrame, true>(p, num_sms, stream) : launch_bm<kModeFrame, false>(p, num_sms, stream);
}

}  // namespace vxb
