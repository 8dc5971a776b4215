// rt_device.cuh — device helpers shared by the layout-A kernels: slab / Moller-Trumbore tests, primary-ray generation,
// surface evaluation at a hit, work-counter flush.  Arithmetic follows the reference expression for expression
// (tests/regression/raycast/geometry.h:1416-1468, render.h:8-21,190-208,233-250); compiled with -fmad=false.
#pragma once
#include "rt_math.cuh"
#include "rt_raycast.h"

namespace vxb {
namespace dev {

constexpr int kBlockThreads = 256;
constexpr uint32_t kTileW = 8, kTileH = 4;   // one warp
constexpr uint32_t kIndirect = 15u;          // stack-entry kind: payload is a node index that must be read
constexpr uint32_t kPayloadMask = 0x0FFFFFFFu;

struct Ray {
  f3 o, d;
};
struct Hit {
  float dist;
  float bx, by, bz;
  uint32_t blas, tri;
};
struct Local {   // per-thread work counters
  uint32_t rays, box, tri, blas, shaded, err;
  uint32_t lane_steps, warp_steps;   // traversal iterations of this lane / issued by this lane as its warp's leader
};

// k-th tile handed out by the work counter -> tile of the image and its cost group (0xFFFFFFFF = none)
__device__ __forceinline__ uint32_t sched_tile(const TileSchedule &s, uint32_t k, uint32_t ntiles, uint32_t &group) {
  group = 0xFFFFFFFFu;
  if (!s.order) return k;
  if (k < s.lead) return ntiles - s.lead + k;
  const uint32_t q = k - s.lead, pos = q >> s.group_shift;
  group = __ldg(s.order + pos);
  return (group << s.group_shift) | (q & ((1u << s.group_shift) - 1u));
}

// One traversal iteration (a node fetch + its box tests, or one leaf) is about to run on the lanes converged here.
// lane_steps / (32 * warp_steps) is the SIMT efficiency the simulator reports per RT unit (sim/simx/rt_unit.h:15-45,
// rt_total_simt_efficiency / rt_total_warps).
template <bool STATS>
__device__ __forceinline__ void simt_step(Local &lc) {
  if (STATS) {
    unsigned m = __activemask();
    lc.lane_steps++;
    if ((threadIdx.x & 31u) == (unsigned)(__ffs(m) - 1)) lc.warp_steps++;
  }
}

__device__ __forceinline__ float4 ldg4(const void *p) { return __ldg(reinterpret_cast<const float4 *>(p)); }
// the same load pinned where it is written: the compiler otherwise sinks a node's box loads behind the branch on its header
// word, which turns one round trip per node into two dependent ones (measured on the layout-B walk)
__device__ __forceinline__ float4 ldg4_pinned(const void *p) {
  float4 v;
  asm volatile("ld.global.nc.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}

// ---- ray_t::intersect(aabbMin, aabbMax) (geometry.h:1447-1468) with the reciprocal hoisted ----
// `nan_safe` is set for rays with an infinite reciprocal (a zero/denormal direction component): only then can
// 0*inf = NaN appear, and only then do std::min/std::max differ from the hardware min/max.
__device__ __forceinline__ float slab(float4 lo, float4 hi, f3 o, f3 id, bool nan_safe) {
  float tx1 = (lo.x - o.x) * id.x, tx2 = (hi.x - o.x) * id.x;
  float ty1 = (lo.y - o.y) * id.y, ty2 = (hi.y - o.y) * id.y;
  float tz1 = (lo.z - o.z) * id.z, tz2 = (hi.z - o.z) * id.z;
  float tmin, tmax;
  if (!nan_safe) {
    tmin = fminf(tx1, tx2), tmax = fmaxf(tx1, tx2);
    tmin = fmaxf(tmin, fminf(ty1, ty2)), tmax = fminf(tmax, fmaxf(ty1, ty2));
    tmin = fmaxf(tmin, fminf(tz1, tz2)), tmax = fminf(tmax, fmaxf(tz1, tz2));
  } else {
    tmin = std_min(tx1, tx2), tmax = std_max(tx1, tx2);
    tmin = std_max(tmin, std_min(ty1, ty2)), tmax = std_min(tmax, std_max(ty1, ty2));
    tmin = std_max(tmin, std_min(tz1, tz2)), tmax = std_min(tmax, std_max(tz1, tz2));
  }
  if (tmax < tmin || tmax <= 0) return kLargeFloat;
  return tmin;
}

// Same test, split into (tmin, passed): the reference's "d != LARGE_FLOAT && d < hit.dist" (render.h:109-110) equals
// "passed && tmin < hit.dist" because hit.dist never exceeds LARGE_FLOAT.
__device__ __forceinline__ bool slab_hit(float4 lo, float4 hi, f3 o, f3 id, bool nan_safe, float limit, float &tmin_out) {
  float tx1 = (lo.x - o.x) * id.x, tx2 = (hi.x - o.x) * id.x;
  float ty1 = (lo.y - o.y) * id.y, ty2 = (hi.y - o.y) * id.y;
  float tz1 = (lo.z - o.z) * id.z, tz2 = (hi.z - o.z) * id.z;
  float tmin, tmax;
  if (!nan_safe) {
    tmin = fmaxf(fmaxf(fminf(tx1, tx2), fminf(ty1, ty2)), fminf(tz1, tz2));
    tmax = fminf(fminf(fmaxf(tx1, tx2), fmaxf(ty1, ty2)), fmaxf(tz1, tz2));
  } else {
    tmin = std_min(tx1, tx2), tmax = std_max(tx1, tx2);
    tmin = std_max(tmin, std_min(ty1, ty2)), tmax = std_min(tmax, std_max(ty1, ty2));
    tmin = std_max(tmin, std_min(tz1, tz2)), tmax = std_min(tmax, std_max(tz1, tz2));
  }
  tmin_out = tmin;
  return !(tmax < tmin || tmax <= 0) && (tmin < limit);
}

// A/B variant (-DVXB_SLAB_FMA=1, fast order's 4-wide records only; NOT the default): the slab planes as one FFMA each,
// t = plane * id - o * id, on boxes padded outward by kSlabPadUlps ulps when the records are built.  It is NOT the reference's
// arithmetic (one rounding instead of two, a different cancellation), so it can accept or reject a grazing box the reference
// decides the other way: the hit then still comes from the exact Moller-Trumbore test, but the set of triangles reached may
// differ from the reference's own false negatives.  Measured against the exact test in profiles/r02_sweep_slab_fma.txt.
#ifndef VXB_SLAB_FMA
#define VXB_SLAB_FMA 0
#endif
constexpr int kSlabPadUlps = 2;
__device__ __forceinline__ bool slab_hit_fma(float lox, float loy, float loz, float hix, float hiy, float hiz, f3 id, f3 noid, float limit,
                                             float &tmin_out) {
  float tx1 = __fmaf_rn(lox, id.x, noid.x), tx2 = __fmaf_rn(hix, id.x, noid.x);
  float ty1 = __fmaf_rn(loy, id.y, noid.y), ty2 = __fmaf_rn(hiy, id.y, noid.y);
  float tz1 = __fmaf_rn(loz, id.z, noid.z), tz2 = __fmaf_rn(hiz, id.z, noid.z);
  float tmin = fmaxf(fmaxf(fminf(tx1, tx2), fminf(ty1, ty2)), fminf(tz1, tz2));
  float tmax = fminf(fminf(fmaxf(tx1, tx2), fmaxf(ty1, ty2)), fmaxf(tz1, tz2));
  tmin_out = tmin;
  return !(tmax < tmin || tmax <= 0) && (tmin < limit);
}

// ---- ray_t::intersect(tri) (geometry.h:1416-1445), triangle in edge form ----
__device__ __forceinline__ bool moller_trumbore(const Ray &r, f3 v0, f3 e1, f3 e2, float &t, float &w1o, float &w2o) {
  f3 h = cross3(r.d, e2);
  float a = dot3(e1, h);
  if (fabsf(a) < kEpsilon) return false;
  float f = 1.0f / a;
  f3 s = r.o - v0;
  float w1 = f * dot3(s, h);
  if (w1 < 0 || w1 > 1) return false;
  f3 q = cross3(s, e1);
  float w2 = f * dot3(r.d, q);
  if (w2 < 0 || w1 + w2 > 1) return false;
  float tt = f * dot3(e2, q);
  if (tt <= kEpsilon) return false;
  t = tt, w1o = w1, w2o = w2;
  return true;
}

// ---- GenerateRay (render.h:190-208) ----
__device__ __forceinline__ Ray generate_ray(uint32_t x, uint32_t y, const KernelArgA &a) {
  float x_ndc = (x + 0.5f) / (float)a.dst_width - 0.5f;    // "- 0.5" in double then narrowed == one float subtraction (exact difference)
  float y_ndc = (y + 0.5f) / (float)a.dst_height - 0.5f;
  float x_vp = x_ndc * a.viewplane.x;
  float y_vp = y_ndc * a.viewplane.y;
  f3 right = mk3(a.camera_right.x, a.camera_right.y, a.camera_right.z);
  f3 up = mk3(a.camera_up.x, a.camera_up.y, a.camera_up.z);
  f3 fwd = mk3(a.camera_forward.x, a.camera_forward.y, a.camera_forward.z);
  f3 pos = mk3(a.camera_pos.x, a.camera_pos.y, a.camera_pos.z);
  f3 pt_cam = x_vp * right + y_vp * up + fwd;
  f3 pt_w = pt_cam + pos;
  f3 dir = normalize3(pt_w - pos);
  return Ray{pos, dir};
}

struct Surface {   // what shading needs at a hit
  f3 I, N, texColor;
  float reflectivity;
};

__device__ __forceinline__ Surface surface_at(const SceneA &sc, const Ray &ray, const Hit &hit) {
  const BlasNodeA *blas = sc.blas + hit.blas;
  const float *tx = reinterpret_cast<const float *>(sc.triEx + hit.tri);   // 15 floats: N0 N1 N2 uv0 uv1 uv2
  f3 N0 = mk3(__ldg(tx + 0), __ldg(tx + 1), __ldg(tx + 2));
  f3 N1 = mk3(__ldg(tx + 3), __ldg(tx + 4), __ldg(tx + 5));
  f3 N2 = mk3(__ldg(tx + 6), __ldg(tx + 7), __ldg(tx + 8));
  float u0 = __ldg(tx + 9), v0 = __ldg(tx + 10), u1 = __ldg(tx + 11), v1 = __ldg(tx + 12), u2 = __ldg(tx + 13), v2 = __ldg(tx + 14);
  Surface s;
  s.I = ray.o + ray.d * hit.dist;
  // render.h:241 — weights are (N1,N2,N0) x (w1,w2,1-w1-w2)
  f3 N = N1 * hit.bx + N2 * hit.by + N0 * hit.bz;
  // invTransform.transposed() (geometry.h:1141-1147): 3x3 transposed, cells 3,7,11 = 0
  const float *inv = blas->invTransform;
  float T[12] = {__ldg(inv + 0), __ldg(inv + 4), __ldg(inv + 8), 0.0f, __ldg(inv + 1), __ldg(inv + 5), __ldg(inv + 9), 0.0f,
                 __ldg(inv + 2), __ldg(inv + 6), __ldg(inv + 10), 0.0f};
  s.N = normalize3(xform_vec(N, T));
  float uvx = u1 * hit.bx + u2 * hit.by + u0 * hit.bz;
  float uvy = v1 * hit.bx + v2 * hit.by + v0 * hit.bz;
  // texSample (render.h:8-21)
  uint32_t tw = __ldg(&blas->tex_width), th = __ldg(&blas->tex_height);
  uint32_t iu = f2u_x86(uvx * tw), iv = f2u_x86(uvy * th);
  iu %= tw, iv %= th;
  const uint32_t *pixels = reinterpret_cast<const uint32_t *>(sc.tex + __ldg(&blas->tex_offset));
  s.texColor = rgb8_to_rgb32f(__ldg(pixels + (iu + iv * tw)));
  s.reflectivity = __ldg(&blas->reflectivity);
  return s;
}

__device__ __forceinline__ void flush_counters(WorkCounters *c, const Local &lc, bool stats) {
  unsigned rays = __reduce_add_sync(0xffffffffu, lc.rays);
  unsigned err = __reduce_or_sync(0xffffffffu, lc.err);
  unsigned box = 0, tri = 0, blas = 0, shaded = 0, lsteps = 0, wsteps = 0;
  if (stats) {
    lsteps = __reduce_add_sync(0xffffffffu, lc.lane_steps);
    wsteps = __reduce_add_sync(0xffffffffu, lc.warp_steps);
    box = __reduce_add_sync(0xffffffffu, lc.box);
    tri = __reduce_add_sync(0xffffffffu, lc.tri);
    blas = __reduce_add_sync(0xffffffffu, lc.blas);
    shaded = __reduce_add_sync(0xffffffffu, lc.shaded);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(&c->rays, (unsigned long long)rays);
    if (err) atomicOr(&c->errors, (unsigned long long)err);
    if (stats) {
      atomicAdd(&c->box_tests, (unsigned long long)box);
      atomicAdd(&c->tri_tests, (unsigned long long)tri);
      atomicAdd(&c->blas_entered, (unsigned long long)blas);
      atomicAdd(&c->shaded_hits, (unsigned long long)shaded);
      atomicAdd(&c->lane_steps, (unsigned long long)lsteps);
      atomicAdd(&c->warp_steps, (unsigned long long)wsteps);
      atomicAdd(&c->warps, 1ull);
    }
  }
}

}  // namespace dev
}  // namespace vxb
