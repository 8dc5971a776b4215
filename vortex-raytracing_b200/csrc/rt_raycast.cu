// rt_raycast.cu — sm_100a kernels for the layout-A hot path: primary-ray generation, TLAS/BLAS (BVH2)
// traversal, slab and Moller-Trumbore tests, closest-hit shading with the bounce loop, RGB8 pack.
//
// Reference behaviour being reproduced (tests/regression/raycast/):
//   GenerateRay render.h:190-208 | TLASIntersect :143-188 | BLASIntersect :128-141 | BVHIntersect :74-126
//   Trace :210-276 | texSample :8-21 | diffuseLighting :58-70 | ray_t::intersect geometry.h:1416-1468
//   pixel loop tracer.cpp:249-263 / kernel.cpp:9-32
//
// B200 design (not a translation of the above):
//   * one ray per lane, one 8x4-pixel screen tile per warp, warps pull tiles from an atomic counter
//     (persistent CTAs, grid = resident CTAs x SMs);
//   * BVH nodes are consumed in place: the two children of an inner node are adjacent 32-byte nodes, i.e.
//     one 64-byte record fetched with four 128-bit read-only loads; the parent is never re-read because the
//     traversal stack carries (leftFirst,triCount) of the child, not its index;
//   * triangles are re-flattened once per scene into 48-byte edge-form records in leaf order (gather kernel),
//     removing the triIdx indirection and the unaligned 36-byte stride: three 128-bit loads per test;
//   * 1/dir is hoisted out of the slab test (bit-identical: same IEEE divide, done once per ray and BLAS);
//   * fast order: near child first + re-cull at pop against the current closest hit; strict order: the
//     reference's far-first BLAS order with no re-cull (used to prove arithmetic parity);
//   * FP32 only, compiled with -fmad=false so that every result rounds like the reference's x86 build.
#include <cstdio>
#include "rt_device.cuh"

namespace vxb {

namespace {
using namespace dev;

__device__ __forceinline__ uint32_t encode_child(uint32_t leftFirst, uint32_t triCount, uint32_t nodeIdx) {
  if (triCount < kIndirect && leftFirst <= kPayloadMask) return (triCount << 28) | leftFirst;
  return (kIndirect << 28) | nodeIdx;   // big leaf (or huge index): re-read the node when popped
}

// ---- BLASIntersect + BVHIntersect (render.h:74-141) ----
// Returns true as soon as a hit is accepted when ANY (shadow rays); otherwise updates `hit` to the closest.
template <bool STRICT, bool STATS, bool ANY>
__device__ __forceinline__ bool blas_intersect(const SceneA &sc, uint32_t blasIdx, const Ray &wray, Hit &hit, Local &lc,
                                               uint32_t *stk_e, float *stk_d) {
  const BlasNodeA *blas = sc.blas + blasIdx;
  const float *inv = blas->invTransform;
  // ray_t::transform (geometry.h:1411-1414): direction first (w=0), then origin (w=1); no renormalisation
  float M[12];
#pragma unroll
  for (int i = 0; i < 12; ++i) M[i] = __ldg(inv + i);
  Ray r;
  r.d = xform_vec(wray.d, M);
  r.o = xform_pos(wray.o, M);
  const f3 id = mk3(1.0f / r.d.x, 1.0f / r.d.y, 1.0f / r.d.z);
  const bool nan_safe = isinf(id.x) || isinf(id.y) || isinf(id.z) || isnan(id.x) || isnan(id.y) || isnan(id.z);
  const uint8_t *nodes = sc.nodes + (size_t)__ldg(&blas->bvh_offset) * sizeof(BvhNodeA);
  if (STATS) lc.blas++;

  uint32_t sp = 0;
  uint32_t cur = (kIndirect << 28) | 0u;   // root: node 0 of this BLAS, read on entry
  for (;;) {
    simt_step<STATS>(lc);
    uint32_t kind = cur >> 28, first = cur & kPayloadMask, count = kind;
    if (kind == kIndirect) {
      const uint8_t *n = nodes + (size_t)first * sizeof(BvhNodeA);
      float4 a = ldg4(n), b = ldg4(n + 16);
      first = __float_as_uint(a.w);
      count = __float_as_uint(b.w);
    }
    if (count == 0) {
      // inner node: both children are one 64-byte record
      const uint8_t *pair = nodes + (size_t)first * sizeof(BvhNodeA);
      float4 l0 = ldg4(pair), l1 = ldg4(pair + 16), r0 = ldg4(pair + 32), r1 = ldg4(pair + 48);
      float dL = slab(l0, l1, r.o, id, nan_safe);
      float dR = slab(r0, r1, r.o, id, nan_safe);
      if (STATS) lc.box += 2;
      bool hitL = (dL != kLargeFloat) && (dL < hit.dist);
      bool hitR = (dR != kLargeFloat) && (dR < hit.dist);
      uint32_t eL = encode_child(__float_as_uint(l0.w), __float_as_uint(l1.w), first);
      uint32_t eR = encode_child(__float_as_uint(r0.w), __float_as_uint(r1.w), first + 1);
      if (hitL && hitR) {
        // STRICT: render.h:113-118 pushes the nearer child first, so the FARTHER one is visited first
        bool right_next = STRICT ? (dL < dR) : (dR < dL);
        if (sp < kStackSize) {
          stk_e[sp] = right_next ? eL : eR;
          if (!STRICT) stk_d[sp] = right_next ? dL : dR;
          ++sp;
        } else {
          lc.err |= 1u;   // deeper than the reference's own stack[64] (render.h:80): the run is reported as failed
        }
        cur = right_next ? eR : eL;
        continue;
      } else if (hitL) {
        cur = eL;
        continue;
      } else if (hitR) {
        cur = eR;
        continue;
      }
    } else {
      const GpuTri *tp = sc.gtri + first;
      for (uint32_t i = 0; i < count; ++i) {
        float4 q0 = ldg4(tp + i), q1 = ldg4(reinterpret_cast<const uint8_t *>(tp + i) + 16),
               q2 = ldg4(reinterpret_cast<const uint8_t *>(tp + i) + 32);
        float t, w1, w2;
        if (STATS) lc.tri++;
        if (moller_trumbore(r, mk3(q0.x, q0.y, q0.z), mk3(q1.x, q1.y, q1.z), mk3(q2.x, q2.y, q2.z), t, w1, w2) && t < hit.dist) {
          hit.dist = t;
          if (ANY) return true;
          hit.bx = w1, hit.by = w2, hit.bz = 1 - w1 - w2;
          hit.blas = blasIdx;
          hit.tri = __float_as_uint(q0.w);
        }
      }
    }
    // pop
    if (STRICT) {
      if (sp == 0) break;
      cur = stk_e[--sp];
    } else {
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
  return false;
}

// ---- TLASIntersect (render.h:143-188): near-first over instance boxes ----
template <bool STRICT, bool STATS, bool ANY>
__device__ __forceinline__ bool tlas_intersect(const SceneA &sc, const Ray &ray, Hit &hit, Local &lc, uint32_t *stk_e, float *stk_d) {
  lc.rays++;
  uint32_t tstack[kStackSize];
  uint32_t sp = 0;
  uint32_t nodeIdx = sc.tlas_root;
  const f3 id = mk3(1.0f / ray.d.x, 1.0f / ray.d.y, 1.0f / ray.d.z);
  const bool nan_safe = isinf(id.x) || isinf(id.y) || isinf(id.z) || isnan(id.x) || isnan(id.y) || isnan(id.z);
  for (;;) {
    simt_step<STATS>(lc);
    const uint8_t *n = reinterpret_cast<const uint8_t *>(sc.tlas + nodeIdx);
    float4 a = ldg4(n), b = ldg4(n + 16);
    uint32_t leftRight = __float_as_uint(a.w);
    if (leftRight == 0) {
      if (blas_intersect<STRICT, STATS, ANY>(sc, __float_as_uint(b.w), ray, hit, lc, stk_e, stk_d)) return true;
    } else {
      uint32_t left = leftRight & 0xFFFFu, right = leftRight >> 16;
      const uint8_t *pl = reinterpret_cast<const uint8_t *>(sc.tlas + left);
      const uint8_t *pr = reinterpret_cast<const uint8_t *>(sc.tlas + right);
      float dL = slab(ldg4(pl), ldg4(pl + 16), ray.o, id, nan_safe);
      float dR = slab(ldg4(pr), ldg4(pr + 16), ray.o, id, nan_safe);
      if (STATS) lc.box += 2;
      bool hitL = (dL != kLargeFloat) && (dL < hit.dist);
      bool hitR = (dR != kLargeFloat) && (dR < hit.dist);
      if (hitL && hitR) {
        if (dL > dR) {
          uint32_t t = left;
          left = right, right = t;
        }
        if (sp < kStackSize) tstack[sp++] = right;
        else lc.err |= 1u;
        nodeIdx = left;
        continue;
      } else if (hitL) {
        nodeIdx = left;
        continue;
      } else if (hitR) {
        nodeIdx = right;
        continue;
      }
    }
    if (sp == 0) break;
    nodeIdx = tstack[--sp];
  }
  return false;
}

// ---- Trace (render.h:210-276): the reference's mirror-bounce closest-hit shading ----
template <bool STRICT, bool STATS>
__device__ __forceinline__ f3 trace_ref(const SceneA &sc, const KernelArgA &a, Ray cur, Local &lc, uint32_t *stk_e, float *stk_d) {
  const f3 light_pos = mk3(a.light_pos.x, a.light_pos.y, a.light_pos.z);
  const f3 light_color = mk3(a.light_color.x, a.light_color.y, a.light_color.z);
  const f3 ambient = mk3(a.ambient_color.x, a.ambient_color.y, a.ambient_color.z);
  const f3 background = mk3(a.background_color.x, a.background_color.y, a.background_color.z);
  f3 radiance = mk3(0, 0, 0);
  float throughput = 1.0f;
  for (uint32_t bounce = 0; bounce < a.max_depth; ++bounce) {
    Hit hit{kLargeFloat, 0, 0, 0, 0, 0};
    tlas_intersect<STRICT, STATS, false>(sc, cur, hit, lc, stk_e, stk_d);
    if (hit.dist == kLargeFloat) {
      radiance = radiance + background * throughput;
      break;
    }
    if (STATS) lc.shaded++;
    Surface s = surface_at(sc, cur, hit);
    // diffuseLighting (render.h:58-70)
    f3 L = light_pos - s.I;
    float dist = length3(L);
    L = L * (1.0f / dist);
    float att = 1.0f / (1.0f + dist * 0.1f);
    float NdotL = std_max(0.0f, dot3(s.N, L));
    f3 diffuse = s.texColor * (ambient + att * light_color * NdotL);
    radiance = radiance + throughput * diffuse * (1 - s.reflectivity);
    throughput *= s.reflectivity;
    if (s.reflectivity > 0.0f && bounce + 1 < a.max_depth) {
      f3 R = normalize3(cur.d - 2.0f * s.N * dot3(s.N, cur.d));
      cur.o = s.I + R * 0.001f;
      cur.d = R;
      continue;
    }
    radiance = radiance + throughput * background;
    break;
  }
  return radiance;
}

// ---- extended Trace: shadow rays and/or diffuse bounces (no reference counterpart —
//      kept expression for expression in lock-step with the CPU oracles' TraceExt) ----
template <uint32_t MODE, bool STRICT, bool STATS>
__device__ __forceinline__ f3 trace_ext(const SceneA &sc, const KernelArgA &a, Ray cur, uint32_t &seed, Local &lc, uint32_t *stk_e,
                                        float *stk_d) {
  constexpr bool SHADOW = (MODE & 1u) != 0, DIFFUSE = (MODE & 2u) != 0;
  const f3 light_pos = mk3(a.light_pos.x, a.light_pos.y, a.light_pos.z);
  const f3 light_color = mk3(a.light_color.x, a.light_color.y, a.light_color.z);
  const f3 ambient = mk3(a.ambient_color.x, a.ambient_color.y, a.ambient_color.z);
  const f3 background = mk3(a.background_color.x, a.background_color.y, a.background_color.z);
  f3 radiance = mk3(0, 0, 0);
  f3 thr = mk3(1, 1, 1);
  for (uint32_t bounce = 0; bounce < a.max_depth; ++bounce) {
    Hit hit{kLargeFloat, 0, 0, 0, 0, 0};
    tlas_intersect<STRICT, STATS, false>(sc, cur, hit, lc, stk_e, stk_d);
    if (hit.dist == kLargeFloat) {
      radiance = radiance + background * thr;
      break;
    }
    if (STATS) lc.shaded++;
    Surface s = surface_at(sc, cur, hit);
    f3 Ns = (dot3(s.N, cur.d) > 0.0f) ? neg(s.N) : s.N;
    f3 O = s.I + Ns * 0.001f;
    f3 L = light_pos - s.I;
    float dist = length3(L);
    L = L * (1.0f / dist);
    float att = 1.0f / (1.0f + dist * 0.1f);
    float NdotL = std_max(0.0f, dot3(s.N, L));
    if (SHADOW && NdotL > 0.0f) {
      f3 Lv = light_pos - O;
      float d2 = length3(Lv);
      f3 D = Lv * (1.0f / d2);
      Hit sh{d2, 0, 0, 0, 0, 0};
      // any-hit early-out: same boolean as the oracle's "closest hit below d2 exists" 
      if (tlas_intersect<STRICT, STATS, true>(sc, Ray{O, D}, sh, lc, stk_e, stk_d)) NdotL = 0.0f;
    }
    f3 diffuse = s.texColor * (ambient + att * light_color * NdotL);
    if (DIFFUSE) {
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

// ---- the pixel loop (tracer.cpp:249-263): persistent warps pulling 8x4 tiles ----
template <uint32_t MODE, bool STRICT, bool STATS>
__global__ void __launch_bounds__(kBlockThreads) k_render_a(const __grid_constant__ RenderA p) {
  uint32_t stk_e[kStackSize];
  float stk_d[kStackSize];
  const KernelArgA &a = p.arg;
  const uint32_t W = a.dst_width, H = a.dst_height;
  const uint32_t lane = threadIdx.x & 31;
  const uint32_t tcols = (W + kTileW - 1) / kTileW;
  const uint32_t band_rows = p.part.band_rows, world = p.part.world, rank = p.part.rank;
  const uint32_t nbands_total = (H + band_rows - 1) / band_rows;
  const uint32_t nbands = nbands_total > rank ? (nbands_total - rank + world - 1) / world : 0;   // bands j with j % world == rank
  const uint32_t tiles_per_band = (band_rows / kTileH) * tcols;
  const uint32_t ntiles = nbands * tiles_per_band;
  Local lc{0, 0, 0, 0, 0};
  for (;;) {
    uint32_t tile = 0;
    if (lane == 0) tile = atomicAdd(p.tile_counter, 1u);
    tile = __shfl_sync(0xffffffffu, tile, 0);
    if (tile >= ntiles) break;
    uint32_t band_local = tile / tiles_per_band, r = tile % tiles_per_band;
    uint32_t trow = r / tcols, tcol = r % tcols;
    uint32_t x = tcol * kTileW + (lane & (kTileW - 1));
    uint32_t y = (band_local * world + rank) * band_rows + trow * kTileH + (lane / kTileW);
    if (x < W && y < H) {
      uint32_t out_idx = y * W + x;
      f3 color = mk3(0, 0, 0);
      for (uint32_t s = 0; s < a.samples_per_pixel; ++s) {
        Ray ray = generate_ray(x, y, a);
        if (MODE == 0) {
          color = color + trace_ref<STRICT, STATS>(p.scene, a, ray, lc, stk_e, stk_d);
        } else {
          uint32_t seed = wang_hash(out_idx * a.samples_per_pixel + s + 1);
          if (seed == 0) seed = 1;
          color = color + trace_ext<MODE, STRICT, STATS>(p.scene, a, ray, seed, lc, stk_e, stk_d);
        }
      }
      p.out[out_idx] = rgb32f_to_rgb8(color);
    }
  }
  flush_counters(p.counters, lc, STATS);
}

// ---- closest-hit records for explicit or primary rays ----
template <bool STRICT, bool STATS>
__global__ void __launch_bounds__(kBlockThreads) k_trace_a(const __grid_constant__ TraceA p) {
  uint32_t stk_e[kStackSize];
  float stk_d[kStackSize];
  Local lc{0, 0, 0, 0, 0};
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  // whole warps iterate together so that the counter flush below is convergent
  const uint64_t n_round = (p.count + 31) / 32 * 32;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += stride) {
    if (i < p.count) {
      Ray ray;
      if (p.rays) {
        const float *q = p.rays + 6 * i;
        ray.o = mk3(__ldg(q), __ldg(q + 1), __ldg(q + 2));
        ray.d = mk3(__ldg(q + 3), __ldg(q + 4), __ldg(q + 5));
      } else {
        ray = generate_ray((uint32_t)(i % p.arg.dst_width), (uint32_t)(i / p.arg.dst_width), p.arg);
      }
      Hit hit{kLargeFloat, 0, 0, 0, 0, 0};
      tlas_intersect<STRICT, STATS, false>(p.scene, ray, hit, lc, stk_e, stk_d);
      HitA o{hit.dist, hit.bx, hit.by, hit.bz, hit.blas, hit.tri};
      p.hits[i] = o;
    }
  }
  __syncwarp();
  flush_counters(p.counters, lc, STATS);
}

__global__ void k_gather_tris(const float *__restrict__ tri, const uint32_t *__restrict__ triIdx, GpuTri *__restrict__ gtri,
                              uint64_t n_slots, uint64_t n_tris) {
  uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n_slots) return;
  uint32_t id = triIdx[p];
  GpuTri g;
  if (id < n_tris) {
    const float *t = tri + (uint64_t)id * 9;
    float v0x = t[0], v0y = t[1], v0z = t[2];
    g.v0x = v0x, g.v0y = v0y, g.v0z = v0z;
    g.e1x = t[3] - v0x, g.e1y = t[4] - v0y, g.e1z = t[5] - v0z;   // edge1 = v1 - v0 (geometry.h:1417)
    g.e2x = t[6] - v0x, g.e2y = t[7] - v0y, g.e2z = t[8] - v0z;   // edge2 = v2 - v0 (geometry.h:1418)
  } else {
    g.v0x = g.v0y = g.v0z = 0.f;   // out-of-range index: degenerate triangle, never hit (|a| < EPSILON)
    g.e1x = g.e1y = g.e1z = 0.f;
    g.e2x = g.e2y = g.e2z = 0.f;
  }
  g.triIdx = id;
  g._p1 = g._p2 = 0;
  gtri[p] = g;
}

// ---- BVH2 -> 4-wide records (see GpuNode4) ----
__device__ __forceinline__ uint32_t wide_desc(uint32_t leftFirst, uint32_t triCount, uint32_t nodeIdx) {
  if (triCount == 0) return nodeIdx;                                   // inner: kind 0, payload = its own (instance-local) index
  if (triCount < 15u && leftFirst < (1u << 27)) return (triCount << 27) | leftFirst;
  return (15u << 27) | nodeIdx;                                        // big leaf: read the node when visited
}
__global__ void k_build_wide(const uint8_t *__restrict__ nodes, uint64_t n_nodes, const uint32_t *__restrict__ bases, uint32_t n_bases,
                             GpuNode4 *__restrict__ wide) {
  uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= n_nodes) return;
  const uint8_t *n = nodes + g * sizeof(BvhNodeA);
  float4 a = ldg4(n), b = ldg4(n + 16);
  if (__float_as_uint(b.w) != 0) return;                               // a leaf has no wide record
  // instance base: last entry of `bases` that is <= g
  uint32_t lo = 0, hi = n_bases;
  while (hi - lo > 1) {
    uint32_t mid = (lo + hi) / 2;
    if (bases[mid] <= g) lo = mid;
    else hi = mid;
  }
  const uint64_t base = n_bases ? bases[lo] : 0;
  const uint64_t end = (lo + 1 < n_bases) ? bases[lo + 1] : n_nodes;
  GpuNode4 w;
  const float qnan = __uint_as_float(0x7FC00000u);
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    w.lox[k] = w.loy[k] = w.loz[k] = w.hix[k] = w.hiy[k] = w.hiz[k] = qnan;
    w.desc[k] = 0xFFFFFFFFu;
    w._pad[k] = 0;
  }
  int slot = 0;
  auto put = [&](uint32_t local) {                                       // child `local` (instance-local index) becomes a slot
    const uint8_t *c = nodes + (base + local) * sizeof(BvhNodeA);
    float4 c0 = ldg4(c), c1 = ldg4(c + 16);
#if VXB_SLAB_FMA   // A/B variant: boxes padded outward for the one-rounding slab test (rt_device.cuh)
    auto down = [](float v) { for (int k = 0; k < kSlabPadUlps; ++k) v = nextafterf(v, -INFINITY); return v; };
    auto up = [](float v) { for (int k = 0; k < kSlabPadUlps; ++k) v = nextafterf(v, INFINITY); return v; };
    c0.x = down(c0.x), c0.y = down(c0.y), c0.z = down(c0.z), c1.x = up(c1.x), c1.y = up(c1.y), c1.z = up(c1.z);
#endif
    w.lox[slot] = c0.x, w.loy[slot] = c0.y, w.loz[slot] = c0.z;
    w.hix[slot] = c1.x, w.hiy[slot] = c1.y, w.hiz[slot] = c1.z;
    w.desc[slot] = wide_desc(__float_as_uint(c0.w), __float_as_uint(c1.w), local);
    ++slot;
  };
  const uint32_t first = __float_as_uint(a.w);
  for (uint32_t c = first; c < first + 2; ++c) {
    if (base + c + 0 >= end) continue;                                  // malformed index: leave the slot unused
    const uint8_t *cn = nodes + (base + c) * sizeof(BvhNodeA);
    float4 c0 = ldg4(cn), c1 = ldg4(cn + 16);
    uint32_t cfirst = __float_as_uint(c0.w), ccount = __float_as_uint(c1.w);
    if (ccount != 0 || base + cfirst + 1 >= end) put(c);                 // leaf child (or malformed grandchildren): one slot
    else put(cfirst), put(cfirst + 1);                                   // inner child: its two children
  }
  wide[g] = w;
}

template <uint32_t MODE, bool STRICT, bool STATS>
int launch_render_t(const RenderA &p, int num_sms, cudaStream_t stream) {
  auto kern = k_render_a<MODE, STRICT, STATS>;
  int per_sm = 0;
  cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kBlockThreads, 0);
  if (e != cudaSuccess) return -(int)e;
  if (per_sm < 1) per_sm = 1;
  kern<<<num_sms * per_sm, kBlockThreads, 0, stream>>>(p);
  e = cudaGetLastError();
  return e == cudaSuccess ? 1 : -(int)e;
}

template <bool STRICT, bool STATS>
int launch_render_m(const RenderA &p, int num_sms, cudaStream_t stream) {
  switch (p.mode & 3u) {
    case 0: return launch_render_t<0, STRICT, STATS>(p, num_sms, stream);
    case 1: return launch_render_t<1, STRICT, STATS>(p, num_sms, stream);
    case 2: return launch_render_t<2, STRICT, STATS>(p, num_sms, stream);
    default: return launch_render_t<3, STRICT, STATS>(p, num_sms, stream);
  }
}

}  // namespace

cudaError_t launch_gather_tris(const TriA *tri, const uint32_t *triIdx, GpuTri *gtri, uint64_t n_slots, uint64_t n_tris,
                               cudaStream_t stream) {
  if (n_slots == 0) return cudaSuccess;
  unsigned blocks = (unsigned)((n_slots + 255) / 256);
  k_gather_tris<<<blocks, 256, 0, stream>>>(reinterpret_cast<const float *>(tri), triIdx, gtri, n_slots, n_tris);
  return cudaGetLastError();
}

cudaError_t launch_build_wide(const uint8_t *nodes, uint64_t n_nodes, const uint32_t *bases, uint32_t n_bases, GpuNode4 *wide,
                              cudaStream_t stream) {
  if (n_nodes == 0) return cudaSuccess;
  k_build_wide<<<(unsigned)((n_nodes + 127) / 128), 128, 0, stream>>>(nodes, n_nodes, bases, n_bases, wide);
  return cudaGetLastError();
}

int launch_render_a(const RenderA &p, int num_sms, cudaStream_t stream) {
  if (p.strict_order) return p.stats ? launch_render_m<true, true>(p, num_sms, stream) : launch_render_m<true, false>(p, num_sms, stream);
  return p.stats ? launch_render_m<false, true>(p, num_sms, stream) : launch_render_m<false, false>(p, num_sms, stream);
}

int launch_trace_a(const TraceA &p, int num_sms, cudaStream_t stream) {
  if (p.count == 0) return 0;
  uint64_t want = (p.count + kBlockThreads - 1) / kBlockThreads;
  uint64_t cap = (uint64_t)num_sms * 8;
  unsigned blocks = (unsigned)(want < cap ? want : cap);
  if (p.strict_order) {
    if (p.stats) k_trace_a<true, true><<<blocks, kBlockThreads, 0, stream>>>(p);
    else k_trace_a<true, false><<<blocks, kBlockThreads, 0, stream>>>(p);
  } else {
    if (p.stats) k_trace_a<false, true><<<blocks, kBlockThreads, 0, stream>>>(p);
    else k_trace_a<false, false><<<blocks, kBlockThreads, 0, stream>>>(p);
  }
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 1 : -(int)e;
}

}  // namespace vxb
