// rt_rtu_programs.cuh — shader programs written against vx_raytrace_b200.cuh, and the kernel that runs one.
// Included by rt_rtu.cu inside its anonymous namespace.
//
// A program is a struct with static device members
//     any_hit(rt, rayID, arg) / closest_hit(rt, rayID, arg) / miss(rt, rayID, arg)
// and is dispatched through sbt_call<Program>() by the work word's type, the way the reference kernel indexes its shader
// binding table (kernel.cpp:85-91).  Two programs are built in:
//   0  ShippedShaders   the three shaders of tests/regression/raytracing/shaders restated through the API — frames are
//                       identical to the fused kernel k_rtu_render (tests/test_gpu_rtu.py), which proves the shim;
//   1  SkipOddTriangles the same closest-hit / miss, with an any-hit shader that answers VX_RT_COMMIT_CONT for triangles
//                       of odd index (a cut-out): exercises the path the shipped any-hit never takes.
#pragma once

#ifndef VXB_PROGRAM_MIN_BLOCKS
#define VXB_PROGRAM_MIN_BLOCKS 3
#endif

struct RayPayload {     // shaders/shader.h:7-12
  f3 color;
  uint32_t bounce;
  bool done;
};

template <class P, class RT>
__device__ void sbt_call(RT &rt, uint32_t work, const KernelArgB &arg);

template <bool SKIP_ODD>
struct ShaderSet {
  // shaders/anyhit.cpp:8-36: alpha is the constant 1, so the shipped shader always accepts
  template <class RT>
  static __device__ void any_hit(RT &rt, uint32_t rayID, const KernelArgB &) {
    if (rayID == 0) return;
    uint32_t tri = rt.getAttr(rayID, VX_RT_HIT_TRI_IDX);
    rt.commit(rayID, (SKIP_ODD && (tri & 1u)) ? VX_RT_COMMIT_CONT : VX_RT_COMMIT_ACCEPT);
  }
  // shaders/miss.cpp:7-14
  template <class RT>
  static __device__ void miss(RT &rt, uint32_t rayID, const KernelArgB &arg) {
    if (rayID == 0) return;
    RayPayload *payload = static_cast<RayPayload *>(rt.getPayload(rayID));
    payload->color = mk3(arg.background_color.x, arg.background_color.y, arg.background_color.z);
    payload->done = true;
    rt.commit(rayID, VX_RT_COMMIT_TERM);
  }
  // shaders/closest.cpp:11-130.  Not inlined: it re-enters itself through sbt_call for the mirror ray.
  template <class RT>
  static __device__ __noinline__ void closest_hit(RT &rt, uint32_t rayID, const KernelArgB &arg) {
    if (rayID == 0) return;
    Ray ray;
    ray.o = mk3(__uint_as_float(rt.getAttr(rayID, VX_RT_RAY_RO_X)), __uint_as_float(rt.getAttr(rayID, VX_RT_RAY_RO_Y)),
                __uint_as_float(rt.getAttr(rayID, VX_RT_RAY_RO_Z)));
    ray.d = mk3(__uint_as_float(rt.getAttr(rayID, VX_RT_RAY_RD_X)), __uint_as_float(rt.getAttr(rayID, VX_RT_RAY_RD_Y)),
                __uint_as_float(rt.getAttr(rayID, VX_RT_RAY_RD_Z)));
    Hit hit;
    hit.dist = __uint_as_float(rt.getAttr(rayID, VX_RT_HIT_DIST));
    hit.bx = __uint_as_float(rt.getAttr(rayID, VX_RT_HIT_BX));
    hit.by = __uint_as_float(rt.getAttr(rayID, VX_RT_HIT_BY));
    hit.bz = __uint_as_float(rt.getAttr(rayID, VX_RT_HIT_BZ));
    hit.blas = rt.getAttr(rayID, VX_RT_HIT_BLAS_IDX);
    hit.tri = rt.getAttr(rayID, VX_RT_HIT_TRI_IDX);
    RayPayload *payload = static_cast<RayPayload *>(rt.getPayload(rayID));
    const f3 background = mk3(arg.background_color.x, arg.background_color.y, arg.background_color.z);
    SurfaceB s = surface_b(rt.sc, ray, hit);
    f3 L = mk3(arg.light_pos.x, arg.light_pos.y, arg.light_pos.z) - s.I;      // diffuseLighting (rtx_shading.h:55-67)
    float dist = length3(L);
    L = L * (1.0f / dist);
    float att = 1.0f / (1.0f + dist * 0.1f);
    float NdotL = std_max(0.0f, dot3(s.N, L));
    f3 diffuse = s.texColor * (mk3(arg.ambient_color.x, arg.ambient_color.y, arg.ambient_color.z) +
                               att * mk3(arg.light_color.x, arg.light_color.y, arg.light_color.z) * NdotL);
    f3 radiance = mk3(0, 0, 0);
    float throughput = 1.0f;
    radiance = radiance + throughput * diffuse * (1 - s.reflectivity);
    throughput *= s.reflectivity;
    if (s.reflectivity > 0.0f && payload->bounce + 1 < arg.max_depth && payload->bounce + 1 < kMaxBounce) {
      f3 R = normalize3(ray.d - 2.0f * s.N * dot3(s.N, ray.d));
      f3 o = s.I + R * 0.001f;
      RayPayload sec;
      sec.color = mk3(0, 0, 0), sec.bounce = payload->bounce + 1, sec.done = false;
      uint32_t secID;
      rt.traceRay(o.x, o.y, o.z, R.x, R.y, R.z, &sec, secID);
      uint32_t work;
      while ((work = rt.getWork()) != 0) sbt_call<ShaderSet<SKIP_ODD>>(rt, work, arg);   // closest.cpp:109-115
      radiance = radiance + sec.color * throughput;
    } else {
      radiance = radiance + background * throughput;
    }
    payload->color = radiance;
    payload->done = true;
    rt.commit(rayID, VX_RT_COMMIT_TERM);
  }
};

template <class P, class RT>
__device__ void sbt_call(RT &rt, uint32_t work, const KernelArgB &arg) {
  const uint32_t type = (uint32_t)__ffs((int)(work >> 28)) - 1u;   // __builtin_ctz(ret >> 28), kernel.cpp:87
  const uint32_t id = work & 0x0FFFFFFFu;
  if (type == kShaderAny) P::any_hit(rt, id, arg);
  else if (type == kShaderClosest) P::closest_hit(rt, id, arg);
  else if (type == kShaderMiss) P::miss(rt, id, arg);
}

// kernel.cpp:41-133 with one lane in both roles (producer: traceRay per pixel; consumer: getWork + SBT dispatch until the
// pixel's payload is done), on the tile scheduler of k_rtu_render.
template <class P, bool STRICT, bool STATS>
__global__ void __launch_bounds__(kBlockThreads, VXB_PROGRAM_MIN_BLOCKS) k_rtu_program(const __grid_constant__ RtuRender p) {
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
  RtUnit<STRICT, STATS> rt(p.scene, lc, stk_e, stk_d);
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
      Ray ray = generate_ray_b(x, y, a, p.camera);
      RayPayload payload;
      payload.color = mk3(0, 0, 0), payload.bounce = 0, payload.done = false;
      uint32_t rayID;
      rt.traceRay(ray.o.x, ray.o.y, ray.o.z, ray.d.x, ray.d.y, ray.d.z, &payload, rayID);
      uint32_t work;
      while (!payload.done && (work = rt.getWork()) != 0) sbt_call<P>(rt, work, a);
      p.out[y * W + x] = rgb32f_to_rgb8(payload.color);
    }
    __syncwarp();
    if (lane == 0 && grp != 0xFFFFFFFFu && p.sched.cost) atomicMax(p.sched.cost + grp, (uint32_t)clock() - t0);
  }
  flush_counters(p.counters, lc, STATS);
}
