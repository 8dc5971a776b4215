// vx_raytrace_b200.cuh — the RT unit's programming model on the GPU: traceRay / getWork / getAttr / commit
// with the reference's names, attribute ids and work-word encoding, so that any-hit / closest-hit / miss shaders written
// for the Vortex RT unit (kernel/include/vx_raytrace.h:21-67, sim/simx/rt_unit.cpp:98-213) compile as CUDA device code.
//
// Included by rt_rtu.cu inside its anonymous namespace (it drives that file's traversal cores); not a standalone header.
//
// Mapping.  The simulated unit is shared by the lanes of a core and keeps rays in maps keyed by a global rayID; here every
// CUDA lane owns a private unit (RtUnit) with a small LIFO of ray slots, which is all the reference's shaders need: a shader
// traces a nested ray and drains getWork() until it is finished before it returns (closest.cpp:104-117), so the newest
// ray is always the one being traversed.
//   traceRay  -> takes a slot, rayID = slot + 1 (0 stays the invalid id, rt_unit.cpp:81), the ray is queued
//   getWork   -> advances the newest queued ray to its next event and returns (1 << (28 + type)) | rayID with
//                type = MISS 0 / CLOSEST 1 / ANY 3 — the index into the shader binding table (tracer.cpp:244-248); 0 = no work
//   getAttr   -> VX_RT_RAY_* / VX_RT_HIT_* of that ray (while an any-hit candidate is pending: the candidate's barycentrics
//                and ids next to the committed distance, as rt_traversal.cpp:135-149 leaves them)
//   commit    -> ACCEPT: candidate becomes the closest hit, traversal resumes;  CONT: candidate ignored, traversal resumes;
//                TERM: the slot is released
// Strict order (VX_DCR_B200_RT_ORDER=1) runs the reference's trail/short-stack algorithm; there a CONT would re-find the
// same candidate forever (the reference restarts from the root with an unchanged hit distance), so it is reported as a
// run error (bit 2) instead of spinning.  Fast order resumes behind the ignored triangle.
#pragma once

constexpr uint32_t VX_RT_RAY_RO_X = 0xFD0, VX_RT_RAY_RO_Y = 0xFD1, VX_RT_RAY_RO_Z = 0xFD2;      // hw/VX_types.toml [rtu_attr]
constexpr uint32_t VX_RT_RAY_RD_X = 0xFD3, VX_RT_RAY_RD_Y = 0xFD4, VX_RT_RAY_RD_Z = 0xFD5;
constexpr uint32_t VX_RT_HIT_DIST = 0xFD6, VX_RT_HIT_BX = 0xFD7, VX_RT_HIT_BY = 0xFD8, VX_RT_HIT_BZ = 0xFD9;
constexpr uint32_t VX_RT_HIT_BLAS_IDX = 0xFDA, VX_RT_HIT_TRI_IDX = 0xFDB, VX_RT_RAY_PAYLOAD_ADDR = 0xFDC;
constexpr uint32_t VX_RT_COMMIT_CONT = 0xFDD, VX_RT_COMMIT_ACCEPT = 0xFDE, VX_RT_COMMIT_TERM = 0xFDF;
constexpr uint32_t kShaderMiss = 0, kShaderClosest = 1, kShaderAny = 3;                           // rt_unit.cpp:10
constexpr uint32_t kUnitRays = kMaxBounce + 1;   // primary + nested rays alive at once per lane
constexpr uint32_t kNoNode = 0xFFFFFFFFu;

// ---- fast order as a resumable walk: trace_fast (rt_rtu.cu) cut at every candidate triangle ----
struct FastWalk {
  ObjRay world;
  Instance in;
  uint32_t cur, sp, leaf_k, leaf_end;
};
__device__ __forceinline__ void fast_begin(FastWalk &w, const Ray &wray) {
  w.world.o = wray.o, w.world.d = wray.d;
  set_inv(w.world);
  w.in.base = 0, w.in.blasIdx = 0, w.in.ray = w.world;
  w.cur = kTlasBit | 0u, w.sp = 0, w.leaf_k = 0, w.leaf_end = 0;
}
// true = traversal complete; false = `cand` holds a triangle closer than hit.dist, waiting for a commit
template <bool STATS>
__device__ bool fast_advance(const RtuScene &sc, const Ray &wray, FastWalk &w, const Hit &hit, Hit &cand, Local &lc, uint32_t *stk_e,
                             float *stk_d) {
  for (;;) {
    while (w.leaf_k < w.leaf_end) {
      uint32_t triIdx = w.leaf_k++;
      if (triIdx >= sc.n_tri) {
        w.leaf_k = w.leaf_end;
        break;
      }
      const uint8_t *tp = reinterpret_cast<const uint8_t *>(sc.gtri + triIdx);
      float4 q0 = ldg4(tp), q1 = ldg4(tp + 16), q2 = ldg4(tp + 32);
      float t, w1, w2;
      if (STATS) lc.tri++;
      Ray rr{w.in.ray.o, w.in.ray.d};
      if (moller_trumbore(rr, mk3(q0.x, q0.y, q0.z), mk3(q1.x, q1.y, q1.z), mk3(q2.x, q2.y, q2.z), t, w1, w2) && t < hit.dist) {
        cand = Hit{t, w1, w2, 1 - w1 - w2, w.in.blasIdx, triIdx};
        return false;
      }
    }
    if (w.cur == kNoNode) {
      bool found = false;
      while (w.sp != 0) {
        --w.sp;
        if (stk_d[w.sp] < hit.dist) {
          w.cur = stk_e[w.sp];
          found = true;
          break;
        }
      }
      if (!found) return true;
    }
    const uint32_t cur = w.cur;
    w.cur = kNoNode;
    simt_step<STATS>(lc);
    const bool top = (cur & kTlasBit) != 0;
    const uint32_t idx = cur & ~kTlasBit;
    if (!(top ? (idx < sc.n_tlas) : (idx < sc.n_bvh))) continue;
    const GpuNodeB *n = (top ? sc.gtlas : sc.gbvh) + idx;
    uint4 info;
    NodeBoxes b;
    load_node(n, info, b);
    if (STATS) lc.box++;
    if (!(info.z & kFlagLeaf)) {
      const ObjRay &r = top ? w.world : w.in.ray;
      const uint32_t child0 = (top ? 0u : w.in.base) + info.x;
      const float inf = __int_as_float(0x7F800000);
      float d[4];
      uint32_t e[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float dk = child_box(b, k, r);
        bool h = ((info.z >> (4 + k)) & 1u) && (dk < hit.dist);
        d[k] = h ? dk : inf;
        e[k] = (top ? kTlasBit : 0u) | (child0 + k);
      }
      // same order as trace_fast: distance ascending, later child first among equals
#define VXB_PSWAP(i, j)                                   \
  if (d[j] < d[i] || (d[j] == d[i] && e[j] > e[i])) {     \
    float td = d[i]; d[i] = d[j]; d[j] = td;              \
    uint32_t te = e[i]; e[i] = e[j]; e[j] = te;           \
  }
      VXB_PSWAP(0, 1) VXB_PSWAP(2, 3) VXB_PSWAP(0, 2) VXB_PSWAP(1, 3) VXB_PSWAP(1, 2)
#undef VXB_PSWAP
      if (d[0] < inf) {
        if (w.sp + 3 <= kStackB) {
          if (d[3] < inf) { stk_e[w.sp] = e[3]; stk_d[w.sp] = d[3]; ++w.sp; }
          if (d[2] < inf) { stk_e[w.sp] = e[2]; stk_d[w.sp] = d[2]; ++w.sp; }
          if (d[1] < inf) { stk_e[w.sp] = e[1]; stk_d[w.sp] = d[1]; ++w.sp; }
        } else {
          lc.err |= 1u;
        }
        w.cur = e[0];
      }
    } else if (top) {
      if (enter_instance(sc, wray, info.y, w.in)) {
        if (STATS) lc.blas++;
        w.cur = w.in.base;
      }
    } else {
      w.leaf_k = info.x, w.leaf_end = info.x + info.y;
    }
  }
}

// ---- the per-lane unit ----
template <bool STRICT, bool STATS>
struct RtUnit {
  enum : uint32_t { kQueued = 1, kAwaitCommit = 2, kReported = 3 };
  struct Slot {
    Ray ray;
    Hit hit;
    void *payload;
    uint32_t state;
  };
  const RtuScene &sc;
  Local &lc;
  uint32_t *stk_e;
  float *stk_d;
  Slot slot[kUnitRays];
  uint32_t top = 0;
  // traversal state of the newest ray
  Hit cand;
  FastWalk walk;                    // fast order
  ObjRay world;                     // strict order: trail + ring stack + pending distance (rt_unit.cpp:87-116)
  uint8_t trail[STRICT ? kTrailLevels : 1];
  Ring stack;
  float pending;

  __device__ RtUnit(const RtuScene &s, Local &l, uint32_t *e, float *d) : sc(s), lc(l), stk_e(e), stk_d(d) {}

  __device__ void traceRay(float ox, float oy, float oz, float dx, float dy, float dz, void *payload, uint32_t &rayID) {
    if (top >= kUnitRays) {   // deeper nesting than the unit keeps: reported as a run error
      lc.err |= 2u;
      rayID = 0;
      return;
    }
    Slot &s = slot[top];
    s.ray = Ray{mk3(ox, oy, oz), mk3(dx, dy, dz)};
    s.hit = Hit{kLargeFloat, 0, 0, 0, 0, 0};
    s.payload = payload;
    s.state = kQueued;
    rayID = ++top;
    lc.rays++;
    if (STRICT) {
      world.o = s.ray.o, world.d = s.ray.d;
      set_inv(world);
      for (uint32_t i = 0; i < kTrailLevels; ++i) trail[STRICT ? i : 0] = 0;
      stack.head = 0, stack.count = 0;
      pending = 0.0f;
    } else {
      fast_begin(walk, s.ray);
    }
  }

  __device__ uint32_t getWork() {
    if (top == 0) return 0;
    Slot &s = slot[top - 1];
    if (s.state != kQueued) return 0;
    bool done;
    if (STRICT) done = strict_traverse<STATS>(sc, s.ray, world, s.hit, pending, trail, stack, lc);
    else done = fast_advance<STATS>(sc, s.ray, walk, s.hit, cand, lc, stk_e, stk_d);
    uint32_t type;
    if (!done) {
      s.state = kAwaitCommit;
      type = kShaderAny;
    } else {
      s.state = kReported;
      type = (s.hit.dist == kLargeFloat) ? kShaderMiss : kShaderClosest;
      if (STATS && type == kShaderClosest) lc.shaded++;
    }
    return (1u << (28 + type)) | (top & 0x0FFFFFFFu);
  }

  __device__ uint32_t getAttr(uint32_t rayID, uint32_t attrID) const {
    if (rayID == 0 || rayID > top) return 0;
    const Slot &s = slot[rayID - 1];
    // strict order writes a candidate's barycentrics / ids straight into the hit record like the reference; fast order keeps
    // the candidate aside until it is accepted
    const bool c = !STRICT && s.state == kAwaitCommit && rayID == top;
    switch (attrID) {
      case VX_RT_RAY_RO_X: return __float_as_uint(s.ray.o.x);
      case VX_RT_RAY_RO_Y: return __float_as_uint(s.ray.o.y);
      case VX_RT_RAY_RO_Z: return __float_as_uint(s.ray.o.z);
      case VX_RT_RAY_RD_X: return __float_as_uint(s.ray.d.x);
      case VX_RT_RAY_RD_Y: return __float_as_uint(s.ray.d.y);
      case VX_RT_RAY_RD_Z: return __float_as_uint(s.ray.d.z);
      case VX_RT_HIT_DIST: return __float_as_uint(s.hit.dist);
      case VX_RT_HIT_BX: return __float_as_uint(c ? cand.bx : s.hit.bx);
      case VX_RT_HIT_BY: return __float_as_uint(c ? cand.by : s.hit.by);
      case VX_RT_HIT_BZ: return __float_as_uint(c ? cand.bz : s.hit.bz);
      case VX_RT_HIT_BLAS_IDX: return c ? cand.blas : s.hit.blas;
      case VX_RT_HIT_TRI_IDX: return c ? cand.tri : s.hit.tri;
      default: return 0;
    }
  }
  __device__ void *getPayload(uint32_t rayID) const {   // VX_RT_RAY_PAYLOAD_ADDR is a 32-bit address there, a pointer here
    return (rayID == 0 || rayID > top) ? nullptr : slot[rayID - 1].payload;
  }

  __device__ void commit(uint32_t rayID, uint32_t actionID) {
    if (rayID == 0 || rayID > top) return;
    Slot &s = slot[rayID - 1];
    if (actionID == VX_RT_COMMIT_TERM) {
      s.state = 0;
      while (top != 0 && slot[top - 1].state == 0) --top;
      return;
    }
    if (s.state != kAwaitCommit || rayID != top) return;
    if (actionID == VX_RT_COMMIT_ACCEPT) {
      if (STRICT) s.hit.dist = pending;   // hits_[rayID].dist = pending_dist (rt_unit.cpp:199-202)
      else s.hit = cand;
      s.state = kQueued;
    } else if (actionID == VX_RT_COMMIT_CONT) {
      if (STRICT) {
        lc.err |= 4u;                     // see the header comment; the candidate is accepted so that the ray still finishes
        s.hit.dist = pending;
      }
      s.state = kQueued;
    }
  }
};
