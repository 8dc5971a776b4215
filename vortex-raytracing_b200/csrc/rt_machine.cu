// rt_machine.cu — the layout-A render kernel as a persistent, warp-scheduled ray state machine (sm_100a).
//
// Why: ncu on the one-pixel-per-lane kernel (profiles/r01_*) shows 5.5-8.8 active threads per warp instruction: rays of one
// warp finish at very different times (silhouettes, grazing rays, any-hit shadow rays) and inner-node, leaf and shading code
// serialise against each other.  Here every lane owns a ray *slot*; a warp repeatedly runs the one operation most of its
// lanes are waiting for:
//     GEN   empty lanes pull the next pixels from the tile counter (ballot-ranked, so the warp stays compact)
//     TRAV  while-while traversal: all lanes descend inner nodes together; leaves/instance entries are postponed until fewer
//           than kInnerQuorum lanes still have an inner node, then tested together
//     POST  lanes whose traversal finished shade their hit (or resolve their shadow ray) and either queue the next ray of the
//           path in the same slot (shadow ray, mirror/diffuse bounce, next sample) or write the pixel and become empty
// Path state that must survive a traversal (accumulated colour, throughput, RNG seed, the saved shading context of a
// shadow ray, the world-space ray while inside an instance) lives in shared memory, one word column per thread; registers
// hold only the traversal state.  Per-ray arithmetic is unchanged from rt_raycast.cu — same slab and Moller-Trumbore
// expressions, same shading expressions, -fmad=false — so strict order stays bit-identical to the reference
// (tests/regression/raycast/render.h:74-276) and fast order differs from it only in equal-t ties.
#include <cub/device/device_radix_sort.cuh>
#include "rt_device.cuh"

namespace vxb {
namespace {
using namespace dev;

constexpr uint32_t kTlasTag = 0x80000000u;   // stack entry / cursor: payload is a TLAS node index
constexpr uint32_t kKindShift = 27;          // bits 27..30: 0 = inner node pair, 1..14 = leaf with that many triangles, 15 = read the node
constexpr uint32_t kPay = (1u << kKindShift) - 1;
constexpr uint32_t kReadNode = 15u;

// One array serves both levels; the reference keeps 64 entries per level (render.h:80,151), so 128 here never overflows on a
// tree the reference can walk.  Every push is bounds-checked: overflow sets error bit 0 and vx_ready_wait fails the run.
constexpr uint32_t kMachineStack = 2 * kStackSize;

constexpr uint32_t S_EMPTY = 0, S_TRAV = 1, S_DONE = 2, S_MASK = 3;
constexpr uint32_t F_ANY = 4, F_NAN = 8, F_INBLAS = 16, F_OCCL = 32;

// Scheduling parameters come with the launch (RenderA::quorum_refill / quorum_inner / chunk; defaults 12 / 8 / 32 chosen by
// the sweep in profiles/r01_sweep_quorum_chunk.txt): GEN/POST run when fewer than quorum_refill lanes are traversing, the
// inner-node loop is left when fewer than quorum_inner lanes hold an inner node, a warp reserves `chunk` pixels (one 8x4
// tile) per atomicAdd — 256-pixel chunks left SMs idle 32-39 % of the kernel (ncu sm__cycles_active avg vs elapsed).
#ifndef VXB_SORT_MINMAX
#define VXB_SORT_MINMAX 0
#endif
#ifndef VXB_MIN_BLOCKS
#define VXB_MIN_BLOCKS 4
#endif
// A/B variant (-DVXB_SMEM_TOP=1, NOT the default; north_star lists "BVH top levels staged in shared memory" as a candidate):
// every CTA copies the first three levels of the 4-wide tree of a single-instance scene (<= 21 records, 2.7 KB) into shared
// memory and re-labels the cursors between them, so that visits of those records read shared memory instead of L1.
// Measured in profiles/r02_sweep_smem_top.txt.
#ifndef VXB_SMEM_TOP
#define VXB_SMEM_TOP 0
#endif
// ... measured: frames lose 2 % (the staged records are the lines L1 keeps anyway, and generic loads replace ld.global.nc),
// incoherent rays gain 6 % without the stack ring below and lose 1 % with it (profiles/r02_sweep_trace_ring_top.txt): off
#ifndef VXB_TRACE_TOP
#define VXB_TRACE_TOP 0
#endif
constexpr uint32_t kTopTag = 1u << 26;   // inner cursor whose payload is a slot of the staged tree top
constexpr uint32_t kTopSlots = 21;

// shared-memory columns (one 32-bit word per thread each)
enum : uint32_t {
  C_PIX, C_COLX, C_COLY, C_COLZ, C_RADX, C_RADY, C_RADZ, C_THRX, C_THRY, C_THRZ, C_SEED, C_PACK,   // PACK: bounce | sample<<8 | stage<<31
  C_WOX, C_WOY, C_WOZ, C_WDX, C_WDY, C_WDZ,                                                         // world-space ray of the slot
  C_T0, C_GRP,                                                                                      // schedule feedback: clock at GEN, cost group
  C_BASE_END,
  C_TEXX = C_BASE_END, C_TEXY, C_TEXZ, C_ATT, C_NDL, C_NX, C_NY, C_NZ, C_IX, C_IY, C_IZ, C_REFL, C_PDX, C_PDY, C_PDZ,   // shadow continuation
  C_SHADOW_END
};

struct Lane {
  f3 o, d, id;          // active ray (world space at TLAS level, object space inside an instance) and its reciprocal direction
  float dist, bx, by;   // closest hit so far
  uint32_t hblas, htri;
  uint32_t cur, sp, st;
  uint32_t blas;
  uint32_t steps;         // statistics builds: traversal steps of the ray in this slot
  uint32_t base;          // -DVXB_SMEM_STACK: entries below this index are in local memory, [base, sp) in the shared-memory ring
  const uint8_t *nodes;   // first node of the instance being traversed
  const uint8_t *wide;    // its 4-wide records (fast order), same node indexing
};

// cursor of a child node: small leaves and inner nodes carry (triCount, leftFirst) so the node is never re-read; payloads
// fit 27 bits because the launch refuses scenes with >= 2^27 nodes or triangle slots (kMachineMaxNodes)
__device__ __forceinline__ uint32_t enc_child(uint32_t leftFirst, uint32_t triCount, uint32_t nodeIdx) {
  bool direct = triCount < kReadNode;
  return ((direct ? triCount : kReadNode) << kKindShift) | (direct ? leftFirst : nodeIdx);
}

__device__ __forceinline__ void set_recip(Lane &L) {
  L.id = mk3(1.0f / L.d.x, 1.0f / L.d.y, 1.0f / L.d.z);
  bool bad = isinf(L.id.x) || isinf(L.id.y) || isinf(L.id.z) || isnan(L.id.x) || isnan(L.id.y) || isnan(L.id.z);
  L.st = bad ? (L.st | F_NAN) : (L.st & ~F_NAN);
}

struct Smem {
  float *base;
  uint32_t tid;
  __device__ __forceinline__ float &f(uint32_t col) const { return base[col * kBlockThreads + tid]; }
  __device__ __forceinline__ uint32_t &u(uint32_t col) const { return reinterpret_cast<uint32_t *>(base)[col * kBlockThreads + tid]; }
  __device__ __forceinline__ f3 get3(uint32_t col) const { return mk3(f(col), f(col + 1), f(col + 2)); }
  __device__ __forceinline__ void put3(uint32_t col, f3 v) const { f(col) = v.x, f(col + 1) = v.y, f(col + 2) = v.z; }
};

// queue a ray in this lane's slot: TLASIntersect entry (render.h:143-153)
__device__ __forceinline__ void start_ray(Lane &L, const Smem &sm, const SceneA &sc, f3 o, f3 d, float tmax, bool any, Local &lc) {
  sm.put3(C_WOX, o);
  sm.put3(C_WDX, d);
  L.o = o, L.d = d;
  L.st = S_TRAV | (any ? F_ANY : 0u);
  set_recip(L);
  L.dist = tmax, L.bx = 0.f, L.by = 0.f, L.hblas = 0, L.htri = 0;
  L.cur = kTlasTag | sc.tlas_root;
  L.sp = 0;
  L.base = 0;
  L.steps = 0;
  lc.rays++;
}

// statistics builds: one lane-iteration of the latency table (WorkCounters::latency): row = what the warp is doing (0 GEN,
// 1 POST, 2 TRAV), column = what this lane is doing meanwhile.  Called with the warp converged.
template <bool STATS>
__device__ __forceinline__ void lat_note(uint32_t *lat, uint32_t row, uint32_t st, bool active, bool exhausted) {
  if (STATS) {
    const uint32_t s = st & S_MASK;
    const uint32_t col = active ? 3u : (s == S_TRAV ? 2u : (s == S_DONE ? 0u : (exhausted ? 4u : 1u)));
    lat[row * 5u + col]++;
  }
}

// The traversal stack: entries (cursor, entry distance) in local memory.  A/B variant -DVXB_SMEM_STACK=K (NOT the default): the
// newest K entries live in a shared-memory ring (one uint2 column per thread, conflict-free), older ones are written back to
// the local-memory arrays when the ring is full and read from there once the ring has drained — the stack's hot end then no
// longer competes with node records for L1 lines.  Measured in profiles/r02_sweep_smem_stack.txt.
#ifndef VXB_SMEM_STACK
#define VXB_SMEM_STACK 0
#endif
// ... measured: frames lose 3-6 % with K = 8 (shared memory taken from L1, coherent rays keep their stack lines hot anyway),
// incoherent rays gain 13 % (K = 8) / 10 % (K = 16): the hit-record mode uses the ring by default
#ifndef VXB_TRACE_RING
#define VXB_TRACE_RING 8
#endif
struct Stk {
  uint32_t *e;
  float *d;
  uint2 *ring;   // [VXB_SMEM_STACK][kBlockThreads], this thread's column already applied
};
template <int RING>
__device__ __forceinline__ void stk_push(Lane &L, const Stk &s, uint32_t e, float d) {
  if (RING > 0) {
    constexpr uint32_t K = RING > 0 ? RING : 1;
    if (L.sp - L.base == K) {   // ring full: its oldest entry goes to local memory
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
__device__ __forceinline__ void stk_pop(Lane &L, const Stk &s, uint32_t &e, float &d) {   // L.sp > 0
  --L.sp;
  if (RING > 0) {
    constexpr uint32_t K = RING > 0 ? RING : 1;
    if (L.sp < L.base) {   // the ring has drained: older entries are in local memory
      L.base = L.sp;
      e = s.e[L.sp], d = s.d[L.sp];
    } else {
      const uint2 o = s.ring[(L.sp % K) * kBlockThreads];
      e = o.x, d = __uint_as_float(o.y);
    }
  } else {
    e = s.e[L.sp], d = s.d[L.sp];
  }
}

template <bool STRICT, int RING>
__device__ __forceinline__ void pop(Lane &L, const Stk &stk) {
  uint32_t e;
  float d;
  if (STRICT) {   // the reference never re-tests a pushed node (render.h:82-83, :151-152)
    if (L.sp == 0) { L.st = (L.st & ~S_MASK) | S_DONE; return; }
    stk_pop<RING>(L, stk, e, d);
    L.cur = e;
  } else {
    for (;;) {
      if (L.sp == 0) { L.st = (L.st & ~S_MASK) | S_DONE; return; }
      stk_pop<RING>(L, stk, e, d);
      if (d < L.dist) { L.cur = e; return; }
    }
  }
}

// one inner-node visit inside an instance: both children are one 64-byte record (render.h:100-124)
template <bool STRICT, bool STATS, int RING>
__device__ __forceinline__ void step_inner(Lane &L, const SceneA &sc, bool careful, const Stk &stk, Local &lc) {
  uint32_t first = L.cur;   // kind bits are zero for an inner pair
  const uint8_t *pair = L.nodes + (size_t)first * sizeof(BvhNodeA);
  float4 l0 = ldg4(pair), l1 = ldg4(pair + 16), r0 = ldg4(pair + 32), r1 = ldg4(pair + 48);
  float dL, dR;
  bool hitL = slab_hit(l0, l1, L.o, L.id, careful, L.dist, dL);
  bool hitR = slab_hit(r0, r1, L.o, L.id, careful, L.dist, dR);
  if (STATS) lc.box += 2;
  uint32_t eL = enc_child(__float_as_uint(l0.w), __float_as_uint(l1.w), first);
  uint32_t eR = enc_child(__float_as_uint(r0.w), __float_as_uint(r1.w), first + 1);
  if (hitL && hitR) {
    bool right_next = STRICT ? (dL < dR) : (dR < dL);   // STRICT: the reference visits the FARTHER child first (render.h:113-118)
    if (L.sp < kMachineStack) {
      stk_push<RING>(L, stk, right_next ? eL : eR, right_next ? dL : dR);
    } else {
      lc.err |= 1u;
    }
    L.cur = right_next ? eR : eL;
  } else if (hitL) {
    L.cur = eL;
  } else if (hitR) {
    L.cur = eR;
  } else {
    pop<STRICT, RING>(L, stk);
  }
}

// one visit of a 4-wide record (GpuNode4): the cursor's payload is the inner BVH2 node whose grandchildren are tested.
// Nearest slot next, the others pushed farthest first; every pushed entry carries its entry distance for the re-cull at pop.
template <bool STATS, int RING, bool TOP>
__device__ __forceinline__ void step_wide(Lane &L, bool careful, const Stk &stk, Local &lc, const uint8_t *top) {
  float4 lox, loy, loz, hix, hiy, hiz;
  uint4 ds;
  if (TOP) {   // generic loads: lanes of one warp may read shared memory and global memory in the same instruction
    const uint8_t *w = (L.cur & kTopTag) ? top + (size_t)(L.cur & 31u) * sizeof(GpuNode4) : L.wide + (size_t)L.cur * sizeof(GpuNode4);
    const float4 *wv = reinterpret_cast<const float4 *>(w);
    lox = wv[0], loy = wv[1], loz = wv[2], hix = wv[3], hiy = wv[4], hiz = wv[5];
    ds = *reinterpret_cast<const uint4 *>(w + 96);
  } else {
    const uint8_t *w = L.wide + (size_t)L.cur * sizeof(GpuNode4);
    lox = ldg4(w), loy = ldg4(w + 16), loz = ldg4(w + 32), hix = ldg4(w + 48), hiy = ldg4(w + 64), hiz = ldg4(w + 80);
    ds = __ldg(reinterpret_cast<const uint4 *>(w + 96));
  }
  float d[4];
  uint32_t e[4] = {ds.x, ds.y, ds.z, ds.w};
  const float kInf = __int_as_float(0x7F800000);
#if VXB_SLAB_FMA
  if (!careful) {
    const f3 noid = mk3(-(L.o.x * L.id.x), -(L.o.y * L.id.y), -(L.o.z * L.id.z));
    float t;
    bool h;
    h = slab_hit_fma(lox.x, loy.x, loz.x, hix.x, hiy.x, hiz.x, L.id, noid, L.dist, t);
    d[0] = h ? t : kInf;
    h = slab_hit_fma(lox.y, loy.y, loz.y, hix.y, hiy.y, hiz.y, L.id, noid, L.dist, t);
    d[1] = h ? t : kInf;
    h = slab_hit_fma(lox.z, loy.z, loz.z, hix.z, hiy.z, hiz.z, L.id, noid, L.dist, t);
    d[2] = h ? t : kInf;
    h = slab_hit_fma(lox.w, loy.w, loz.w, hix.w, hiy.w, hiz.w, L.id, noid, L.dist, t);
    d[3] = h ? t : kInf;
  } else
#endif
  {
    float t;
    bool h;
    h = slab_hit(make_float4(lox.x, loy.x, loz.x, 0.f), make_float4(hix.x, hiy.x, hiz.x, 0.f), L.o, L.id, careful, L.dist, t);
    d[0] = h ? t : kInf;
    h = slab_hit(make_float4(lox.y, loy.y, loz.y, 0.f), make_float4(hix.y, hiy.y, hiz.y, 0.f), L.o, L.id, careful, L.dist, t);
    d[1] = h ? t : kInf;
    h = slab_hit(make_float4(lox.z, loy.z, loz.z, 0.f), make_float4(hix.z, hiy.z, hiz.z, 0.f), L.o, L.id, careful, L.dist, t);
    d[2] = h ? t : kInf;
    h = slab_hit(make_float4(lox.w, loy.w, loz.w, 0.f), make_float4(hix.w, hiy.w, hiz.w, 0.f), L.o, L.id, careful, L.dist, t);
    d[3] = h ? t : kInf;   // unused slots hold NaN boxes: every comparison of the slab test fails
  }
  if (STATS) lc.box += 4;
#if VXB_SORT_MINMAX   // compare-exchange without moves: distances through min/max (never NaN here), cursors through two selects
#define VXB_CSWAP(i, j)                                     \
  {                                                         \
    const bool sw = d[j] < d[i];                            \
    const float dl = fminf(d[i], d[j]), dh = fmaxf(d[i], d[j]); \
    const uint32_t el = sw ? e[j] : e[i], eh = sw ? e[i] : e[j]; \
    d[i] = dl, d[j] = dh, e[i] = el, e[j] = eh;             \
  }
#else
#define VXB_CSWAP(i, j)                       \
  if (d[j] < d[i]) {                          \
    float td = d[i]; d[i] = d[j]; d[j] = td;  \
    uint32_t te = e[i]; e[i] = e[j]; e[j] = te; \
  }
#endif
  VXB_CSWAP(0, 1) VXB_CSWAP(2, 3) VXB_CSWAP(0, 2) VXB_CSWAP(1, 3) VXB_CSWAP(1, 2)
#undef VXB_CSWAP
  if (d[0] < kInf) {
    if (L.sp + 3 <= kMachineStack) {
      if (d[3] < kInf) stk_push<RING>(L, stk, e[3], d[3]);
      if (d[2] < kInf) stk_push<RING>(L, stk, e[2], d[2]);
      if (d[1] < kInf) stk_push<RING>(L, stk, e[1], d[1]);
    } else {
      lc.err |= 1u;   // stack overflow: vx_ready_wait reports the run as failed
    }
    L.cur = e[0];
  } else {
    pop<false, RING>(L, stk);
  }
}

// everything that is not an inner-node pair: leaves, nodes that must be read first, TLAS nodes / instance entry
template <bool STRICT, bool STATS, int RING, bool TOP>
__device__ __forceinline__ void step_other(Lane &L, const Smem &sm, const SceneA &sc, const Stk &stk, Local &lc, bool top_staged) {
  if (L.cur & kTlasTag) {
    if (L.st & F_INBLAS) {   // back from an instance: the active ray becomes the world-space ray again
      L.o = sm.get3(C_WOX), L.d = sm.get3(C_WDX);
      L.st &= ~F_INBLAS;
      set_recip(L);
    }
    const uint8_t *n = reinterpret_cast<const uint8_t *>(sc.tlas + (L.cur & kPay));
    float4 a = ldg4(n), b = ldg4(n + 16);
    uint32_t leftRight = __float_as_uint(a.w);
    if (leftRight == 0) {
      // BLASIntersect (render.h:128-141): ray_t::transform = direction (w=0) then origin (w=1), not renormalised
      uint32_t bi = __float_as_uint(b.w);
      const BlasNodeA *blas = sc.blas + bi;
      float M[12];
#pragma unroll
      for (int i = 0; i < 12; ++i) M[i] = __ldg(blas->invTransform + i);
      f3 od = xform_vec(L.d, M), oo = xform_pos(L.o, M);
      L.d = od, L.o = oo;
      L.st |= F_INBLAS;
      set_recip(L);
      const size_t off = __ldg(&blas->bvh_offset);
      L.nodes = sc.nodes + off * sizeof(BvhNodeA);
      L.wide = reinterpret_cast<const uint8_t *>(sc.wide) + off * sizeof(GpuNode4);
      L.blas = bi;
      L.cur = kReadNode << kKindShift;   // node 0 of the instance
      if (STATS) lc.blas++;
    } else {
      uint32_t left = leftRight & 0xFFFFu, right = leftRight >> 16;
      const uint8_t *pl = reinterpret_cast<const uint8_t *>(sc.tlas + left);
      const uint8_t *pr = reinterpret_cast<const uint8_t *>(sc.tlas + right);
      bool careful = (L.st & F_NAN) != 0;
      float dL, dR;
      bool hitL = slab_hit(ldg4(pl), ldg4(pl + 16), L.o, L.id, careful, L.dist, dL);
      bool hitR = slab_hit(ldg4(pr), ldg4(pr + 16), L.o, L.id, careful, L.dist, dR);
      if (STATS) lc.box += 2;
      if (hitL && hitR) {   // near child first at this level (render.h:175-180)
        bool swap = dL > dR;
        if (L.sp < kMachineStack) {
          stk_push<RING>(L, stk, kTlasTag | (swap ? left : right), swap ? dL : dR);
        } else {
          lc.err |= 1u;
        }
        L.cur = kTlasTag | (swap ? right : left);
      } else if (hitL) {
        L.cur = kTlasTag | left;
      } else if (hitR) {
        L.cur = kTlasTag | right;
      } else {
        pop<STRICT, RING>(L, stk);
      }
    }
    return;
  }
  uint32_t kind = (L.cur >> kKindShift) & 15u, first = L.cur & kPay, count = kind;
  if (kind == kReadNode) {
    const uint8_t *n = L.nodes + (size_t)first * sizeof(BvhNodeA);
    float4 a = ldg4(n), b = ldg4(n + 16);
    if (!STRICT && first == 0) {
      // instance root: the reference never tests this box, but every box below it lies inside it and IEEE rounding is
      // monotonic, so a ray that fails here fails both children too — rejecting it now skips only failing tests
      float t0;
      if (!slab_hit(a, b, L.o, L.id, (L.st & F_NAN) != 0, L.dist, t0)) {
        pop<STRICT, RING>(L, stk);
        return;
      }
    }
    first = __float_as_uint(a.w);
    count = __float_as_uint(b.w);
    if (count == 0) {   // an inner node after all: visited by the inner loop (its child pair, or its 4-wide record)
      L.cur = STRICT ? (first & kPay) : (L.cur & kPay);
      if (TOP && !STRICT && top_staged && L.cur == 0u) L.cur = kTopTag;   // the root's record sits in slot 0
      return;
    }
  }
  const GpuTri *tp = sc.gtri + first;
  Ray r{L.o, L.d};
  for (uint32_t i = 0; i < count; ++i) {
    const uint8_t *q = reinterpret_cast<const uint8_t *>(tp + i);
    float4 q0 = ldg4(q), q1 = ldg4(q + 16), q2 = ldg4(q + 32);
    float t, w1, w2;
    if (STATS) lc.tri++;
    if (moller_trumbore(r, mk3(q0.x, q0.y, q0.z), mk3(q1.x, q1.y, q1.z), mk3(q2.x, q2.y, q2.z), t, w1, w2) && t < L.dist) {
      L.dist = t;
      if (L.st & F_ANY) {   // shadow ray: any accepted hit below tmax decides
        L.st = (L.st & ~S_MASK) | S_DONE | F_OCCL;
        return;
      }
      L.bx = w1, L.by = w2;
      L.hblas = L.blas;
      L.htri = __float_as_uint(q0.w);
    }
  }
  pop<STRICT, RING>(L, stk);
}

__device__ __forceinline__ bool is_inner(const Lane &L) { return (L.st & S_MASK) == S_TRAV && (L.cur >> kKindShift) == 0u; }

// TRAV: runs until fewer than `quorum` lanes of the warp are still traversing
template <bool STRICT, bool STATS, bool COST, int RING, bool TOP>
__device__ __forceinline__ void traverse(Lane &L, const Smem &sm, const SceneA &sc, uint32_t quorum, uint32_t inner_quorum, const Stk &stk,
                                         Local &lc, uint32_t *lat, bool exhausted, const uint8_t *top) {
  for (;;) {
    // warp-uniform choice of the NaN-exact min/max path (no divergence); F_NAN only changes in step_other / start_ray
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
        if (STRICT) step_inner<STRICT, STATS, RING>(L, sc, careful, stk, lc);   // the reference's own BVH2 walk
        else step_wide<STATS, RING, TOP>(L, careful, stk, lc, top);                   // 4-wide records (the launch guarantees sc.wide)
      }
    }
    lat_note<STATS>(lat, 2, L.st, (L.st & S_MASK) == S_TRAV && !is_inner(L), exhausted);
    if ((L.st & S_MASK) == S_TRAV && !is_inner(L)) {
      simt_step<STATS>(lc);
      if (STATS) L.steps++;
      if (COST) sm.u(C_T0) += 1u;
      step_other<STRICT, STATS, RING, TOP>(L, sm, sc, stk, lc, top != nullptr);
    }
    unsigned t = __ballot_sync(0xffffffffu, (L.st & S_MASK) == S_TRAV);
    if ((uint32_t)__popc(t) < quorum) return;
  }
}

struct Frame {   // warp-uniform view of the launch
  uint32_t W, H, spp, max_depth;
  uint32_t tcols, tiles_per_band, band_rows, world, rank;
  uint32_t total;   // work items (pixel slots of whole 8x4 tiles) of this rank
};

// begin sample `s` of pixel `pix`: GenerateRay + the per-sample seed of the extended modes (tracer.cpp:254-257)
template <uint32_t MODE>
__device__ __forceinline__ void begin_sample(Lane &L, const Smem &sm, const RenderA &p, uint32_t pix, uint32_t s, Local &lc) {
  const KernelArgA &a = p.arg;
  sm.put3(C_RADX, mk3(0, 0, 0));
  sm.put3(C_THRX, mk3(1, 1, 1));
  sm.u(C_PACK) = (s << 8);   // bounce 0, stage = closest
  if (MODE != 0) {
    uint32_t seed = wang_hash(pix * a.samples_per_pixel + s + 1);
    sm.u(C_SEED) = seed ? seed : 1u;
  }
  Ray r = generate_ray(pix % a.dst_width, pix / a.dst_width, a);
  start_ray(L, sm, p.scene, r.o, r.d, kLargeFloat, false, lc);
}

constexpr uint32_t kModeTrace = 4;   // work items are rays, POST writes the ray_hit_t record (vx_b200_trace_*)

// hit-record mode: item `i` is ray i of the caller's buffer (or primary ray i)
__device__ __forceinline__ void begin_trace(Lane &L, const Smem &sm, const RenderA &p, uint32_t i, Local &lc) {
  Ray r;
  if (p.rays) {
    const float *q = p.rays + 6ull * i;
    r.o = mk3(__ldg(q), __ldg(q + 1), __ldg(q + 2));
    r.d = mk3(__ldg(q + 3), __ldg(q + 4), __ldg(q + 5));
  } else {
    r = generate_ray(i % p.arg.dst_width, i / p.arg.dst_width, p.arg);
  }
  start_ray(L, sm, p.scene, r.o, r.d, kLargeFloat, false, lc);
}

// POST: closest-hit shading / shadow resolution and the bounce loop of Trace (render.h:210-276) and of the extended
// modes (shadow rays, diffuse bounces; same expressions as rt_raycast.cu:trace_ext and the CPU oracles' TraceExt)
template <uint32_t MODE, bool STATS, bool COST>
__device__ __forceinline__ void post(Lane &L, const Smem &sm, const RenderA &p, Local &lc) {
  constexpr bool SHADOW = (MODE & 1u) != 0, DIFFUSE = (MODE & 2u) != 0;
  const KernelArgA &a = p.arg;
  if (MODE == kModeTrace) {
    const bool hit = L.dist != kLargeFloat;   // misses keep the all-zero record of render.h's ray_hit_t initialiser
    p.hits[sm.u(C_PIX)] = HitA{L.dist, hit ? L.bx : 0.f, hit ? L.by : 0.f, hit ? 1 - L.bx - L.by : 0.f, L.hblas, L.htri};
    L.st = S_EMPTY;
    return;
  }
  const f3 light_pos = mk3(a.light_pos.x, a.light_pos.y, a.light_pos.z);
  const f3 light_color = mk3(a.light_color.x, a.light_color.y, a.light_color.z);
  const f3 ambient = mk3(a.ambient_color.x, a.ambient_color.y, a.ambient_color.z);
  const f3 background = mk3(a.background_color.x, a.background_color.y, a.background_color.z);
  uint32_t pack = sm.u(C_PACK);
  uint32_t bounce = pack & 255u, sample = (pack >> 8) & 0x7FFFFFu;
  bool shadow_stage = SHADOW && (pack >> 31);
  f3 radiance = sm.get3(C_RADX), thr = sm.get3(C_THRX);
  f3 texColor, N, I, prev_d;
  float att, NdotL, refl;
  bool end_path = false;
  if (shadow_stage) {
    texColor = sm.get3(C_TEXX), N = sm.get3(C_NX), I = sm.get3(C_IX), prev_d = sm.get3(C_PDX);
    att = sm.f(C_ATT), NdotL = sm.f(C_NDL), refl = sm.f(C_REFL);
    if (L.st & F_OCCL) NdotL = 0.0f;
  } else {
    if (L.dist == kLargeFloat) {
      radiance = radiance + background * thr;
      end_path = true;
    } else {
      if (STATS) lc.shaded++;
      Ray wr{sm.get3(C_WOX), sm.get3(C_WDX)};
      Hit h{L.dist, L.bx, L.by, 1 - L.bx - L.by, L.hblas, L.htri};
      Surface s = surface_at(p.scene, wr, h);
      texColor = s.texColor, N = s.N, I = s.I, refl = s.reflectivity, prev_d = wr.d;
      f3 Lv = light_pos - I;                       // diffuseLighting (render.h:58-70)
      float dist = length3(Lv);
      Lv = Lv * (1.0f / dist);
      att = 1.0f / (1.0f + dist * 0.1f);
      NdotL = std_max(0.0f, dot3(N, Lv));
      if (SHADOW && NdotL > 0.0f) {
        f3 Ns = (dot3(N, prev_d) > 0.0f) ? neg(N) : N;
        f3 O = I + Ns * 0.001f;
        f3 Sv = light_pos - O;
        float d2 = length3(Sv);
        f3 D = Sv * (1.0f / d2);
        sm.put3(C_TEXX, texColor), sm.put3(C_NX, N), sm.put3(C_IX, I), sm.put3(C_PDX, prev_d);
        sm.f(C_ATT) = att, sm.f(C_NDL) = NdotL, sm.f(C_REFL) = refl;
        sm.u(C_PACK) = pack | 0x80000000u;
        start_ray(L, sm, p.scene, O, D, d2, true, lc);
        return;
      }
    }
  }
  if (!end_path) {
    f3 diffuse = texColor * (ambient + att * light_color * NdotL);
    if (DIFFUSE) {
      radiance = radiance + thr * diffuse;
      thr = thr * texColor;
      if (bounce + 1 < a.max_depth) {
        uint32_t seed = sm.u(C_SEED);
        float px, py, pz, l2;
        do {
          px = random_float(seed) * 2.0f - 1.0f;
          py = random_float(seed) * 2.0f - 1.0f;
          pz = random_float(seed) * 2.0f - 1.0f;
          l2 = px * px + py * py + pz * pz;
        } while (l2 > 1.0f || l2 < 1e-6f);
        sm.u(C_SEED) = seed;
        float il = 1.0f / sqrtf(l2);
        f3 Ns = (dot3(N, prev_d) > 0.0f) ? neg(N) : N;
        f3 O = I + Ns * 0.001f;
        f3 D = Ns + mk3(px * il, py * il, pz * il);
        float dl2 = dot3(D, D);
        if (dl2 < 1e-12f) D = Ns;
        else D = D * (1.0f / sqrtf(dl2));
        sm.put3(C_RADX, radiance), sm.put3(C_THRX, thr);
        sm.u(C_PACK) = (sample << 8) | (bounce + 1);
        start_ray(L, sm, p.scene, O, D, kLargeFloat, false, lc);
        return;
      }
    } else {
      radiance = radiance + thr * diffuse * (1 - refl);
      thr = thr * refl;
      if (refl > 0.0f && bounce + 1 < a.max_depth) {
        f3 R = normalize3(prev_d - 2.0f * N * dot3(N, prev_d));
        sm.put3(C_RADX, radiance), sm.put3(C_THRX, thr);
        sm.u(C_PACK) = (sample << 8) | (bounce + 1);
        start_ray(L, sm, p.scene, I + R * 0.001f, R, kLargeFloat, false, lc);
        return;
      }
      radiance = radiance + thr * background;
    }
  }
  // end of this sample: color += Trace(...) (tracer.cpp:255-258); no jitter, no division by spp
  uint32_t pix = sm.u(C_PIX);
  f3 color = sm.get3(C_COLX) + radiance;
  if (sample + 1 < a.samples_per_pixel) {
    sm.put3(C_COLX, color);
    begin_sample<MODE>(L, sm, p, pix, sample + 1, lc);
    return;
  }
  p.out[pix] = rgb32f_to_rgb8(color);
  // most traversal steps spent on one pixel of this group (one pixel in eight reports; a reduction without return value, so
  // the warp does not wait for it)
  if (COST && sm.u(C_GRP) != 0xFFFFFFFFu) atomicMax(p.sched.cost + sm.u(C_GRP), sm.u(C_T0));
  L.st = S_EMPTY;
}

template <uint32_t MODE, bool STRICT, bool STATS, bool COST>
__global__ void __launch_bounds__(kBlockThreads, VXB_MIN_BLOCKS) k_render_m(const __grid_constant__ RenderA p) {
  constexpr uint32_t kCols = (MODE == kModeTrace) ? C_BASE_END : ((MODE & 1u) ? C_SHADOW_END : C_BASE_END);
  __shared__ float s_cols[kCols * kBlockThreads];
  uint32_t stk_e[kMachineStack];
  float stk_d[kMachineStack];
  constexpr int RING = (MODE == kModeTrace) ? VXB_TRACE_RING : VXB_SMEM_STACK;
  constexpr bool TOP = !STRICT && ((MODE == kModeTrace) ? VXB_TRACE_TOP : VXB_SMEM_TOP) != 0;
  extern __shared__ uint2 s_ring[];   // RING * kBlockThreads entries (dynamic: beyond the 48 KB static limit)
  const Stk stk{stk_e, stk_d, RING > 0 ? s_ring + threadIdx.x : nullptr};
  const Smem sm{s_cols, threadIdx.x};
  const KernelArgA &a = p.arg;
  const uint32_t lane = threadIdx.x & 31;
  Frame f;
  f.W = a.dst_width, f.H = a.dst_height;
  f.tcols = (f.W + kTileW - 1) / kTileW;
  f.band_rows = p.part.band_rows, f.world = p.part.world, f.rank = p.part.rank;
  const uint32_t nbands_total = (f.H + f.band_rows - 1) / f.band_rows;
  const uint32_t nbands = nbands_total > f.rank ? (nbands_total - f.rank + f.world - 1) / f.world : 0;
  f.tiles_per_band = (f.band_rows / kTileH) * f.tcols;
  f.total = nbands * f.tiles_per_band * 32u;
  if (MODE != kModeTrace && p.tile_end) f.total = (min(p.tile_end, nbands * f.tiles_per_band) - p.tile_first) * 32u;   // one strip
  if (MODE == kModeTrace) f.total = (uint32_t)((p.count + 31u) / 32u * 32u);

  Lane L;
  L.st = S_EMPTY, L.sp = 0, L.base = 0, L.cur = 0, L.dist = kLargeFloat, L.bx = L.by = 0.f, L.hblas = L.htri = 0, L.nodes = p.scene.nodes, L.wide = reinterpret_cast<const uint8_t *>(p.scene.wide), L.blas = 0;
  L.o = L.d = L.id = mk3(0, 0, 0);
  L.steps = 0;
  Local lc{0, 0, 0, 0, 0};
  uint32_t lat[STATS ? 15 : 1];
  if (STATS)
    for (int k = 0; k < 15; ++k) lat[k] = 0;
  const uint8_t *top = nullptr;
  __shared__ __align__(128) GpuNode4 s_top[TOP ? kTopSlots : 1];
  if (TOP && p.scene.wide && p.n_blas == 1) {   // warp-uniform: launch parameters only
    const uint32_t t = threadIdx.x;
    auto inner = [](uint32_t d) { return d != 0xFFFFFFFFu && (d >> kKindShift) == 0u; };
    if (t == 0) s_top[0] = p.scene.wide[0];
    __syncthreads();
    uint32_t d1 = t < 4 ? s_top[0].desc[t] : 0xFFFFFFFFu;
    if (t < 4 && inner(d1)) s_top[1 + t] = p.scene.wide[d1];
    __syncthreads();
    if (t < 4 && inner(d1)) s_top[0].desc[t] = kTopTag | (1 + t);
    uint32_t d2 = 0xFFFFFFFFu;
    __syncthreads();
    if (t < 16 && (s_top[0].desc[t / 4] & kTopTag) && (s_top[0].desc[t / 4] >> kKindShift) == 0u) d2 = s_top[1 + t / 4].desc[t % 4];
    if (t < 16 && inner(d2) && !(d2 & kTopTag)) s_top[5 + t] = p.scene.wide[d2];
    __syncthreads();
    if (t < 16 && inner(d2) && !(d2 & kTopTag)) s_top[1 + t / 4].desc[t % 4] = kTopTag | (5 + t);
    __syncthreads();
    top = reinterpret_cast<const uint8_t *>(s_top);
  }
  uint32_t chunk_next = 0, chunk_end = 0;
  bool exhausted = (f.total == 0);

  for (;;) {
    unsigned trav = __ballot_sync(0xffffffffu, (L.st & S_MASK) == S_TRAV);
    unsigned done = __ballot_sync(0xffffffffu, (L.st & S_MASK) == S_DONE);
    const uint32_t quorum = exhausted ? 1u : p.quorum_refill;
    if (trav && ((uint32_t)__popc(trav) >= quorum || (!done && exhausted))) {
      traverse<STRICT, STATS, COST, RING, TOP>(L, sm, p.scene, quorum, exhausted ? 1u : p.quorum_inner, stk, lc, lat, exhausted, top);
      continue;
    }
    if (done) {
      lat_note<STATS>(lat, 1, L.st, (L.st & S_MASK) == S_DONE, exhausted);
      if (STATS && (L.st & S_MASK) == S_DONE)   // the ray in this slot is complete: its step count goes into the histogram
        atomicAdd(&p.counters->steps_hist[L.steps ? min(15, 32 - __clz(L.steps)) : 0], 1ull);
      if ((L.st & S_MASK) == S_DONE) post<MODE, STATS, COST>(L, sm, p, lc);
      continue;   // POST may have queued follow-up rays; re-evaluate before pulling new pixels
    }
    if (exhausted) {
      if (!trav) break;
      continue;
    }
    // GEN: empty lanes take the next pixels of the warp's chunk, compactly (ballot rank)
    unsigned need = __ballot_sync(0xffffffffu, (L.st & S_MASK) == S_EMPTY);
    while (need) {
      if (chunk_next == chunk_end) {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(p.tile_counter, p.chunk);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= f.total) {
          exhausted = true;
          break;
        }
        chunk_next = base;
        chunk_end = min(base + p.chunk, f.total);
      }
      uint32_t avail = chunk_end - chunk_next;
      bool mine = (need >> lane) & 1u;
      uint32_t my_rank = __popc(need & ((1u << lane) - 1u));
      bool take = mine && my_rank < avail;
      if (take) {
        uint32_t item = chunk_next + my_rank;
        if (MODE == kModeTrace) {
          if (item < p.count) {
            const uint32_t src = p.perm ? __ldg(p.perm + item) : item;   // coherence-ordered batch: item k is ray perm[k]
            sm.u(C_PIX) = src;
            begin_trace(L, sm, p, src, lc);
          }
        } else {
        uint32_t grp, in_tile = item & 31u;
        uint32_t tile = sched_tile(p.sched, (item >> 5) + p.tile_first, f.total >> 5, grp);   // strips run without a schedule order
        uint32_t band_local = tile / f.tiles_per_band, r = tile % f.tiles_per_band;
        uint32_t trow = r / f.tcols, tcol = r % f.tcols;
        uint32_t x = tcol * kTileW + (in_tile & (kTileW - 1));
        uint32_t y = (band_local * f.world + f.rank) * f.band_rows + trow * kTileH + (in_tile / kTileW);
        if (x < f.W && y < f.H) {
          uint32_t pix = y * f.W + x;
          sm.u(C_PIX) = pix;
          if (COST) sm.u(C_GRP) = (in_tile & 7u) ? 0xFFFFFFFFu : grp, sm.u(C_T0) = 0u;
          sm.put3(C_COLX, mk3(0, 0, 0));
          begin_sample<MODE>(L, sm, p, pix, 0, lc);
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

// ---- frames, one pixel per lane, over the same 4-wide records and step functions (fast order; RenderA::direct) ----
// The A/B of round 2 on layout B (rt_machine_b.cu) showed that for coherent rays the warp scheduling costs more than the
// divergence it removes; this kernel is that experiment for layout A: persistent warps pull 8x4 tiles, every lane runs its
// pixel's rays (primary, shadow, bounces, samples) to completion through step_wide / step_other / post.
template <uint32_t MODE, bool STATS>
__global__ void __launch_bounds__(kBlockThreads, VXB_MIN_BLOCKS) k_render_direct(const __grid_constant__ RenderA p) {
  constexpr uint32_t kCols = (MODE & 1u) ? C_SHADOW_END : C_BASE_END;
  __shared__ float s_cols[kCols * kBlockThreads];
  uint32_t stk_e[kMachineStack];
  float stk_d[kMachineStack];
  const Stk stk{stk_e, stk_d, nullptr};
  const Smem sm{s_cols, threadIdx.x};
  const KernelArgA &a = p.arg;
  const uint32_t lane = threadIdx.x & 31;
  const uint32_t W = a.dst_width, H = a.dst_height;
  const uint32_t tcols = (W + kTileW - 1) / kTileW;
  const uint32_t band_rows = p.part.band_rows, world = p.part.world, rank = p.part.rank;
  const uint32_t nbands_total = (H + band_rows - 1) / band_rows;
  const uint32_t nbands = nbands_total > rank ? (nbands_total - rank + world - 1) / world : 0;
  const uint32_t tiles_per_band = (band_rows / kTileH) * tcols;
  const uint32_t all_tiles = nbands * tiles_per_band;
  const uint32_t ntiles = p.tile_end ? min(p.tile_end, all_tiles) - p.tile_first : all_tiles;
  Lane L;
  L.st = S_EMPTY, L.sp = 0, L.base = 0, L.cur = 0, L.dist = kLargeFloat, L.bx = L.by = 0.f, L.hblas = L.htri = 0, L.nodes = p.scene.nodes, L.wide = reinterpret_cast<const uint8_t *>(p.scene.wide), L.blas = 0;
  L.o = L.d = L.id = mk3(0, 0, 0);
  L.steps = 0;
  Local lc{0, 0, 0, 0, 0};
  for (;;) {
    uint32_t t = 0;
    if (lane == 0) t = atomicAdd(p.tile_counter, 1u);
    t = __shfl_sync(0xffffffffu, t, 0);
    if (t >= ntiles) break;
    uint32_t grp;
    const uint32_t tile = sched_tile(p.sched, t + p.tile_first, all_tiles, grp);
    const uint32_t band_local = tile / tiles_per_band, r = tile % tiles_per_band;
    const uint32_t trow = r / tcols, tcol = r % tcols;
    const uint32_t x = tcol * kTileW + (lane & (kTileW - 1));
    const uint32_t y = (band_local * world + rank) * band_rows + trow * kTileH + (lane / kTileW);
    uint32_t steps = 0;
    if (x < W && y < H) {
      const uint32_t pix = y * W + x;
      sm.u(C_PIX) = pix;
      sm.put3(C_COLX, mk3(0, 0, 0));
      begin_sample<MODE>(L, sm, p, pix, 0, lc);
      do {
        while ((L.st & S_MASK) == S_TRAV) {
          simt_step<STATS>(lc);
          ++steps;
          if (is_inner(L)) step_wide<STATS, 0, false>(L, (L.st & F_NAN) != 0, stk, lc, nullptr);
          else step_other<false, STATS, 0, false>(L, sm, p.scene, stk, lc, false);
        }
        post<MODE, STATS, false>(L, sm, p, lc);
      } while ((L.st & S_MASK) != S_EMPTY);
    }
    if (p.sched.cost && grp != 0xFFFFFFFFu) {   // most traversal steps spent on one pixel of this group
      steps = __reduce_max_sync(0xffffffffu, steps);
      if (lane == 0) atomicMax(p.sched.cost + grp, steps);
    }
  }
  __syncwarp();
  flush_counters(p.counters, lc, STATS);
}

template <uint32_t MODE, bool STATS>
int launch_d(const RenderA &p, int num_sms, cudaStream_t stream) {
  auto kern = k_render_direct<MODE, STATS>;
  int per_sm = 0;
  cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kBlockThreads, 0);
  if (e != cudaSuccess) return -(int)e;
  if (per_sm < 1) per_sm = 1;
  kern<<<num_sms * per_sm, kBlockThreads, 0, stream>>>(p);
  e = cudaGetLastError();
  return e == cudaSuccess ? 1 : -(int)e;
}

template <uint32_t MODE, bool STRICT, bool STATS>
int launch_m(const RenderA &p, int num_sms, cudaStream_t stream) {
  // the cost-recording build runs only on the frames that re-measure the schedule (p.sched.cost set), and never for rays
  auto kern = (MODE != kModeTrace && p.sched.cost) ? k_render_m<MODE, STRICT, STATS, MODE != kModeTrace> : k_render_m<MODE, STRICT, STATS, false>;
  int per_sm = 0;
  const size_t dyn = (size_t)((MODE == kModeTrace) ? VXB_TRACE_RING : VXB_SMEM_STACK) * kBlockThreads * sizeof(uint2);
  cudaError_t e = cudaSuccess;
  if (dyn) e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
  if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kBlockThreads, dyn);
  if (e != cudaSuccess) return -(int)e;
  if (per_sm < 1) per_sm = 1;
  if (p.blocks_per_sm_cap && per_sm > (int)p.blocks_per_sm_cap) per_sm = (int)p.blocks_per_sm_cap;
  kern<<<num_sms * per_sm, kBlockThreads, dyn, stream>>>(p);
  e = cudaGetLastError();
  return e == cudaSuccess ? 1 : -(int)e;
}

template <bool STRICT, bool STATS>
int launch_mm(const RenderA &p, int num_sms, cudaStream_t stream) {
  switch (p.mode & 3u) {
    case 0: return launch_m<0, STRICT, STATS>(p, num_sms, stream);
    case 1: return launch_m<1, STRICT, STATS>(p, num_sms, stream);
    case 2: return launch_m<2, STRICT, STATS>(p, num_sms, stream);
    default: return launch_m<3, STRICT, STATS>(p, num_sms, stream);
  }
}

}  // namespace

// ---------------------------------------------------------------- coherence ordering of ray batches
namespace {
__device__ __forceinline__ uint32_t spread10(uint32_t v) {   // 10 bits -> every third bit
  v &= 0x3FFu;
  v = (v | (v << 16)) & 0x030000FFu;
  v = (v | (v << 8)) & 0x0300F00Fu;
  v = (v | (v << 4)) & 0x030C30C3u;
  v = (v | (v << 2)) & 0x09249249u;
  return v;
}
__global__ void k_ray_keys(const TlasNodeA *__restrict__ tlas, uint32_t root, const float *__restrict__ rays, uint32_t n, uint32_t bits,
                           uint32_t *__restrict__ keys, uint32_t *__restrict__ idx, uint32_t *__restrict__ jumps) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  const float *r = reinterpret_cast<const float *>(tlas + root);   // aabbMin.xyz, leftRight, aabbMax.xyz, blas
  float lo[3] = {__ldg(r), __ldg(r + 1), __ldg(r + 2)}, hi[3] = {__ldg(r + 4), __ldg(r + 5), __ldg(r + 6)};
  uint32_t key = 0;
  if (i < n) {
    const float *q = rays + 6ull * i;
    uint32_t c[3];
    const float cells = (float)(1u << bits);
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      float ext = hi[a] - lo[a];
      float u = ext > 0.0f ? (__ldg(q + a) - lo[a]) / ext : 0.0f;
      u = fminf(fmaxf(u, 0.0f), 1.0f);          // origins outside the scene fall into the border cells (NaN -> 0)
      c[a] = min((uint32_t)(u * cells), (1u << bits) - 1u);
    }
    uint32_t morton = (spread10(c[0]) << 2) | (spread10(c[1]) << 1) | spread10(c[2]);
    uint32_t octant = (__ldg(q + 3) < 0.0f ? 4u : 0u) | (__ldg(q + 4) < 0.0f ? 2u : 0u) | (__ldg(q + 5) < 0.0f ? 1u : 0u);
    key = (octant << (3 * bits)) | morton;
    keys[i] = key, idx[i] = i;
  }
  // incoherence probe: neighbours in different coarse cells (octant + the top two Morton bits per axis)
  const uint32_t coarse = key >> (3 * bits >= 6 ? 3 * bits - 6 : 0);
  uint32_t next = __shfl_down_sync(0xffffffffu, coarse, 1);
  bool jump = (threadIdx.x & 31u) != 31u && i + 1 < n && next != coarse;
  unsigned m = __ballot_sync(0xffffffffu, jump);
  // one block in eight reports (a single counter: every warp reporting serialises 2^18 atomics per 2^23 rays)
  if ((blockIdx.x & 7u) == 0 && (threadIdx.x & 31u) == 0 && m) atomicAdd(jumps, (uint32_t)__popc(m));
}
}  // namespace

size_t ray_order_tmp_bytes(uint64_t n) {
  size_t bytes = 0;
  cub::DoubleBuffer<uint32_t> k(nullptr, nullptr), v(nullptr, nullptr);
  cub::DeviceRadixSort::SortPairs(nullptr, bytes, k, v, (int)n, 0, 32);
  return bytes;
}
cudaError_t ray_order_keys(const SceneA &scene, const float *rays, uint64_t n, uint32_t bits, const RayOrderBuffers &b, cudaStream_t stream) {
  cudaError_t e = cudaMemsetAsync(b.jumps, 0, 4, stream);
  if (e != cudaSuccess) return e;
  k_ray_keys<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(scene.tlas, scene.tlas_root, rays, (uint32_t)n, bits, b.keys[0], b.idx[0], b.jumps);
  return cudaGetLastError();
}
cudaError_t ray_order_sort(uint64_t n, uint32_t bits, const RayOrderBuffers &b, const uint32_t **sorted_idx, cudaStream_t stream) {
  cub::DoubleBuffer<uint32_t> k(b.keys[0], b.keys[1]), v(b.idx[0], b.idx[1]);
  size_t bytes = b.tmp_bytes;
  cudaError_t e = cub::DeviceRadixSort::SortPairs(b.tmp, bytes, k, v, (int)n, 0, (int)(3 * bits + 3), stream);
  *sorted_idx = v.Current();
  return e;
}

int launch_trace_machine(const RenderA &p, int num_sms, cudaStream_t stream) {
  if (p.count == 0) return 0;
  if (!p.strict_order && !p.scene.wide) return -(int)cudaErrorInvalidValue;   // fast order walks the 4-wide records
  if (p.strict_order) return p.stats ? launch_m<kModeTrace, true, true>(p, num_sms, stream) : launch_m<kModeTrace, true, false>(p, num_sms, stream);
  return p.stats ? launch_m<kModeTrace, false, true>(p, num_sms, stream) : launch_m<kModeTrace, false, false>(p, num_sms, stream);
}

int launch_render_machine(const RenderA &p, int num_sms, cudaStream_t stream) {
  if (!p.strict_order && !p.scene.wide) return -(int)cudaErrorInvalidValue;   // fast order walks the 4-wide records
  if (!p.strict_order && p.direct) {
    switch (p.mode & 3u) {
      case 0: return p.stats ? launch_d<0, true>(p, num_sms, stream) : launch_d<0, false>(p, num_sms, stream);
      case 1: return p.stats ? launch_d<1, true>(p, num_sms, stream) : launch_d<1, false>(p, num_sms, stream);
      case 2: return p.stats ? launch_d<2, true>(p, num_sms, stream) : launch_d<2, false>(p, num_sms, stream);
      default: return p.stats ? launch_d<3, true>(p, num_sms, stream) : launch_d<3, false>(p, num_sms, stream);
    }
  }
  if (p.strict_order) return p.stats ? launch_mm<true, true>(p, num_sms, stream) : launch_mm<true, false>(p, num_sms, stream);
  return p.stats ? launch_mm<false, true>(p, num_sms, stream) : launch_mm<false, false>(p, num_sms, stream);
}

}  // namespace vxb
