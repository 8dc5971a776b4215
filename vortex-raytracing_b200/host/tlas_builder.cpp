// tlas_builder.cpp — SAH TLAS over instances (reference: tests/regression/raycast/bvh.cpp:181-346) and the 4x4 matrix
// helpers the instance records need (geometry.h:1149-1231).  Instance counts are small (the TLAS child fields are 16-bit,
// common.h:72-82), so this is plain sequential code; what matters is byte-identical output.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <numeric>

#include "bvh2_builder.h"

namespace vxh {
namespace {

constexpr int kBins = 8;
constexpr float kLarge = 1e30f;
inline float lo_of(float a, float b) { return a < b ? a : b; }
inline float hi_of(float a, float b) { return a > b ? a : b; }
inline float ax(const Vec3 &v, int a) { return a == 0 ? v.x : (a == 1 ? v.y : v.z); }
inline Vec3 vlo(const Vec3 &a, const Vec3 &b) { return {lo_of(a.x, b.x), lo_of(a.y, b.y), lo_of(a.z, b.z)}; }
inline Vec3 vhi(const Vec3 &a, const Vec3 &b) { return {hi_of(a.x, b.x), hi_of(a.y, b.y), hi_of(a.z, b.z)}; }
inline float area2(const Vec3 &lo, const Vec3 &hi) {   // surfaceArea (geometry.h:971-974)
  float ex = hi.x - lo.x, ey = hi.y - lo.y, ez = hi.z - lo.z;
  return 2 * (ex * ey + ey * ez + ez * ex);
}

struct TlasBuild {
  TlasNode *nodes;
  const uint32_t *tri_counts;
  std::vector<uint32_t> order;   // nodeIndices_
  uint32_t next_inner;

  float mid(uint32_t slot, int a) const { return (ax(nodes[order[slot]].lo, a) + ax(nodes[order[slot]].hi, a)) / 2; }

  // TLAS::partition (bvh.cpp:308-346)
  uint32_t split(int start, int end, int a, float pos) {
    int l = start, r = end;
    while (l <= r) {
      while (l <= end && mid((uint32_t)l, a) < pos) ++l;
      while (r >= start && !(mid((uint32_t)r, a) < pos)) --r;
      if (l < r) std::swap(order[l++], order[r--]);
    }
    if (r < start) return (uint32_t)end;
    if (l > end) return (uint32_t)start;
    return (uint32_t)r;
  }

  // TLAS::buildRecursive (bvh.cpp:214-306)
  uint32_t build(uint32_t start, uint32_t end) {
    if (start == end) return order[start];
    Vec3 lo = nodes[order[start]].lo, hi = nodes[order[start]].hi;
    for (uint32_t i = start + 1; i <= end; ++i) lo = vlo(lo, nodes[order[i]].lo), hi = vhi(hi, nodes[order[i]].hi);
    int best_axis = 0;
    float best_pos = 0.0f, best_cost = std::numeric_limits<float>::infinity();
    Vec3 ext{hi.x - lo.x, hi.y - lo.y, hi.z - lo.z};
    for (int a = 0; a < 3; ++a) {
      if (ax(ext, a) <= 0) continue;
      float width = ax(ext, a) / kBins;
      for (uint32_t i = 1; i < (uint32_t)kBins; ++i) {
        float pos = ax(lo, a) + i * width;
        uint32_t ln = 0, rn = 0;
        Vec3 llo{kLarge, kLarge, kLarge}, lhi{-kLarge, -kLarge, -kLarge}, rlo = llo, rhi = lhi;
        for (uint32_t j = start; j <= end; ++j) {
          const TlasNode &n = nodes[order[j]];
          if (mid(j, a) < pos) llo = vlo(llo, n.lo), lhi = vhi(lhi, n.hi), ln += tri_counts[n.blasIdx];
          else rlo = vlo(rlo, n.lo), rhi = vhi(rhi, n.hi), rn += tri_counts[n.blasIdx];
        }
        if (ln == 0 || rn == 0) continue;
        float la = area2(llo, lhi), ra = area2(rlo, rhi);
        float cost = la * ln + ra * rn;
        if (cost < best_cost) best_cost = cost, best_axis = a, best_pos = pos;
      }
    }
    if (best_cost == std::numeric_limits<float>::infinity()) {   // median fallback (bvh.cpp:273-284)
      best_axis = (ext.x > ext.y) ? ((ext.x > ext.z) ? 0 : 2) : ((ext.y > ext.z) ? 1 : 2);
      std::vector<float> mids;
      for (uint32_t i = start; i <= end; ++i) mids.push_back((ax(nodes[order[i]].lo, best_axis) + ax(nodes[order[i]].hi, best_axis)) * 0.5f);
      std::sort(mids.begin(), mids.end());
      best_pos = mids[mids.size() / 2];
    }
    uint32_t m = split((int)start, (int)end, best_axis, best_pos);
    if (m == start || m == end) m = (start + end) / 2;
    uint32_t l = build(start, m);
    uint32_t r = build(m + 1, end);
    uint32_t me = next_inner++;
    nodes[me].leftRight = (r << 16) | l;
    nodes[me].lo = lo, nodes[me].hi = hi;
    return me;
  }
};

// cofactor table of the 4x4 inverse: entry e = sum over six signed products m[i]*m[j]*m[k], in this order
// (geometry.h:1151-1182).  Encoded as sign | i<<8 | j<<4 | k.
#define T_(s, i, j, k) (uint16_t)(((s) < 0 ? 0x8000u : 0u) | ((i) << 8) | ((j) << 4) | (k))
constexpr uint16_t kInv[16][6] = {
    {T_(+1, 5, 10, 15), T_(-1, 5, 11, 14), T_(-1, 9, 6, 15), T_(+1, 9, 7, 14), T_(+1, 13, 6, 11), T_(-1, 13, 7, 10)},
    {T_(-1, 1, 10, 15), T_(+1, 1, 11, 14), T_(+1, 9, 2, 15), T_(-1, 9, 3, 14), T_(-1, 13, 2, 11), T_(+1, 13, 3, 10)},
    {T_(+1, 1, 6, 15), T_(-1, 1, 7, 14), T_(-1, 5, 2, 15), T_(+1, 5, 3, 14), T_(+1, 13, 2, 7), T_(-1, 13, 3, 6)},
    {T_(-1, 1, 6, 11), T_(+1, 1, 7, 10), T_(+1, 5, 2, 11), T_(-1, 5, 3, 10), T_(-1, 9, 2, 7), T_(+1, 9, 3, 6)},
    {T_(-1, 4, 10, 15), T_(+1, 4, 11, 14), T_(+1, 8, 6, 15), T_(-1, 8, 7, 14), T_(-1, 12, 6, 11), T_(+1, 12, 7, 10)},
    {T_(+1, 0, 10, 15), T_(-1, 0, 11, 14), T_(-1, 8, 2, 15), T_(+1, 8, 3, 14), T_(+1, 12, 2, 11), T_(-1, 12, 3, 10)},
    {T_(-1, 0, 6, 15), T_(+1, 0, 7, 14), T_(+1, 4, 2, 15), T_(-1, 4, 3, 14), T_(-1, 12, 2, 7), T_(+1, 12, 3, 6)},
    {T_(+1, 0, 6, 11), T_(-1, 0, 7, 10), T_(-1, 4, 2, 11), T_(+1, 4, 3, 10), T_(+1, 8, 2, 7), T_(-1, 8, 3, 6)},
    {T_(+1, 4, 9, 15), T_(-1, 4, 11, 13), T_(-1, 8, 5, 15), T_(+1, 8, 7, 13), T_(+1, 12, 5, 11), T_(-1, 12, 7, 9)},
    {T_(-1, 0, 9, 15), T_(+1, 0, 11, 13), T_(+1, 8, 1, 15), T_(-1, 8, 3, 13), T_(-1, 12, 1, 11), T_(+1, 12, 3, 9)},
    {T_(+1, 0, 5, 15), T_(-1, 0, 7, 13), T_(-1, 4, 1, 15), T_(+1, 4, 3, 13), T_(+1, 12, 1, 7), T_(-1, 12, 3, 5)},
    {T_(-1, 0, 5, 11), T_(+1, 0, 7, 9), T_(+1, 4, 1, 11), T_(-1, 4, 3, 9), T_(-1, 8, 1, 7), T_(+1, 8, 3, 5)},
    {T_(-1, 4, 9, 14), T_(+1, 4, 10, 13), T_(+1, 8, 5, 14), T_(-1, 8, 6, 13), T_(-1, 12, 5, 10), T_(+1, 12, 6, 9)},
    {T_(+1, 0, 9, 14), T_(-1, 0, 10, 13), T_(-1, 8, 1, 14), T_(+1, 8, 2, 13), T_(+1, 12, 1, 10), T_(-1, 12, 2, 9)},
    {T_(-1, 0, 5, 14), T_(+1, 0, 6, 13), T_(+1, 4, 1, 14), T_(-1, 4, 2, 13), T_(-1, 12, 1, 6), T_(+1, 12, 2, 5)},
    {T_(+1, 0, 5, 10), T_(-1, 0, 6, 9), T_(-1, 4, 1, 10), T_(+1, 4, 2, 9), T_(+1, 8, 1, 6), T_(-1, 8, 2, 5)},
};
#undef T_

}  // namespace

uint32_t build_tlas(const BvhNode *bvh, const BlasNode *blas, const uint32_t *tri_counts, uint32_t n, TlasNode *tlas) {
  if (n == 0) return 0;
  TlasBuild tb{tlas, tri_counts, std::vector<uint32_t>(n), n};
  for (uint32_t i = 0; i < n; ++i) {   // leaves: world box = box of the 8 transformed corners of the instance root (bvh.cpp:186-207)
    const BvhNode &root = bvh[blas[i].bvh_offset];
    const float *M = blas[i].transform;
    Vec3 lo{kLarge, kLarge, kLarge}, hi{-kLarge, -kLarge, -kLarge};
    for (int c = 0; c < 8; ++c) {
      float x = (c & 1) ? root.hi.x : root.lo.x, y = (c & 2) ? root.hi.y : root.lo.y, z = (c & 4) ? root.hi.z : root.lo.z;
      Vec3 p{M[0] * x + M[1] * y + M[2] * z + M[3] * 1.0f, M[4] * x + M[5] * y + M[6] * z + M[7] * 1.0f,
             M[8] * x + M[9] * y + M[10] * z + M[11] * 1.0f};   // TransformPosition (geometry.h:1287-1297)
      lo = vlo(lo, p), hi = vhi(hi, p);
    }
    tlas[i].lo = lo, tlas[i].hi = hi, tlas[i].blasIdx = i, tlas[i].leftRight = 0;
    tb.order[i] = i;
  }
  return tb.build(0, n - 1);
}

void mat4_mul(const float *a, const float *b, float *out) {
  float r[16];
  for (int i = 0; i < 16; i += 4)
    for (int j = 0; j < 4; ++j)
      r[i + j] = (a[i + 0] * b[j + 0]) + (a[i + 1] * b[j + 4]) + (a[i + 2] * b[j + 8]) + (a[i + 3] * b[j + 12]);
  std::memcpy(out, r, sizeof(r));
}

void mat4_inverse(const float *m, float *out) {
  float inv[16];
  for (int e = 0; e < 16; ++e) {
    float acc = 0.0f;
    for (int t = 0; t < 6; ++t) {
      uint16_t q = kInv[e][t];
      float a = m[(q >> 8) & 15];
      float term = ((q & 0x8000u) ? -a : a) * m[(q >> 4) & 15] * m[q & 15];
      acc = t ? acc + term : term;
    }
    inv[e] = acc;
  }
  const float det = m[0] * inv[0] + m[1] * inv[4] + m[2] * inv[8] + m[3] * inv[12];
  float r[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};   // singular -> identity (geometry.h:1184-1190)
  if (det != 0) {
    const float invdet = 1.0f / det;
    for (int i = 0; i < 16; ++i) r[i] = inv[i] * invdet;
  }
  std::memcpy(out, r, sizeof(r));
}

}  // namespace vxh
