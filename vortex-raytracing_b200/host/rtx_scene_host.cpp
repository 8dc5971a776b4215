// rtx_scene_host.cpp — host-side scene preparation for the RT-unit ("raytracing", layout B) path: what Scene::init, BVH
// (build + quantize) and TLAS (build + quantize) do in the reference raytracing host
// (tests/regression/raytracing/scene.cpp:17-119, bvh.cpp:12-260, bvh.cpp:262-623, tracer.cpp:195-218), emitting the SAME
// byte layouts (raytracing/common.h:20-99,164-195): triangles physically reordered into leaf order, 52-byte quantized
// BVH4 nodes with consecutive children, quantized TLAS, blas_node_t, kernel_arg_t.
//
// The reference's BVH4 builder is not reproducible (it reads uninitialised cluster bounds, bvh.cpp:79-86), so byte
// identity with it is not a meaningful goal; this builder is deterministic and many-core instead:
//   1. the task-parallel binned-SAH BVH2 of bvh2_builder.cpp (shared with layout A) is built per mesh,
//   2. collapsed to a BVH4 by repeatedly opening the child with the largest surface area (children written consecutively,
//      breadth-first per node, as rt_traversal.cpp:92 addresses them: leftFirst + k),
//   3. triangles / triEx are permuted into leaf order (the RT unit has no index indirection, rt_traversal.cpp:127-130),
//   4. nodes are quantized with the reference's formulas (bvh.cpp:209-245: e = ceil(log2(extent / 255)), mins floor, maxes
//      ceil) and then VERIFIED against the decode the RT unit performs (origin + ldexp(q, e), rt_traversal.cpp:61-67):
//      any box that float rounding left non-conservative is widened by one step, so child boxes always contain their
//      geometry (the reference does not check this).
#include <algorithm>
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>

#include "../../include/vortex_b200_host.h"
#include "bvh2_builder.h"

#define VXH_EXPORT extern "C" __attribute__((visibility("default")))

namespace {
using namespace vxh;

#pragma pack(push, 1)
struct QChild { uint8_t meta; uint8_t q[6]; };
struct QNode {                       // bvh_quantized_node_t (raytracing/common.h:58-67) == simx BVHNode — 52 B
  Vec3 origin; int8_t ex, ey, ez; uint8_t imask;
  uint32_t leftFirst, leafIdx;
  QChild children[4];
};
struct TriExB { float n[9]; float uv[6]; uint32_t texId; };   // 64 B
struct BlasB {                       // blas_node_t (raytracing/common.h:86-99) — 160 B
  uint32_t bvh_offset; float invTransform[16]; float transform[16]; uint32_t pad0;
  uint64_t mat_offset; uint32_t tex_width, tex_height; float reflectivity; uint32_t pad1;
};
struct Material { uint8_t bytes[88]; };
struct KernelArgB {                  // kernel_arg_t (raytracing/common.h:164-195) — 216 B
  uint32_t dst_width, dst_height; uint64_t dst_addr;
  uint64_t tri_addr, triEx_addr, triIdx_addr, mat_addr, tex_addr, bvh_addr, qBvh_addr, blas_addr, tlas_addr;
  uint32_t tlas_root;
  Vec3 camera_pos, camera_forward, camera_right, camera_up; float viewplane[2];
  uint32_t samples_per_pixel, max_depth;
  Vec3 light_pos, light_color, ambient_color, background_color;
  uint32_t pad; uint64_t sbt_addr;
};
#pragma pack(pop)
static_assert(sizeof(QNode) == 52 && sizeof(TriExB) == 64 && sizeof(BlasB) == 160 && sizeof(KernelArgB) == 216, "layout B records");

struct Box { Vec3 lo, hi; };
inline float area(const Box &b) {
  float ex = b.hi.x - b.lo.x, ey = b.hi.y - b.lo.y, ez = b.hi.z - b.lo.z;
  return ex * ey + ey * ez + ez * ex;
}

// a BVH4 node before quantization
struct Wide {
  Box box;
  uint32_t leftFirst = 0;   // inner: index of first child; leaf: first triangle (mesh-local slot)
  uint32_t triCount = 0;    // 0 = inner
  uint32_t childCount = 0;
};

struct RtxScene {
  std::vector<uint32_t> tri_counts, node_counts;
  std::vector<Tri> tri;
  std::vector<TriExB> triEx;
  std::vector<uint32_t> triIdx;
  std::vector<Material> mat;
  std::vector<uint8_t> tex;
  std::vector<BvhNode> bvh2;        // uploaded as the (unused by the RT unit) bvh buffer
  std::vector<QNode> qbvh;
  std::vector<Box> root_box;        // per mesh, object space
  std::vector<BlasB> blas;
  std::vector<QNode> tlas;
  uint32_t tlas_nodes = 0;
  Box world{{1e30f, 1e30f, 1e30f}, {-1e30f, -1e30f, -1e30f}};   // union of the instances' world boxes (set by build)
};

// ---- quantization (bvh.cpp:200-258), verified against the RT unit's decode ----
inline int8_t exponent_for(float extent) {
  if (!(extent > 0.0f)) return 0;    // the reference casts ceil(log2(0)) = -inf to int8_t (UB; 0 on x86-64)
  float e = std::ceil(std::log2(extent / 255.0f));
  if (e < -126.0f) e = -126.0f;
  if (e > 127.0f) e = 127.0f;
  return (int8_t)e;
}
inline bool quantize_axis(float origin, int8_t e, float lo, float hi, uint8_t &qlo, uint8_t &qhi) {
  float scale = std::exp2f((float)e);
  float a = std::floor((lo - origin) / scale), b = std::ceil((hi - origin) / scale);
  if (a < 0.0f) a = 0.0f;
  // widen until the decoded box (origin + ldexp(q, e), one rounding each) really contains [lo, hi]
  while (a > 0.0f && origin + std::ldexp(a, e) > lo) a -= 1.0f;
  while (b <= 255.0f && origin + std::ldexp(b, e) < hi) b += 1.0f;
  if (b > 255.0f || a > 255.0f) return false;
  qlo = (uint8_t)a, qhi = (uint8_t)b;
  return true;
}
void quantize_node(QNode &q, const Box &self, const Box *child, uint32_t n_children) {
  q.origin = self.lo;
  int8_t e[3] = {exponent_for(self.hi.x - self.lo.x), exponent_for(self.hi.y - self.lo.y), exponent_for(self.hi.z - self.lo.z)};
  const float o[3] = {self.lo.x, self.lo.y, self.lo.z};
  for (int axis = 0; axis < 3; ++axis) {
    for (;;) {
      bool ok = true;
      for (uint32_t k = 0; k < n_children && ok; ++k) {
        const float lo = axis == 0 ? child[k].lo.x : (axis == 1 ? child[k].lo.y : child[k].lo.z);
        const float hi = axis == 0 ? child[k].hi.x : (axis == 1 ? child[k].hi.y : child[k].hi.z);
        ok = quantize_axis(o[axis], e[axis], lo, hi, q.children[k].q[axis], q.children[k].q[3 + axis]);
      }
      if (ok || e[axis] == 127) break;
      e[axis]++;   // a ceil landed on 256: one coarser step always fits
    }
  }
  q.ex = e[0], q.ey = e[1], q.ez = e[2];
  for (uint32_t k = 0; k < 4; ++k) {
    q.children[k].meta = k < n_children ? 1 : 0;
    if (k >= n_children) std::memset(q.children[k].q, 0, 6);
  }
}

// ---- BVH2 -> BVH4 collapse for one mesh; returns nodes in the reference's addressing (children consecutive) ----
void collapse(const BvhNode *b2, std::vector<Wide> &out) {
  struct Item { uint32_t wide, src; };
  out.clear();
  out.push_back(Wide{});
  std::vector<Item> queue{{0u, 0u}};
  for (size_t qi = 0; qi < queue.size(); ++qi) {
    const Item it = queue[qi];
    const BvhNode &n = b2[it.src];
    out[it.wide].box = Box{n.lo, n.hi};
    if (n.triCount != 0) {
      out[it.wide].leftFirst = n.leftFirst, out[it.wide].triCount = n.triCount, out[it.wide].childCount = 0;
      continue;
    }
    uint32_t kids[4] = {n.leftFirst, n.leftFirst + 1, 0, 0};
    uint32_t nk = 2;
    while (nk < 4) {   // open the inner child with the largest surface area
      int best = -1;
      float best_area = -1.0f;
      for (uint32_t k = 0; k < nk; ++k) {
        const BvhNode &c = b2[kids[k]];
        if (c.triCount != 0) continue;
        float a = area(Box{c.lo, c.hi});
        if (a > best_area) best_area = a, best = (int)k;
      }
      if (best < 0) break;
      const BvhNode &c = b2[kids[best]];
      kids[best] = c.leftFirst;
      kids[nk++] = c.leftFirst + 1;
    }
    const uint32_t first = (uint32_t)out.size();
    out[it.wide].leftFirst = first, out[it.wide].triCount = 0, out[it.wide].childCount = nk;
    for (uint32_t k = 0; k < nk; ++k) {
      out.push_back(Wide{});
      queue.push_back({first + k, kids[k]});
    }
  }
}

inline Vec3 xform_pos(Vec3 p, const float *M) {   // TransformPosition (geometry.h:1280-1293, w = 1)
  return {M[0] * p.x + M[1] * p.y + M[2] * p.z + M[3] * 1.0f, M[4] * p.x + M[5] * p.y + M[6] * p.z + M[7] * 1.0f,
          M[8] * p.x + M[9] * p.y + M[10] * p.z + M[11] * 1.0f};
}

}  // namespace

// ---------------------------------------------------------------- C ABI
VXH_EXPORT vxh_rtx_scene_h vxh_rtx_scene_create(uint32_t n_meshes, const uint32_t *tri_counts, const float *tris, const uint32_t *triex16,
                                                const uint32_t *n_mats, const void *mats, const uint8_t *tex, uint64_t tex_bytes,
                                                const float *reflectivity, const float *transforms, int threads) {
  if (n_meshes == 0 || !tri_counts || !tris || !triex16 || !n_mats) return nullptr;
  if (threads <= 0) threads = (int)std::thread::hardware_concurrency();
  auto sc = new RtxScene;
  uint64_t num_tris = 0, total_mats = 0;
  for (uint32_t i = 0; i < n_meshes; ++i) num_tris += tri_counts[i], total_mats += n_mats[i];
  if (num_tris == 0 || num_tris > 0x7FFFFFFFull) { delete sc; return nullptr; }
  sc->tri_counts.assign(tri_counts, tri_counts + n_meshes);
  sc->tri.resize(num_tris), sc->triEx.resize(num_tris), sc->triIdx.resize(num_tris);
  sc->bvh2.assign(num_tris * 2, BvhNode{});
  sc->qbvh.assign(num_tris * 2, QNode{});            // scene.cpp:36-37: 2N nodes allocated
  sc->mat.resize(total_mats);
  if (total_mats && mats) std::memcpy(sc->mat.data(), mats, total_mats * sizeof(Material));
  if (tex && tex_bytes) sc->tex.assign(tex, tex + tex_bytes);
  sc->blas.assign(n_meshes, BlasB{});
  sc->root_box.resize(n_meshes), sc->node_counts.resize(n_meshes);
  const Tri *src_tri = reinterpret_cast<const Tri *>(tris);
  const TriExB *src_ex = reinterpret_cast<const TriExB *>(triex16);
  std::vector<Vec3> centroids(num_tris);
  std::vector<uint32_t> idx(num_tris);
  for (uint64_t j = 0; j < num_tris; ++j) {
    const Tri &t = src_tri[j];
    idx[j] = (uint32_t)j;
    centroids[j] = {(t.v0.x + t.v1.x + t.v2.x) / 3, (t.v0.y + t.v1.y + t.v2.y) / 3, (t.v0.z + t.v1.z + t.v2.z) / 3};   // scene.cpp:86
  }
  uint32_t node_offset = 0, tri_offset = 0, mat_offset = 0;
  std::vector<Wide> wide;
  for (uint32_t i = 0; i < n_meshes; ++i) {
    const uint32_t count = tri_counts[i];
    BvhNode *b2 = sc->bvh2.data() + node_offset;
    uint32_t used2 = build_bvh2(src_tri, centroids.data(), count, idx.data() + tri_offset, b2, threads);
    (void)used2;
    // leaf order -> physical order (bvh.cpp:120-131 swaps the triangles themselves)
    for (uint32_t p = 0; p < count; ++p) {
      uint32_t g = idx[tri_offset + p];
      sc->tri[tri_offset + p] = src_tri[g];
      sc->triEx[tri_offset + p] = src_ex[g];
      sc->triEx[tri_offset + p].texId += mat_offset;           // scene.cpp:56-58
      sc->triIdx[tri_offset + p] = tri_offset + p;             // scene.cpp:84 (never permuted by the layout-B builder)
    }
    // the BVH2 leaves hold mesh-local slots of idx[]: slot s of this mesh is now triangle tri_offset + s
    collapse(b2, wide);
    if (node_offset + wide.size() > sc->qbvh.size()) { delete sc; return nullptr; }
    QNode *q = sc->qbvh.data() + node_offset;
    for (size_t k = 0; k < wide.size(); ++k) {
      const Wide &w = wide[k];
      QNode &n = q[k];
      std::memset(&n, 0, sizeof(n));
      n.imask = 0;
      Box kids[4];
      for (uint32_t c = 0; c < w.childCount; ++c) kids[c] = wide[w.leftFirst + c].box;
      quantize_node(n, w.box, kids, w.childCount);
      if (w.triCount != 0) n.leftFirst = w.leftFirst + tri_offset, n.leafIdx = w.triCount;   // bvh.cpp:254: += tri_offset_
      else n.leftFirst = w.leftFirst, n.leafIdx = 0;
    }
    sc->root_box[i] = wide[0].box;
    sc->node_counts[i] = (uint32_t)wide.size();
    BlasB &b = sc->blas[i];
    b.bvh_offset = node_offset;
    b.mat_offset = mat_offset;
    for (int c = 0; c < 16; ++c) b.transform[c] = b.invTransform[c] = (c % 5 == 0) ? 1.0f : 0.0f;
    b.reflectivity = reflectivity ? reflectivity[i] : 0.0f;   // the reference scene forces 0 (scene.cpp:96)
    if (transforms) {
      mat4_mul(transforms + 16 * i, b.transform, b.transform);
      mat4_inverse(b.transform, b.invTransform);
    }
    node_offset += (uint32_t)wide.size();
    tri_offset += count;
    mat_offset += n_mats[i];
  }
  return sc;
}

VXH_EXPORT void vxh_rtx_scene_apply_transform(vxh_rtx_scene_h h, const float *m16) {
  if (!h || !m16) return;
  for (auto &b : ((RtxScene *)h)->blas) {
    mat4_mul(m16, b.transform, b.transform);
    mat4_inverse(b.transform, b.invTransform);
  }
}

// TLAS::build + quantize (bvh.cpp:292-340,560-620): a BVH4 over the instances' world boxes, node 0 = root
VXH_EXPORT void vxh_rtx_scene_build(vxh_rtx_scene_h h) {
  if (!h) return;
  auto sc = (RtxScene *)h;
  const uint32_t n = (uint32_t)sc->blas.size();
  sc->world = Box{{1e30f, 1e30f, 1e30f}, {-1e30f, -1e30f, -1e30f}};
  std::vector<Box> leaf(n);
  for (uint32_t i = 0; i < n; ++i) {   // world box of the 8 transformed corners (bvh.cpp:303-312)
    const Box &o = sc->root_box[i];
    Box w{{1e30f, 1e30f, 1e30f}, {-1e30f, -1e30f, -1e30f}};
    for (int c = 0; c < 8; ++c) {
      Vec3 p = xform_pos({c & 1 ? o.hi.x : o.lo.x, c & 2 ? o.hi.y : o.lo.y, c & 4 ? o.hi.z : o.lo.z}, sc->blas[i].transform);
      w.lo = {std::min(w.lo.x, p.x), std::min(w.lo.y, p.y), std::min(w.lo.z, p.z)};
      w.hi = {std::max(w.hi.x, p.x), std::max(w.hi.y, p.y), std::max(w.hi.z, p.z)};
    }
    leaf[i] = w;
    sc->world.lo = {std::min(sc->world.lo.x, w.lo.x), std::min(sc->world.lo.y, w.lo.y), std::min(sc->world.lo.z, w.lo.z)};
    sc->world.hi = {std::max(sc->world.hi.x, w.hi.x), std::max(sc->world.hi.y, w.hi.y), std::max(sc->world.hi.z, w.hi.z)};
  }
  struct Range { uint32_t lo, hi; };   // [lo, hi) over `order`
  std::vector<uint32_t> order(n);
  for (uint32_t i = 0; i < n; ++i) order[i] = i;
  auto bounds = [&](Range r) {
    Box b{{1e30f, 1e30f, 1e30f}, {-1e30f, -1e30f, -1e30f}};
    for (uint32_t i = r.lo; i < r.hi; ++i) {
      const Box &l = leaf[order[i]];
      b.lo = {std::min(b.lo.x, l.lo.x), std::min(b.lo.y, l.lo.y), std::min(b.lo.z, l.lo.z)};
      b.hi = {std::max(b.hi.x, l.hi.x), std::max(b.hi.y, l.hi.y), std::max(b.hi.z, l.hi.z)};
    }
    return b;
  };
  struct Pending { uint32_t node; Range r; };
  std::vector<QNode> nodes(1);
  std::vector<Box> boxes(1);
  std::vector<Pending> queue{{0u, {0u, n}}};
  for (size_t qi = 0; qi < queue.size(); ++qi) {
    Pending p = queue[qi];
    Box self = bounds(p.r);
    boxes[p.node] = self;
    QNode q;
    std::memset(&q, 0, sizeof(q));
    q.imask = 1;
    if (p.r.hi - p.r.lo == 1) {
      q.leafIdx = order[p.r.lo];                  // blasIdx
      quantize_node(q, self, nullptr, 0);
      nodes[p.node] = q;
      continue;
    }
    // up to four groups: split the most populous group at the median centroid of its longest axis
    Range groups[4] = {p.r};
    uint32_t ng = 1;
    while (ng < 4) {
      int best = -1;
      uint32_t best_n = 1;
      for (uint32_t g = 0; g < ng; ++g)
        if (groups[g].hi - groups[g].lo > best_n) best_n = groups[g].hi - groups[g].lo, best = (int)g;
      if (best < 0) break;
      Range r = groups[best];
      Box b = bounds(r);
      float ex = b.hi.x - b.lo.x, ey = b.hi.y - b.lo.y, ez = b.hi.z - b.lo.z;
      int axis = (ex >= ey && ex >= ez) ? 0 : (ey >= ez ? 1 : 2);
      auto key = [&](uint32_t i) {
        const Box &l = leaf[i];
        return axis == 0 ? l.lo.x + l.hi.x : (axis == 1 ? l.lo.y + l.hi.y : l.lo.z + l.hi.z);
      };
      std::stable_sort(order.begin() + r.lo, order.begin() + r.hi, [&](uint32_t a, uint32_t c) { return key(a) < key(c); });
      uint32_t mid = r.lo + (r.hi - r.lo) / 2;
      groups[best] = {r.lo, mid};
      groups[ng++] = {mid, r.hi};
    }
    const uint32_t first = (uint32_t)nodes.size();
    Box kids[4];
    for (uint32_t g = 0; g < ng; ++g) {
      kids[g] = bounds(groups[g]);
      nodes.push_back(QNode{});
      boxes.push_back(kids[g]);
      queue.push_back({first + g, groups[g]});
    }
    q.leftFirst = first, q.leafIdx = 0xFFFFFFFFu;   // inner TLAS node (bvh.cpp:575: blasIdx == UINT32_MAX)
    quantize_node(q, self, kids, ng);
    nodes[p.node] = q;
  }
  sc->tlas_nodes = (uint32_t)nodes.size();
  nodes.resize(std::max<size_t>(nodes.size(), 2 * (size_t)n - 1));   // 2M-1 allocated (bvh.cpp:268-272)
  sc->tlas = std::move(nodes);
}

VXH_EXPORT void vxh_rtx_scene_view(vxh_rtx_scene_h h, vxh_rtx_scene_view_t *v) {
  if (!h || !v) return;
  auto sc = (RtxScene *)h;
  v->tri = sc->tri.data(), v->tri_bytes = sc->tri.size() * sizeof(Tri);
  v->triEx = sc->triEx.data(), v->triEx_bytes = sc->triEx.size() * sizeof(TriExB);
  v->triIdx = sc->triIdx.data(), v->triIdx_bytes = sc->triIdx.size() * 4;
  v->mat = sc->mat.data(), v->mat_bytes = sc->mat.size() * sizeof(Material);
  v->tex = sc->tex.data(), v->tex_bytes = sc->tex.size();
  v->bvh = sc->bvh2.data(), v->bvh_bytes = sc->bvh2.size() * sizeof(BvhNode);
  v->qbvh = sc->qbvh.data(), v->qbvh_bytes = sc->qbvh.size() * sizeof(QNode);
  v->blas = sc->blas.data(), v->blas_bytes = sc->blas.size() * sizeof(BlasB);
  v->tlas = sc->tlas.data(), v->tlas_bytes = sc->tlas.size() * sizeof(QNode);
  v->tlas_root = 0;
  uint32_t used = 0;
  for (auto c : sc->node_counts) used += c;
  v->bvh_nodes_used = used, v->tlas_nodes_used = sc->tlas_nodes, v->_pad = 0;
}

VXH_EXPORT void vxh_rtx_scene_free(vxh_rtx_scene_h h) { delete (RtxScene *)h; }

// Tracer::setup of the raytracing host (tracer.cpp:195-235): TLAS build, framing camera, viewplane, lights
VXH_EXPORT void vxh_rtx_setup(vxh_rtx_scene_h h, void *kernel_arg_216, uint32_t width, uint32_t height, uint32_t spp, uint32_t max_depth,
                              float vfov_deg, float zoom, const float *light_pos, const float *light_color, const float *ambient,
                              const float *background) {
  if (!h || !kernel_arg_216) return;
  auto sc = (RtxScene *)h;
  KernelArgB a;
  std::memset(&a, 0, sizeof(a));
  a.dst_width = width, a.dst_height = height, a.samples_per_pixel = spp, a.max_depth = max_depth;
  vxh_rtx_scene_build(h);
  a.tlas_root = 0;
  {   // Scene::computeFramingCamera (raytracing/scene.cpp:163-197) over the world boxes of all instances
    Vec3 lo = sc->world.lo, hi = sc->world.hi;
    float vfov = vfov_deg * 0.017453292519943295769236907684886f;
    Vec3 center = {(lo.x + hi.x) * 0.5f, (lo.y + hi.y) * 0.5f, (lo.z + hi.z) * 0.5f};
    Vec3 d = {hi.x - center.x, hi.y - center.y, hi.z - center.z};
    float radius = sqrtf(d.x * d.x + d.y * d.y + d.z * d.z);
    float distance = radius / tanf(vfov) * zoom;
    Vec3 pos = {center.x, center.y, center.z - distance};
    Vec3 f = {center.x - pos.x, center.y - pos.y, center.z - pos.z};
    float il = 1.0f / sqrtf(f.x * f.x + f.y * f.y + f.z * f.z);
    Vec3 fwd = {f.x * il, f.y * il, f.z * il};
    Vec3 r = {fwd.y * 0.0f - fwd.z * 1.0f, fwd.z * 0.0f - fwd.x * 0.0f, fwd.x * 1.0f - fwd.y * 0.0f};   // cross(fwd, (0,1,0))
    float ir = 1.0f / sqrtf(r.x * r.x + r.y * r.y + r.z * r.z);
    Vec3 right = {r.x * ir, r.y * ir, r.z * ir};
    Vec3 up = {right.y * fwd.z - right.z * fwd.y, right.z * fwd.x - right.x * fwd.z, right.x * fwd.y - right.y * fwd.x};
    a.camera_pos = pos, a.camera_forward = fwd, a.camera_right = right, a.camera_up = up;
  }
  float aspect = float(width) / height;               // tracer.cpp:221-225 (tan of the RAW degree value, as the reference)
  float vh = 2.0f * tanf(vfov_deg * 0.5f);
  a.viewplane[0] = vh * aspect, a.viewplane[1] = vh;
  a.light_pos = {light_pos[0], light_pos[1], light_pos[2]};
  a.light_color = {light_color[0], light_color[1], light_color[2]};
  a.ambient_color = {ambient[0], ambient[1], ambient[2]};
  a.background_color = {background[0], background[1], background[2]};
  std::memcpy(kernel_arg_216, &a, sizeof(a));
}
