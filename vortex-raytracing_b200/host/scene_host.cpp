// scene_host.cpp — C ABI of libvortex-b200-host.so (include/vortex_b200_host.h): scene assembly (Scene::init,
// tests/regression/raycast/scene.cpp:17-82), instance transforms (common.h:56-59), TLAS build and Tracer::setup's
// camera / viewplane maths (tracer.cpp:152-196, scene.cpp:138-172), emitting the reference's layout-A byte buffers.
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>

#include "../../include/vortex_b200_host.h"
#include "bvh2_builder.h"

#define VXH_EXPORT extern "C" __attribute__((visibility("default")))

namespace {

using namespace vxh;

struct KernelArg {   // kernel_arg_t (raycast/common.h:120-147), 192 B
  uint32_t dst_width, dst_height;
  uint64_t dst_addr, tri_addr, triEx_addr, triIdx_addr, tex_addr, bvh_addr, blas_addr, tlas_addr;
  uint32_t tlas_root;
  Vec3 camera_pos, camera_forward, camera_right, camera_up;
  float viewplane[2];
  uint32_t samples_per_pixel, max_depth;
  Vec3 light_pos, light_color, ambient_color, background_color;
  uint32_t tail;
};
static_assert(sizeof(KernelArg) == 192, "kernel_arg_t");

struct SceneData {
  std::vector<uint32_t> tri_counts, bvh_counts;
  std::vector<Tri> tri;
  std::vector<TriEx> triEx;
  std::vector<uint32_t> triIdx;
  std::vector<BvhNode> bvh;
  std::vector<BlasNode> blas;
  std::vector<TlasNode> tlas;
  std::vector<uint8_t> tex;
  uint32_t tlas_root = 0;
};

inline Vec3 sub(Vec3 a, Vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline Vec3 add(Vec3 a, Vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline Vec3 mul(Vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline float dot(Vec3 a, Vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline Vec3 cross(Vec3 a, Vec3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline Vec3 unit(Vec3 v) { return mul(v, 1.0f / sqrtf(dot(v, v))); }   // normalize (geometry.h:913-916)

void apply(BlasNode &b, const float *T) {   // blas_node_t::applyTransform
  mat4_mul(T, b.transform, b.transform);
  mat4_inverse(b.transform, b.invTransform);
}

}  // namespace

VXH_EXPORT vxh_scene_h vxh_scene_create(uint32_t n_meshes, const uint32_t *tri_counts, const float *tris, const float *triex,
                                        const uint32_t *tex_w, const uint32_t *tex_h, const uint32_t *texels, const float *reflectivity,
                                        const float *transforms, int threads) {
  if (n_meshes == 0 || !tri_counts || !tris || !triex || !tex_w || !tex_h || !texels || !reflectivity) return nullptr;
  if (threads <= 0) threads = (int)std::thread::hardware_concurrency();
  auto sc = new SceneData;
  uint64_t n = 0, texels_total = 0;
  for (uint32_t i = 0; i < n_meshes; ++i) n += tri_counts[i], texels_total += (uint64_t)tex_w[i] * tex_h[i];
  sc->tri_counts.assign(tri_counts, tri_counts + n_meshes);
  sc->bvh_counts.assign(n_meshes, 0);
  sc->tri.resize(n);
  sc->triEx.resize(n);
  sc->triIdx.resize(n);
  sc->bvh.assign(n * 2, BvhNode{{0, 0, 0}, 0, {0, 0, 0}, 0});
  sc->blas.resize(n_meshes);
  sc->tlas.assign(2 * (size_t)n_meshes - 1, TlasNode{{0, 0, 0}, 0, {0, 0, 0}, 0});
  sc->tex.resize(texels_total * 4);
  std::memcpy(sc->tri.data(), tris, n * sizeof(Tri));
  std::memcpy(sc->triEx.data(), triex, n * sizeof(TriEx));
  std::memcpy(sc->tex.data(), texels, texels_total * 4);
  std::vector<Vec3> centroids(n);
  uint32_t node_at = 0, tri_at = 0;
  uint64_t tex_at = 0;
  static const float kIdentity[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
  for (uint32_t m = 0; m < n_meshes; ++m) {
    for (uint32_t j = 0; j < tri_counts[m]; ++j) {   // identity index + centroid = (v0 + v1 + v2) / 3 (scene.cpp:48-53)
      const Tri &t = sc->tri[tri_at + j];
      sc->triIdx[tri_at + j] = tri_at + j;
      Vec3 s = add(add(t.v0, t.v1), t.v2);
      centroids[tri_at + j] = {s.x / 3, s.y / 3, s.z / 3};
    }
    BlasNode &b = sc->blas[m];
    std::memset(&b, 0, sizeof(b));
    std::memcpy(b.transform, kIdentity, 64);
    std::memcpy(b.invTransform, kIdentity, 64);
    b.bvh_offset = node_at, b.tex_offset = tex_at;
    b.tex_width = tex_w[m], b.tex_height = tex_h[m], b.reflectivity = reflectivity[m];
    sc->bvh_counts[m] = build_bvh2(sc->tri.data(), centroids.data(), tri_counts[m], sc->triIdx.data() + tri_at, sc->bvh.data() + node_at, threads);
    node_at += sc->bvh_counts[m];
    tri_at += tri_counts[m];
    tex_at += (uint64_t)tex_w[m] * tex_h[m] * 4;
  }
  if (transforms)
    for (uint32_t m = 0; m < n_meshes; ++m) apply(sc->blas[m], transforms + 16 * m);
  return sc;
}

VXH_EXPORT void vxh_scene_apply_transform(vxh_scene_h h, const float *m16) {
  if (!h || !m16) return;
  for (auto &b : ((SceneData *)h)->blas) apply(b, m16);
}

VXH_EXPORT void vxh_scene_build(vxh_scene_h h) {
  if (!h) return;
  auto sc = (SceneData *)h;
  sc->tlas_root = build_tlas(sc->bvh.data(), sc->blas.data(), sc->tri_counts.data(), (uint32_t)sc->blas.size(), sc->tlas.data());
}

VXH_EXPORT void vxh_scene_view(vxh_scene_h h, vxh_scene_view_t *v) {
  if (!h || !v) return;
  auto sc = (SceneData *)h;
  v->tri = sc->tri.data(), v->tri_bytes = sc->tri.size() * sizeof(Tri);
  v->triEx = sc->triEx.data(), v->triEx_bytes = sc->triEx.size() * sizeof(TriEx);
  v->triIdx = sc->triIdx.data(), v->triIdx_bytes = sc->triIdx.size() * 4;
  v->bvh = sc->bvh.data(), v->bvh_bytes = sc->bvh.size() * sizeof(BvhNode);
  v->blas = sc->blas.data(), v->blas_bytes = sc->blas.size() * sizeof(BlasNode);
  v->tlas = sc->tlas.data(), v->tlas_bytes = sc->tlas.size() * sizeof(TlasNode);
  v->tex = sc->tex.data(), v->tex_bytes = sc->tex.size();
  v->tlas_root = sc->tlas_root;
  uint32_t used = 0;
  for (auto c : sc->bvh_counts) used += c;
  v->bvh_nodes_used = used;
}

VXH_EXPORT void vxh_scene_free(vxh_scene_h h) { delete (SceneData *)h; }

VXH_EXPORT void vxh_setup(vxh_scene_h h, void *kernel_arg_192, uint32_t width, uint32_t height, uint32_t spp, uint32_t max_depth,
                          float vfov_deg, float zoom, int rotate_scene, const float *light_pos, const float *light_color,
                          const float *ambient, const float *background) {
  if (!h || !kernel_arg_192) return;
  auto sc = (SceneData *)h;
  KernelArg a;
  std::memset(&a, 0, sizeof(a));
  a.dst_width = width, a.dst_height = height, a.samples_per_pixel = spp, a.max_depth = max_depth;
  if (rotate_scene) {   // RotateX(-PI/4) * RotateY(PI/4) (tracer.cpp:157-158, geometry.h:1023-1039)
    const float pi = 3.14159265358979323846264f;
    float ax = -pi / 4, ay = pi / 4;
    float rx[16] = {1, 0, 0, 0, 0, cosf(ax), -sinf(ax), 0, 0, sinf(ax), cosf(ax), 0, 0, 0, 0, 1};
    float ry[16] = {cosf(ay), 0, sinf(ay), 0, 0, 1, 0, 0, -sinf(ay), 0, cosf(ay), 0, 0, 0, 0, 1};
    float T[16];
    mat4_mul(rx, ry, T);
    vxh_scene_apply_transform(h, T);
  }
  vxh_scene_build(h);
  a.tlas_root = sc->tlas_root;
  {   // Scene::computeFramingCamera (scene.cpp:138-172): bounding sphere of all TLAS nodes, dolly back along -Z
    float vfov = vfov_deg * 0.017453292519943295769236907684886f;
    Vec3 lo{1e30f, 1e30f, 1e30f}, hi{-1e30f, -1e30f, -1e30f};
    for (auto &n : sc->tlas) {
      lo = {std::min(lo.x, n.lo.x), std::min(lo.y, n.lo.y), std::min(lo.z, n.lo.z)};
      hi = {std::max(hi.x, n.hi.x), std::max(hi.y, n.hi.y), std::max(hi.z, n.hi.z)};
    }
    Vec3 center = mul(add(lo, hi), 0.5f);
    Vec3 d = sub(hi, center);
    float radius = sqrtf(dot(d, d));
    float distance = radius / tanf(vfov);
    distance *= zoom;
    Vec3 pos = sub(center, mul(Vec3{0.0f, 0.0f, 1.0f}, distance));
    Vec3 fwd = unit(sub(center, pos));
    Vec3 right = unit(cross(fwd, Vec3{0.0f, 1.0f, 0.0f}));
    Vec3 up = cross(right, fwd);
    a.camera_pos = pos, a.camera_forward = fwd, a.camera_right = right, a.camera_up = up;
  }
  float aspect = float(width) / height;               // tracer.cpp:179-183 (tan of the RAW degree value, as the reference)
  float vh = 2.0f * tanf(vfov_deg * 0.5f);
  a.viewplane[0] = vh * aspect, a.viewplane[1] = vh;
  a.light_pos = {light_pos[0], light_pos[1], light_pos[2]};
  a.light_color = {light_color[0], light_color[1], light_color[2]};
  a.ambient_color = {ambient[0], ambient[1], ambient[2]};
  a.background_color = {background[0], background[1], background[2]};
  std::memcpy(kernel_arg_192, &a, sizeof(a));
}
