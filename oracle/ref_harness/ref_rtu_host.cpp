// oracle/ref_harness/ref_rtu_host.cpp — TEST INFRASTRUCTURE ONLY.
//
// Host side of the RT-unit ("raytracing", layout B) oracle: the REFERENCE's own Scene / BVH (BVH4 + quantizer) /
// TLAS classes from tests/regression/raytracing/{scene,bvh,treelet}.cpp, compiled where they lie, driven from C.
// No reference code is copied: this file #includes scene.h / rtx_shading.h and calls into them.
//
//   rtuh_scene_create/build/view : buffers exactly as raytracing/tracer.cpp:156-258 would upload them, from caller
//                                  triangles (the glue is Scene::init, scene.cpp:17-119, minus the Mesh/OBJ loader).
//                                  NOTE the reference BVH4 builder reads uninitialised cluster bounds
//                                  (bvh.cpp:79-86 -> findBestSplitPlane :140), so its output may vary run to run;
//                                  every comparison therefore uses ONE set of buffers for all implementations.
//   rtuh_camera                  : Tracer::setup camera/viewplane maths (tracer.cpp:197-218).
//   rtuh_render                  : the per-pixel program of kernel.cpp:28-39,41-133 + shaders/{closest,miss}.cpp, run on
//                                  a flat memory image.  Those files are RISC-V only (vx_raytrace.h inline .insn), so
//                                  their few lines of control flow are restated here with getAttr/commit replaced by
//                                  struct reads; ALL arithmetic is the reference's own (geometry.h operators,
//                                  rtx_shading.h texSample/diffuseLighting, common.h RGB32FtoRGB8) and traversal is
//                                  the verbatim sim/simx/rt_traversal.cpp (ref_rtu.cpp).
#include <vector>
#include <thread>
#include <atomic>
#include <cstring>
#include <cstdio>
#include <cmath>
#include <iostream>
#include <sstream>

#define private public
#include "scene.h"
#undef private
#include "rtx_shading.h"

extern "C" {
struct rtu_hit_t { float dist, bx, by, bz; uint32_t blasIdx, triIdx; };
void rtu_trace_rays(const uint8_t *mem, uint64_t base, uint64_t size, uint32_t tlas, uint32_t blas, uint32_t qbvh, uint32_t tri,
                    const float *rays, uint64_t n, rtu_hit_t *hits, uint32_t threads, uint64_t *work);
}

static_assert(sizeof(kernel_arg_t) == 216 && sizeof(bvh_quantized_node_t) == 52 && sizeof(blas_node_t) == 160, "layout B");
static_assert(sizeof(tri_ex_t) == 64 && sizeof(material_info_t) == 88 && sizeof(tri_t) == 36, "layout B");

namespace {
struct Quiet {   // the reference builders print progress to std::cout; keep test logs clean
  std::streambuf *old;
  std::ostringstream sink;
  Quiet() : old(std::cout.rdbuf(sink.rdbuf())) {}
  ~Quiet() { std::cout.rdbuf(old); }
};
}

extern "C" {

struct rtuh_view_t {
  const void *tri;    uint64_t tri_bytes;
  const void *triEx;  uint64_t triEx_bytes;
  const void *triIdx; uint64_t triIdx_bytes;
  const void *mat;    uint64_t mat_bytes;
  const void *tex;    uint64_t tex_bytes;
  const void *bvh;    uint64_t bvh_bytes;
  const void *qbvh;   uint64_t qbvh_bytes;
  const void *blas;   uint64_t blas_bytes;
  const void *tlas;   uint64_t tlas_bytes;     // QUANTIZED TLAS nodes (tracer.cpp:172,243)
  uint32_t tlas_root, bvh_nodes_used, tlas_nodes_used, _pad;
};

// triex: 16 words per triangle (15 floats + texId LOCAL to its mesh); mats: n_mats[i] material_info_t per mesh with
// tex_offset/tex_width/tex_height already final (byte offsets into `tex`).
void *rtuh_scene_create(uint32_t n_meshes, const uint32_t *tri_counts, const float *tris, const uint32_t *triex,
                        const uint32_t *n_mats, const void *mats, const uint8_t *tex, uint64_t tex_bytes,
                        const float *transforms /*16 per mesh or null*/) {
  Quiet q;
  auto sc = new Scene(std::vector<Mesh *>());
  uint64_t num_tris = 0, total_mats = 0;
  for (uint32_t i = 0; i < n_meshes; ++i) num_tris += tri_counts[i], total_mats += n_mats[i];
  sc->tex_buf_.assign(tex, tex + tex_bytes);
  sc->mat_buf_.resize(total_mats);
  if (total_mats) memcpy(sc->mat_buf_.data(), mats, total_mats * sizeof(material_info_t));
  sc->tri_buf_.resize(num_tris);
  sc->triEx_buf_.resize(num_tris);
  sc->triIdx_buf_.resize(num_tris);
  sc->centroids_.resize(num_tris);
  sc->bvh_nodes_.resize(num_tris * 2);
  sc->bvh_quantized_nodes_.resize(num_tris * 2);
  memset((void *)sc->bvh_quantized_nodes_.data(), 0, num_tris * 2 * sizeof(bvh_quantized_node_t));
  sc->blas_nodes_.resize(n_meshes);
  sc->bvh_list_.assign(n_meshes, nullptr);
  memcpy((void *)sc->tri_buf_.data(), tris, num_tris * sizeof(tri_t));
  memcpy((void *)sc->triEx_buf_.data(), triex, num_tris * sizeof(tri_ex_t));
  uint32_t bvh_offset = 0, tri_offset = 0, mat_offset = 0;
  for (uint32_t i = 0; i < n_meshes; ++i) {
    for (uint32_t j = 0; j < tri_counts[i]; ++j) {
      auto &tri = sc->tri_buf_[tri_offset + j];
      sc->triEx_buf_[tri_offset + j].texId += mat_offset;            // scene.cpp:56-58
      sc->triIdx_buf_[tri_offset + j] = tri_offset + j;              // scene.cpp:83-87
      sc->centroids_[tri_offset + j] = (tri.v0 + tri.v1 + tri.v2) / 3;
    }
    auto &b = sc->blas_nodes_[i];                                    // scene.cpp:90-97
    b.bvh_offset = bvh_offset;
    b.mat_offset = mat_offset;
    b.transform = mat4_t::Identity();
    b.invTransform = mat4_t::Identity();
    b.reflectivity = 0.0;
    auto bvh = new BVH(sc->tri_buf_.data() + tri_offset, sc->centroids_.data() + tri_offset, tri_counts[i],
                       sc->bvh_nodes_.data() + bvh_offset, sc->bvh_quantized_nodes_.data() + bvh_offset,
                       sc->triIdx_buf_.data() + tri_offset, sc->triEx_buf_.data() + tri_offset, tri_offset);
    bvh_offset += bvh->nodeCount();
    tri_offset += tri_counts[i];
    mat_offset += n_mats[i];
    sc->bvh_list_[i] = bvh;
  }
  sc->tlas_ = new TLAS(sc->bvh_list_, sc->blas_nodes_.data());
  if (transforms) {
    for (uint32_t i = 0; i < n_meshes; ++i) {
      mat4_t T;
      memcpy(T.cell, transforms + 16 * i, 64);
      sc->blas_nodes_[i].applyTransform(T);
    }
  }
  return sc;
}

void rtuh_scene_set_reflectivity(void *h, uint32_t blas, float r) { ((Scene *)h)->blas_nodes_[blas].reflectivity = r; }

void rtuh_scene_build(void *h) {
  Quiet q;
  ((Scene *)h)->build();
}

void rtuh_scene_view(void *h, rtuh_view_t *v) {
  auto sc = (Scene *)h;
  uint32_t used = 0;
  for (auto b : sc->bvh_list_) used += b->nodeCount();
  v->tri = sc->tri_buf_.data(), v->tri_bytes = sc->tri_buf_.size() * sizeof(tri_t);
  v->triEx = sc->triEx_buf_.data(), v->triEx_bytes = sc->triEx_buf_.size() * sizeof(tri_ex_t);
  v->triIdx = sc->triIdx_buf_.data(), v->triIdx_bytes = sc->triIdx_buf_.size() * 4;
  v->mat = sc->mat_buf_.data(), v->mat_bytes = sc->mat_buf_.size() * sizeof(material_info_t);
  v->tex = sc->tex_buf_.data(), v->tex_bytes = sc->tex_buf_.size();
  v->bvh = sc->bvh_nodes_.data(), v->bvh_bytes = sc->bvh_nodes_.size() * sizeof(bvh_node_t);
  v->qbvh = sc->bvh_quantized_nodes_.data(), v->qbvh_bytes = sc->bvh_quantized_nodes_.size() * sizeof(bvh_quantized_node_t);
  v->blas = sc->blas_nodes_.data(), v->blas_bytes = sc->blas_nodes_.size() * sizeof(blas_node_t);
  v->tlas = sc->tlas_->qnodes().data(), v->tlas_bytes = sc->tlas_->qnodes().size() * sizeof(bvh_quantized_node_t);
  v->tlas_root = sc->tlas_root();
  v->bvh_nodes_used = used;
  v->tlas_nodes_used = sc->tlas_->nodeIndex_;
  v->_pad = 0;
}

void rtuh_scene_destroy(void *h) {
  auto sc = (Scene *)h;
  sc->meshes_.clear();
  delete sc;
}

// Tracer::setup (tracer.cpp:197-218) camera + viewplane + lights into a kernel_arg_t
void rtuh_camera(void *h, uint32_t W, uint32_t H, float camera_vfov, float zoom, void *karg) {
  auto sc = (Scene *)h;
  auto a = (kernel_arg_t *)karg;
  float3_t camera_pos, camera_target, camera_up;
  sc->computeFramingCamera(camera_vfov * DEG2RAD, zoom, &camera_pos, &camera_target, &camera_up);
  float3_t forward = normalize(camera_target - camera_pos);
  float3_t right = normalize(cross(forward, camera_up));
  float3_t up = cross(right, forward);
  a->camera_pos = camera_pos, a->camera_forward = forward, a->camera_right = right, a->camera_up = up;
  float aspect_ratio = float(W) / H;
  float viewport_height = 2.0f * tan(camera_vfov * 0.5f);
  float viewport_width = viewport_height * aspect_ratio;
  a->viewplane = {viewport_width, viewport_height};
  a->dst_width = W, a->dst_height = H;
  a->tlas_root = sc->tlas_root();
}

}  // extern "C"

namespace {

// kernel.cpp:28-39 — the hard-coded camera the shipped RTU kernel uses (camera 0)
ray_t GenerateRayFixed(uint32_t x, uint32_t y, const kernel_arg_t *arg) {
  auto pos = float3_t(0.0, 100.0, 0.0);
  auto front = float3_t(1.0, 0.0, 0.0);
  float FOV = 1.0;
  float u = (x * 2.0 - arg->dst_width) / arg->dst_height;
  float v = (y * 2.0 - arg->dst_height) / arg->dst_height;
  auto right = cross(front, float3_t(0.0, 1.0, 0.0));
  auto up = cross(right, front);
  auto dir = normalize(u * right + v * up + FOV * front);
  return ray_t{pos, dir};
}

// kernel.cpp:13-26 — the kernel_arg_t pinhole camera (commented out in the reference kernel; camera 1)
ray_t GenerateRayPinhole(uint32_t x, uint32_t y, const kernel_arg_t *arg) {
  float x_ndc = (x + 0.5f) / arg->dst_width - 0.5;
  float y_ndc = (y + 0.5f) / arg->dst_height - 0.5;
  float x_vp = x_ndc * arg->viewplane.x;
  float y_vp = y_ndc * arg->viewplane.y;
  auto pt_cam = x_vp * arg->camera_right + y_vp * arg->camera_up + arg->camera_forward;
  auto pt_w = pt_cam + arg->camera_pos;
  auto camera_dir = normalize(pt_w - arg->camera_pos);
  return ray_t{arg->camera_pos, camera_dir};
}

struct Image {
  const uint8_t *mem;
  uint64_t base, size;
  uint32_t dcr[4];
  template <class T> const T *at(uint64_t addr) const { return reinterpret_cast<const T *>(mem + (addr - base)); }
};

// shaders/closest.cpp:11-130 and shaders/miss.cpp:7-14; `bounce` is ray_payload_t::bounce
float3_t Shade(const Image &im, const kernel_arg_t *arg, const ray_t &ray, uint32_t bounce) {
  rtu_hit_t h;
  float r6[6] = {ray.orig.x, ray.orig.y, ray.orig.z, ray.dir.x, ray.dir.y, ray.dir.z};
  rtu_trace_rays(im.mem, im.base, im.size, im.dcr[0], im.dcr[1], im.dcr[2], im.dcr[3], r6, 1, &h, 1, nullptr);
  if (h.dist == LARGE_FLOAT) return arg->background_color;            // miss.cpp:12
  auto blas_ptr = im.at<blas_node_t>(arg->blas_addr);
  auto triEx_ptr = im.at<tri_ex_t>(arg->triEx_addr);
  auto mat_ptr = im.at<material_info_t>(arg->mat_addr);
  auto tex_ptr = im.at<uint8_t>(arg->tex_addr);
  float3_t radiance = {0, 0, 0};
  float throughput = 1.0f;
  ray_hit_t hit;
  hit.dist = h.dist, hit.bcoords = float3_t(h.bx, h.by, h.bz), hit.blasIdx = h.blasIdx, hit.triIdx = h.triIdx;
  auto &blas = blas_ptr[hit.blasIdx];
  const tri_ex_t &triEx = triEx_ptr[hit.triIdx];
  const material_info_t &mat = mat_ptr[triEx.texId];
  float3_t I = ray.orig + ray.dir * hit.dist;
  float3_t N = triEx.N1 * hit.bcoords.x + triEx.N2 * hit.bcoords.y + triEx.N0 * hit.bcoords.z;
  mat4_t invTranspose = blas.invTransform.transposed();
  N = normalize(TransformVector(N, invTranspose));
  float2_t uv = triEx.uv1 * hit.bcoords.x + triEx.uv2 * hit.bcoords.y + triEx.uv0 * hit.bcoords.z;
  float3_t texColor;
  if (mat.diffuse_tex_id >= 0) {
    auto tex_pixels = reinterpret_cast<const uint32_t *>(tex_ptr + mat.tex_offset);
    texColor = texSample(uv, tex_pixels, mat.tex_width, mat.tex_height);
  } else {
    texColor = mat.diffuse;
  }
  float3_t diffuse = diffuseLighting(I, N, texColor, arg->ambient_color, arg->light_color, arg->light_pos);
  auto reflectivity = blas.reflectivity;
  radiance += throughput * diffuse * (1 - reflectivity);
  throughput *= reflectivity;
  if (reflectivity > 0.0f && bounce + 1 < arg->max_depth) {
    float3_t R = normalize(ray.dir - 2.0f * N * dot(N, ray.dir));
    ray_t sec_ray;
    sec_ray.orig = I + R * 0.001f;
    sec_ray.dir = R;
    float3_t sec = Shade(im, arg, sec_ray, bounce + 1);                // nested traceRay + getWork loop (closest.cpp:104-117)
    radiance += sec * throughput;
  } else {
    radiance += arg->background_color * throughput;
  }
  return radiance;
}

}  // namespace

extern "C" {

// Renders arg->dst_width x dst_height into out (row-major 0x00RRGGBB) from a flat memory image.
// camera: 0 = kernel.cpp hard-coded camera, 1 = kernel_arg_t pinhole.  dcr[4] = RTU DCRs 0x006..0x009.
void rtuh_render(const uint8_t *mem, uint64_t base, uint64_t size, const void *karg, const uint32_t *dcr, uint32_t camera,
                 uint32_t *out, uint32_t threads) {
  auto arg = (const kernel_arg_t *)karg;
  Image im{mem, base, size, {dcr[0], dcr[1], dcr[2], dcr[3]}};
  std::atomic<uint32_t> next{0};
  auto worker = [&]() {
    for (;;) {
      uint32_t y = next.fetch_add(1);
      if (y >= arg->dst_height) break;
      for (uint32_t x = 0; x < arg->dst_width; ++x) {
        ray_t ray = camera ? GenerateRayPinhole(x, y, arg) : GenerateRayFixed(x, y, arg);
        // kernel.cpp:66-81: every sample re-traces the same ray into the same payload; the last one is what is written
        float3_t color = Shade(im, arg, ray, 0);
        out[x + y * arg->dst_width] = RGB32FtoRGB8(color);
      }
    }
  };
  if (threads <= 1) worker();
  else {
    std::vector<std::thread> pool;
    for (uint32_t t = 0; t < threads; ++t) pool.emplace_back(worker);
    for (auto &t : pool) t.join();
  }
}

// primary rays of the chosen camera, 6 floats each (so hit records can be compared ray by ray)
void rtuh_primary_rays(const void *karg, uint32_t camera, float *rays) {
  auto arg = (const kernel_arg_t *)karg;
  for (uint32_t y = 0; y < arg->dst_height; ++y)
    for (uint32_t x = 0; x < arg->dst_width; ++x) {
      ray_t r = camera ? GenerateRayPinhole(x, y, arg) : GenerateRayFixed(x, y, arg);
      float *q = rays + 6 * (uint64_t(y) * arg->dst_width + x);
      q[0] = r.orig.x, q[1] = r.orig.y, q[2] = r.orig.z, q[3] = r.dir.x, q[4] = r.dir.y, q[5] = r.dir.z;
    }
}

}  // extern "C"

// The OBJ/MTL/PNG loader (mesh.cpp, surface.cpp) is not part of this harness: scenes arrive as triangle arrays.
// Scene::~Scene deletes its (empty) mesh list, which needs this symbol to exist.
Mesh::~Mesh() {}
