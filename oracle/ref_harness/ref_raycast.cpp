// oracle/ref_harness/ref_raycast.cpp — TEST INFRASTRUCTURE ONLY.
//
// A thin C-callable harness around the REFERENCE's own raycast sources, compiled where they lie
// under /root/reference (see oracle/Makefile, target libref_raycast.so).  No reference code is
// copied: this file #includes tests/regression/raycast/{render.h,scene.h,bvh.h,mesh.h} and links
// bvh.cpp / scene.cpp / mesh.cpp / surface.cpp unmodified.
//
// What it gives the tests and bench.py (cpu_baseline.kind == "reference"):
//   * scenes built by the reference's BVH (bvh.cpp:23-66) and TLAS (bvh.cpp:181-212) builders,
//     either from the shipped OBJ/PNG assets (Mesh + Scene::init, scene.cpp:17-82) or from
//     caller-supplied triangle arrays (procedural scenes; same glue as Scene::init);
//   * Tracer::setup's camera/viewplane maths (tracer.cpp:152-196) on those scenes;
//   * the hot loop itself: GenerateRay / TLASIntersect / Trace (render.h:143-276) called per ray,
//     single-threaded as written or row-parallel over std::thread (no arithmetic change);
//   * work counters are NOT taken here (that needs source edits); see oracle/port for those.
//
// Extensions that do not exist in the reference (shadow rays, diffuse path tracing)
// are driven from here with the reference's TLASIntersect as the only intersection primitive,
// mirroring oracle/port/rc_port.c expression for expression.

#include <vector>
#include <thread>
#include <atomic>
#include <string>
#include <cstring>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <limits>
#include <algorithm>

#define private public   // harness needs Scene's buffers to inject procedural meshes
#include "scene.h"
#undef private
#define __UNIFORM__
#include "render.h"

extern "C" {

struct ref_scene_view_t {
  const void *tri;    uint64_t tri_bytes;
  const void *triEx;  uint64_t triEx_bytes;
  const void *triIdx; uint64_t triIdx_bytes;
  const void *bvh;    uint64_t bvh_bytes;     // full allocation (2*N nodes, as the reference uploads)
  const void *blas;   uint64_t blas_bytes;
  const void *tlas;   uint64_t tlas_bytes;
  const void *tex;    uint64_t tex_bytes;
  uint32_t tlas_root;
  uint32_t bvh_nodes_used;
};

struct RefScene {
  Scene *scene = nullptr;
  std::vector<uint32_t> tri_counts;   // per mesh
};

static_assert(sizeof(kernel_arg_t) == 192, "layout A kernel_arg_t");
static_assert(sizeof(bvh_node_t) == 32 && sizeof(tlas_node_t) == 32 && sizeof(blas_node_t) == 160, "layout A nodes");
static_assert(sizeof(tri_t) == 36 && sizeof(tri_ex_t) == 60 && sizeof(ray_hit_t) == 24, "layout A tris");

// ---- scene from the shipped assets: exactly Tracer::init (tracer.cpp:89-100) ----
void *ref_scene_load_assets(const char *model_file, uint32_t mesh_count) {
  static const struct { const char *texture; float reflectivity; } materials[] =
      {{"ceramic.png", 0}, {"red.png", 0.5}, {"flower.png", 0.3}};   // tracer.cpp:13
  if (mesh_count == 0 || mesh_count > 3) return nullptr;
  std::vector<Mesh *> meshes(mesh_count);
  std::string base = std::string(ASSETS_PATHS) + "/assets/";
  for (uint32_t i = 0; i < mesh_count; ++i)
    meshes[i] = new Mesh((base + model_file).c_str(), (base + materials[i].texture).c_str(), materials[i].reflectivity);
  auto rs = new RefScene;
  rs->scene = new Scene(meshes);
  if (rs->scene->init() != 0) return nullptr;
  for (auto m : meshes) rs->tri_counts.push_back(m->tri().size());
  return rs;
}

// ---- scene from caller triangles: the glue of Scene::init (scene.cpp:17-82) around the
//      reference BVH/TLAS classes.  Textures are 0x00RRGGBB uint32 texels (surface.cpp:47-50).
void *ref_scene_create(uint32_t n_meshes, const uint32_t *tri_counts,
                       const float *tris /*9 per tri*/, const float *triex /*15 per tri*/,
                       const uint32_t *tex_w, const uint32_t *tex_h, const uint32_t *texels,
                       const float *reflectivity, const float *transforms /*16 per mesh, row-major*/) {
  auto rs = new RefScene;
  auto sc = new Scene(std::vector<Mesh *>());
  rs->scene = sc;
  uint64_t num_tris = 0, tex_px = 0;
  for (uint32_t i = 0; i < n_meshes; ++i) { num_tris += tri_counts[i]; tex_px += uint64_t(tex_w[i]) * tex_h[i]; }
  sc->tex_buf_.resize(tex_px * 4);
  sc->tri_buf_.resize(num_tris);
  sc->triEx_buf_.resize(num_tris);
  sc->triIdx_buf_.resize(num_tris);
  sc->centroids_.resize(num_tris);
  sc->bvh_nodes_.resize(num_tris * 2);
  sc->blas_nodes_.resize(n_meshes);
  sc->bvh_list_.assign(n_meshes, nullptr);
  memcpy(sc->tri_buf_.data(), tris, num_tris * sizeof(tri_t));
  memcpy(sc->triEx_buf_.data(), triex, num_tris * sizeof(tri_ex_t));
  memcpy(sc->tex_buf_.data(), texels, tex_px * 4);
  uint32_t bvh_offset = 0, tri_offset = 0; uint64_t tex_offset = 0;
  for (uint32_t i = 0; i < n_meshes; ++i) {
    for (uint32_t j = 0; j < tri_counts[i]; ++j) {
      auto &tri = sc->tri_buf_[tri_offset + j];
      sc->triIdx_buf_[tri_offset + j] = tri_offset + j;
      sc->centroids_[tri_offset + j] = (tri.v0 + tri.v1 + tri.v2) / 3;
    }
    auto &b = sc->blas_nodes_[i];
    b.bvh_offset = bvh_offset;
    b.tex_offset = tex_offset;
    b.transform = mat4_t::Identity();
    b.invTransform = mat4_t::Identity();
    b.tex_width = tex_w[i];
    b.tex_height = tex_h[i];
    b.reflectivity = reflectivity[i];
    auto bvh = new BVH(sc->tri_buf_.data(), sc->centroids_.data(), tri_counts[i],
                       sc->bvh_nodes_.data() + bvh_offset, sc->triIdx_buf_.data() + tri_offset);
    bvh_offset += bvh->nodeCount();
    tri_offset += tri_counts[i];
    tex_offset += uint64_t(tex_w[i]) * tex_h[i] * 4;
    sc->bvh_list_[i] = bvh;
    rs->tri_counts.push_back(tri_counts[i]);
  }
  sc->tlas_ = new TLAS(sc->bvh_list_, sc->blas_nodes_.data());
  if (transforms) {
    for (uint32_t i = 0; i < n_meshes; ++i) {
      mat4_t T;
      memcpy(T.cell, transforms + 16 * i, 64);
      sc->blas_nodes_[i].applyTransform(T);      // common.h:56-59
    }
  }
  return rs;
}

void ref_scene_apply_transform(void *h, const float *m16) {
  mat4_t T; memcpy(T.cell, m16, 64);
  ((RefScene *)h)->scene->applyTransform(T);
}

void ref_scene_build(void *h) { ((RefScene *)h)->scene->build(); }

void ref_scene_view(void *h, ref_scene_view_t *v) {
  auto sc = ((RefScene *)h)->scene;
  v->tri = sc->tri_buf_.data();       v->tri_bytes = sc->tri_buf_.size() * sizeof(tri_t);
  v->triEx = sc->triEx_buf_.data();   v->triEx_bytes = sc->triEx_buf_.size() * sizeof(tri_ex_t);
  v->triIdx = sc->triIdx_buf_.data(); v->triIdx_bytes = sc->triIdx_buf_.size() * 4;
  v->bvh = sc->bvh_nodes_.data();     v->bvh_bytes = sc->bvh_nodes_.size() * sizeof(bvh_node_t);
  v->blas = sc->blas_nodes_.data();   v->blas_bytes = sc->blas_nodes_.size() * sizeof(blas_node_t);
  v->tlas = sc->tlas_->nodes().data(); v->tlas_bytes = sc->tlas_->nodes().size() * sizeof(tlas_node_t);
  v->tex = sc->tex_buf_.data();       v->tex_bytes = sc->tex_buf_.size();
  v->tlas_root = sc->tlas_root();
  uint32_t used = 0;
  for (auto b : sc->bvh_list_) used += b->nodeCount();
  v->bvh_nodes_used = used;
}

void ref_scene_free(void *h) {
  auto rs = (RefScene *)h;
  delete rs->scene;
  delete rs;
}

// ---- Tracer::setup (tracer.cpp:152-196): scene rotation, TLAS build, framing camera, viewplane, lights.
//      Address fields are filled with HOST pointers, as the reference's use_cpu_ path does (tracer.cpp:102-109).
void ref_setup(void *h, void *arg_out, uint32_t width, uint32_t height, uint32_t spp, uint32_t max_depth,
               float camera_vfov, float zoom, int rotate_scene,
               const float *light_pos, const float *light_color, const float *ambient, const float *background) {
  auto sc = ((RefScene *)h)->scene;
  kernel_arg_t a = {};
  a.dst_width = width; a.dst_height = height;
  a.samples_per_pixel = spp; a.max_depth = max_depth;
  if (rotate_scene) {
    auto T = mat4_t::RotateX(-PI / 4) * mat4_t::RotateY(PI / 4);
    sc->applyTransform(T);
  }
  sc->build();
  a.tlas_root = sc->tlas_root();
  {
    float3_t camera_pos, camera_target, camera_up;
    sc->computeFramingCamera(camera_vfov * DEG2RAD, zoom, &camera_pos, &camera_target, &camera_up);
    float3_t forward = normalize(camera_target - camera_pos);
    float3_t right = normalize(cross(forward, camera_up));
    float3_t up = cross(right, forward);
    a.camera_pos = camera_pos; a.camera_forward = forward; a.camera_right = right; a.camera_up = up;
  }
  {
    float aspect_ratio = float(width) / height;
    float viewport_height = 2.0f * tan(camera_vfov * 0.5f);   // raw degrees, as tracer.cpp:181 does
    float viewport_width = viewport_height * aspect_ratio;
    a.viewplane = {viewport_width, viewport_height};
  }
  a.light_pos = float3_t(light_pos[0], light_pos[1], light_pos[2]);
  a.light_color = float3_t(light_color[0], light_color[1], light_color[2]);
  a.ambient_color = float3_t(ambient[0], ambient[1], ambient[2]);
  a.background_color = float3_t(background[0], background[1], background[2]);
  a.tri_addr = (uint64_t)sc->tri_buf_.data();
  a.triEx_addr = (uint64_t)sc->triEx_buf_.data();
  a.triIdx_addr = (uint64_t)sc->triIdx_buf_.data();
  a.bvh_addr = (uint64_t)sc->bvh_nodes_.data();
  a.tlas_addr = (uint64_t)sc->tlas_->nodes().data();
  a.blas_addr = (uint64_t)sc->blas_nodes_.data();
  a.tex_addr = (uint64_t)sc->tex_buf_.data();
  memcpy(arg_out, &a, sizeof(a));
}

// ---- extended Trace (SURVEY §8d extensions; no counterpart in the reference).  mode bit0 = shadow
//      rays, bit1 = diffuse bounces.  mode == 0 never reaches here: the reference's Trace is called.
#define RT_MODE_SHADOW  1u
#define RT_MODE_DIFFUSE 2u

static inline bool occluded(const float3_t &O, const float3_t &D, float maxd, const kernel_arg_t *arg) {
  ray_hit_t sh;
  sh.dist = maxd;
  ray_t r{O, D};
  TLASIntersect(r, arg->tlas_root, (const tlas_node_t *)arg->tlas_addr, (const blas_node_t *)arg->blas_addr,
                (const bvh_node_t *)arg->bvh_addr, (const uint32_t *)arg->triIdx_addr, (const tri_t *)arg->tri_addr, &sh);
  return sh.dist < maxd;
}

static float3_t TraceExt(const ray_t &ray, const kernel_arg_t *arg, uint32_t mode, uint32_t *seed, uint64_t *nrays) {
  auto tri_ptr = (const tri_t *)arg->tri_addr;
  auto bvh_ptr = (const bvh_node_t *)arg->bvh_addr;
  auto idx_ptr = (const uint32_t *)arg->triIdx_addr;
  auto tlas_ptr = (const tlas_node_t *)arg->tlas_addr;
  auto blas_ptr = (const blas_node_t *)arg->blas_addr;
  auto triEx_ptr = (const tri_ex_t *)arg->triEx_addr;
  auto tex_ptr = (const uint8_t *)arg->tex_addr;
  ray_t cur_ray = ray;
  float3_t radiance = {0, 0, 0};
  float3_t thr = {1, 1, 1};
  for (uint32_t bounce = 0; bounce < arg->max_depth; ++bounce) {
    ray_hit_t hit;
    TLASIntersect(cur_ray, arg->tlas_root, tlas_ptr, blas_ptr, bvh_ptr, idx_ptr, tri_ptr, &hit);
    ++*nrays;
    if (hit.dist == LARGE_FLOAT) {
      radiance += arg->background_color * thr;
      break;
    }
    auto &blas = blas_ptr[hit.blasIdx];
    const tri_ex_t &triEx = triEx_ptr[hit.triIdx];
    float3_t I = cur_ray.orig + cur_ray.dir * hit.dist;
    float3_t N = triEx.N1 * hit.bcoords.x + triEx.N2 * hit.bcoords.y + triEx.N0 * hit.bcoords.z;
    mat4_t invTranspose = blas.invTransform.transposed();
    N = normalize(TransformVector(N, invTranspose));
    float2_t uv = triEx.uv1 * hit.bcoords.x + triEx.uv2 * hit.bcoords.y + triEx.uv0 * hit.bcoords.z;
    auto tex_pixels = (const uint32_t *)(tex_ptr + blas.tex_offset);
    float3_t texColor = texSample(uv, tex_pixels, blas.tex_width, blas.tex_height);
    // shading normal on the side the ray arrived from (offsets + bounce hemisphere)
    float3_t Ns = (dot(N, cur_ray.dir) > 0.0f) ? -N : N;
    float3_t O = I + Ns * 0.001f;
    // diffuseLighting (render.h:58-70) with a visibility factor folded into NdotL
    float3_t L = arg->light_pos - I;
    float dist = length(L);
    L *= 1.0f / dist;
    float att = 1.0f / (1.0f + dist * 0.1f);
    float NdotL = std::max(0.0f, dot(N, L));
    if ((mode & RT_MODE_SHADOW) && NdotL > 0.0f) {
      float3_t Lv = arg->light_pos - O;
      float d2 = length(Lv);
      float3_t D = Lv * (1.0f / d2);
      ++*nrays;
      if (occluded(O, D, d2, arg)) NdotL = 0.0f;
    }
    float3_t diffuse = texColor * (arg->ambient_color + att * arg->light_color * NdotL);
    if (mode & RT_MODE_DIFFUSE) {
      radiance += thr * diffuse;
      thr = thr * texColor;
      if (bounce + 1 < arg->max_depth) {
        float px, py, pz, l2;
        do {
          px = RandomFloat(seed) * 2.0f - 1.0f;
          py = RandomFloat(seed) * 2.0f - 1.0f;
          pz = RandomFloat(seed) * 2.0f - 1.0f;
          l2 = px * px + py * py + pz * pz;
        } while (l2 > 1.0f || l2 < 1e-6f);
        float il = 1.0f / sqrtf(l2);
        float3_t D = Ns + float3_t(px * il, py * il, pz * il);
        float dl2 = dot(D, D);
        if (dl2 < 1e-12f) D = Ns; else D = D * (1.0f / sqrtf(dl2));
        cur_ray.orig = O;
        cur_ray.dir = D;
        continue;
      }
      break;
    }
    // mirror path of the reference (render.h:253-272), scalar reflectivity
    auto reflectivity = blas.reflectivity;
    radiance += thr * diffuse * (1 - reflectivity);
    thr = thr * reflectivity;
    if (reflectivity > 0.0f && bounce + 1 < arg->max_depth) {
      float3_t R = normalize(cur_ray.dir - 2.0f * N * dot(N, cur_ray.dir));
      cur_ray.orig = I + R * 0.001f;
      cur_ray.dir = R;
      continue;
    }
    radiance += thr * arg->background_color;
    break;
  }
  return radiance;
}

}  // extern "C"
template <class F> static void parallel_rows(uint32_t rows, int threads, F fn) {
  if (threads <= 1) { for (uint32_t y = 0; y < rows; ++y) fn(y); return; }
  std::atomic<uint32_t> next{0};
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t)
    pool.emplace_back([&] { for (;;) { uint32_t y = next.fetch_add(1); if (y >= rows) break; fn(y); } });
  for (auto &th : pool) th.join();
}

extern "C" {
// ---- Tracer::render (tracer.cpp:249-263): the timed CPU baseline.  rows [y0,y1). Returns rays traced.
uint64_t ref_render(const void *argp, uint32_t *out, uint32_t mode, uint32_t y0, uint32_t y1, int threads) {
  auto arg = (const kernel_arg_t *)argp;
  std::atomic<uint64_t> total{0};
  parallel_rows(y1 - y0, threads, [&](uint32_t r) {
    uint32_t y = y0 + r;
    uint64_t nrays = 0;
    for (uint32_t x = 0; x < arg->dst_width; ++x) {
      uint32_t out_idx = y * arg->dst_width + x;
      float3_t color = float3_t(0, 0, 0);
      for (uint32_t s = 0; s < arg->samples_per_pixel; ++s) {
        auto ray = GenerateRay(x, y, arg);
        if (mode == 0) {
          color += Trace(ray, arg);
        } else {
          uint32_t seed = WangHash(out_idx * arg->samples_per_pixel + s + 1);
          if (seed == 0) seed = 1;
          color += TraceExt(ray, arg, mode, &seed, &nrays);
        }
      }
      out[out_idx] = RGB32FtoRGB8(color);
    }
    total += nrays;
  });
  return total.load();
}

// ---- per-ray dumps: the closest-hit record of each ray, as TLASIntersect leaves it (render.h:226-227).
//      hits: 24 B ray_hit_t per ray {dist, bx, by, bz, blasIdx, triIdx}.
void ref_trace_rays(const void *argp, const float *rays /*6 per ray: o, d*/, uint64_t n, void *hits, int threads) {
  auto arg = (const kernel_arg_t *)argp;
  auto out = (ray_hit_t *)hits;
  const uint64_t chunk = 4096;
  uint32_t nchunks = (uint32_t)((n + chunk - 1) / chunk);
  parallel_rows(nchunks, threads, [&](uint32_t c) {
    uint64_t e = std::min<uint64_t>(n, (c + 1) * chunk);
    for (uint64_t i = c * chunk; i < e; ++i) {
      ray_t r{float3_t(rays[6 * i], rays[6 * i + 1], rays[6 * i + 2]), float3_t(rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5])};
      ray_hit_t hit;
      TLASIntersect(r, arg->tlas_root, (const tlas_node_t *)arg->tlas_addr, (const blas_node_t *)arg->blas_addr,
                    (const bvh_node_t *)arg->bvh_addr, (const uint32_t *)arg->triIdx_addr, (const tri_t *)arg->tri_addr, &hit);
      out[i] = hit;
    }
  });
}

void ref_trace_primary(const void *argp, void *hits, float *rays_out /*optional, 6 per pixel*/, int threads) {
  auto arg = (const kernel_arg_t *)argp;
  auto out = (ray_hit_t *)hits;
  parallel_rows(arg->dst_height, threads, [&](uint32_t y) {
    for (uint32_t x = 0; x < arg->dst_width; ++x) {
      uint64_t i = uint64_t(y) * arg->dst_width + x;
      auto ray = GenerateRay(x, y, arg);
      if (rays_out) {
        rays_out[6 * i + 0] = ray.orig.x; rays_out[6 * i + 1] = ray.orig.y; rays_out[6 * i + 2] = ray.orig.z;
        rays_out[6 * i + 3] = ray.dir.x;  rays_out[6 * i + 4] = ray.dir.y;  rays_out[6 * i + 5] = ray.dir.z;
      }
      ray_hit_t hit;
      TLASIntersect(ray, arg->tlas_root, (const tlas_node_t *)arg->tlas_addr, (const blas_node_t *)arg->blas_addr,
                    (const bvh_node_t *)arg->bvh_addr, (const uint32_t *)arg->triIdx_addr, (const tri_t *)arg->tri_addr, &hit);
      out[i] = hit;
    }
  });
}

int ref_hw_threads() { return (int)std::thread::hardware_concurrency(); }

}  // extern "C"
