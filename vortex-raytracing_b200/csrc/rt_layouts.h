// rt_layouts.h — byte layouts of the buffers a Vortex ray-tracing host uploads, as this backend reads them.
//
// These are declarations of an ABI (offsets and sizes), written for this backend; every struct is
// static_asserted against the sizes/offsets SURVEY.md §8(a) measured on the reference headers:
//   layout A = tests/regression/raycast/common.h:19-147  (software ray-caster, BVH2, index indirection)
//   layout B = tests/regression/raytracing/common.h:20-195 (RT-unit path, quantized BVH4)
#pragma once
#include <cstddef>
#include <cstdint>

namespace vxb {

struct vec2 { float x, y; };
struct vec3 { float x, y, z; };

// ---------------------------------------------------------------- layout A (raycast)
struct TriA {            // geometry.h:1401-1405  — 36 B, 4-byte aligned only
  vec3 v0, v1, v2;
};
struct TriExA {          // raycast/common.h:19-22 — 60 B
  vec3 N0, N1, N2;
  vec2 uv0, uv1, uv2;
};
struct HitA {            // raycast/common.h:24-29 — 24 B
  float dist;
  float bx, by, bz;
  uint32_t blasIdx;
  uint32_t triIdx;
};
struct BvhNodeA {        // raycast/common.h:32-44 — 32 B; leaf iff triCount != 0
  vec3 aabbMin;
  uint32_t leftFirst;    // inner: children at leftFirst, leftFirst+1 (BLAS-local); leaf: first slot in triIdx (GLOBAL buffer)
  vec3 aabbMax;
  uint32_t triCount;
};
struct BlasNodeA {       // raycast/common.h:47-60 — 160 B; row-major 4x4, translation in cells 3,7,11
  float transform[16];
  float invTransform[16];
  uint32_t bvh_offset;
  uint32_t _pad0;
  uint64_t tex_offset;   // BYTE offset into the texture buffer
  uint32_t tex_width;
  uint32_t tex_height;
  float reflectivity;
  uint32_t _pad1;
};
struct TlasNodeA {       // raycast/common.h:63-83 — 32 B; leaf iff leftRight == 0; children = low/high 16 bits
  vec3 aabbMin;
  uint32_t leftRight;
  vec3 aabbMax;
  uint32_t blasIdx;
};
struct KernelArgA {      // raycast/common.h:120-147 — 192 B
  uint32_t dst_width;
  uint32_t dst_height;
  uint64_t dst_addr;
  uint64_t tri_addr;
  uint64_t triEx_addr;
  uint64_t triIdx_addr;
  uint64_t tex_addr;
  uint64_t bvh_addr;
  uint64_t blas_addr;
  uint64_t tlas_addr;
  uint32_t tlas_root;
  vec3 camera_pos;
  vec3 camera_forward;
  vec3 camera_right;
  vec3 camera_up;
  vec2 viewplane;
  uint32_t samples_per_pixel;
  uint32_t max_depth;
  vec3 light_pos;
  vec3 light_color;
  vec3 ambient_color;
  vec3 background_color;
  uint32_t _tail_pad;
};

static_assert(sizeof(TriA) == 36 && sizeof(TriExA) == 60 && sizeof(HitA) == 24, "layout A triangles");
static_assert(sizeof(BvhNodeA) == 32 && offsetof(BvhNodeA, leftFirst) == 12 && offsetof(BvhNodeA, triCount) == 28, "layout A bvh node");
static_assert(sizeof(TlasNodeA) == 32 && offsetof(TlasNodeA, blasIdx) == 28, "layout A tlas node");
static_assert(sizeof(BlasNodeA) == 160 && offsetof(BlasNodeA, invTransform) == 64 && offsetof(BlasNodeA, bvh_offset) == 128 &&
              offsetof(BlasNodeA, tex_offset) == 136 && offsetof(BlasNodeA, tex_width) == 144 &&
              offsetof(BlasNodeA, reflectivity) == 152, "layout A blas node");
static_assert(sizeof(KernelArgA) == 192 && offsetof(KernelArgA, tlas_root) == 72 && offsetof(KernelArgA, camera_pos) == 76 &&
              offsetof(KernelArgA, viewplane) == 124 && offsetof(KernelArgA, samples_per_pixel) == 132 &&
              offsetof(KernelArgA, light_pos) == 140 && offsetof(KernelArgA, background_color) == 176, "layout A kernel_arg_t");

constexpr float kLargeFloat = 1e30f;   // geometry.h:15
constexpr float kEpsilon = 1e-6f;      // geometry.h:17
constexpr uint32_t kStackSize = 64;    // render.h:5

// ---------------------------------------------------------------- layout B (raytracing / simx RT unit)
struct MaterialB {       // raytracing/common.h:20-36 — 88 B; shading reads diffuse, diffuse_tex_id, tex_* only
  vec3 ambient, diffuse, specular, emissive;
  float shininess, ior, dissolve, reflectivity;
  int32_t diffuse_tex_id;   // -1 = untextured
  int32_t illum;
  uint32_t tex_width, tex_height;
  uint64_t tex_offset;      // BYTE offset into the texture buffer
};
struct TriExB {          // raytracing/common.h:39-43 — 64 B
  vec3 N0, N1, N2;
  vec2 uv0, uv1, uv2;
  uint32_t texId;           // global material index
};
struct QChildB {         // raytracing/common.h:53-56 — 7 B
  uint8_t meta;             // non-zero = valid child
  uint8_t qaabb[6];         // qlo.xyz, qhi.xyz
};
struct QNodeB {          // raytracing/common.h:58-67 == simx BVHNode (sim/simx/rt_traversal.h:19-33) — 52 B
  vec3 origin;
  int8_t ex, ey, ez;
  uint8_t imask;            // 1 = TLAS node, 0 = BLAS node
  uint32_t leftFirst;       // inner: first child node index (children consecutive); BLAS leaf: first triangle
  uint32_t leafIdx;         // BLAS node: triCount (0 => inner); TLAS node: blasIdx (UINT32_MAX => inner)
  QChildB children[4];
};
struct BlasNodeB {       // raytracing/common.h:86-99 — 160 B; simx reads bvh_offset + 12 floats of invTransform, stride 160
  uint32_t bvh_offset;
  float invTransform[16];
  float transform[16];
  uint32_t _pad0;
  uint64_t mat_offset;
  uint32_t tex_width, tex_height;
  float reflectivity;
  uint32_t _pad1;
};
struct KernelArgB {      // raytracing/common.h:164-195 — 216 B
  uint32_t dst_width, dst_height;
  uint64_t dst_addr;
  uint64_t tri_addr, triEx_addr, triIdx_addr, mat_addr, tex_addr, bvh_addr, qBvh_addr, blas_addr, tlas_addr;
  uint32_t tlas_root;
  vec3 camera_pos, camera_forward, camera_right, camera_up;
  vec2 viewplane;
  uint32_t samples_per_pixel, max_depth;
  vec3 light_pos, light_color, ambient_color, background_color;
  uint32_t _pad;
  uint64_t sbt_addr;
};
static_assert(sizeof(MaterialB) == 88 && offsetof(MaterialB, diffuse_tex_id) == 64 && offsetof(MaterialB, tex_offset) == 80, "layout B material");
static_assert(sizeof(TriExB) == 64 && offsetof(TriExB, texId) == 60, "layout B triEx");
static_assert(sizeof(QChildB) == 7 && sizeof(QNodeB) == 52 && offsetof(QNodeB, leftFirst) == 16 && offsetof(QNodeB, children) == 24, "layout B qnode");
static_assert(sizeof(BlasNodeB) == 160 && offsetof(BlasNodeB, invTransform) == 4 && offsetof(BlasNodeB, transform) == 68 &&
              offsetof(BlasNodeB, mat_offset) == 136 && offsetof(BlasNodeB, reflectivity) == 152, "layout B blas");
static_assert(sizeof(KernelArgB) == 216 && offsetof(KernelArgB, qBvh_addr) == 64 && offsetof(KernelArgB, tlas_root) == 88 &&
              offsetof(KernelArgB, camera_pos) == 92 && offsetof(KernelArgB, viewplane) == 140 && offsetof(KernelArgB, samples_per_pixel) == 148 &&
              offsetof(KernelArgB, light_pos) == 156 && offsetof(KernelArgB, sbt_addr) == 208, "layout B kernel_arg_t");

// ---------------------------------------------------------------- GPU-side triangle record
// One per slot of the triIdx buffer (leaf order): removes the index indirection and the 36-byte
// unaligned stride.  e1/e2 are the same single IEEE subtractions ray_t::intersect performs
// (geometry.h:1417-1418), so precomputing them is bit-identical.
struct alignas(16) GpuTri {
  float v0x, v0y, v0z; uint32_t triIdx;
  float e1x, e1y, e1z; uint32_t _p1;
  float e2x, e2y, e2z; uint32_t _p2;
};
static_assert(sizeof(GpuTri) == 48, "GpuTri");

}  // namespace vxb
