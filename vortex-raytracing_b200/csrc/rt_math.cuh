// rt_math.cuh — FP32 helpers whose operand order and rounding follow the reference's geometry.h
// expression for expression (tests/regression/raycast/geometry.h).  This translation unit is compiled
// with -fmad=false, IEEE division and square root (nvcc defaults -prec-div=true -prec-sqrt=true), so
// every operation rounds exactly as the x86-64 SSE build of the reference does (g++ -O2, no -march:
// no FMA contraction).  Nothing here is copied: these are the arithmetic definitions restated.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace vxb {

struct f3 {
  float x, y, z;
};

__device__ __forceinline__ f3 mk3(float x, float y, float z) { return f3{x, y, z}; }
__device__ __forceinline__ f3 operator+(f3 a, f3 b) { return f3{a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ f3 operator-(f3 a, f3 b) { return f3{a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ f3 operator*(f3 a, f3 b) { return f3{a.x * b.x, a.y * b.y, a.z * b.z}; }
__device__ __forceinline__ f3 operator*(f3 a, float s) { return f3{a.x * s, a.y * s, a.z * s}; }
__device__ __forceinline__ f3 operator*(float s, f3 a) { return f3{s * a.x, s * a.y, s * a.z}; }
__device__ __forceinline__ f3 neg(f3 a) { return f3{-a.x, -a.y, -a.z}; }

// geometry.h:888 — (x*x + y*y) + z*z
__device__ __forceinline__ float dot3(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
// geometry.h:952
__device__ __forceinline__ f3 cross3(f3 a, f3 b) {
  return f3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
// geometry.h:902
__device__ __forceinline__ float length3(f3 v) { return sqrtf(dot3(v, v)); }
// geometry.h:913-916 with rsqrtf(x) := 1.0f / sqrtf(x) (geometry.h:180) — NOT the approximate intrinsic
__device__ __forceinline__ f3 normalize3(f3 v) {
  float invLen = 1.0f / sqrtf(dot3(v, v));
  return v * invLen;
}

// std::min / std::max as the reference calls them (geometry.h:1453-1462): they return the FIRST
// argument when the comparison is false, which matters only when a NaN is involved.
__device__ __forceinline__ float std_min(float a, float b) { return (b < a) ? b : a; }
__device__ __forceinline__ float std_max(float a, float b) { return (a < b) ? b : a; }

// float4(a, w) * M, row-major M (geometry.h:1280-1293): every product and the trailing "+ m3*w" are kept,
// because x + 0 is not the identity for x = -0 and 0*inf is NaN.
__device__ __forceinline__ f3 xform_vec(f3 a, const float *__restrict__ M) {  // TransformVector, w = 0
  return f3{M[0] * a.x + M[1] * a.y + M[2] * a.z + M[3] * 0.0f,
            M[4] * a.x + M[5] * a.y + M[6] * a.z + M[7] * 0.0f,
            M[8] * a.x + M[9] * a.y + M[10] * a.z + M[11] * 0.0f};
}
__device__ __forceinline__ f3 xform_pos(f3 a, const float *__restrict__ M) {  // TransformPosition, w = 1
  return f3{M[0] * a.x + M[1] * a.y + M[2] * a.z + M[3] * 1.0f,
            M[4] * a.x + M[5] * a.y + M[6] * a.z + M[7] * 1.0f,
            M[8] * a.x + M[9] * a.y + M[10] * a.z + M[11] * 1.0f};
}

// raycast/common.h:85-103
__device__ __forceinline__ uint32_t wang_hash(uint32_t s) {
  s = (s ^ 61u) ^ (s >> 16);
  s *= 9u, s = s ^ (s >> 4);
  s *= 0x27d4eb2du;
  s = s ^ (s >> 15);
  return s;
}
__device__ __forceinline__ uint32_t random_int(uint32_t &s) {
  s ^= s << 13;
  s ^= s >> 17;
  s ^= s << 5;
  return s;
}
__device__ __forceinline__ float random_float(uint32_t &s) { return (float)random_int(s) * 2.3283064365387e-10f; }

// raycast/common.h:105-118
__device__ __forceinline__ uint32_t rgb32f_to_rgb8(f3 c) {
  int r = (int)(std_min(c.x, 1.f) * 255);
  int g = (int)(std_min(c.y, 1.f) * 255);
  int b = (int)(std_min(c.z, 1.f) * 255);
  return (uint32_t)((r << 16) + (g << 8) + b);
}
__device__ __forceinline__ f3 rgb8_to_rgb32f(uint32_t c) {
  const float s = 1 / 256.0f;
  int r = (c >> 16) & 255, g = (c >> 8) & 255, b = c & 255;
  return f3{r * s, g * s, b * s};
}

// uint32_t(float) as x86-64 g++ lowers it (cvttss2si to 64 bits, low word kept): negative inputs wrap
// instead of saturating to 0 as CUDA's cvt.rzi.u32.f32 would.
__device__ __forceinline__ uint32_t f2u_x86(float f) { return (uint32_t)(long long)f; }

}  // namespace vxb
