// rt_rtu.h — host-side launch interface of the layout-B (simx RT-unit) CUDA path.
#pragma once
#include <cuda_runtime.h>
#include "rt_layouts.h"
#include "rt_raycast.h"

// b200-private DCR: which primary-ray generator the layout-B kernel uses.
//   0 = the hard-coded Sponza camera of tests/regression/raytracing/kernel.cpp:28-39 (what the reference runs)
//   1 = the kernel_arg_t pinhole camera (render.h-style GenerateRay), for scenes that are not Sponza
#ifndef VX_DCR_B200_RT_CAMERA
#define VX_DCR_B200_RT_CAMERA 0x0F3
#endif

namespace vxb {

struct RtuScene {
  const QNodeB *tlas;       // quantized TLAS nodes (DCR 0x006)
  const BlasNodeB *blas;    // DCR 0x007
  const QNodeB *qbvh;       // quantized BLAS nodes (DCR 0x008)
  const TriA *tri;          // reordered triangles (DCR 0x009)
  const TriExB *triEx;
  const MaterialB *mat;
  const uint8_t *tex;
  uint32_t tlas_root;
};

struct RtuRender {
  RtuScene scene;
  KernelArgB arg;
  uint32_t *out;
  Partition part;
  uint32_t camera;
  WorkCounters *counters;
  uint32_t *tile_counter;
  int stats;
};

struct RtuTrace {
  RtuScene scene;
  KernelArgB arg;
  const float *rays;
  uint64_t count;
  HitA *hits;
  uint32_t camera;
  WorkCounters *counters;
  int stats;
};

int launch_rtu_render(const RtuRender &p, int num_sms, cudaStream_t stream);
int launch_rtu_trace(const RtuTrace &p, int num_sms, cudaStream_t stream);
inline void rtu_release(void *) {}

}  // namespace vxb
