// rt_rtu.h — host-side launch interface of the layout-B (simx RT-unit, quantized BVH4) CUDA path.
#pragma once
#include <cuda_runtime.h>
#include "rt_layouts.h"
#include "rt_raycast.h"

// b200-private DCR: which primary-ray generator the layout-B kernel uses.
//   0 = the hard-coded Sponza camera of tests/regression/raytracing/kernel.cpp:28-39 (what the reference runs)
//   1 = the kernel_arg_t pinhole camera (kernel.cpp:13-26, commented out there), for scenes that are not Sponza
#ifndef VX_DCR_B200_RT_CAMERA
#define VX_DCR_B200_RT_CAMERA 0x0F3
#endif

namespace vxb {

// One per quantized node (TLAS or BLAS), same index as in the uploaded buffer: the four child boxes decoded once
// (origin + ldexp(q, e), rt_traversal.cpp:61-67 — bit-identical, the decode is exact arithmetic + one rounding add),
// SoA so that a visit is seven 128-bit loads of one 128-byte line.
struct alignas(128) GpuNodeB {
  float lox[4], loy[4], loz[4], hix[4], hiy[4], hiz[4];   // child boxes; NaN for invalid children (meta == 0)
  uint32_t leftFirst;   // QNodeB.leftFirst
  uint32_t leafData;    // QNodeB.leafIdx
  uint32_t flags;       // bit 0: TLAS node (imask == 1); bit 1: leaf; bits 4-7: valid-child mask
  uint32_t _pad[5];     // [1..4] (bytes 112-127): one cursor per child, written by launch_rtu_link for the ray state machine —
                        // bit 31 TLAS space | kind << 27 | payload: kind 0 inner node (absolute index in its buffer), 1..13 BLAS
                        // leaf with that many triangles (payload = first triangle), 14 instance (blasIdx), 15 "read the node";
                        // 0xFFFFFFFF = no child
};
static_assert(sizeof(GpuNodeB) == 128, "GpuNodeB");

struct RtuScene {
  const GpuNodeB *gtlas;    // decoded TLAS nodes (DCR 0x006 buffer)
  const GpuNodeB *gbvh;     // decoded BLAS nodes (DCR 0x008 buffer)
  const GpuTri *gtri;       // edge-form triangles, same order as the tri buffer (DCR 0x009); triIdx = position
  const BlasNodeB *blas;    // DCR 0x007, consumed in place
  const TriExB *triEx;
  const MaterialB *mat;
  const uint8_t *tex;
  uint32_t n_tlas, n_bvh, n_tri, n_blas;   // element counts of the buffers above (bounds for hostile indices)
  uint32_t n_mat, _pad;
  uint64_t tex_bytes;
};

struct RtuRender {
  RtuScene scene;
  KernelArgB arg;
  uint32_t *out;
  Partition part;
  uint32_t camera;
  uint32_t strict_order;    // 1 = the reference's restart-trail / 5-entry short-stack order, visit for visit
  WorkCounters *counters;
  uint32_t *tile_counter;
  int stats;
  TileSchedule sched;       // frame-to-frame schedule feedback (rt_raycast.h)
  uint32_t mode;            // VX_B200_MODE_* bits: shadow rays / diffuse bounces (builder-defined extension, as for layout A)
};

struct RtuTrace {
  RtuScene scene;
  KernelArgB arg;
  const float *rays;        // 6 floats per ray, or nullptr => primary rays of `arg` / `camera`
  uint64_t count;
  HitA *hits;
  uint32_t camera;
  uint32_t strict_order;
  WorkCounters *counters;
  int stats;
};

// the ray state machine over layout B (rt_machine_b.cu): frames (hits == nullptr) or hit records
struct RtuMachine {
  RtuScene scene;
  KernelArgB arg;
  uint32_t *out;
  Partition part;
  uint32_t camera;
  WorkCounters *counters;
  uint32_t *tile_counter;
  int stats;
  TileSchedule sched;
  uint32_t quorum_refill, quorum_inner, chunk;   // as RenderA
  uint32_t tile_first, tile_end;                 // frames: tiles [tile_first, tile_end) of the rank's enumeration (0, 0 = all)
  const float *rays;                             // hit records: 6 floats per ray, or nullptr => primary rays
  HitA *hits;
  uint64_t count;
};
// child cursors of the decoded records: nodes = TLAS records then BLAS records (as decoded), bases = ascending bvh_offset values
cudaError_t launch_rtu_link(GpuNodeB *nodes, uint64_t n_tlas, uint64_t n_bvh, const uint32_t *bases, uint32_t n_bases, cudaStream_t stream);
bool rtu_machine_supports(const KernelArgB &a);   // nested mirror rays the slot columns can hold
int launch_rtu_machine(const RtuMachine &p, int num_sms, cudaStream_t stream);
// frames with one ray per lane over the same cursor-carrying records (no warp scheduling); p.sched.cost in SM clocks per tile
int launch_rtu_direct(const RtuMachine &p, int num_sms, cudaStream_t stream);

// scene re-flatten (one launch each): decoded nodes and edge-form triangles
cudaError_t launch_rtu_decode_nodes(const QNodeB *nodes, GpuNodeB *out, uint64_t n, cudaStream_t stream);
cudaError_t launch_rtu_gather_tris(const TriA *tri, GpuTri *gtri, uint64_t n, cudaStream_t stream);
// return the number of kernels launched (>0) or a negative cudaError_t
int launch_rtu_render(const RtuRender &p, int num_sms, cudaStream_t stream);
int launch_rtu_program(const RtuRender &p, uint32_t program, int num_sms, cudaStream_t stream);   // returns launches or -cudaError
int launch_rtu_trace(const RtuTrace &p, int num_sms, cudaStream_t stream);

}  // namespace vxb
