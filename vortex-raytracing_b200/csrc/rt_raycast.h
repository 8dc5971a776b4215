// rt_raycast.h — host-side launch interface of the layout-A (software ray-caster) CUDA path.
#pragma once
#include <cuda_runtime.h>
#include "rt_layouts.h"

namespace vxb {

// GPU-side 4-wide node derived from the uploaded BVH2 (fast order only; strict order walks the BVH2 in place): wide[n] holds,
// for the inner BVH2 node n of an instance, the boxes of its grandchildren (a child that is a leaf stays one slot) — the very
// float boxes of the uploaded bvh_node_t records, SoA, plus one descriptor per slot.  One 128-byte line per visit, half as
// many visits, pops and dependent loads per ray as the 2-wide walk; same closest hit (a superset/subset of the same exact
// box tests never changes the minimum-t triangle, only which of several equal-t triangles is met first).
struct alignas(128) GpuNode4 {
  float lox[4], loy[4], loz[4], hix[4], hiy[4], hiz[4];   // NaN boxes in unused slots
  uint32_t desc[4];     // slot cursor: kind<<27 | payload (rt_machine.cu: 0 = inner node index, 1..14 = leaf with that many
                        // triangles at payload, 15 = read node payload); 0xFFFFFFFF = unused slot
  uint32_t _pad[4];
};
static_assert(sizeof(GpuNode4) == 128, "GpuNode4");

// Device pointers of one uploaded layout-A scene (already translated from Vortex addresses).
struct SceneA {
  const uint8_t *nodes;     // BvhNodeA[]  — consumed in place (a node pair is one 64-byte, 32-byte-aligned record)
  const GpuTri *gtri;       // one per triIdx slot (built by launch_gather_tris)
  const GpuNode4 *wide;     // one per bvh_node_t slot (built by launch_build_wide), or nullptr
  const TlasNodeA *tlas;
  const BlasNodeA *blas;
  const TriExA *triEx;
  const uint8_t *tex;
  uint32_t tlas_root;
};

struct Partition {
  uint32_t rank = 0, world = 1, band_rows = 8;
};

// Frame-to-frame schedule feedback: the tiles of a frame are cut into groups of consecutive tiles; every group
// records the longest time one of its pixels (layout A: a lane of the state machine) or tiles (layout B: a warp) was occupied,
// and the next frame hands the groups out in descending order of that latency, so that the slowest rays of the image (rays
// grazing the geometry near a horizon take 50-100x the median) start first instead of forming the tail of the launch.
// order == nullptr: row-major order.
struct TileSchedule {
  const uint32_t *order;    // device: schedule position -> group
  uint32_t *cost;           // device: per group, the largest per-pixel (A: traversal steps) / per-tile (B: SM clocks) cost seen by
                            // this launch (zeroed by the host); nullptr on frames that only apply the order
  uint32_t group_shift, n_groups;   // a group = 2^group_shift consecutive tiles
  uint32_t lead;            // tiles beyond n_groups << group_shift (the last tiles of the image): handed out first, no cost kept
};

struct WorkCounters {       // device-resident, zeroed before each run
  unsigned long long rays, box_tests, tri_tests, blas_entered, shaded_hits;
  unsigned long long errors;   // bit 0: a traversal stack overflowed (the run is reported as failed)
  unsigned long long lane_steps, warp_steps, warps;   // statistics mode: SIMT efficiency = lane_steps / (32 * warp_steps)
  // statistics mode, ray state machine: what the lanes were doing while their warp fetched work / shaded / traversed
  // (vx_b200_stats_t.latency_dist, the shape of RTUnit::PerfStats::rt_latency_dist) and traversal steps per ray in log2 buckets
  unsigned long long latency[15], steps_hist[16];
};

struct RenderA {
  SceneA scene;
  KernelArgA arg;           // camera, lights, spp, depth, dimensions (addresses inside are NOT used by the kernel)
  uint32_t *out;            // framebuffer W*H
  Partition part;
  uint32_t mode;            // VX_B200_MODE_* bits
  uint32_t strict_order;    // 1 = reference far-first BLAS order, no re-cull at pop
  WorkCounters *counters;
  uint32_t *tile_counter;   // device word, zeroed before each launch
  int stats;                // gather box/tri/blas counters
  uint32_t quorum_refill;   // ray state machine: run GEN/POST when fewer lanes than this are traversing
  uint32_t chunk;           //                    pixels a warp reserves per atomicAdd on the work counter
  uint32_t blocks_per_sm_cap;   //                resident blocks per SM to launch (0 = as many as fit)
  TileSchedule sched;       // frame modes
  uint32_t quorum_inner;    //                    leave the inner-node loop when fewer lanes than this hold an inner node
  // hit-record mode of the state machine (launch_trace_machine): closest-hit records instead of pixels
  const float *rays;        // 6 floats per ray, or nullptr => primary rays of `arg`
  HitA *hits;
  uint64_t count;
  const uint32_t *perm;     // hit-record mode: work item k traces ray perm[k] (launch_ray_order), or nullptr = identity
  uint32_t n_blas;          // instances in the blas buffer (the -DVXB_SMEM_TOP=1 A/B variant stages the tree top only for one)
  uint32_t direct;          // frames, fast order: 1 = one pixel per lane over the 4-wide records (k_render_direct), 0 = the
                            // warp-scheduled state machine
  uint32_t tile_first, tile_end;   // frame modes: this launch renders tiles [tile_first, tile_end) of the rank's enumeration
                                   // (0, 0 = all of them); strips of one frame are launched back to back so that the finished
                                   // bands of one strip travel to the host while the next strip is traced
};

struct TraceA {
  SceneA scene;
  KernelArgA arg;
  const float *rays;        // 6 floats per ray, or nullptr => primary rays of `arg`
  uint64_t count;
  HitA *hits;
  uint32_t strict_order;
  WorkCounters *counters;
  int stats;
};

// gtri[p] = triangle triIdx[p] in edge form, p in [0, n_slots).  Slots whose index is >= n_tris get a degenerate triangle.
cudaError_t launch_gather_tris(const TriA *tri, const uint32_t *triIdx, GpuTri *gtri, uint64_t n_slots, uint64_t n_tris,
                               cudaStream_t stream);
// wide[g] for every inner node g in [0, n_nodes); bases = ascending bvh_offset values of the instances (child indices are
// instance-local: node g belongs to the last base <= g)
cudaError_t launch_build_wide(const uint8_t *nodes, uint64_t n_nodes, const uint32_t *bases, uint32_t n_bases, GpuNode4 *wide,
                              cudaStream_t stream);
// returns the number of kernels launched (>0) or a negative cudaError_t
int launch_render_a(const RenderA &p, int num_sms, cudaStream_t stream);
int launch_trace_a(const TraceA &p, int num_sms, cudaStream_t stream);
// the warp-scheduled ray state machine (rt_machine.cu); scenes must have < 2^27 BVH nodes (else use launch_render_a)
int launch_render_machine(const RenderA &p, int num_sms, cudaStream_t stream);
// same machine, work items = rays p.rays[0..count) (or primary rays), output = p.hits; count < 2^32
// Coherence ordering of caller-supplied rays (hit-record mode): keys = (direction octant, Morton code of the origin inside the
// scene's bounds, `bits` per axis).  ray_order_keys fills keys/idx (idx = identity) and counts in *jumps how often neighbouring
// rays fall into different coarse cells (a measure of how incoherent the batch already is); ray_order_sort sorts idx by key
// (CUB radix sort over 3*bits+3 bits; tmp sized by ray_order_tmp_bytes).  Returns cudaError_t.
struct RayOrderBuffers {
  uint32_t *keys[2];
  uint32_t *idx[2];
  void *tmp;
  size_t tmp_bytes;
  uint32_t *jumps;   // device word
};
size_t ray_order_tmp_bytes(uint64_t n);
cudaError_t ray_order_keys(const SceneA &scene, const float *rays, uint64_t n, uint32_t bits, const RayOrderBuffers &b, cudaStream_t stream);
cudaError_t ray_order_sort(uint64_t n, uint32_t bits, const RayOrderBuffers &b, const uint32_t **sorted_idx, cudaStream_t stream);
int launch_trace_machine(const RenderA &p, int num_sms, cudaStream_t stream);
constexpr uint64_t kMachineMaxNodes = 1ull << 27;

}  // namespace vxb
