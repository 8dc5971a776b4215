// rt_xfer.h — device-side data movement of the multi-GPU device group (vx_group.inc).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace vxb {

constexpr int kMaxPeers = 15;

// Replication of an uploaded slice over NVLink: the GPU that received bytes [src, src+bytes) through its own PCIe link stores
// them at the same arena offset of up to kMaxPeers peer GPUs (peer-mapped pointers), reading its HBM once.  Stores are
// fire-and-forget, so one pass keeps every outgoing NVLink lane busy; `bytes` must be a multiple of 16, pointers 16-byte aligned.
struct PushArgs {
  const uint8_t *src;
  uint8_t *dst[kMaxPeers];
  int n_dst;
  uint64_t bytes;
};
cudaError_t launch_push_slice(const PushArgs &a, int num_sms, cudaStream_t stream);

}  // namespace vxb
