// rt_xfer.cu — the push kernel behind DeviceGroup's scene replication (see rt_xfer.h).
#include "rt_xfer.h"

namespace vxb {
namespace {

__global__ void __launch_bounds__(256) k_push_slice(const __grid_constant__ PushArgs a) {
  const uint64_t n16 = a.bytes / 16;
  const uint4 *src = reinterpret_cast<const uint4 *>(a.src);
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride * 4) {
    uint4 v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k)
      if (i + k * stride < n16) v[k] = __ldcs(src + i + k * stride);   // streamed: this copy of the data is read once here
    for (int d = 0; d < a.n_dst; ++d) {
      uint4 *dst = reinterpret_cast<uint4 *>(a.dst[d]);
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (i + k * stride < n16) dst[i + k * stride] = v[k];
    }
  }
}

}  // namespace

cudaError_t launch_push_slice(const PushArgs &a, int num_sms, cudaStream_t stream) {
  if (a.bytes == 0 || a.n_dst == 0) return cudaSuccess;
  const uint64_t per_block = 256ull * 16 * 4;
  uint64_t blocks = (a.bytes + per_block - 1) / per_block;
  const uint64_t cap = (uint64_t)num_sms * 4;
  if (blocks > cap) blocks = cap;
  k_push_slice<<<(unsigned)blocks, 256, 0, stream>>>(a);
  return cudaGetLastError();
}

}  // namespace vxb
