// L2 -> SM gather-bandwidth probe (see rt_probe.h).  The traversal kernels read one 32-byte box (reference layout A), one
// 52-byte node (layout B) or one 128-byte 4-wide record per visit from wherever the walk goes next; this measures how many
// such records per second the memory system can deliver to the SMs when nothing else limits the kernel.
#include "rt_probe.h"

namespace vxb {
namespace {

__device__ __forceinline__ uint32_t mix(uint32_t x) {
  x ^= x >> 16, x *= 0x7feb352du, x ^= x >> 15, x *= 0x846ca68bu, x ^= x >> 16;
  return x;
}

// VEC = 16-byte loads per record; four independent records in flight per lane and iteration
template <int VEC>
__global__ void __launch_bounds__(256) k_gather_probe(const uint4 *__restrict__ buf, uint32_t rec_mask, uint32_t iters, uint32_t *sink) {
  uint32_t s = mix(blockIdx.x * 256u + threadIdx.x + 1u);
  uint32_t acc = 0;
  for (uint32_t it = 0; it < iters; ++it) {
    uint4 v[4][VEC];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      s = s * 1664525u + 1013904223u;
      const uint4 *rec = buf + (size_t)(mix(s) & rec_mask) * VEC;
#pragma unroll
      for (int j = 0; j < VEC; ++j) v[k][j] = __ldg(rec + j);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
      for (int j = 0; j < VEC; ++j) acc ^= v[k][j].x ^ v[k][j].y ^ v[k][j].z ^ v[k][j].w;
  }
  if (acc == 0x9e3779b9u) *sink = acc;   // never true for the zero-filled buffer: keeps the loads alive
}

}  // namespace

cudaError_t probe_gather(uint64_t buffer_bytes, uint32_t record_bytes, int num_sms, cudaStream_t stream, double *gbs) {
  if (!gbs || (record_bytes != 32 && record_bytes != 128) || buffer_bytes < (1u << 20)) return cudaErrorInvalidValue;
  uint64_t records = 1;
  while (records * 2 * record_bytes <= buffer_bytes && records < (1ull << 31)) records *= 2;   // power of two: index by mask
  void *buf = nullptr;
  uint32_t *sink = nullptr;
  cudaError_t e = cudaMalloc(&buf, records * record_bytes);
  if (e != cudaSuccess) return e;
  e = cudaMalloc(&sink, 4);
  if (e != cudaSuccess) { cudaFree(buf); return e; }
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  cudaEventCreate(&e0), cudaEventCreate(&e1);
  cudaMemsetAsync(buf, 0, records * record_bytes, stream);
  cudaMemsetAsync(sink, 0, 4, stream);
  const int blocks = num_sms * 8;
  const uint32_t iters = record_bytes == 32 ? 256 : 96;
  const double bytes = (double)blocks * 256.0 * iters * 4.0 * record_bytes;
  double best = 0.0;
  for (int rep = 0; rep < 4 && e == cudaSuccess; ++rep) {   // launch 0 warms the buffer into L2
    cudaEventRecord(e0, stream);
    if (record_bytes == 32) k_gather_probe<2><<<blocks, 256, 0, stream>>>((const uint4 *)buf, (uint32_t)(records - 1), iters, sink);
    else k_gather_probe<8><<<blocks, 256, 0, stream>>>((const uint4 *)buf, (uint32_t)(records - 1), iters, sink);
    cudaEventRecord(e1, stream);
    e = cudaEventSynchronize(e1);
    if (e == cudaSuccess) e = cudaGetLastError();
    float ms = 0.f;
    if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, e0, e1);
    if (e == cudaSuccess && rep > 0 && ms > 0.f && bytes / (ms * 1e6) > best) best = bytes / (ms * 1e6);
  }
  cudaEventDestroy(e0), cudaEventDestroy(e1);
  cudaFree(sink), cudaFree(buf);
  *gbs = best;
  return e;
}

}  // namespace vxb
