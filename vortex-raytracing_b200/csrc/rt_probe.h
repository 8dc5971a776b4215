// L2 -> SM gather-bandwidth probe: the second of the three roofline denominators SURVEY.md §8(d) asks for (HBM copy bandwidth
// comes from MEASURED_PEAKS.json, the FP32 peak from SM count x clock).  No reference counterpart.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace vxb {
// Random gathers of `record_bytes` (32 or 128, naturally aligned) over `buffer_bytes` of device memory, every lane its own
// record, read as 16-byte ld.global.nc like the traversal kernels read theirs (L1 only merges the loads of one record).
// Returns the best of three timed launches in *gbs (bytes gathered / second / 1e9).
cudaError_t probe_gather(uint64_t buffer_bytes, uint32_t record_bytes, int num_sms, cudaStream_t stream, double *gbs);
}  // namespace vxb
