// placeholder until the layout-B kernels land
#include "rt_rtu.h"
namespace vxb {
int launch_rtu_render(const RtuRender &, int, cudaStream_t) { return -(int)cudaErrorNotSupported; }
int launch_rtu_trace(const RtuTrace &, int, cudaStream_t) { return -(int)cudaErrorNotSupported; }
}
