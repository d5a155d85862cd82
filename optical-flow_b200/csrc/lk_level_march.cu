// Strip-marching fused level kernel (throughput path).  Placeholder: not instantiated for
// any radius yet, so the generic tile kernel handles every level.
#include "of_device.cuh"
#include "of_kernels.h"

namespace ofb {

cudaError_t launch_lk_level_march(const LevelArgs&, int, cudaStream_t, unsigned long long*, bool* handled)
{
    *handled = false;
    return cudaSuccess;
}

}  // namespace ofb
