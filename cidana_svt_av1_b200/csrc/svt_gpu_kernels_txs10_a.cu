// residual_kernel<W, H, QM = false, DIST = true, FULL = 210> for size group 0 (see kernel_launch.cuh)
#include "kernel_launch.cuh"
namespace svt {
SVT_DEFINE_PART(false, true, 210, 0)
}  // namespace svt
