// residual_kernel<W, H, QM = false, DIST = true, FULL = 110> for size group 1 (see kernel_launch.cuh)
#include "kernel_launch.cuh"
namespace svt {
SVT_DEFINE_PART(false, true, 110, 1)
}  // namespace svt
