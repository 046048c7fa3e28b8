// residual_kernel<W, H, QM = false, DIST = true, FULL = 108> for size group 0 (see kernel_launch.cuh)
#include "kernel_launch.cuh"
namespace svt {
SVT_DEFINE_PART(false, true, 108, 0)
}  // namespace svt
