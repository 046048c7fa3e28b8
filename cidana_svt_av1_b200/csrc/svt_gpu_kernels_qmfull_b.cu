// residual_kernel<W, H, QM = true, DIST = false, FULL = true> for size group 1 (see kernel_launch.cuh)
#include "kernel_launch.cuh"
namespace svt {
SVT_DEFINE_PART(true, false, true, 1)
}  // namespace svt
