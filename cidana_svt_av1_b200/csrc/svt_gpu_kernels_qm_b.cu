// residual_kernel<W, H, QM = true, DIST = false, FULL = false> for size group 1 (see kernel_launch.cuh)
#include "kernel_launch.cuh"
namespace svt {
SVT_DEFINE_PART(true, false, false, 1)
}  // namespace svt
