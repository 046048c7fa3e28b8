// residual_kernel<W, H, QM = true, DIST = false, FULL = false> for size group 0 (see kernel_launch.cuh)
#include "kernel_launch.cuh"
namespace svt {
SVT_DEFINE_PART(true, false, false, 0)
}  // namespace svt
