// residual_kernel<W, H, QM = true, DIST = false, FULL = 0> for size group 0 (see kernel_launch.cuh)
#include "kernel_launch.cuh"
namespace svt {
SVT_DEFINE_PART(true, false, 0, 0)
}  // namespace svt
