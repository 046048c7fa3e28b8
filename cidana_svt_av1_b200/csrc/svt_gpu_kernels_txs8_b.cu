// residual_kernel<W, H, QM = false, DIST = true, FULL = 208> for size group 1 (see kernel_launch.cuh)
#include "kernel_launch.cuh"
namespace svt {
SVT_DEFINE_PART(false, true, 208, 1)
}  // namespace svt
