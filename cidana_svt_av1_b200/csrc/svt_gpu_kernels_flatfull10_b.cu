// residual_kernel<W, H, QM = false, DIST = false, FULL = 10> for size group 1 (see kernel_launch.cuh)
#include "kernel_launch.cuh"
namespace svt {
SVT_DEFINE_PART(false, false, 10, 1)
}  // namespace svt
