// residual_kernel<W, H, QM = false, DIST = false, FULL = 10> for size group 0 (see kernel_launch.cuh)
#include "kernel_launch.cuh"
namespace svt {
SVT_DEFINE_PART(false, false, 10, 0)
}  // namespace svt
