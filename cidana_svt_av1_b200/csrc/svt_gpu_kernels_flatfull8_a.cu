// residual_kernel<W, H, QM = false, DIST = false, FULL = 8> for size group 0 (see kernel_launch.cuh)
#include "kernel_launch.cuh"
namespace svt {
SVT_DEFINE_PART(false, false, 8, 0)
}  // namespace svt
