// residual_kernel<W, H, QM = true, DIST = false> for size group 1 (see kernel_launch.cuh)
#include "kernel_launch.cuh"
namespace svt {
#undef SVT_DECLARE_PART
SVT_DEFINE_PART(true, false, 1)
}  // namespace svt
