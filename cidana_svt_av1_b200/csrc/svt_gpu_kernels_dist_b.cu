// residual_kernel<W, H, QM = false, DIST = true> for size group 1 (see kernel_launch.cuh)
#include "kernel_launch.cuh"
namespace svt {
#undef SVT_DECLARE_PART
SVT_DEFINE_PART(false, true, 1)
}  // namespace svt
