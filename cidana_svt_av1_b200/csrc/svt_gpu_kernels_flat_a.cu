// residual_kernel<W, H, QM = false, DIST = false> for size group 0 (see kernel_launch.cuh)
#include "kernel_launch.cuh"
namespace svt {
#undef SVT_DECLARE_PART
SVT_DEFINE_PART(false, false, 0)
}  // namespace svt
