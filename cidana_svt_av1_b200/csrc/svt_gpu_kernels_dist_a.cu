// residual_kernel<W, H, QM = false, DIST = true> for size group 0 (see kernel_launch.cuh)
#include "kernel_launch.cuh"
namespace svt {
#undef SVT_DECLARE_PART
SVT_DEFINE_PART(false, true, 0)
}  // namespace svt
