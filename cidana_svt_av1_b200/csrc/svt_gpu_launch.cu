// launch_bucket: picks the kernel variant (flat / qmatrix / distortion) and the translation unit
// that holds the size (kernel_launch.cuh).
#include "kernel_launch.cuh"
#include "svt_gpu_internal.h"

namespace svt {
template <bool QM, bool DIST, bool FULL>
static cudaError_t launch_v(int tx_size, const KParams& p, int first, int count, cudaStream_t st) {
    return part_of(tx_size) ? launch_part<QM, DIST, FULL, 1>(tx_size, p, first, count, st)
                            : launch_part<QM, DIST, FULL, 0>(tx_size, p, first, count, st);
}
cudaError_t launch_bucket(int tx_size, const KParams& p, int first, int count, cudaStream_t st) {
    // the qmatrix variant carries the 64-bit quantizer; keep it out of the common instruction stream
    if (p.dist) return launch_v<false, true, false>(tx_size, p, first, count, st);  // flat quantizer only (checked by run)
    // the headline configuration (all stages, q + dq out) has its switches compiled in (CfgFull)
    if (CfgFull::matches(p))
        return p.qm_wt ? launch_v<true, false, true>(tx_size, p, first, count, st)
                       : launch_v<false, false, true>(tx_size, p, first, count, st);
    return p.qm_wt ? launch_v<true, false, false>(tx_size, p, first, count, st)
                   : launch_v<false, false, false>(tx_size, p, first, count, st);
}
}  // namespace svt
