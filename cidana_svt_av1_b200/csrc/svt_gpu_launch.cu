// launch_bucket: picks the kernel variant (flat / qmatrix / distortion) and the translation unit
// that holds the size (kernel_launch.cuh).
#include "kernel_launch.cuh"
#include "svt_gpu_internal.h"

namespace svt {
template <bool QM, bool DIST, int FULL>
static cudaError_t launch_v(int tx_size, const KParams& p, int first, int count, cudaStream_t st) {
    return part_of(tx_size) ? launch_part<QM, DIST, FULL, 1>(tx_size, p, first, count, st)
                            : launch_part<QM, DIST, FULL, 0>(tx_size, p, first, count, st);
}
cudaError_t launch_bucket(int tx_size, const KParams& p, int first, int count, cudaStream_t st) {
    // the qmatrix variant carries the 64-bit quantizer; keep it out of the common instruction stream
    if (p.dist) {  // flat quantizer only (checked by run); the two mode-decision shapes are compiled in
        if (CfgFull<108>::matches(p)) return launch_v<false, true, 108>(tx_size, p, first, count, st);
        if (CfgFull<110>::matches(p)) return launch_v<false, true, 110>(tx_size, p, first, count, st);
        if (CfgFull<208>::matches(p)) return launch_v<false, true, 208>(tx_size, p, first, count, st);
        if (CfgFull<210>::matches(p)) return launch_v<false, true, 210>(tx_size, p, first, count, st);
        return launch_v<false, true, 0>(tx_size, p, first, count, st);
    }
    // the headline configuration (all stages, q + dq out, one bit depth) has its switches compiled
    // in (CfgFull); the qmatrix path exists for 10-bit only
    if (CfgFull<8>::matches(p) && !p.qm_wt) return launch_v<false, false, 8>(tx_size, p, first, count, st);
    if (CfgFull<10>::matches(p))
        return p.qm_wt ? launch_v<true, false, 10>(tx_size, p, first, count, st)
                       : launch_v<false, false, 10>(tx_size, p, first, count, st);
    return p.qm_wt ? launch_v<true, false, 0>(tx_size, p, first, count, st)
                   : launch_v<false, false, 0>(tx_size, p, first, count, st);
}
}  // namespace svt
