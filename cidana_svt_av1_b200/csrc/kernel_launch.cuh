// Kernel instantiation + launch, split over several translation units so that they compile in
// parallel: svt_gpu_kernels_{flat,flatfull8,flatfull10,qm,qmfull10,dist,md8,md10,txs8,txs10}_{a,b}.cu each instantiate launch_part<QM, DIST, FULL, PART>
// for half of the 19 transform sizes; svt_gpu_api.cu only sees launch_bucket().
#pragma once
#include <cuda_runtime.h>
#include <stdlib.h>
#include "residual_kernel.cuh"

namespace svt {

template <int W, int H, bool QM, bool DIST, int FULL>
static cudaError_t launch_size(const KParams& p, int first, int count, cudaStream_t st) {
    using G = Geo<W, H>;
    static bool configured[64] = {};  // per device: opt in to > 48 KB of dynamic shared memory
    int dev = 0;
    cudaGetDevice(&dev);
    if (!configured[dev & 63]) {
        cudaError_t e = cudaFuncSetAttribute(residual_kernel<W, H, QM, DIST, FULL>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM_BYTES);
        if (e != cudaSuccess) return e;
        configured[dev & 63] = true;
    }
    const int grid = (count + G::NB - 1) / G::NB;
    residual_kernel<W, H, QM, DIST, FULL><<<grid, G::NT, G::SMEM_BYTES, st>>>(p, first, count);
    return cudaGetLastError();
}

// PART 0: the sizes with a 16-point or shorter longest side that are squares or 2:1; PART 1: the rest
template <bool QM, bool DIST, int FULL, int PART>
cudaError_t launch_part(int tx_size, const KParams& p, int first, int count, cudaStream_t st) {
#define CASE(id, w, h) case id: return launch_size<w, h, QM, DIST, FULL>(p, first, count, st);
    if constexpr (PART == 0) {
        switch (tx_size) {
            CASE(0, 4, 4) CASE(1, 8, 8) CASE(2, 16, 16) CASE(3, 32, 32) CASE(5, 4, 8) CASE(6, 8, 4)
            CASE(7, 8, 16) CASE(8, 16, 8) CASE(13, 4, 16) CASE(14, 16, 4)
        }
    } else {
        switch (tx_size) {
            CASE(4, 64, 64) CASE(9, 16, 32) CASE(10, 32, 16) CASE(11, 32, 64) CASE(12, 64, 32)
            CASE(15, 8, 32) CASE(16, 32, 8) CASE(17, 16, 64) CASE(18, 64, 16)
        }
    }
#undef CASE
    return cudaErrorInvalidValue;
}
constexpr bool part_of(int tx_size) {  // 1 for the sizes of PART 1
    return tx_size == 4 || tx_size == 9 || tx_size == 10 || tx_size == 11 || tx_size == 12 ||
           tx_size >= 15;
}

#define SVT_DECLARE_PART(QM, DIST, FULL, PART) \
    extern template cudaError_t launch_part<QM, DIST, FULL, PART>(int, const KParams&, int, int, cudaStream_t);
SVT_DECLARE_PART(false, false, 0, 0) SVT_DECLARE_PART(false, false, 0, 1)
SVT_DECLARE_PART(false, false, 8, 0) SVT_DECLARE_PART(false, false, 8, 1)
SVT_DECLARE_PART(false, false, 10, 0) SVT_DECLARE_PART(false, false, 10, 1)
SVT_DECLARE_PART(true, false, 0, 0) SVT_DECLARE_PART(true, false, 0, 1)
SVT_DECLARE_PART(true, false, 10, 0) SVT_DECLARE_PART(true, false, 10, 1)
SVT_DECLARE_PART(false, true, 0, 0) SVT_DECLARE_PART(false, true, 0, 1)
SVT_DECLARE_PART(false, true, 108, 0) SVT_DECLARE_PART(false, true, 108, 1)
SVT_DECLARE_PART(false, true, 110, 0) SVT_DECLARE_PART(false, true, 110, 1)
SVT_DECLARE_PART(false, true, 208, 0) SVT_DECLARE_PART(false, true, 208, 1)
SVT_DECLARE_PART(false, true, 210, 0) SVT_DECLARE_PART(false, true, 210, 1)
#define SVT_DEFINE_PART(QM, DIST, FULL, PART) \
    template cudaError_t launch_part<QM, DIST, FULL, PART>(int, const KParams&, int, int, cudaStream_t);

}  // namespace svt
