// svtgpu_quantize_b_noalloc: the stand-alone quantizer (quant_kernel, residual_kernel.cuh; the
// arithmetic of aom_highbd_quantize_b*, Source/Lib/Common/ASM_AVX2/highbd_quantize_intrin_avx2.c:80-125)
// for callers that must not allocate: code under CUDA graph capture, tight serving loops, harnesses
// that forbid cudaMalloc.  It has no counterpart in the
// reference.
//
// The level-A quantizer slots stage through per-thread device buffers they cudaMalloc on first
// use.  This entry point instead takes DEVICE pointers for its inputs and outputs, passes the ten
// table values by value, spreads the coefficients over several CTAs and keeps the only working
// memory it needs -- the cross-CTA eob reduction -- in a file-scope __device__ array of fixed size.
// It calls no cudaMalloc / cudaFree / allocating library routine: one cudaMemsetAsync of the
// ticket word and one kernel launch on the caller's stream.
//
// Scratch rules:
//   * the array is part of the library image: every GPU that loads libsvtgpu.so holds its 256 bytes
//     until the process exits;
//   * in host code the array's NAME is a host-side stand-in, not a device address; the device
//     address comes from cudaGetSymbolAddress, once per device, with that device current;
//   * every call on a device shares the array, and it keeps the last call's contents: the ticket
//     word is cleared by each call before its launch (and again by the kernel's last CTA), the
//     per-CTA partials are overwritten before they are read;
//   * calls that OVERLAP on one GPU (two streams, two host threads) race on it.  Serialise them.
#include <cuda_runtime.h>
#include <stdio.h>
#include <mutex>

#include "../../include/svt_gpu.h"
#include "residual_kernel.cuh"
#include "svt_gpu_internal.h"

using namespace svt;

namespace {

constexpr int kThreads = 256, kPerThread = 4, kPerCta = kThreads * kPerThread;  // 1024 coefficients per CTA
constexpr int kMaxCtas = SVTGPU_NOALLOC_MAX_COEFFS / kPerCta;                  // 32
constexpr int kScratchWords = 64;  // [0] ticket, [1] unused, [2 .. 2 + kMaxCtas) per-CTA eob partials
static_assert(2 + kMaxCtas <= kScratchWords, "scratch too small for the stated maximum");

static __device__ __align__(256) unsigned int g_noalloc_scratch[kScratchWords];

struct QTab { int16_t v[10]; };  // zbin, round, quant, quant_shift, dequant x {DC, AC}

template <int LOG_SCALE>
__global__ void __launch_bounds__(kThreads)
quant_scratch_kernel(const int32_t* __restrict__ coeff, int n, const QTab tab, const int16_t* __restrict__ iscan,
                     int32_t* __restrict__ qcoeff, int32_t* __restrict__ dqcoeff, uint16_t* eob,
                     unsigned int* scratch) {
    const QSet dc = make_qset<LOG_SCALE, true>(tab.v[0], tab.v[2], tab.v[4], tab.v[6], tab.v[8]);
    const QSet ac = make_qset<LOG_SCALE, true>(tab.v[1], tab.v[3], tab.v[5], tab.v[7], tab.v[9]);
    __shared__ int s_eob;
    __shared__ bool s_last;
    if (threadIdx.x == 0) s_eob = 0;
    __syncthreads();
    int e = 0;
    const int base = blockIdx.x * kPerCta;
#pragma unroll
    for (int k = 0; k < kPerThread; ++k) {
        const int i = base + k * kThreads + threadIdx.x;  // consecutive lanes, consecutive words
        if (i < n) {
            int qc, dqc;
            const int aq = quant_flat<LOG_SCALE, true>(coeff[i], i == 0 ? dc : ac, 0x7fffffff, qc, dqc);
            qcoeff[i] = qc;
            dqcoeff[i] = dqc;
            if (aq != 0) { const int s = int(iscan[i]) + 1; e = s > e ? s : e; }
        }
    }
    if (e) atomicMax(&s_eob, e);
    __syncthreads();
    if (threadIdx.x == 0) {
        volatile unsigned int* part = scratch + 2;
        part[blockIdx.x] = unsigned(s_eob);
        __threadfence();
        s_last = atomicAdd(scratch, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (s_last && threadIdx.x == 0) {  // the CTA that arrives last folds the partials
        __threadfence();
        const volatile unsigned int* part = scratch + 2;
        unsigned m = 0;
        for (unsigned b = 0; b < gridDim.x; ++b) m = part[b] > m ? part[b] : m;
        *eob = uint16_t(m);
        scratch[0] = 0;  // leave the ticket clear for the next call
    }
}

std::mutex g_mu;
unsigned int* g_scratch_ptr[64] = {};  // per device, resolved once

// the device address of the scratch array on the CURRENT device
int scratch_of_current_device(unsigned int** out) {
    int dev = 0;
    CK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) {
        svtgpu_set_error("%s%s (line %d)", "svtgpu_quantize_b_noalloc: device ordinal beyond 63", "", __LINE__);
        return SVTGPU_ERR_ARG;
    }
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratch_ptr[dev]) {
        void* p = nullptr;
        CK(cudaGetSymbolAddress(&p, g_noalloc_scratch));  // NOT &g_noalloc_scratch: that is a host shadow
        if (!p) {
            svtgpu_set_error("%s%s (line %d)", "cudaGetSymbolAddress returned a null scratch address", "", __LINE__);
            return SVTGPU_ERR_CUDA;
        }
        g_scratch_ptr[dev] = static_cast<unsigned int*>(p);
    }
    *out = g_scratch_ptr[dev];
    return SVTGPU_OK;
}

}  // namespace

extern "C" int svtgpu_quantize_b_noalloc(const int32_t* d_coeff, intptr_t n_coeffs, int log_scale,
                                         const int16_t* zbin, const int16_t* round, const int16_t* quant,
                                         const int16_t* quant_shift, const int16_t* dequant,
                                         int32_t* d_qcoeff, int32_t* d_dqcoeff, uint16_t* d_eob,
                                         const int16_t* d_iscan, void* cuda_stream) {
    // argument checks first: they need no device
    if (n_coeffs > SVTGPU_NOALLOC_MAX_COEFFS) {
        char got[32], cap[32];
        snprintf(got, sizeof(got), "%lld", (long long)n_coeffs);
        snprintf(cap, sizeof(cap), "%d", SVTGPU_NOALLOC_MAX_COEFFS);
        static thread_local char msg[160];
        snprintf(msg, sizeof(msg), "svtgpu_quantize_b_noalloc: n_coeffs %s exceeds SVTGPU_NOALLOC_MAX_COEFFS = %s "
                 "(the fixed scratch array holds one eob partial per 1024 coefficients)", got, cap);
        svtgpu_set_error("%s%s (line %d)", msg, "", __LINE__);
        return SVTGPU_ERR_ARG;
    }
    if (n_coeffs <= 0 || log_scale < 0 || log_scale > 2 || !d_coeff || !zbin || !round || !quant || !quant_shift ||
        !dequant || !d_qcoeff || !d_dqcoeff || !d_eob || !d_iscan) {
        svtgpu_set_error("%s%s (line %d)", "svtgpu_quantize_b_noalloc: null pointer, n_coeffs <= 0 or log_scale outside 0..2",
                         "", __LINE__);
        return SVTGPU_ERR_ARG;
    }
    unsigned int* scratch = nullptr;
    const int rc = scratch_of_current_device(&scratch);
    if (rc != SVTGPU_OK) return rc;
    QTab tab;
    const int16_t* rows[5] = { zbin, round, quant, quant_shift, dequant };
    for (int i = 0; i < 5; ++i) { tab.v[2 * i] = rows[i][0]; tab.v[2 * i + 1] = rows[i][1]; }
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    const int n = int(n_coeffs), grid = (n + kPerCta - 1) / kPerCta;
    // the ticket must start at zero whatever an earlier (possibly failed) call left behind
    CK(cudaMemsetAsync(scratch, 0, sizeof(unsigned int), st));
    switch (log_scale) {
    case 0: quant_scratch_kernel<0><<<grid, kThreads, 0, st>>>(d_coeff, n, tab, d_iscan, d_qcoeff, d_dqcoeff, d_eob, scratch); break;
    case 1: quant_scratch_kernel<1><<<grid, kThreads, 0, st>>>(d_coeff, n, tab, d_iscan, d_qcoeff, d_dqcoeff, d_eob, scratch); break;
    default: quant_scratch_kernel<2><<<grid, kThreads, 0, st>>>(d_coeff, n, tab, d_iscan, d_qcoeff, d_dqcoeff, d_eob, scratch); break;
    }
    CK(cudaGetLastError());
    ++g_total_launches;
    return SVTGPU_OK;
}
