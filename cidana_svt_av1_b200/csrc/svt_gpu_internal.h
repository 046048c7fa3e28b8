// Internal definitions shared by the translation units of libsvtgpu.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/svt_gpu.h"
#include "residual_kernel.cuh"

void svtgpu_set_error(const char* fmt, const char* a, const char* b, int line);

#define CK(call)                                                                              \
    do {                                                                                      \
        cudaError_t e_ = (call);                                                              \
        if (e_ != cudaSuccess) {                                                              \
            svtgpu_set_error("CUDA error: %s at %s (line %d)", cudaGetErrorString(e_), #call, \
                             __LINE__);                                                       \
            return SVTGPU_ERR_CUDA;                                                           \
        }                                                                                     \
    } while (0)

// One lane of the host-buffer pipeline (svtgpu_submit_host): a stream, device buffers grown on
// demand and page-locked staging for the sorted block descriptors.
struct HostSlot {
    cudaStream_t st = nullptr;
    cudaEvent_t done = nullptr;
    bool busy = false;
    int ticket = -1;
    svt::DevBlock* d_blocks = nullptr;
    svt::DevBlock* h_blocks = nullptr;  // page-locked
    size_t cap_hblocks = 0;
    size_t cap_blocks = 0;
    int16_t* d_res = nullptr;
    size_t cap_res = 0;
    void* d_pix = nullptr;  // prediction in, reconstruction out (in place)
    size_t cap_pix = 0;     // bytes
    // one capacity per buffer: a failed allocation can only leave ITS buffer empty
    int32_t* d_q = nullptr;
    size_t cap_q = 0;
    int32_t* d_dq = nullptr;
    size_t cap_dq = 0;
    uint16_t* d_eob = nullptr;
    size_t cap_eob = 0;
    uint64_t* d_energy = nullptr;
    size_t cap_energy = 0;
    uint64_t* d_dist = nullptr;
    size_t cap_dist = 0;
    void* d_src = nullptr;  // source pixels (SVTGPU_STAGE_RESIDUAL)
    size_t cap_src = 0;     // bytes
};

struct SvtGpuContext {
    static constexpr int kSide = 4;
    static constexpr int kHostSlots = 3;
    HostSlot host[kHostSlots];
    int host_next = 0;
    int host_tickets = 0;
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t side[kSide] = {};
    cudaEvent_t side_done[kSide] = {};
    cudaEvent_t fork = nullptr;
    int16_t* d_iscan = nullptr;
    int iscan_off[19][3] = {};
    int16_t* d_qtables = nullptr;
    int n_qrows = 0;
    uint8_t* d_qm_wt = nullptr;
    uint8_t* d_qm_iwt = nullptr;
    int qm_off[2][19] = {};
    long long launches = 0;
};

struct SvtGpuPlan {
    SvtGpuContext* ctx = nullptr;
    int n = 0;
    svt::DevBlock* d_blocks = nullptr;
    int first[19] = {};
    int count[19] = {};
    uint32_t max_qrow = 0;
    bool uses_qm = false;
    uint32_t qm_sizes[2] = {};  // [plane > 0]: bit s = some USE_QM block of tx_size s
    bool any_clamp16 = false;
    int bd_uniform = 0;  // 8 / 10 when every block has that bit depth (selects the compiled-in pixel type)
    // extents the block list reaches (elements): what a host frame must provide
    size_t need_res = 0, need_pix = 0, need_coeff = 0;
};

#include <atomic>
extern std::atomic<long long> g_total_launches;  // svt_gpu_api.cu

// svtgpu_run plus the internal STAGE_FWD_FULL flag (used by the RTCD-level forward slots)
int svtgpu_run_internal(SvtGpuContext* c, const SvtGpuPlan* pl, const SvtGpuBuffers* b, int stages,
                        void* cuda_stream);

// one launch of the fused kernel for one size bucket (svt_gpu_launch.cu)
namespace svt {
cudaError_t launch_bucket(int tx_size, const KParams& p, int first, int count, cudaStream_t st);
}

// frees the lanes of the host-buffer pipeline (svt_gpu_host.cu)
void svtgpu_host_release(SvtGpuContext* c);
