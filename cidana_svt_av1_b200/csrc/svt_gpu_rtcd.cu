// libsvtgpu.so, levels A and B of include/svt_gpu.h: the reference's own call signatures with HOST
// pointers.  Each call stages one transform block through pinned memory, runs the same sm_100a
// kernels as the batch API and returns when the result is back in the caller's buffers
// (the encoder reads eob / recon immediately after the call, Source/Lib/Common/Codec/
// EbCodingLoop.c:697-731).  State is thread-local: up to MIN(40, cores) EncDec threads call
// concurrently (Source/Lib/Encoder/Codec/EbEncHandle.c:576) and none of them takes a lock here.
// A CUDA failure prints and abort()s -- the slots return void and there is no CPU fallback.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/svt_gpu.h"
#include "host_plan.h"
#include "residual_kernel.cuh"
#include "svt_gpu_internal.h"

using namespace svt;

namespace {

[[noreturn]] void die(const char* what) {
    fprintf(stderr, "libsvtgpu: %s: %s\n", what, svtgpu_last_error());
    abort();
}
#define CKA(call)                                                                            \
    do {                                                                                     \
        cudaError_t e_ = (call);                                                             \
        if (e_ != cudaSuccess) {                                                             \
            fprintf(stderr, "libsvtgpu: CUDA error %s at %s:%d\n", cudaGetErrorString(e_),  \
                    __FILE__, __LINE__);                                                     \
            abort();                                                                         \
        }                                                                                    \
    } while (0)

constexpr int kMaxPix = 64 * 64;

// Per-thread staging: one context, one single-block plan whose device descriptor is rewritten
// per call, pinned host mirrors and device buffers sized for the largest block.
struct Tls {
    SvtGpuContext* ctx = nullptr;
    SvtGpuPlan plan;
    // pinned host staging
    int16_t* h_res = nullptr;
    uint16_t* h_pix = nullptr;   // also viewed as uint8_t
    int32_t* h_coeff = nullptr;  // 3 x kMaxPix: in/coeff, q, dq
    int16_t* h_tab = nullptr;    // 40 (table row) + kMaxPix (iscan)
    DevBlock* h_blk = nullptr;
    uint64_t* h_small = nullptr; // [0] energy, [1] eob
    // device
    int16_t* d_res = nullptr;
    uint16_t* d_pix = nullptr;
    int32_t* d_coeff = nullptr;  // 3 x kMaxPix
    int16_t* d_tab = nullptr;
    uint64_t* d_small = nullptr;
    // tx-type search of one block (svtgpu_tx_search_block): up to 16 candidates in one launch
    SvtGpuPlan plan16;
    DevBlock* h_blk16 = nullptr;       // pinned, 16 descriptors
    int32_t* d_q16 = nullptr;          // 16 x 1024 quantised coefficients
    int32_t* h_q16 = nullptr;          // pinned
    uint64_t* d_out16 = nullptr;       // [0..15] energy, [16..47] distortion pairs, then 16 x u16 eob
    uint64_t* h_out16 = nullptr;       // pinned

    void init() {
        if (ctx) return;
        const char* e = getenv("SVTGPU_DEVICE");
        if (svtgpu_create(&ctx, e ? atoi(e) : 0) != SVTGPU_OK) die("svtgpu_create");
        CKA(cudaMallocHost(&h_res, kMaxPix * sizeof(int16_t)));
        CKA(cudaMallocHost(&h_pix, kMaxPix * sizeof(uint16_t)));
        CKA(cudaMallocHost(&h_coeff, 3 * kMaxPix * sizeof(int32_t)));
        CKA(cudaMallocHost(&h_tab, (40 + kMaxPix) * sizeof(int16_t)));
        CKA(cudaMallocHost(&h_blk, sizeof(DevBlock)));
        CKA(cudaMallocHost(&h_small, 2 * sizeof(uint64_t)));
        CKA(cudaMalloc(&d_res, kMaxPix * sizeof(int16_t)));
        CKA(cudaMalloc(&d_pix, kMaxPix * sizeof(uint16_t)));
        CKA(cudaMalloc(&d_coeff, 3 * kMaxPix * sizeof(int32_t)));
        CKA(cudaMalloc(&d_tab, (40 + kMaxPix) * sizeof(int16_t)));
        CKA(cudaMalloc(&d_small, 2 * sizeof(uint64_t)));
        CKA(cudaMalloc(&plan.d_blocks, sizeof(DevBlock)));
        plan.ctx = ctx;
        plan.n = 1;
        CKA(cudaMalloc(&plan16.d_blocks, 16 * sizeof(DevBlock)));
        plan16.ctx = ctx;
        CKA(cudaMallocHost(&h_blk16, 16 * sizeof(DevBlock)));
        CKA(cudaMalloc(&d_q16, 16 * 1024 * sizeof(int32_t)));
        CKA(cudaMallocHost(&h_q16, 16 * 1024 * sizeof(int32_t)));
        CKA(cudaMalloc(&d_out16, 52 * sizeof(uint64_t)));
        CKA(cudaMallocHost(&h_out16, 52 * sizeof(uint64_t)));
        // the quantizer table of this context is the single row staged in d_tab
        ctx->d_qtables = d_tab;
        ctx->n_qrows = 1;
    }
    ~Tls() {
        if (!ctx) return;
        ctx->d_qtables = nullptr;  // owned here, not by the context
        cudaStreamSynchronize(ctx->stream);
        cudaFree(plan16.d_blocks); cudaFree(d_q16); cudaFree(d_out16);
        cudaFreeHost(h_blk16); cudaFreeHost(h_q16); cudaFreeHost(h_out16);
        cudaFree(plan.d_blocks); cudaFree(d_res); cudaFree(d_pix); cudaFree(d_coeff);
        cudaFree(d_tab); cudaFree(d_small);
        cudaFreeHost(h_res); cudaFreeHost(h_pix); cudaFreeHost(h_coeff); cudaFreeHost(h_tab);
        cudaFreeHost(h_blk); cudaFreeHost(h_small);
        svtgpu_destroy(ctx);
    }
};
thread_local Tls g_tls;

inline void check_pair(int s, int t) {
    if (!tx_pair_valid(s, t)) {
        fprintf(stderr, "libsvtgpu: illegal (tx_size %d, tx_type %d)\n", s, t);
        abort();
    }
}

// Stage the descriptor and launch `stages` for one block held densely in the staging buffers.
void run_one(Tls& t, int s, int tx_type, int bd, int stages, bool clamp16) {
    for (int i = 0; i < 19; ++i) { t.plan.first[i] = 0; t.plan.count[i] = 0; }
    t.plan.count[s] = 1;
    t.plan.max_qrow = 0;
    DevBlock& d = *t.h_blk;
    d.res_off = 0; d.res_stride = kTxW[s]; d.rec_off = 0; d.rec_stride = kTxW[s];
    d.coeff_off = 0; d.qrow = 0; d.id = 0;
    d.meta = make_meta(tx_type, bd, 0, clamp16, false);
    cudaStream_t st = t.ctx->stream;
    CKA(cudaMemcpyAsync(t.plan.d_blocks, t.h_blk, sizeof(DevBlock), cudaMemcpyHostToDevice, st));
    SvtGpuBuffers b;
    memset(&b, 0, sizeof(b));
    b.residual = t.d_res; b.pred = t.d_pix; b.recon = t.d_pix;
    b.coeff_in = t.d_coeff; b.coeff = t.d_coeff;
    b.qcoeff = t.d_coeff + kMaxPix; b.dqcoeff = t.d_coeff + 2 * kMaxPix;
    b.eob = reinterpret_cast<uint16_t*>(t.d_small + 1); b.energy = t.d_small;
    if ((stages & STAGE_INV) && !(stages & STAGE_QUANT)) b.coeff_in = t.d_coeff + 2 * kMaxPix;
    if (svtgpu_run_internal(t.ctx, &t.plan, &b, stages, st) != SVTGPU_OK) die("svtgpu_run");
}

}  // namespace

// ---- forward ---------------------------------------------------------------------------------------
// full = RTCD-level result (all w*h coefficients at stride w); otherwise av1_estimate_transform's
// packed result + three_quad_energy
static void fwd_host(int s, int16_t* input, uint32_t stride, int32_t* output, int tx_type, int bd,
                     bool full, uint64_t* energy) {
    check_pair(s, tx_type);
    Tls& t = g_tls;
    t.init();
    const int w = kTxW[s], h = kTxH[s];
    for (int r = 0; r < h; ++r) memcpy(t.h_res + r * w, input + size_t(r) * stride, w * sizeof(int16_t));
    cudaStream_t st = t.ctx->stream;
    CKA(cudaMemcpyAsync(t.d_res, t.h_res, w * h * sizeof(int16_t), cudaMemcpyHostToDevice, st));
    run_one(t, s, tx_type, bd, STAGE_FWD | (full ? STAGE_FWD_FULL : 0), false);
    const int n = full ? w * h : max_eob(s);
    CKA(cudaMemcpyAsync(t.h_coeff, t.d_coeff, n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    CKA(cudaMemcpyAsync(t.h_small, t.d_small, sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
    CKA(cudaStreamSynchronize(st));
    memcpy(output, t.h_coeff, n * sizeof(int32_t));
    if (energy) *energy = t.h_small[0];
}

#define SVTGPU_FWD_DEF(WxH, ID)                                                                   \
    extern "C" void svtgpu_av1_fwd_txfm2d_##WxH(int16_t* input, int32_t* output,                 \
                                                uint32_t input_stride, uint8_t transform_type,   \
                                                uint8_t bit_depth) {                              \
        fwd_host(ID, input, input_stride, output, transform_type, bit_depth, true, nullptr);     \
    }
SVTGPU_FWD_DEF(4x4, 0) SVTGPU_FWD_DEF(8x8, 1) SVTGPU_FWD_DEF(16x16, 2) SVTGPU_FWD_DEF(32x32, 3)
SVTGPU_FWD_DEF(64x64, 4) SVTGPU_FWD_DEF(4x8, 5) SVTGPU_FWD_DEF(8x4, 6) SVTGPU_FWD_DEF(8x16, 7)
SVTGPU_FWD_DEF(16x8, 8) SVTGPU_FWD_DEF(16x32, 9) SVTGPU_FWD_DEF(32x16, 10) SVTGPU_FWD_DEF(32x64, 11)
SVTGPU_FWD_DEF(64x32, 12) SVTGPU_FWD_DEF(4x16, 13) SVTGPU_FWD_DEF(16x4, 14) SVTGPU_FWD_DEF(8x32, 15)
SVTGPU_FWD_DEF(32x8, 16) SVTGPU_FWD_DEF(16x64, 17) SVTGPU_FWD_DEF(64x16, 18)

extern "C" int svtgpu_av1_estimate_transform(int16_t* residual_buffer, uint32_t residual_stride,
                                             int32_t* coeff_buffer, uint32_t coeff_stride,
                                             uint8_t transform_size, uint64_t* three_quad_energy,
                                             int16_t* transform_inner_array_ptr,
                                             uint32_t bit_increment, uint8_t transform_type,
                                             int asm_type, int component_type,
                                             int trans_coeff_shape) {
    (void)coeff_stride; (void)transform_inner_array_ptr; (void)asm_type; (void)component_type;
    (void)trans_coeff_shape;  // unused by the reference too, EbTransforms.c:4934-4939
    uint64_t e = 0;
    fwd_host(transform_size, residual_buffer, residual_stride, coeff_buffer, transform_type,
             bit_increment ? 10 : 8, false, &e);
    // the reference only writes the energy for the 64-point sizes
    if (three_quad_energy && (kTxW[transform_size] == 64 || kTxH[transform_size] == 64))
        *three_quad_energy = e;
    return 0;  // EB_ErrorNone
}

// ---- quantizer -------------------------------------------------------------------------------------
template <int LS>
static void quant_slot(const int32_t* coeff_ptr, intptr_t n_coeffs, int32_t skip_block,
                       const int16_t* zbin, const int16_t* round, const int16_t* quant,
                       const int16_t* quant_shift, int32_t* qcoeff, int32_t* dqcoeff,
                       const int16_t* dequant, uint16_t* eob, const int16_t* iscan) {
    const int n = int(n_coeffs);
    if (n <= 0 || n > kMaxPix) { fprintf(stderr, "libsvtgpu: bad n_coeffs %d\n", n); abort(); }
    if (skip_block) {  // EbFullLoop.c:253-256
        memset(qcoeff, 0, n * sizeof(int32_t));
        memset(dqcoeff, 0, n * sizeof(int32_t));
        *eob = 0;
        return;
    }
    Tls& t = g_tls;
    t.init();
    const int16_t* rows[5] = { zbin, round, quant, quant_shift, dequant };
    for (int i = 0; i < 5; ++i) { t.h_tab[2 * i] = rows[i][0]; t.h_tab[2 * i + 1] = rows[i][1]; }
    memcpy(t.h_tab + 40, iscan, n * sizeof(int16_t));
    memcpy(t.h_coeff, coeff_ptr, n * sizeof(int32_t));
    cudaStream_t st = t.ctx->stream;
    CKA(cudaMemcpyAsync(t.d_tab, t.h_tab, (40 + n) * sizeof(int16_t), cudaMemcpyHostToDevice, st));
    CKA(cudaMemcpyAsync(t.d_coeff, t.h_coeff, n * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    uint16_t* d_eob = reinterpret_cast<uint16_t*>(t.d_small + 1);
    quant_kernel<LS><<<1, 256, 0, st>>>(t.d_coeff, n, t.d_tab, t.d_tab + 40, t.d_coeff + kMaxPix,
                                        t.d_coeff + 2 * kMaxPix, d_eob);
    CKA(cudaGetLastError());
    t.ctx->launches++;
    ++g_total_launches;
    CKA(cudaMemcpyAsync(t.h_coeff + kMaxPix, t.d_coeff + kMaxPix, n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    CKA(cudaMemcpyAsync(t.h_coeff + 2 * kMaxPix, t.d_coeff + 2 * kMaxPix, n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    CKA(cudaMemcpyAsync(t.h_small + 1, t.d_small + 1, sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
    CKA(cudaStreamSynchronize(st));
    memcpy(qcoeff, t.h_coeff + kMaxPix, n * sizeof(int32_t));
    memcpy(dqcoeff, t.h_coeff + 2 * kMaxPix, n * sizeof(int32_t));
    *eob = *reinterpret_cast<uint16_t*>(t.h_small + 1);
}

#define SVTGPU_QUANT_DEF(name, LS)                                                                 \
    extern "C" void name(const int32_t* coeff_ptr, intptr_t n_coeffs, int32_t skip_block,         \
                         const int16_t* zbin_ptr, const int16_t* round_ptr,                       \
                         const int16_t* quant_ptr, const int16_t* quant_shift_ptr,                \
                         int32_t* qcoeff_ptr, int32_t* dqcoeff_ptr, const int16_t* dequant_ptr,   \
                         uint16_t* eob_ptr, const int16_t* scan, const int16_t* iscan) {          \
        (void)scan;                                                                                \
        quant_slot<LS>(coeff_ptr, n_coeffs, skip_block, zbin_ptr, round_ptr, quant_ptr,           \
                       quant_shift_ptr, qcoeff_ptr, dqcoeff_ptr, dequant_ptr, eob_ptr, iscan);    \
    }
SVTGPU_QUANT_DEF(svtgpu_aom_quantize_b, 0) SVTGPU_QUANT_DEF(svtgpu_aom_quantize_b_32x32, 1)
SVTGPU_QUANT_DEF(svtgpu_aom_quantize_b_64x64, 2) SVTGPU_QUANT_DEF(svtgpu_aom_highbd_quantize_b, 0)
SVTGPU_QUANT_DEF(svtgpu_aom_highbd_quantize_b_32x32, 1)
SVTGPU_QUANT_DEF(svtgpu_aom_highbd_quantize_b_64x64, 2)

extern "C" void svtgpu_av1_quantize_inv_quantize(const int16_t* zbin, const int16_t* round,
                                                 const int16_t* quant, const int16_t* quant_shift,
                                                 const int16_t* dequant, int32_t* coeff,
                                                 int32_t* quant_coeff, int32_t* recon_coeff,
                                                 uint8_t txsize, uint16_t* eob,
                                                 uint32_t* y_count_non_zero_coeffs,
                                                 uint32_t bit_increment, uint8_t tx_type,
                                                 int clamp16) {
    check_pair(txsize, tx_type);
    Tls& t = g_tls;
    t.init();
    const int n = max_eob(txsize);
    const int16_t* rows[5] = { zbin, round, quant, quant_shift, dequant };
    for (int i = 0; i < 5; ++i) memcpy(t.h_tab + 8 * i, rows[i], 2 * sizeof(int16_t));
    memcpy(t.h_coeff, coeff, n * sizeof(int32_t));
    cudaStream_t st = t.ctx->stream;
    CKA(cudaMemcpyAsync(t.d_tab, t.h_tab, 40 * sizeof(int16_t), cudaMemcpyHostToDevice, st));
    CKA(cudaMemcpyAsync(t.d_coeff, t.h_coeff, n * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    run_one(t, txsize, tx_type, bit_increment ? 10 : 8, STAGE_QUANT, clamp16 && !bit_increment);
    CKA(cudaMemcpyAsync(t.h_coeff + kMaxPix, t.d_coeff + kMaxPix, n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    CKA(cudaMemcpyAsync(t.h_coeff + 2 * kMaxPix, t.d_coeff + 2 * kMaxPix, n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    CKA(cudaMemcpyAsync(t.h_small + 1, t.d_small + 1, sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
    CKA(cudaStreamSynchronize(st));
    memcpy(quant_coeff, t.h_coeff + kMaxPix, n * sizeof(int32_t));
    memcpy(recon_coeff, t.h_coeff + 2 * kMaxPix, n * sizeof(int32_t));
    *eob = *reinterpret_cast<uint16_t*>(t.h_small + 1);
    if (y_count_non_zero_coeffs) *y_count_non_zero_coeffs = *eob;  // EbFullLoop.c:650
}

// ---- inverse + reconstruction ----------------------------------------------------------------------
// pix16: the caller's pixels are uint16_t (highbd slots: even for bd 8, as av1_inv_txfm_add_c's
// widened copy, EbTransforms.c:8882); otherwise uint8_t.
static void inv_host(int s, const int32_t* input, void* output, int32_t stride, int tx_type, int bd,
                     bool pix16) {
    check_pair(s, tx_type);
    if (bd != 8 && bd != 10) { fprintf(stderr, "libsvtgpu: bit depth %d not supported\n", bd); abort(); }
    Tls& t = g_tls;
    t.init();
    const int w = kTxW[s], h = kTxH[s], n = max_eob(s);
    uint8_t* h8 = reinterpret_cast<uint8_t*>(t.h_pix);
    for (int r = 0; r < h; ++r) {
        if (bd == 10) {
            memcpy(t.h_pix + r * w, static_cast<uint16_t*>(output) + size_t(r) * stride, w * sizeof(uint16_t));
        } else if (pix16) {
            const uint16_t* src = static_cast<uint16_t*>(output) + size_t(r) * stride;
            for (int c = 0; c < w; ++c) h8[r * w + c] = uint8_t(src[c]);
        } else {
            memcpy(h8 + r * w, static_cast<uint8_t*>(output) + size_t(r) * stride, w);
        }
    }
    memcpy(t.h_coeff + 2 * kMaxPix, input, n * sizeof(int32_t));
    cudaStream_t st = t.ctx->stream;
    const size_t pbytes = size_t(w) * h * (bd == 10 ? 2 : 1);
    CKA(cudaMemcpyAsync(t.d_pix, t.h_pix, pbytes, cudaMemcpyHostToDevice, st));
    CKA(cudaMemcpyAsync(t.d_coeff + 2 * kMaxPix, t.h_coeff + 2 * kMaxPix, n * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    run_one(t, s, tx_type, bd, STAGE_INV, false);
    CKA(cudaMemcpyAsync(t.h_pix, t.d_pix, pbytes, cudaMemcpyDeviceToHost, st));
    CKA(cudaStreamSynchronize(st));
    for (int r = 0; r < h; ++r) {
        if (bd == 10) {
            memcpy(static_cast<uint16_t*>(output) + size_t(r) * stride, t.h_pix + r * w, w * sizeof(uint16_t));
        } else if (pix16) {
            uint16_t* dst = static_cast<uint16_t*>(output) + size_t(r) * stride;
            for (int c = 0; c < w; ++c) dst[c] = h8[r * w + c];
        } else {
            memcpy(static_cast<uint8_t*>(output) + size_t(r) * stride, h8 + r * w, w);
        }
    }
}

#define SVTGPU_INV_SQ_DEF(WxH, ID)                                                                \
    extern "C" void svtgpu_av1_inv_txfm2d_add_##WxH(const int32_t* input, uint16_t* output,      \
                                                    int32_t stride, uint8_t tx_type, int32_t bd) { \
        inv_host(ID, input, output, stride, tx_type, bd, true);                                   \
    }
#define SVTGPU_INV_RECT_DEF(WxH, ID)                                                              \
    extern "C" void svtgpu_av1_inv_txfm2d_add_##WxH(const int32_t* input, uint16_t* output,      \
                                                    int32_t stride, uint8_t tx_type,              \
                                                    uint8_t tx_size, int32_t eob, int32_t bd) {   \
        (void)tx_size; (void)eob; /* eob only selects shortcuts in the reference */               \
        inv_host(ID, input, output, stride, tx_type, bd, true);                                   \
    }
#define SVTGPU_INV_4N_DEF(WxH, ID)                                                                \
    extern "C" void svtgpu_av1_inv_txfm2d_add_##WxH(const int32_t* input, uint16_t* output,      \
                                                    int32_t stride, uint8_t tx_type,              \
                                                    uint8_t tx_size, int32_t bd) {                \
        (void)tx_size;                                                                            \
        inv_host(ID, input, output, stride, tx_type, bd, true);                                   \
    }
SVTGPU_INV_SQ_DEF(4x4, 0) SVTGPU_INV_SQ_DEF(8x8, 1) SVTGPU_INV_SQ_DEF(16x16, 2)
SVTGPU_INV_SQ_DEF(32x32, 3) SVTGPU_INV_SQ_DEF(64x64, 4)
SVTGPU_INV_RECT_DEF(8x16, 7) SVTGPU_INV_RECT_DEF(16x8, 8) SVTGPU_INV_RECT_DEF(16x32, 9)
SVTGPU_INV_RECT_DEF(32x16, 10) SVTGPU_INV_RECT_DEF(32x8, 16) SVTGPU_INV_RECT_DEF(8x32, 15)
SVTGPU_INV_RECT_DEF(32x64, 11) SVTGPU_INV_RECT_DEF(64x32, 12) SVTGPU_INV_RECT_DEF(16x64, 17)
SVTGPU_INV_RECT_DEF(64x16, 18)
SVTGPU_INV_4N_DEF(4x8, 5) SVTGPU_INV_4N_DEF(8x4, 6) SVTGPU_INV_4N_DEF(4x16, 13)
SVTGPU_INV_4N_DEF(16x4, 14)

extern "C" void svtgpu_av1_inv_txfm_add(const int32_t* dqcoeff, uint8_t* dst, int32_t stride,
                                        const SvtGpuTxfmParam* p) {
    inv_host(p->tx_size, dqcoeff, dst, stride, p->tx_type, 8, false);
}

extern "C" int svtgpu_av1_inv_transform_recon8bit(int32_t* coeff_buffer, uint8_t* recon_buffer,
                                                  uint32_t recon_stride, uint8_t txsize,
                                                  uint8_t transform_type, int component_type,
                                                  uint32_t eob) {
    (void)component_type; (void)eob;
    inv_host(txsize, coeff_buffer, recon_buffer, int32_t(recon_stride), transform_type, 8, false);
    return 0;
}
extern "C" int svtgpu_av1_inv_transform_recon(int32_t* coeff_buffer, uint8_t* recon_buffer,
                                              uint32_t recon_stride, uint8_t txsize,
                                              uint32_t bit_increment, uint8_t transform_type,
                                              int component_type, uint32_t eob) {
    (void)component_type; (void)eob;
    // the encoder passes CONVERT_TO_BYTEPTR(uint16_t*) (EbDefinitions.h:700-701): undo the shift
    uint16_t* px = reinterpret_cast<uint16_t*>(uintptr_t(recon_buffer) << 1);
    inv_host(txsize, coeff_buffer, px, int32_t(recon_stride), transform_type,
             bit_increment ? 10 : 8, true);
    return 0;
}

// ---- tx-type search of one block ------------------------------------------------------------------
// The per-type loop of encode_pass_tx_search / ProductFullLoopTxSearch (EbFullLoop.c:1127-1318,
// :927-1100) runs av1_estimate_transform -> av1_quantize_inv_quantize ->
// picture_full_distortion32_bits once per candidate type.  Here the residual is staged once and
// all candidate types of the block go through ONE launch of the fused path (forward + quantise +
// distortion pair): the caller keeps its own rate estimate and cost, candidate by candidate.
extern "C" int svtgpu_tx_search_block(const int16_t* residual, uint32_t residual_stride, uint8_t tx_size,
                                      const uint8_t* tx_types, int n_types, const int16_t* zbin,
                                      const int16_t* round, const int16_t* quant, const int16_t* quant_shift,
                                      const int16_t* dequant, uint32_t bit_increment, int32_t* quant_coeff,
                                      uint16_t* eob, uint64_t* dist, uint64_t* three_quad_energy) {
    if (!residual || !tx_types || n_types < 1 || n_types > 16 || tx_size > 18 || !quant_coeff || !eob || !dist)
        return SVTGPU_ERR_ARG;
    for (int k = 0; k < n_types; ++k) check_pair(tx_size, tx_types[k]);
    Tls& t = g_tls;
    t.init();
    const int s = tx_size, w = kTxW[s], h = kTxH[s], n = max_eob(s), bd = bit_increment ? 10 : 8;
    for (int r = 0; r < h; ++r) memcpy(t.h_res + r * w, residual + size_t(r) * residual_stride, w * sizeof(int16_t));
    const int16_t* rows[5] = { zbin, round, quant, quant_shift, dequant };
    for (int i = 0; i < 5; ++i) memcpy(t.h_tab + 8 * i, rows[i], 2 * sizeof(int16_t));
    for (int k = 0; k < n_types; ++k) {
        DevBlock& d = t.h_blk16[k];
        d.res_off = 0; d.res_stride = w; d.rec_off = 0; d.rec_stride = w;
        d.coeff_off = uint32_t(k * n); d.qrow = 0; d.id = uint32_t(k);
        d.meta = make_meta(tx_types[k], bd, 0, false, false);
    }
    cudaStream_t st = t.ctx->stream;
    CKA(cudaMemcpyAsync(t.d_res, t.h_res, w * h * sizeof(int16_t), cudaMemcpyHostToDevice, st));
    CKA(cudaMemcpyAsync(t.d_tab, t.h_tab, 40 * sizeof(int16_t), cudaMemcpyHostToDevice, st));
    CKA(cudaMemcpyAsync(t.plan16.d_blocks, t.h_blk16, n_types * sizeof(DevBlock), cudaMemcpyHostToDevice, st));
    for (int i = 0; i < 19; ++i) { t.plan16.first[i] = 0; t.plan16.count[i] = 0; }
    t.plan16.count[s] = n_types;
    t.plan16.n = n_types;
    t.plan16.max_qrow = 0;
    t.plan16.bd_uniform = bd;
    SvtGpuBuffers b;
    memset(&b, 0, sizeof(b));
    b.residual = t.d_res;
    b.qcoeff = t.d_q16;
    b.energy = t.d_out16;
    b.dist = t.d_out16 + 16;
    b.eob = reinterpret_cast<uint16_t*>(t.d_out16 + 48);
    if (svtgpu_run_internal(t.ctx, &t.plan16, &b, STAGE_FWD | STAGE_QUANT, st) != SVTGPU_OK) die("svtgpu_run (tx search)");
    CKA(cudaMemcpyAsync(t.h_q16, t.d_q16, size_t(n_types) * n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    CKA(cudaMemcpyAsync(t.h_out16, t.d_out16, 52 * sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
    CKA(cudaStreamSynchronize(st));
    memcpy(quant_coeff, t.h_q16, size_t(n_types) * n * sizeof(int32_t));
    const uint16_t* e16 = reinterpret_cast<const uint16_t*>(t.h_out16 + 48);
    for (int k = 0; k < n_types; ++k) {
        eob[k] = e16[k];
        dist[2 * k] = t.h_out16[16 + 2 * k];
        dist[2 * k + 1] = t.h_out16[16 + 2 * k + 1];
        if (three_quad_energy) three_quad_energy[k] = t.h_out16[k];
    }
    return SVTGPU_OK;
}

// ---- the slot table an installer walks ------------------------------------------
#define E(slot) { #slot, (void*)svtgpu_##slot }
static const SvtGpuRtcdEntry k_rtcd[] = {
    E(av1_fwd_txfm2d_4x4), E(av1_fwd_txfm2d_8x8), E(av1_fwd_txfm2d_16x16), E(av1_fwd_txfm2d_32x32),
    E(av1_fwd_txfm2d_64x64), E(av1_fwd_txfm2d_4x8), E(av1_fwd_txfm2d_8x4), E(av1_fwd_txfm2d_8x16),
    E(av1_fwd_txfm2d_16x8), E(av1_fwd_txfm2d_16x32), E(av1_fwd_txfm2d_32x16),
    E(av1_fwd_txfm2d_32x64), E(av1_fwd_txfm2d_64x32), E(av1_fwd_txfm2d_4x16),
    E(av1_fwd_txfm2d_16x4), E(av1_fwd_txfm2d_8x32), E(av1_fwd_txfm2d_32x8),
    E(av1_fwd_txfm2d_16x64), E(av1_fwd_txfm2d_64x16),
    E(aom_quantize_b), E(aom_quantize_b_32x32), E(aom_quantize_b_64x64), E(aom_highbd_quantize_b),
    E(aom_highbd_quantize_b_32x32), E(aom_highbd_quantize_b_64x64),
    E(av1_inv_txfm2d_add_4x4), E(av1_inv_txfm2d_add_8x8), E(av1_inv_txfm2d_add_16x16),
    E(av1_inv_txfm2d_add_32x32), E(av1_inv_txfm2d_add_64x64), E(av1_inv_txfm2d_add_4x8),
    E(av1_inv_txfm2d_add_8x4), E(av1_inv_txfm2d_add_8x16), E(av1_inv_txfm2d_add_16x8),
    E(av1_inv_txfm2d_add_16x32), E(av1_inv_txfm2d_add_32x16), E(av1_inv_txfm2d_add_32x64),
    E(av1_inv_txfm2d_add_64x32), E(av1_inv_txfm2d_add_4x16), E(av1_inv_txfm2d_add_16x4),
    E(av1_inv_txfm2d_add_8x32), E(av1_inv_txfm2d_add_32x8), E(av1_inv_txfm2d_add_16x64),
    E(av1_inv_txfm2d_add_64x16), E(av1_inv_txfm_add),
};
#undef E
extern "C" int svtgpu_rtcd_table(const SvtGpuRtcdEntry** entries) {
    if (entries) *entries = k_rtcd;
    return int(sizeof(k_rtcd) / sizeof(k_rtcd[0]));
}
