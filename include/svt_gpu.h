/*
 * svt_gpu.h -- C ABI of libsvtgpu.so: the B200 (sm_100a) implementation of SVT-AV1's
 * residual-coding path (forward 2-D transform, quantize/dequantize/eob, inverse transform +
 * reconstruction).  Plain pointers and sizes only; no CUDA or torch types appear here.
 *
 * Three levels, mirroring SURVEY.md 8b.  Reference paths are relative to
 * Source/Lib/Common/Codec/ of Cidana-SVT-AV1.
 *
 *   C. batch API (svtgpu_create .. svtgpu_run)  -- device-resident, asynchronous, what a batched
 *      caller and the benchmarks use; one call = every transform block of a frame batch.
 *   B. entry points  svtgpu_av1_estimate_transform / svtgpu_av1_quantize_inv_quantize /
 *      svtgpu_av1_inv_transform_recon[8bit] -- same arguments as EbTransforms.h:3896-3987 minus
 *      the PictureControlSet (the five quantizer table rows are passed instead); host pointers,
 *      synchronous.
 *   A. RTCD slots  svtgpu_av1_fwd_txfm2d_WxH, svtgpu_aom_[highbd_]quantize_b[_32x32|_64x64],
 *      svtgpu_av1_inv_txfm2d_add_WxH, svtgpu_av1_inv_txfm_add -- byte-for-byte the prototypes of
 *      the function pointers at aom_dsp_rtcd.h:158-255, 321-342, 347-423, so that
 *      `av1_fwd_txfm2d_8x8 = svtgpu_av1_fwd_txfm2d_8x8;` compiles in the reference.  Host
 *      pointers, synchronous, re-entrant (thread-local staging), void return: a CUDA failure
 *      prints and abort()s -- there is no CPU fallback.
 *
 * TxType / TxSize are 1-byte packed enums in the reference (EbDefinitions.h:621, :728), hence
 * uint8_t below.
 */
#ifndef SVT_GPU_H
#define SVT_GPU_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------------------------------ */
/* Level C: batch API                                                                          */
/* ------------------------------------------------------------------------------------------ */
typedef struct SvtGpuContext SvtGpuContext;
typedef struct SvtGpuPlan SvtGpuPlan;

enum { SVTGPU_OK = 0, SVTGPU_ERR_ARG = -1, SVTGPU_ERR_CUDA = -2, SVTGPU_ERR_NOMEM = -3 };
enum { SVTGPU_STAGE_FWD = 1, SVTGPU_STAGE_QUANT = 2, SVTGPU_STAGE_INV = 4, SVTGPU_STAGE_ALL = 7,
       /* with STAGE_FWD: residual = src - pred computed on the fly (ResidualKernel,
        * aom_dsp_rtcd.h:2372; residual_kernel16bit, EbPictureOperators.c:134) instead of read */
       SVTGPU_STAGE_RESIDUAL = 16 };
enum { SVTGPU_BLK_CLAMP16 = 1, /* 8-bit C quantizer variant, EbFullLoop.c:86-88 */
       SVTGPU_BLK_USE_QM = 2 /* quantization matrices, EbFullLoop.c:239-296 */ };

/* One transform block.  Offsets index the batch-wide arrays of SvtGpuBuffers. */
typedef struct SvtGpuBlock {
    uint8_t tx_size;     /* TxSize 0..18 */
    uint8_t tx_type;     /* TxType 0..15; must be legal for tx_size (AV1 5.11) */
    uint8_t bit_depth;   /* 8 or 10 */
    uint8_t plane;       /* 0 Y, 1 U, 2 V (selects the qmatrix set) */
    uint32_t res_off;    /* int16 elements into residual */
    uint32_t res_stride; /* int16 elements */
    uint32_t rec_off;    /* pixels into pred / recon */
    uint32_t rec_stride; /* pixels */
    uint32_t coeff_off;  /* int32 elements into coeff / qcoeff / dqcoeff; multiple of 4 */
    uint32_t qrow;       /* row of the uploaded quantizer tables */
    uint32_t flags;      /* SVTGPU_BLK_* */
} SvtGpuBlock;

/* Device pointers (from svtgpu_malloc or any CUDA allocation of the same device). */
typedef struct SvtGpuBuffers {
    const int16_t *residual; /* STAGE_FWD input */
    const void *pred;        /* STAGE_INV input: uint8 (bd 8) / uint16 (bd 10) pixels */
    void *recon;             /* STAGE_INV output; may equal pred (in place, like the reference) */
    const int32_t *coeff_in; /* input of QUANT / INV when STAGE_FWD is not requested */
    int32_t *coeff;          /* optional output: transform coefficients, packed to min(w,32) */
    int32_t *qcoeff;         /* STAGE_QUANT outputs */
    int32_t *dqcoeff;        /* optional (NULL: not written; the inverse stage does not need it) */
    uint16_t *eob;           /* [n_blocks], indexed like the caller's block list */
    uint64_t *energy;        /* [n_blocks] three_quad_energy (0 unless a 64-point size) */
    /* SURVEY 8f-1, what mode decision consumes (EbFullLoop.c:763-835): */
    const void *src;         /* STAGE_RESIDUAL input: source pixels, same type and geometry as pred */
    uint64_t *dist;          /* optional output [n_blocks][2]: tuFullDistortion {DIST_CALC_RESIDUAL,
                              * DIST_CALC_PREDICTION} = picture_full_distortion32_bits over the coded
                              * region (cbf-zero rule on eob), + three_quad_energy, scaled by the tx
                              * size (EbFullLoop.c:808-835).  Flat quantizer only.  When dist is
                              * given, qcoeff and/or dqcoeff may be NULL and are then not written. */
    /* SURVEY 8f-4, the input of the coefficient-rate estimate (Av1TuEstimateCoeffBits,
     * EbRateDistortionCost.c:450): */
    uint8_t *levels;         /* optional STAGE_QUANT output: per block the padded level map exactly as
                              * av1_txb_init_levels (EbRateDistortionCost.c:125) fills `levels_buf`:
                              * TX_PAD_TOP = 2 zero rows, then h rows of {w levels min(|qcoeff|, 127),
                              * TX_PAD_HOR = 4 zeros}, then TX_PAD_BOTTOM = 4 zero rows + TX_PAD_END =
                              * 16 zeros, with w x h the coded size (64 -> 32): svtgpu_levels_size()
                              * bytes starting at levels + levels_off[i] */
    const uint32_t *levels_off; /* [n_blocks] byte offsets, multiples of 4 (required with levels) */
} SvtGpuBuffers;
/* bytes of one block's padded level map: (w + 4) * (h + 6) + 16 over the coded size */
size_t svtgpu_levels_size(int tx_size);

int svtgpu_create(SvtGpuContext **ctx, int device);
void svtgpu_destroy(SvtGpuContext *ctx);
const char *svtgpu_last_error(void);

/* tables[n_rows][5][8] int16: zbin, round, quant, quant_shift, dequant rows exactly as
 * av1_build_quantizer lays them out (EbModeDecisionConfigurationProcess.c:429; [0] DC, [1..7] AC).
 * quant_shift must be non-negative, as invert_quant produces it (1 << (16 - msb), :322-329): the
 * fused kernels multiply by quant_shift << (16 + log_scale).  A row that is not is refused with
 * SVTGPU_ERR_ARG and the tables uploaded before stay in place.  (The stand-alone quantizer slots
 * svtgpu_quantize_b* take their rows per call and accept any int16.) */
int svtgpu_upload_quant_tables(SvtGpuContext *ctx, const int16_t *tables, int n_rows);
/* wt/iwt: concatenated qmatrix arrays; off[plane>0][tx_size] = start of that size's matrix or -1
 * (what av1_qm_init binds into gqmatrix/giqmatrix, :524-554) */
int svtgpu_upload_qmatrix(SvtGpuContext *ctx, const uint8_t *wt, const uint8_t *iwt, int n_bytes,
                          const int32_t *off /* [2][19] */);

/* Validates and buckets the block list by (tx_size, tx_type) and keeps it on the device. */
int svtgpu_plan_create(SvtGpuContext *ctx, const SvtGpuBlock *blocks, int n_blocks,
                       SvtGpuPlan **plan);
void svtgpu_plan_destroy(SvtGpuPlan *plan);

/* Asynchronous on `cuda_stream` (a cudaStream_t, or NULL for the context's own stream). */
int svtgpu_run(SvtGpuContext *ctx, const SvtGpuPlan *plan, const SvtGpuBuffers *dev, int stages,
               void *cuda_stream);
int svtgpu_sync(SvtGpuContext *ctx);
/* kernels launched by this context so far */
long long svtgpu_launch_count(const SvtGpuContext *ctx);
/* kernels launched by every context of this process (levels A and B keep one per thread) */
long long svtgpu_total_launch_count(void);
/* diagnostics: buffer (re)allocations made so far by the host pipeline (0 in steady state) */
long long svtgpu_debug_host_allocs(void);

/* ------------------------------------------------------------------------------------------ */
/* batched tx-type search (ProductFullLoopTxSearch, EbFullLoop.c:927-1100)          */
/* ------------------------------------------------------------------------------------------ */
/* The reference tries the allowed tx types of a block one after another around
 * av1_estimate_transform -> av1_quantize_inv_quantize -> picture_full_distortion32_bits and keeps
 * the cheapest.  Here the candidates of a whole block list go through ONE run of the fused path
 * (STAGE_FWD | STAGE_QUANT [| STAGE_RESIDUAL], `dist` output): the caller gets {eob, distortion
 * pair} (and optionally qcoeff for its rate estimate) per candidate and applies its own cost.
 *
 * svtgpu_tx_search_mask: bit t set = the reference's loop evaluates tx type t for this size
 * (get_ext_tx_set_type / av1_ext_tx_used, EbDefinitions.h:1472-1500; allowed_tx_set_a,
 * EbFullLoop.c:883-902 when tx_search_reduced_set).
 * svtgpu_tx_search_expand: writes one SvtGpuBlock per (block, tried type) into cand[16 * n]
 * (cand index = eob / dist slot), the owning block index into cand_block, and the total
 * coefficient elements (0 unless want_coeff).  Returns the candidate count.  Host only. */
uint32_t svtgpu_tx_search_mask(int tx_size, int is_inter, int reduced_tx_set_used,
                               int tx_search_reduced_set);
int svtgpu_tx_search_expand(const SvtGpuBlock *blocks, int n_blocks, int is_inter,
                            int reduced_tx_set_used, int tx_search_reduced_set, int want_coeff,
                            SvtGpuBlock *cand, int32_t *cand_block, int64_t *coeff_elems);

/* thin memory helpers so that a C caller needs no CUDA headers */
int svtgpu_malloc(SvtGpuContext *ctx, void **dev_ptr, size_t bytes);
int svtgpu_free(SvtGpuContext *ctx, void *dev_ptr);
int svtgpu_malloc_host(SvtGpuContext *ctx, void **pinned_ptr, size_t bytes);
int svtgpu_free_host(SvtGpuContext *ctx, void *pinned_ptr);
int svtgpu_memcpy_h2d(SvtGpuContext *ctx, void *dev, const void *host, size_t bytes);
int svtgpu_memcpy_d2h(SvtGpuContext *ctx, void *host, const void *dev, size_t bytes);
int svtgpu_memset(SvtGpuContext *ctx, void *dev, int value, size_t bytes);

/* ---- host-buffer pipeline: the end-to-end call ------------------------------------------------
 * One picture (or any independent group of blocks) whose inputs and outputs live in HOST memory.
 * svtgpu_submit_host copies the inputs to the device, runs the stages and copies every output
 * back, all asynchronously on one of a small ring of internal streams, so that the upload of
 * picture i+1, the kernels of picture i and the download of picture i-1 overlap (separate copy
 * engines per direction).  The caller's buffers must stay valid and untouched until
 * svtgpu_wait_host returns; they should be page-locked (svtgpu_malloc_host) -- pageable memory
 * works but serialises the copies.  Offsets in `blocks` are relative to the pointers given here.
 * This is the batched counterpart of calling av1_estimate_transform + av1_quantize_inv_quantize +
 * av1_inv_transform_recon for every block of the picture (EbCodingLoop.c:655-731). */
typedef struct SvtGpuHostFrame {
    const SvtGpuBlock *blocks;
    int n_blocks;
    int bit_depth;           /* 8: pixels are uint8; 10: uint16 */
    const int16_t *residual; /* residual_elems int16 */
    size_t residual_elems;
    const void *pred;        /* pixel_elems pixels */
    void *recon;             /* pixel_elems pixels out (may equal pred) */
    size_t pixel_elems;
    int32_t *qcoeff;         /* coeff_elems int32 out */
    int32_t *dqcoeff;        /* coeff_elems int32 out */
    size_t coeff_elems;
    uint16_t *eob;           /* n_blocks out */
    uint64_t *energy;        /* n_blocks out, or NULL */
    const void *src;         /* SVTGPU_STAGE_RESIDUAL: pixel_elems source pixels (residual unused) */
    uint64_t *dist;          /* optional n_blocks x 2 out (see SvtGpuBuffers.dist); qcoeff / dqcoeff
                              * may then be NULL: nothing is computed for or copied into them.
                              * dqcoeff alone may also be NULL (what the encode pass consumes is
                              * qcoeff + eob + recon: EbCodingLoop.c:697-731 feeds recon_coeff only
                              * to the inverse transform, which runs here anyway) */
    const SvtGpuPlan *plan;  /* optional: a plan made once with svtgpu_plan_create for this exact block
                              * list (same context).  `blocks` is then ignored: nothing is validated,
                              * sorted or uploaded per submission -- the per-picture host cost of a
                              * caller whose partitioning repeats (a fixed geometry, a benchmark) */
} SvtGpuHostFrame;
/* returns immediately after queueing; *ticket (optional) identifies the submission */
int svtgpu_submit_host(SvtGpuContext *ctx, const SvtGpuHostFrame *frame, int stages, int *ticket);
/* blocks until submission `ticket` (or all of them when ticket < 0) has fully landed in host memory */
int svtgpu_wait_host(SvtGpuContext *ctx, int ticket);

/* ---- several GPUs from ONE process (svtgpu_create(device_ids[], n)) -------------------
 * For a C caller that cannot run one process per GPU (the encoder).  Pictures are independent on
 * this path, so submission i runs on device i mod n (svtgpu_multi_device_of): whole pictures per
 * device, no collective.  Tables are uploaded to every device; svtgpu_multi_submit_host /
 * svtgpu_multi_wait_host are svtgpu_submit_host / svtgpu_wait_host of the chosen device's context.
 * *ticket = the global submission index.  frame->plan, if given, must have been made on
 * svtgpu_multi_context(m, svtgpu_multi_device_of(n, index of this submission)).  The same device
 * id may be listed more than once (two contexts, twice the lanes, on one GPU). */
#define SVTGPU_MAX_DEVICES 16
typedef struct SvtGpuMulti SvtGpuMulti;
int svtgpu_create_multi(SvtGpuMulti **multi, const int *device_ids, int n_devices);
void svtgpu_multi_destroy(SvtGpuMulti *multi);
int svtgpu_multi_count(const SvtGpuMulti *multi);
SvtGpuContext *svtgpu_multi_context(SvtGpuMulti *multi, int i);
int svtgpu_multi_device_of(int n_devices, long long submission_index); /* host only, no GPU needed */
int svtgpu_multi_upload_quant_tables(SvtGpuMulti *multi, const int16_t *tables, int n_rows);
int svtgpu_multi_upload_qmatrix(SvtGpuMulti *multi, const uint8_t *wt, const uint8_t *iwt, int n_bytes,
                                const int32_t *off /* [2][19] */);
int svtgpu_multi_submit_host(SvtGpuMulti *multi, const SvtGpuHostFrame *frame, int stages,
                             long long *ticket);
/* ticket < 0: every submission on every device; otherwise everything queued so far on the device
 * that submission `ticket` went to */
int svtgpu_multi_wait_host(SvtGpuMulti *multi, long long ticket);

/* ---- allocation-free quantizer ---------------------------------------------------------------
 * No counterpart in the reference: this entry point was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same arithmetic and results as the level-A slots
 * svtgpu_aom_highbd_quantize_b / _32x32 / _64x64 (log_scale 0 / 1 / 2), but:
 *   - d_coeff, d_qcoeff, d_dqcoeff, d_eob (one uint16) and d_iscan are DEVICE pointers of the
 *     current device; the five table rows are host pointers ([0] DC, [1] AC are read, by value);
 *   - asynchronous on `cuda_stream` (a cudaStream_t; NULL = the default stream);
 *   - it calls no cudaMalloc / cudaFree / allocating routine: its only working memory is a 256-byte
 *     file-scope __device__ array (csrc/svt_gpu_noalloc.cu) that every GPU loading the library
 *     holds until exit, whose address is resolved once per device on the first call there (make
 *     that first call outside graph capture);
 *   - n_coeffs above SVTGPU_NOALLOC_MAX_COEFFS is rejected with SVTGPU_ERR_ARG and a message in
 *     svtgpu_last_error();
 *   - every call on a GPU shares that array: calls that overlap on one GPU (different streams or
 *     threads) race on it and must be serialised by the caller. */
#define SVTGPU_NOALLOC_MAX_COEFFS 32768
int svtgpu_quantize_b_noalloc(const int32_t *d_coeff, intptr_t n_coeffs, int log_scale,
                              const int16_t *zbin, const int16_t *round, const int16_t *quant,
                              const int16_t *quant_shift, const int16_t *dequant,
                              int32_t *d_qcoeff, int32_t *d_dqcoeff, uint16_t *d_eob,
                              const int16_t *d_iscan, void *cuda_stream);

/* scan tables the library generates (av1_scan_orders, EbTransforms.h:3349): for tests */
int svtgpu_get_iscan(int tx_size, int scan_class, int16_t *iscan_out /* max_eob entries */);

/* ------------------------------------------------------------------------------------------ */
/* Level B: entry points (host pointers, synchronous)                                          */
/* ------------------------------------------------------------------------------------------ */
/* av1_estimate_transform, EbTransforms.h:3896 / EbTransforms.c:4918 */
int svtgpu_av1_estimate_transform(int16_t *residual_buffer, uint32_t residual_stride,
                                  int32_t *coeff_buffer, uint32_t coeff_stride,
                                  uint8_t transform_size, uint64_t *three_quad_energy,
                                  int16_t *transform_inner_array_ptr, uint32_t bit_increment,
                                  uint8_t transform_type, int asm_type, int component_type,
                                  int trans_coeff_shape);
/* av1_quantize_inv_quantize, EbTransforms.h:3910 / EbFullLoop.c:653, with the table rows that
 * av1_quantize_inv_quantize_ii selects from the PictureControlSet (:548-613) passed directly */
void svtgpu_av1_quantize_inv_quantize(const int16_t *zbin, const int16_t *round,
                                      const int16_t *quant, const int16_t *quant_shift,
                                      const int16_t *dequant, int32_t *coeff, int32_t *quant_coeff,
                                      int32_t *recon_coeff, uint8_t txsize, uint16_t *eob,
                                      uint32_t *y_count_non_zero_coeffs, uint32_t bit_increment,
                                      uint8_t tx_type, int clamp16);
/* av1_inv_transform_recon / av1_inv_transform_recon8bit, EbTransforms.h:3967, :3977 */
int svtgpu_av1_inv_transform_recon(int32_t *coeff_buffer, uint8_t *recon_buffer,
                                   uint32_t recon_stride, uint8_t txsize, uint32_t bit_increment,
                                   uint8_t transform_type, int component_type, uint32_t eob);
int svtgpu_av1_inv_transform_recon8bit(int32_t *coeff_buffer, uint8_t *recon_buffer,
                                       uint32_t recon_stride, uint8_t txsize,
                                       uint8_t transform_type, int component_type, uint32_t eob);

/* The per-type loop of encode_pass_tx_search (EbFullLoop.c:1127-1318) and ProductFullLoopTxSearch
 * (:927-1100) for ONE transform block: the residual is staged once and every candidate type in
 * tx_types[0 .. n_types) (<= 16, each legal for tx_size) goes through one launch of forward
 * transform + quantiser + transform-domain distortion.  Outputs per candidate k:
 * quant_coeff[k * max_eob .. ), eob[k], dist[2k] / dist[2k+1] = tuFullDistortion
 * {DIST_CALC_RESIDUAL, DIST_CALC_PREDICTION} after the energy term and the tx-scale shift
 * (:1231-1257), three_quad_energy[k] (optional).  The caller keeps the rate estimate and the cost
 * (Av1TuEstimateCoeffBits, av1_tu_calc_cost_luma) and applies them candidate by candidate. */
int svtgpu_tx_search_block(const int16_t *residual, uint32_t residual_stride, uint8_t tx_size,
                           const uint8_t *tx_types, int n_types, const int16_t *zbin,
                           const int16_t *round, const int16_t *quant, const int16_t *quant_shift,
                           const int16_t *dequant, uint32_t bit_increment, int32_t *quant_coeff,
                           uint16_t *eob, uint64_t *dist, uint64_t *three_quad_energy);

/* ------------------------------------------------------------------------------------------ */
/* Level A: RTCD slots (aom_dsp_rtcd.h)                                                        */
/* ------------------------------------------------------------------------------------------ */
#define SVTGPU_FWD_DECL(WxH)                                                                  \
    void svtgpu_av1_fwd_txfm2d_##WxH(int16_t *input, int32_t *output, uint32_t input_stride, \
                                     uint8_t transform_type, uint8_t bit_depth);
/* aom_dsp_rtcd.h:158-255 */
SVTGPU_FWD_DECL(4x4) SVTGPU_FWD_DECL(8x8) SVTGPU_FWD_DECL(16x16) SVTGPU_FWD_DECL(32x32)
SVTGPU_FWD_DECL(64x64) SVTGPU_FWD_DECL(4x8) SVTGPU_FWD_DECL(8x4) SVTGPU_FWD_DECL(8x16)
SVTGPU_FWD_DECL(16x8) SVTGPU_FWD_DECL(16x32) SVTGPU_FWD_DECL(32x16) SVTGPU_FWD_DECL(32x64)
SVTGPU_FWD_DECL(64x32) SVTGPU_FWD_DECL(4x16) SVTGPU_FWD_DECL(16x4) SVTGPU_FWD_DECL(8x32)
SVTGPU_FWD_DECL(32x8) SVTGPU_FWD_DECL(16x64) SVTGPU_FWD_DECL(64x16)
#undef SVTGPU_FWD_DECL

/* aom_dsp_rtcd.h:321-342 (the 8-bit and highbd slots share one prototype) */
#define SVTGPU_QUANT_DECL(name)                                                                  \
    void name(const int32_t *coeff_ptr, intptr_t n_coeffs, int32_t skip_block,                  \
              const int16_t *zbin_ptr, const int16_t *round_ptr, const int16_t *quant_ptr,      \
              const int16_t *quant_shift_ptr, int32_t *qcoeff_ptr, int32_t *dqcoeff_ptr,        \
              const int16_t *dequant_ptr, uint16_t *eob_ptr, const int16_t *scan,               \
              const int16_t *iscan);
SVTGPU_QUANT_DECL(svtgpu_aom_quantize_b) SVTGPU_QUANT_DECL(svtgpu_aom_quantize_b_32x32)
SVTGPU_QUANT_DECL(svtgpu_aom_quantize_b_64x64) SVTGPU_QUANT_DECL(svtgpu_aom_highbd_quantize_b)
SVTGPU_QUANT_DECL(svtgpu_aom_highbd_quantize_b_32x32)
SVTGPU_QUANT_DECL(svtgpu_aom_highbd_quantize_b_64x64)
#undef SVTGPU_QUANT_DECL

/* aom_dsp_rtcd.h:347-417: three prototype families */
#define SVTGPU_INV_SQ_DECL(WxH)                                                          \
    void svtgpu_av1_inv_txfm2d_add_##WxH(const int32_t *input, uint16_t *output,        \
                                         int32_t stride, uint8_t tx_type, int32_t bd);
#define SVTGPU_INV_RECT_DECL(WxH)                                                        \
    void svtgpu_av1_inv_txfm2d_add_##WxH(const int32_t *input, uint16_t *output,        \
                                         int32_t stride, uint8_t tx_type, uint8_t tx_size, \
                                         int32_t eob, int32_t bd);
#define SVTGPU_INV_4N_DECL(WxH)                                                          \
    void svtgpu_av1_inv_txfm2d_add_##WxH(const int32_t *input, uint16_t *output,        \
                                         int32_t stride, uint8_t tx_type, uint8_t tx_size, \
                                         int32_t bd);
SVTGPU_INV_SQ_DECL(4x4) SVTGPU_INV_SQ_DECL(8x8) SVTGPU_INV_SQ_DECL(16x16)
SVTGPU_INV_SQ_DECL(32x32) SVTGPU_INV_SQ_DECL(64x64)
SVTGPU_INV_RECT_DECL(8x16) SVTGPU_INV_RECT_DECL(16x8) SVTGPU_INV_RECT_DECL(16x32)
SVTGPU_INV_RECT_DECL(32x16) SVTGPU_INV_RECT_DECL(32x8) SVTGPU_INV_RECT_DECL(8x32)
SVTGPU_INV_RECT_DECL(32x64) SVTGPU_INV_RECT_DECL(64x32) SVTGPU_INV_RECT_DECL(16x64)
SVTGPU_INV_RECT_DECL(64x16)
SVTGPU_INV_4N_DECL(4x8) SVTGPU_INV_4N_DECL(8x4) SVTGPU_INV_4N_DECL(4x16) SVTGPU_INV_4N_DECL(16x4)
#undef SVTGPU_INV_SQ_DECL
#undef SVTGPU_INV_RECT_DECL
#undef SVTGPU_INV_4N_DECL

/* layout-compatible with TxfmParam, EbDefinitions.h:764-776 */
typedef struct SvtGpuTxfmParam {
    uint8_t tx_type;
    uint8_t tx_size;
    int32_t lossless;
    int32_t bd;
    int32_t is_hbd;
    uint8_t tx_set_type;
    int32_t eob;
} SvtGpuTxfmParam;
/* av1_inv_txfm_add (8-bit pixels), aom_dsp_rtcd.h:421-423 */
void svtgpu_av1_inv_txfm_add(const int32_t *dqcoeff, uint8_t *dst, int32_t stride,
                             const SvtGpuTxfmParam *txfm_param);

/* Fills a table of {slot name, function} pairs for the installer shown in INTEGRATION.md. */
typedef struct SvtGpuRtcdEntry { const char *slot; void *fn; } SvtGpuRtcdEntry;
int svtgpu_rtcd_table(const SvtGpuRtcdEntry **entries);

#ifdef __cplusplus
}
#endif
#endif /* SVT_GPU_H */
