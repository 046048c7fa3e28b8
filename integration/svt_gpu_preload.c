/*
 * svt_gpu_preload.c -- LD_PRELOAD installer that puts libsvtgpu.so behind an UNMODIFIED
 * SvtAv1EncApp / libSvtAv1Enc.so (INTEGRATION.md sections 1 and 2).
 *
 *   LD_PRELOAD=libsvtgpu_preload.so SVT_GPU_LIB=/path/libsvtgpu.so SvtAv1EncApp -i ... -b out.ivf
 *
 * Level A (always; needs no reference header): eb_init_encoder is interposed; after the real one
 * has run setup_rtcd_internal (Source/Lib/Encoder/Codec/EbEncHandle.c:917, before any worker
 * thread exists, :1796) every RTCD pointer listed by svtgpu_rtcd_table() is overwritten with its
 * svtgpu_ twin (the pointers are exported globals of the encoder library,
 * Source/Lib/Common/Codec/aom_dsp_rtcd.h:158-423).
 *
 * Level B (built with -DSVT_GPU_LEVEL_B against the reference headers, because
 * av1_quantize_inv_quantize takes a PictureControlSet): the four entry points
 * av1_estimate_transform / av1_quantize_inv_quantize / av1_inv_transform_recon[8bit]
 * (Source/Lib/Common/Codec/EbTransforms.h:3896-3987) are interposed as symbols -- the encoder
 * library is -fPIC without -Bsymbolic, so its own calls go through the PLT -- and forwarded to
 * the svtgpu_ entry points.  This also covers the (size, type) pairs for which
 * av1_estimate_transform hard-wires the _c transform and never consults the RTCD pointer
 * (EbTransforms.c:4951-5235).
 *
 * SVT_GPU_LEVEL = "A", "B" or "AB" (default "AB" when level B is compiled in, else "A").
 * There is no fallback: if the library or a symbol is missing the process aborts.
 */
#define _GNU_SOURCE
#include <dlfcn.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#ifdef SVT_GPU_LEVEL_B
#include "EbDefinitions.h"
#include "EbPictureControlSet.h"
#include "EbTransforms.h"
#include "EbSvtAv1Enc.h"
#include "EbEncDecProcess.h"
#include "EbEntropyCoding.h"
#include "EbRateDistortionCost.h"
#endif
#include "svt_gpu.h"

static void *g_lib;
static int g_level_a, g_level_b, g_level_s;
static long long g_calls[6]; /* estimate_transform, quantize_inv_quantize, recon, recon8bit, tx searches, candidates */

static void die(const char *what, const char *arg) {
    fprintf(stderr, "[svtgpu-preload] %s %s\n", what, arg ? arg : "");
    abort();
}
static void *need(const char *name) {
    void *p = dlsym(g_lib, name);
    if (!p) die("libsvtgpu.so does not export", name);
    return p;
}

static int (*p_estimate)(int16_t *, uint32_t, int32_t *, uint32_t, uint8_t, uint64_t *, int16_t *,
                         uint32_t, uint8_t, int, int, int);
static void (*p_quant)(const int16_t *, const int16_t *, const int16_t *, const int16_t *,
                       const int16_t *, int32_t *, int32_t *, int32_t *, uint8_t, uint16_t *,
                       uint32_t *, uint32_t, uint8_t, int);
static int (*p_recon)(int32_t *, uint8_t *, uint32_t, uint8_t, uint32_t, uint8_t, int, uint32_t);
static int (*p_recon8)(int32_t *, uint8_t *, uint32_t, uint8_t, uint8_t, int, uint32_t);
static int (*p_txsearch)(const int16_t *, uint32_t, uint8_t, const uint8_t *, int, const int16_t *, const int16_t *,
                         const int16_t *, const int16_t *, const int16_t *, uint32_t, int32_t *, uint16_t *,
                         uint64_t *, uint64_t *);

static long long (*p_launches)(void);
static void report(void) {
    if (!g_lib) return;
    fprintf(stderr, "[svtgpu-preload] kernels launched by libsvtgpu.so: %lld; level-B calls served: "
            "estimate_transform %lld, quantize_inv_quantize %lld, inv_transform_recon %lld, "
            "inv_transform_recon8bit %lld; batched tx-type searches %lld (%lld candidates)\n",
            p_launches ? p_launches() : -1, g_calls[0], g_calls[1], g_calls[2], g_calls[3], g_calls[4], g_calls[5]);
}

static void install(void) {
    const char *path = getenv("SVT_GPU_LIB");
    const char *lvl = getenv("SVT_GPU_LEVEL");
#ifdef SVT_GPU_LEVEL_B
    if (!lvl) lvl = "ABS";
#else
    if (!lvl) lvl = "A";
#endif
    g_lib = dlopen(path ? path : "libsvtgpu.so", RTLD_NOW | RTLD_GLOBAL);
    if (!g_lib) die("cannot load libsvtgpu.so:", dlerror());
    p_launches = (long long (*)(void))need("svtgpu_total_launch_count");
    if (strchr(lvl, 'A')) {
        int (*table)(const SvtGpuRtcdEntry **) = (int (*)(const SvtGpuRtcdEntry **))need("svtgpu_rtcd_table");
        const SvtGpuRtcdEntry *e = NULL;
        const int n = table(&e);
        for (int i = 0; i < n; ++i) {
            void **slot = (void **)dlsym(RTLD_DEFAULT, e[i].slot);
            if (!slot) die("the encoder does not export the RTCD pointer", e[i].slot);
            *slot = e[i].fn;
        }
        g_level_a = n;
    }
    if (strchr(lvl, 'B')) {
#ifdef SVT_GPU_LEVEL_B
        p_estimate = need("svtgpu_av1_estimate_transform");
        p_quant = need("svtgpu_av1_quantize_inv_quantize");
        p_recon = need("svtgpu_av1_inv_transform_recon");
        p_recon8 = need("svtgpu_av1_inv_transform_recon8bit");
        g_level_b = 1;
        if (strchr(lvl, 'S')) {
            p_txsearch = need("svtgpu_tx_search_block");
            g_level_s = 1;
        }
#else
        die("built without -DSVT_GPU_LEVEL_B; SVT_GPU_LEVEL =", lvl);
#endif
    }
    fprintf(stderr, "[svtgpu-preload] %d RTCD slots rebound to libsvtgpu.so, level-B entry points %s, tx-type search %s\n",
            g_level_a, g_level_b ? "interposed" : "left to the encoder",
            g_level_s ? "batched (one launch per block)" : "left to the encoder");
}

/* EbSvtAv1Enc.h:439 -- EbErrorType eb_init_encoder(EbComponentType *svt_enc_component) */
#ifdef SVT_GPU_LEVEL_B
EbErrorType eb_init_encoder(EbComponentType *svt_enc_component) {
#else
int eb_init_encoder(void *svt_enc_component) {
#endif
    int (*real)(void *) = (int (*)(void *))dlsym(RTLD_NEXT, "eb_init_encoder");
    if (!real) die("no eb_init_encoder behind the preload", NULL);
    const int rc = real(svt_enc_component);
    if (rc == 0 && !g_lib) install();
    return rc;
}

/* EbSvtAv1Enc.h -- EbErrorType eb_deinit_encoder(EbComponentType *): the application closes its
 * log files right after this call, so the usage report is printed here */
#ifdef SVT_GPU_LEVEL_B
EbErrorType eb_deinit_encoder(EbComponentType *svt_enc_component) {
#else
int eb_deinit_encoder(void *svt_enc_component) {
#endif
    int (*real)(void *) = (int (*)(void *))dlsym(RTLD_NEXT, "eb_deinit_encoder");
    if (!real) die("no eb_deinit_encoder behind the preload", NULL);
    report();
    return real(svt_enc_component);
}

#ifdef SVT_GPU_LEVEL_B
EbErrorType av1_estimate_transform(int16_t *residual_buffer, uint32_t residual_stride,
                                   int32_t *coeff_buffer, uint32_t coeff_stride, TxSize transform_size,
                                   uint64_t *three_quad_energy, int16_t *transform_inner_array_ptr,
                                   uint32_t bit_increment, TxType transform_type, EbAsm asm_type,
                                   PLANE_TYPE component_type, EB_TRANS_COEFF_SHAPE trans_coeff_shape) {
    if (!g_level_b) {
        static EbErrorType (*real)(int16_t *, uint32_t, int32_t *, uint32_t, TxSize, uint64_t *, int16_t *,
                                   uint32_t, TxType, EbAsm, PLANE_TYPE, EB_TRANS_COEFF_SHAPE);
        if (!real) real = dlsym(RTLD_NEXT, "av1_estimate_transform");
        return real(residual_buffer, residual_stride, coeff_buffer, coeff_stride, transform_size,
                    three_quad_energy, transform_inner_array_ptr, bit_increment, transform_type,
                    asm_type, component_type, trans_coeff_shape);
    }
    __atomic_fetch_add(&g_calls[0], 1, __ATOMIC_RELAXED);
    return (EbErrorType)p_estimate(residual_buffer, residual_stride, coeff_buffer, coeff_stride,
                                   transform_size, three_quad_energy, transform_inner_array_ptr,
                                   bit_increment, transform_type, asm_type, component_type,
                                   trans_coeff_shape);
}

void av1_quantize_inv_quantize(PictureControlSet_t *picture_control_set_ptr, int32_t *coeff,
                               const uint32_t coeff_stride, int32_t *quant_coeff, int32_t *recon_coeff,
                               uint32_t qp, uint32_t width, uint32_t height, TxSize txsize, uint16_t *eob,
                               MacroblockPlane candidate_plane, EbAsm asm_type,
                               uint32_t *y_count_non_zero_coeffs,
#if !PF_N2_32X32
                               EbPfMode pf_mode,
#endif
                               uint8_t enable_contouring_qc_update_flag, uint32_t component_type,
                               uint32_t bit_increment, TxType tx_type, EbBool clean_sparse_coeff_flag) {
    if (!g_level_b) {
        static void (*real)(PictureControlSet_t *, int32_t *, const uint32_t, int32_t *, int32_t *, uint32_t,
                            uint32_t, uint32_t, TxSize, uint16_t *, MacroblockPlane, EbAsm, uint32_t *,
#if !PF_N2_32X32
                            EbPfMode,
#endif
                            uint8_t, uint32_t, uint32_t, TxType, EbBool);
        if (!real) real = dlsym(RTLD_NEXT, "av1_quantize_inv_quantize");
        real(picture_control_set_ptr, coeff, coeff_stride, quant_coeff, recon_coeff, qp, width, height,
             txsize, eob, candidate_plane, asm_type, y_count_non_zero_coeffs,
#if !PF_N2_32X32
             pf_mode,
#endif
             enable_contouring_qc_update_flag, component_type, bit_increment, tx_type,
             clean_sparse_coeff_flag);
        return;
    }
    /* The accessor: the only thing the reference needs the PictureControlSet for
     * (av1_quantize_inv_quantize_ii, EbFullLoop.c:548-613): with bit_increment == 0 all seven rows
     * of the plane come from quantsMd / deqMd (the 8-bit tables mode decision uses, also inside a
     * 10-bit encode: EbFullLoop.c:1637 calls chroma with BIT_INCREMENT_8BIT), otherwise from
     * quants / deq (built at encoder_bit_depth, EbModeDecisionConfigurationProcess.c:2176-2202). */
    const PictureParentControlSet_t *p = picture_control_set_ptr->parent_pcs_ptr;
    const uint32_t q = p->base_qindex;
    const Quants *qs = bit_increment ? &p->quants : &p->quantsMd;
    const Dequants *dq = bit_increment ? &p->deq : &p->deqMd;
    const int16_t *zbin, *round, *quant, *shift, *deq;
    if (component_type == COMPONENT_LUMA) {
        zbin = qs->y_zbin[q]; round = qs->y_round[q]; quant = qs->y_quant[q];
        shift = qs->y_quant_shift[q]; deq = dq->y_dequant_QTX[q];
    } else if (component_type == COMPONENT_CHROMA_CB) {
        zbin = qs->u_zbin[q]; round = qs->u_round[q]; quant = qs->u_quant[q];
        shift = qs->u_quant_shift[q]; deq = dq->u_dequant_QTX[q];
    } else {
        zbin = qs->v_zbin[q]; round = qs->v_round[q]; quant = qs->v_quant[q];
        shift = qs->v_quant_shift[q]; deq = dq->v_dequant_QTX[q];
    }
    __atomic_fetch_add(&g_calls[1], 1, __ATOMIC_RELAXED);
    p_quant(zbin, round, quant, shift, deq, coeff, quant_coeff, recon_coeff, txsize, eob,
            y_count_non_zero_coeffs, bit_increment, tx_type, /*clamp16=*/0);
}

EbErrorType av1_inv_transform_recon(int32_t *coeff_buffer, uint8_t *recon_buffer, uint32_t recon_stride,
                                    TxSize txsize, uint32_t bit_increment, TxType transform_type,
                                    PLANE_TYPE component_type, uint32_t eob) {
    if (!g_level_b) {
        static EbErrorType (*real)(int32_t *, uint8_t *, uint32_t, TxSize, uint32_t, TxType, PLANE_TYPE, uint32_t);
        if (!real) real = dlsym(RTLD_NEXT, "av1_inv_transform_recon");
        return real(coeff_buffer, recon_buffer, recon_stride, txsize, bit_increment, transform_type,
                    component_type, eob);
    }
    __atomic_fetch_add(&g_calls[2], 1, __ATOMIC_RELAXED);
    return (EbErrorType)p_recon(coeff_buffer, recon_buffer, recon_stride, txsize, bit_increment,
                                transform_type, component_type, eob);
}

EbErrorType av1_inv_transform_recon8bit(int32_t *coeff_buffer, uint8_t *recon_buffer, uint32_t recon_stride,
                                        TxSize txsize, TxType transform_type, PLANE_TYPE component_type,
                                        uint32_t eob) {
    if (!g_level_b) {
        static EbErrorType (*real)(int32_t *, uint8_t *, uint32_t, TxSize, TxType, PLANE_TYPE, uint32_t);
        if (!real) real = dlsym(RTLD_NEXT, "av1_inv_transform_recon8bit");
        return real(coeff_buffer, recon_buffer, recon_stride, txsize, transform_type, component_type, eob);
    }
    __atomic_fetch_add(&g_calls[3], 1, __ATOMIC_RELAXED);
    return (EbErrorType)p_recon8(coeff_buffer, recon_buffer, recon_stride, txsize, transform_type,
                                 component_type, eob);
}
/* SURVEY 8f-2 wired in: encode_pass_tx_search (EbFullLoop.c:1127-1318), the tx-type search the
 * encode pass runs per luma transform block at enc-mode >= 6 (TX_SEARCH_ENC_DEC,
 * EbPictureDecisionProcess.c:944-960).  The reference walks the candidate types one after another
 * around av1_estimate_transform -> av1_quantize_inv_quantize -> picture_full_distortion32_bits and
 * prices each with Av1TuEstimateCoeffBits + av1_tu_calc_cost_luma.  Here the candidate list is
 * built with the reference's own skip rules, ALL candidates go through one
 * svtgpu_tx_search_block launch, and the reference's rate estimate and cost are then applied to
 * them in the same order -- same winner, 1 launch instead of 2 per candidate.  Everything the loop
 * leaves behind for later code is reproduced: txb_ptr->transform_type, the candidate buffer
 * fields, eob[0] and the quantised coefficients of the last candidate tried (Av1EncodeLoop
 * re-transforms the block with the winning type right after, EbCodingLoop.c:655-731).
 * The 16-bit twin encode_pass_tx_search_hbd is left to the encoder. */
extern uint8_t allowed_tx_set_a[TX_SIZES_ALL][TX_TYPES];
void encode_pass_tx_search(PictureControlSet_t *picture_control_set_ptr, EncDecContext_t *context_ptr,
                           LargestCodingUnit_t *sb_ptr, uint32_t cbQp, EbPictureBufferDesc_t *coeffSamplesTB,
                           EbPictureBufferDesc_t *residual16bit, EbPictureBufferDesc_t *transform16bit,
                           EbPictureBufferDesc_t *inverse_quant_buffer, int16_t *transformScratchBuffer,
                           EbAsm asm_type, uint32_t *count_non_zero_coeffs, uint32_t component_mask,
                           uint32_t use_delta_qp, uint32_t dZoffset, uint16_t *eob,
                           MacroblockPlane *candidate_plane) {
    if (!g_level_s) {
        static void (*real)(PictureControlSet_t *, EncDecContext_t *, LargestCodingUnit_t *, uint32_t,
                            EbPictureBufferDesc_t *, EbPictureBufferDesc_t *, EbPictureBufferDesc_t *,
                            EbPictureBufferDesc_t *, int16_t *, EbAsm, uint32_t *, uint32_t, uint32_t, uint32_t,
                            uint16_t *, MacroblockPlane *);
        if (!real) real = dlsym(RTLD_NEXT, "encode_pass_tx_search");
        real(picture_control_set_ptr, context_ptr, sb_ptr, cbQp, coeffSamplesTB, residual16bit, transform16bit,
             inverse_quant_buffer, transformScratchBuffer, asm_type, count_non_zero_coeffs, component_mask,
             use_delta_qp, dZoffset, eob, candidate_plane);
        return;
    }
    CodingUnit_t *cu_ptr = context_ptr->cu_ptr;
    TransformUnit *txb_ptr = &cu_ptr->transform_unit_array[context_ptr->txb_itr];
    const uint32_t scratchLumaOffset = context_ptr->blk_geom->tx_org_x[context_ptr->txb_itr] +
                                       context_ptr->blk_geom->tx_org_y[context_ptr->txb_itr] * SB_STRIDE_Y;
    const uint32_t coeff1dOffset = context_ptr->coded_area_sb;
    const int32_t is_inter = context_ptr->is_inter;
    const TxSize txSize = context_ptr->blk_geom->txsize[context_ptr->txb_itr];
    const PictureParentControlSet_t *ppcs = picture_control_set_ptr->parent_pcs_ptr;
    const TxSetType tx_set_type = get_ext_tx_set_type(txSize, is_inter, ppcs->reduced_tx_set_used);

    /* the candidate list, by the loop's own rules (EbFullLoop.c:1174-1186) */
    uint8_t types[TX_TYPES];
    int n_types = 0;
    for (int32_t t = DCT_DCT; t < TX_TYPES; ++t) {
        if (ppcs->tx_search_reduced_set)
            if (!allowed_tx_set_a[txSize][t]) continue;
        const int32_t eset = get_ext_tx_set(txSize, is_inter, ppcs->reduced_tx_set_used);
        if (eset <= 0) continue;
        if (av1_ext_tx_used[tx_set_type][t] == 0) continue;
        types[n_types++] = (uint8_t)t;
    }
    TxType best_tx_type = DCT_DCT;
    if (n_types > 0) {
        /* BIT_INCREMENT_8BIT, COMPONENT_LUMA: the *Md tables (EbFullLoop.c:548-557) */
        const uint32_t q = ppcs->base_qindex;
        static __thread int32_t qc[TX_TYPES * 1024];
        uint16_t eobs[TX_TYPES];
        uint64_t dist[TX_TYPES][2], energy[TX_TYPES];
        const int n = av1_get_max_eob(txSize);
        if (p_txsearch(((int16_t *)residual16bit->buffer_y) + scratchLumaOffset, residual16bit->stride_y, (uint8_t)txSize,
                       types, n_types, ppcs->quantsMd.y_zbin[q], ppcs->quantsMd.y_round[q], ppcs->quantsMd.y_quant[q],
                       ppcs->quantsMd.y_quant_shift[q], ppcs->deqMd.y_dequant_QTX[q], 0, qc, eobs, &dist[0][0],
                       energy) != 0)
            die("svtgpu_tx_search_block failed", NULL);
        __atomic_fetch_add(&g_calls[4], 1, __ATOMIC_RELAXED);
        __atomic_fetch_add(&g_calls[5], n_types, __ATOMIC_RELAXED);
        uint64_t bestFullCost = UINT64_MAX, y_full_cost, y_tu_coeff_bits;
        int32_t *qdst = ((int32_t *)coeffSamplesTB->buffer_y) + coeff1dOffset;
        for (int k = 0; k < n_types; ++k) {
            const TxType tx_type = (TxType)types[k];
            context_ptr->three_quad_energy = energy[k];
            y_tu_coeff_bits = 0;
            memcpy(qdst, qc + (size_t)k * n, (size_t)n * sizeof(int32_t)); /* what the quantiser would have left */
            eob[0] = eobs[k];
            const uint32_t yCountNonZeroCoeffsTemp = eobs[k];
            if (yCountNonZeroCoeffsTemp == 0 && tx_type != DCT_DCT) continue;
            txb_ptr->transform_type[PLANE_TYPE_Y] = tx_type;
            ModeDecisionCandidateBuffer_t *candidateBuffer = context_ptr->md_context->candidate_buffer_ptr_array[0];
            candidateBuffer->candidate_ptr->transform_type[PLANE_TYPE_Y] = cu_ptr->transform_unit_array[context_ptr->txb_itr].transform_type[PLANE_TYPE_Y];
            candidateBuffer->candidate_ptr->transform_type[PLANE_TYPE_UV] = cu_ptr->transform_unit_array[context_ptr->txb_itr].transform_type[PLANE_TYPE_UV];
            candidateBuffer->candidate_ptr->type = cu_ptr->prediction_mode_flag;
            candidateBuffer->candidate_ptr->pred_mode = cu_ptr->pred_mode;
            Av1TuEstimateCoeffBits(picture_control_set_ptr, candidateBuffer, context_ptr->cu_ptr, coeff1dOffset, 0,
                                   picture_control_set_ptr->coeff_est_entropy_coder_ptr, coeffSamplesTB,
                                   yCountNonZeroCoeffsTemp, 0, 0, &y_tu_coeff_bits, &y_tu_coeff_bits, &y_tu_coeff_bits,
                                   context_ptr->blk_geom->txsize[context_ptr->txb_itr],
                                   context_ptr->blk_geom->txsize_uv[context_ptr->txb_itr], COMPONENT_LUMA, asm_type);
            av1_tu_calc_cost_luma(context_ptr->cu_ptr->luma_txb_skip_context, candidateBuffer->candidate_ptr,
                                  context_ptr->txb_itr, context_ptr->blk_geom->txsize[context_ptr->txb_itr],
                                  yCountNonZeroCoeffsTemp, dist[k], &y_tu_coeff_bits, &y_full_cost,
                                  context_ptr->full_lambda);
            if (y_full_cost < bestFullCost) {
                bestFullCost = y_full_cost;
                best_tx_type = tx_type;
            }
        }
    }
    txb_ptr->transform_type[PLANE_TYPE_Y] = best_tx_type;
    /* For Inter blocks, transform type of chroma follows luma transfrom type */
    if (is_inter) txb_ptr->transform_type[PLANE_TYPE_UV] = txb_ptr->transform_type[PLANE_TYPE_Y];
    (void)sb_ptr; (void)cbQp; (void)transform16bit; (void)inverse_quant_buffer; (void)transformScratchBuffer;
    (void)count_non_zero_coeffs; (void)component_mask; (void)use_delta_qp; (void)dZoffset; (void)candidate_plane;
}
#endif /* SVT_GPU_LEVEL_B */
