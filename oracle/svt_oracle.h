/*
 * TEST INFRASTRUCTURE ONLY -- CPU restatement (plain C) of the reference's residual-coding path.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this; the product (cidana_svt_av1_b200) never does.
 *
 * Parity: PINNED.  Every function below is checked bit-for-bit against the reference itself
 * compiled from /root/reference (oracle/_ref/libsvtref.so, built by oracle/Makefile `ref`) in
 * tests/test_oracle_vs_ref.py, and against fixtures under tests/golden/ generated from that
 * build by tests/golden/make_golden.py.
 *
 * file:line citations are relative to /root/reference/Source/Lib/Common/Codec/.
 */
#ifndef SVT_ORACLE_H
#define SVT_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* 1-D kernels.  kind: 0 DCT, 1 ADST, 2 identity.  n in {4,8,16,32,64} (ADST: 4,8,16).
 * inverse != 0 applies clamp_value(range_bits) on every add/sub stage (EbTransforms.c:5458).
 * Returns 0, or -1 for an (kind,n) the reference has no legal use of. */
int svto_txfm1d(int kind, int n, int inverse, const int32_t *in, int32_t *out, int cos_bit,
                int range_bits);

/* Av1TranformTwoDCore_c (EbTransforms.c:3701): dense w*h int32 output, stride w. */
int svto_fwd_txfm2d(const int16_t *input, uint32_t stride, int32_t *output, int tx_size,
                    int tx_type, int bd);

/* av1_estimate_transform (EbTransforms.c:4918): as above, then for the 64-point sizes the
 * energy of the discarded region, zero-out and re-pack to stride min(w,32).  coeff must hold
 * w*h int32 (the reference uses it as scratch too). */
int svto_estimate_transform(const int16_t *residual, uint32_t stride, int32_t *coeff, int tx_size,
                            int tx_type, int bd, uint64_t *three_quad_energy);

/* highbd_quantize_b_helper_c (EbFullLoop.c:239) when clamp16 == 0,
 * quantize_b_helper_c_II (EbFullLoop.c:44) when clamp16 != 0.  Tables are the 2-entry
 * {DC, AC} rows of av1_build_quantizer.  qm/iqm may be NULL (flat). */
void svto_quantize(const int32_t *coeff, int n_coeffs, const int16_t *zbin, const int16_t *round,
                   const int16_t *quant, const int16_t *quant_shift, const int16_t *dequant,
                   int32_t *qcoeff, int32_t *dqcoeff, uint16_t *eob, const int16_t *scan,
                   int log_scale, const uint8_t *qm, const uint8_t *iqm, int clamp16);

/* inv_txfm2d_add_c (EbTransforms.c:8180) incl. the 64-point zero re-expansion (:8299-8426):
 * packed dqcoeff in, prediction in / reconstruction out in place, 16-bit pixels. */
int svto_inv_txfm2d_add(const int32_t *dqcoeff, uint16_t *dst, int32_t stride, int tx_size,
                        int tx_type, int bd);
/* av1_inv_txfm_add_c (EbTransforms.c:8882): the same on 8-bit pixels. */
int svto_inv_txfm2d_add8(const int32_t *dqcoeff, uint8_t *dst, int32_t stride, int tx_size,
                         int tx_type);

/* av1_build_quantizer (EbModeDecisionConfigurationProcess.c:429) from caller-supplied
 * dc/ac_qlookup rows: out[256][5][8] = zbin, round, quant, quant_shift, dequant. */
void svto_build_quant_tables(const int16_t *dc_q, const int16_t *ac_q, int bd, int16_t *out);

/* geometry / constant lookups shared with the tests */
int svto_tx_wide(int tx_size);
int svto_tx_high(int tx_size);
int svto_max_eob(int tx_size);       /* av1_get_max_eob, EbTransforms.h:3872 */
int svto_tx_scale(int tx_size);      /* av1_get_tx_scale, EbTransforms.h:325 */
int svto_is_valid(int tx_size, int tx_type); /* is_txfm_allowed, test/TxfmCommon.h:172 */
int svto_scan_class(int tx_size, int tx_type); /* 0 default, 1 mrow, 2 mcol (EbTransforms.h:3349) */
void svto_get_cfg(int tx_size, int tx_type, int32_t *out16);

/* One transform block of a batch (mirrors the product's SvtGpuBlock so that the same arrays
 * feed both). */
typedef struct SvtoBlock {
    uint8_t tx_size, tx_type, bd, plane;
    uint32_t res_off, res_stride; /* int16 elements */
    uint32_t rec_off, rec_stride; /* pixels */
    uint32_t coeff_off;           /* int32 elements */
    uint32_t qrow;                /* row of tables[][5][8] */
    uint32_t flags;               /* product flags (clamp16, ...): unused by the checker */
} SvtoBlock;

/* fused fwd + quant + inverse-recon over blocks [first,last): the CPU baseline "port".
 * recon is uint8 (bd 8) or uint16 (bd 10).  scans[tx_size][class] -> scan table pointer. */
void svto_fused_blocks(const SvtoBlock *blk, int first, int last, const int16_t *residual,
                       void *recon, const int16_t *tables, int32_t *qcoeff, int32_t *dqcoeff,
                       uint16_t *eob, uint64_t *energy, const int16_t *const *scans,
                       int clamp16, const uint8_t *wt, const uint8_t *iwt,
                       const int32_t *qm_off /* [2][19] or NULL */);


/* SURVEY 8f-1 (residual + distortion fusion): see svt_oracle.c */
void svto_residual(const void *input, uint32_t input_stride, const void *pred, uint32_t pred_stride,
                   int16_t *residual, uint32_t residual_stride, int w, int h, int bd);
void svto_tu_distortion(const int32_t *coeff, const int32_t *recon_coeff, int tx_size, int eob,
                        uint64_t three_quad_energy, uint64_t dist[2]);
void svto_fused_blocks_md(const SvtoBlock *blk, int first, int last, const void *src, void *recon,
                          const int16_t *tables, int32_t *qcoeff, int32_t *dqcoeff, uint16_t *eob,
                          uint64_t *energy, uint64_t *dist, const int16_t *const *scans,
                          int clamp16);
/* the same per candidate of a tx-type search: no inverse, `pred` is only read */
void svto_tx_search_blocks(const SvtoBlock *blk, int first, int last, const void *src, const void *pred,
                          const int16_t *tables, int32_t *qcoeff, int32_t *dqcoeff, uint16_t *eob,
                          uint64_t *energy, uint64_t *dist, const int16_t *const *scans,
                          int clamp16);

/* av1_txb_init_levels_c on a whole levels_buf ((width + 4) * (height + 6) + 16 bytes) */
void svto_txb_init_levels(const int32_t *coeff, int width, int height, uint8_t *buf);

#ifdef __cplusplus
}
#endif
#endif
