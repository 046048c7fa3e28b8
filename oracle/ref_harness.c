/*
 * TEST INFRASTRUCTURE ONLY -- a thin C API over the UNMODIFIED reference, compiled against the
 * reference headers where they lie under /root/reference (see oracle/Makefile, target `ref`).
 * Nothing here is product code and nothing in the product links it.
 *
 * It owns the reference's RTCD function-pointer table (the `#define RTCD_C` TU, exactly as
 * Source/Lib/Encoder/Codec/EbEncHandle.c:117-118 does) and calls the reference's own
 * setup_rtcd_internal (Source/Lib/Common/Codec/aom_dsp_rtcd.h:2385) so that "asm" below means
 * precisely what the running encoder binds for ASM_AVX2.
 */
#include <stdint.h>
#include <string.h>
#include <stdlib.h>

#include "EbDefinitions.h"
#include "EbPictureControlSet.h"
#include "EbTransforms.h"
#include "EbModeDecisionConfigurationProcess.h"

#define RTCD_C
#include "aom_dsp_rtcd.h"

/* globals normally defined by EbEncHandle.c:154-161 (that TU is not linked: it would define
 * the RTCD pointers a second time) */
EbMemoryMapEntry *memory_map;
uint32_t *memory_map_index;
uint64_t *total_lib_memory;
uint32_t libMallocCount = 0;
uint32_t lib_thread_count = 0;
uint32_t libSemaphoreCount = 0;
uint32_t libMutexCount = 0;
uint8_t num_groups = 0;

/* prototypes of non-static reference functions that no header exports */
void av1_build_quantizer(aom_bit_depth_t bit_depth, int32_t y_dc_delta_q, int32_t u_dc_delta_q,
    int32_t u_ac_delta_q, int32_t v_dc_delta_q, int32_t v_ac_delta_q,
    Quants *const quants, Dequants *const deq);
void highbd_quantize_b_helper_c(const tran_low_t *coeff_ptr, intptr_t n_coeffs, int32_t skip_block,
    const int16_t *zbin_ptr, const int16_t *round_ptr, const int16_t *quant_ptr,
    const int16_t *quant_shift_ptr, tran_low_t *qcoeff_ptr, tran_low_t *dqcoeff_ptr,
    const int16_t *dequant_ptr, uint16_t *eob_ptr, const int16_t *scan, const int16_t *iscan,
    const qm_val_t *qm_ptr, const qm_val_t *iqm_ptr, const int32_t log_scale);
void quantize_b_helper_c_II(const tran_low_t *coeff_ptr, intptr_t n_coeffs, int32_t skip_block,
    const int16_t *zbin_ptr, const int16_t *round_ptr, const int16_t *quant_ptr,
    const int16_t *quant_shift_ptr, tran_low_t *qcoeff_ptr, tran_low_t *dqcoeff_ptr,
    const int16_t *dequant_ptr, uint16_t *eob_ptr, const int16_t *scan, const int16_t *iscan,
    const qm_val_t *qm_ptr, const qm_val_t *iqm_ptr, const int32_t log_scale);
void av1_inv_txfm_add_c(const tran_low_t *dqcoeff, uint8_t *dst, int32_t stride,
    const TxfmParam *txfm_param);
typedef void(*Txfm1dFn)(const int32_t *input, int32_t *output, int8_t cos_bit,
    const int8_t *stage_range);
#define DECL1D(n) void n(const int32_t *, int32_t *, int8_t, const int8_t *)
DECL1D(av1_fdct4_new); DECL1D(av1_fdct8_new); DECL1D(av1_fdct16_new); DECL1D(av1_fdct32_new);
DECL1D(av1_fdct64_new); DECL1D(av1_fadst4_new); DECL1D(av1_fadst8_new); DECL1D(av1_fadst16_new);
DECL1D(av1_fidentity4_c); DECL1D(av1_fidentity8_c); DECL1D(av1_fidentity16_c);
DECL1D(av1_fidentity32_c); DECL1D(av1_fidentity64_c);
DECL1D(av1_idct4_new); DECL1D(av1_idct8_new); DECL1D(av1_idct16_new); DECL1D(av1_idct32_new);
DECL1D(av1_idct64_new); DECL1D(av1_iadst4_new); DECL1D(av1_iadst8_new); DECL1D(av1_iadst16_new);
DECL1D(av1_iidentity4_c); DECL1D(av1_iidentity8_c); DECL1D(av1_iidentity16_c);
DECL1D(av1_iidentity32_c); DECL1D(av1_iidentity64_c);
extern const int32_t av1_cospi_arr_data[7][64];
extern const int32_t av1_sinpi_arr_data[7][5];

static int g_asm = -1;

/* asm_on = 1: what SvtAv1EncApp binds on an AVX2 host; 0: the pure C table (-asm 0) */
void ref_init(int asm_on) {
    if (g_asm == asm_on) return;
    setup_rtcd_internal(asm_on ? ASM_AVX2 : ASM_NON_AVX2);
    if (!asm_on) {
        /* ASM_NON_AVX2 still enables SSE*; force the plain C bodies for the hot-path slots */
        aom_quantize_b = aom_quantize_b_c_II;
        aom_quantize_b_32x32 = aom_quantize_b_32x32_c_II;
        aom_quantize_b_64x64 = aom_quantize_b_64x64_c_II;
        aom_highbd_quantize_b = aom_highbd_quantize_b_c;
        aom_highbd_quantize_b_32x32 = aom_highbd_quantize_b_32x32_c;
        aom_highbd_quantize_b_64x64 = aom_highbd_quantize_b_64x64_c;
        av1_inv_txfm_add = av1_inv_txfm_add_c;
        av1_fwd_txfm2d_4x4 = Av1TransformTwoD_4x4_c;
        av1_fwd_txfm2d_8x8 = Av1TransformTwoD_8x8_c;
        av1_fwd_txfm2d_16x16 = Av1TransformTwoD_16x16_c;
        av1_fwd_txfm2d_32x32 = Av1TransformTwoD_32x32_c;
        av1_fwd_txfm2d_64x64 = Av1TransformTwoD_64x64_c;
        av1_fwd_txfm2d_4x8 = av1_fwd_txfm2d_4x8_c;
        av1_fwd_txfm2d_8x4 = av1_fwd_txfm2d_8x4_c;
        av1_fwd_txfm2d_8x16 = av1_fwd_txfm2d_8x16_c;
        av1_fwd_txfm2d_16x8 = av1_fwd_txfm2d_16x8_c;
        av1_fwd_txfm2d_16x32 = av1_fwd_txfm2d_16x32_c;
        av1_fwd_txfm2d_32x16 = av1_fwd_txfm2d_32x16_c;
        av1_fwd_txfm2d_32x64 = av1_fwd_txfm2d_32x64_c;
        av1_fwd_txfm2d_64x32 = av1_fwd_txfm2d_64x32_c;
        av1_fwd_txfm2d_4x16 = av1_fwd_txfm2d_4x16_c;
        av1_fwd_txfm2d_16x4 = av1_fwd_txfm2d_16x4_c;
        av1_fwd_txfm2d_8x32 = av1_fwd_txfm2d_8x32_c;
        av1_fwd_txfm2d_32x8 = av1_fwd_txfm2d_32x8_c;
        av1_fwd_txfm2d_16x64 = av1_fwd_txfm2d_16x64_c;
        av1_fwd_txfm2d_64x16 = av1_fwd_txfm2d_64x16_c;
        av1_inv_txfm2d_add_4x4 = av1_inv_txfm2d_add_4x4_c;
        av1_inv_txfm2d_add_8x8 = av1_inv_txfm2d_add_8x8_c;
        av1_inv_txfm2d_add_16x16 = av1_inv_txfm2d_add_16x16_c;
        av1_inv_txfm2d_add_32x32 = av1_inv_txfm2d_add_32x32_c;
        av1_inv_txfm2d_add_64x64 = av1_inv_txfm2d_add_64x64_c;
        av1_inv_txfm2d_add_4x8 = av1_inv_txfm2d_add_4x8_c;
        av1_inv_txfm2d_add_8x4 = av1_inv_txfm2d_add_8x4_c;
        av1_inv_txfm2d_add_8x16 = av1_inv_txfm2d_add_8x16_c;
        av1_inv_txfm2d_add_16x8 = av1_inv_txfm2d_add_16x8_c;
        av1_inv_txfm2d_add_16x32 = av1_inv_txfm2d_add_16x32_c;
        av1_inv_txfm2d_add_32x16 = av1_inv_txfm2d_add_32x16_c;
        av1_inv_txfm2d_add_32x64 = av1_inv_txfm2d_add_32x64_c;
        av1_inv_txfm2d_add_64x32 = av1_inv_txfm2d_add_64x32_c;
        av1_inv_txfm2d_add_4x16 = av1_inv_txfm2d_add_4x16_c;
        av1_inv_txfm2d_add_16x4 = av1_inv_txfm2d_add_16x4_c;
        av1_inv_txfm2d_add_8x32 = av1_inv_txfm2d_add_8x32_c;
        av1_inv_txfm2d_add_32x8 = av1_inv_txfm2d_add_32x8_c;
        av1_inv_txfm2d_add_16x64 = av1_inv_txfm2d_add_16x64_c;
        av1_inv_txfm2d_add_64x16 = av1_inv_txfm2d_add_64x16_c;
    }
    g_asm = asm_on;
}

/* ---- entry point 1: av1_estimate_transform (EbTransforms.c:4918) ---- */
int ref_estimate_transform(int16_t *residual, uint32_t stride, int32_t *coeff, int tx_size,
                           int tx_type, int bd, uint64_t *three_quad_energy) {
    *three_quad_energy = 0;
    return (int)av1_estimate_transform(residual, stride, coeff, 0, (TxSize)tx_size,
        three_quad_energy, NULL, bd == 8 ? 0 : 2, (TxType)tx_type, ASM_AVX2, PLANE_TYPE_Y,
        DEFAULT_SHAPE);
}

/* ---- quantizer: the facades' routing (EbFullLoop.c:335-386, 421-489) without the PCS ----
 * variant 0: the RTCD pointer for this bit depth (what the encoder runs)
 * variant 1: highbd_quantize_b_helper_c (10-bit C, optional qmatrix)
 * variant 2: quantize_b_helper_c_II     (8-bit C with the INT16 clamp) */
void ref_quantize(const int32_t *coeff, int n_coeffs, const int16_t *zbin, const int16_t *round,
                  const int16_t *quant, const int16_t *quant_shift, const int16_t *dequant,
                  int32_t *qcoeff, int32_t *dqcoeff, uint16_t *eob, int tx_size, int tx_type,
                  int bd, int variant, const uint8_t *qm, const uint8_t *iqm) {
    const SCAN_ORDER *so = &av1_scan_orders[tx_size][tx_type];
    const int log_scale = av1_get_tx_scale((TxSize)tx_size);
    if (variant == 1) {
        highbd_quantize_b_helper_c(coeff, n_coeffs, 0, zbin, round, quant, quant_shift, qcoeff,
            dqcoeff, dequant, eob, so->scan, so->iscan, qm, iqm, log_scale);
    } else if (variant == 2) {
        quantize_b_helper_c_II(coeff, n_coeffs, 0, zbin, round, quant, quant_shift, qcoeff,
            dqcoeff, dequant, eob, so->scan, so->iscan, NULL, NULL, log_scale);
    } else {
        void(*fn)(const tran_low_t *, intptr_t, int32_t, const int16_t *, const int16_t *,
            const int16_t *, const int16_t *, tran_low_t *, tran_low_t *, const int16_t *,
            uint16_t *, const int16_t *, const int16_t *);
        if (bd == 8)
            fn = log_scale == 0 ? aom_quantize_b : log_scale == 1 ? aom_quantize_b_32x32
                                                                   : aom_quantize_b_64x64;
        else
            fn = log_scale == 0 ? aom_highbd_quantize_b
                                : log_scale == 1 ? aom_highbd_quantize_b_32x32
                                                 : aom_highbd_quantize_b_64x64;
        fn(coeff, n_coeffs, 0, zbin, round, quant, quant_shift, qcoeff, dqcoeff, dequant, eob,
           so->scan, so->iscan);
    }
}

/* ---- entry point 3: av1_inv_transform_recon / _recon8bit (EbTransforms.c:8905, 8935) ----
 * bd 8: recon is uint8_t; bd 10: recon is uint16_t passed the way the encoder passes it */
int ref_inv_transform_recon(int32_t *dqcoeff, void *recon, uint32_t stride, int tx_size,
                            int tx_type, int bd, uint32_t eob) {
    if (bd == 8)
        return (int)av1_inv_transform_recon8bit(dqcoeff, (uint8_t *)recon, stride,
            (TxSize)tx_size, (TxType)tx_type, PLANE_TYPE_Y, eob);
    return (int)av1_inv_transform_recon(dqcoeff, CONVERT_TO_BYTEPTR((uint16_t *)recon), stride,
        (TxSize)tx_size, 2, (TxType)tx_type, PLANE_TYPE_Y, eob);
}

/* ---- the whole path on a list of blocks, threaded: CPU baseline for bench.py ---- */
typedef struct RefBlock {
    uint8_t tx_size, tx_type, bd, plane;
    uint32_t res_off, res_stride;   /* int16 elements */
    uint32_t rec_off, rec_stride;   /* pixels */
    uint32_t coeff_off;             /* int32 elements */
    uint32_t qrow;                  /* row of the quant table to use */
    uint32_t flags;
} RefBlock;

/* tables: [nrows][5][8] int16: zbin, round, quant, quant_shift, dequant.
 * wt / iwt / off ([2][19], -1 = none): optional quantization matrices; a 10-bit block flagged
 * USE_QM (flags & 2) then goes through highbd_quantize_b_helper_c, the function the reference's
 * facade picks when both matrix pointers are set (EbFullLoop.c:421-489); everything else through
 * the RTCD pointer of its bit depth. */
static void ref_fused_range(const RefBlock *blk, int first, int last, int16_t *residual, void *recon,
                            const int16_t *tables, int32_t *qcoeff, int32_t *dqcoeff, uint16_t *eob,
                            uint64_t *energy, const uint8_t *wt, const uint8_t *iwt, const int32_t *off) {
    int32_t coeff[32 * 32 * 4] __attribute__((aligned(32)));
    for (int i = first; i < last; ++i) {
        const RefBlock *b = &blk[i];
        const int n = av1_get_max_eob((TxSize)b->tx_size);
        const int16_t *t = tables + (size_t)b->qrow * 40;
        ref_estimate_transform(residual + b->res_off, b->res_stride, coeff, b->tx_size,
                               b->tx_type, b->bd, &energy[i]);
        const int o = (wt && (b->flags & 2) && b->bd != 8) ? off[(b->plane > 0) * 19 + b->tx_size] : -1;
        ref_quantize(coeff, n, t, t + 8, t + 16, t + 24, t + 32, qcoeff + b->coeff_off,
                     dqcoeff + b->coeff_off, &eob[i], b->tx_size, b->tx_type, b->bd, o >= 0 ? 1 : 0,
                     o >= 0 ? wt + o : NULL, o >= 0 ? iwt + o : NULL);
        if (eob[i]) {
            void *dst = b->bd == 8 ? (void *)((uint8_t *)recon + b->rec_off)
                                   : (void *)((uint16_t *)recon + b->rec_off);
            ref_inv_transform_recon(dqcoeff + b->coeff_off, dst, b->rec_stride, b->tx_size,
                                    b->tx_type, b->bd, eob[i]);
        }
    }
}
void ref_fused_blocks(const RefBlock *blk, int first, int last, int16_t *residual, void *recon,
                      const int16_t *tables, int32_t *qcoeff, int32_t *dqcoeff, uint16_t *eob,
                      uint64_t *energy) {
    ref_fused_range(blk, first, last, residual, recon, tables, qcoeff, dqcoeff, eob, energy, NULL, NULL, NULL);
}

/* ---- persistent worker pool: the timed CPU arm of bench.py ------------------------------------
 * Workers are created once and park on a condition variable between passes; inside a pass they
 * pull chunks of POOL_CHUNK consecutive blocks from a shared counter, so the pass is balanced
 * whatever the mix of sizes (the 32x16 / 16x32 DCT goes through the C path and is ~3x slower
 * per sample than its neighbours).  ref_pool_run returns the wall time of the pass in seconds,
 * measured here, not in the caller's interpreter. */
#include <pthread.h>
#include <time.h>
#define POOL_MAX 256
#define POOL_CHUNK 64
typedef struct PoolJob {
    const RefBlock *blk; int n;
    int16_t *residual; void *recon; const int16_t *tables;
    int32_t *q, *dq; uint16_t *eob; uint64_t *energy;
    const uint8_t *wt, *iwt; const int32_t *off;
} PoolJob;
static struct {
    pthread_t th[POOL_MAX];
    int n;
    pthread_mutex_t mu;
    pthread_cond_t go, done;
    unsigned long gen;
    int pending, stop;
    volatile int next;
    PoolJob job;
} g_pool = { .mu = PTHREAD_MUTEX_INITIALIZER, .go = PTHREAD_COND_INITIALIZER, .done = PTHREAD_COND_INITIALIZER };

static void pool_drain(const PoolJob *j) {
    for (;;) {
        const int a = __atomic_fetch_add(&g_pool.next, POOL_CHUNK, __ATOMIC_RELAXED);
        if (a >= j->n) break;
        const int b = a + POOL_CHUNK < j->n ? a + POOL_CHUNK : j->n;
        ref_fused_range(j->blk, a, b, j->residual, j->recon, j->tables, j->q, j->dq, j->eob, j->energy,
                        j->wt, j->iwt, j->off);
    }
}
static void *pool_worker(void *arg) {
    unsigned long seen = 0;
    (void)arg;
    for (;;) {
        pthread_mutex_lock(&g_pool.mu);
        while (g_pool.gen == seen && !g_pool.stop) pthread_cond_wait(&g_pool.go, &g_pool.mu);
        if (g_pool.stop) { pthread_mutex_unlock(&g_pool.mu); return NULL; }
        seen = g_pool.gen;
        const PoolJob j = g_pool.job;
        pthread_mutex_unlock(&g_pool.mu);
        pool_drain(&j);
        pthread_mutex_lock(&g_pool.mu);
        if (--g_pool.pending == 0) pthread_cond_signal(&g_pool.done);
        pthread_mutex_unlock(&g_pool.mu);
    }
}
/* threads = total threads working on a pass (the caller's thread is one of them) */
int ref_pool_start(int threads) {
    if (threads < 1) threads = 1;
    if (threads > POOL_MAX) threads = POOL_MAX;
    if (g_pool.n == threads - 1 && !g_pool.stop) return threads;
    pthread_mutex_lock(&g_pool.mu);
    g_pool.stop = 1;
    pthread_cond_broadcast(&g_pool.go);
    pthread_mutex_unlock(&g_pool.mu);
    for (int i = 0; i < g_pool.n; ++i) pthread_join(g_pool.th[i], NULL);
    g_pool.stop = 0;
    g_pool.n = 0;
    for (int i = 0; i < threads - 1; ++i)
        if (pthread_create(&g_pool.th[g_pool.n], NULL, pool_worker, NULL) == 0) ++g_pool.n;
    return g_pool.n + 1;
}
double ref_pool_run(const RefBlock *blk, int n, int16_t *residual, void *recon, const int16_t *tables,
                    int32_t *qcoeff, int32_t *dqcoeff, uint16_t *eob, uint64_t *energy,
                    const uint8_t *wt, const uint8_t *iwt, const int32_t *off) {
    struct timespec t0, t1;
    const PoolJob j = { blk, n, residual, recon, tables, qcoeff, dqcoeff, eob, energy, wt, iwt, off };
    clock_gettime(CLOCK_MONOTONIC, &t0);
    pthread_mutex_lock(&g_pool.mu);
    g_pool.job = j;
    g_pool.next = 0;
    g_pool.pending = g_pool.n;
    ++g_pool.gen;
    pthread_cond_broadcast(&g_pool.go);
    pthread_mutex_unlock(&g_pool.mu);
    pool_drain(&j);
    pthread_mutex_lock(&g_pool.mu);
    while (g_pool.pending) pthread_cond_wait(&g_pool.done, &g_pool.mu);
    pthread_mutex_unlock(&g_pool.mu);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}

/* ---- table dumps (fixtures for tests/golden) ---- */
void ref_build_quant_tables(int bd, int plane, int16_t *out /* [256][5][8] */) {
    Quants *q = (Quants *)malloc(sizeof(Quants));
    Dequants *d = (Dequants *)malloc(sizeof(Dequants));
    av1_build_quantizer((aom_bit_depth_t)bd, 0, 0, 0, 0, 0, q, d);
    for (int qi = 0; qi < 256; ++qi) {
        int16_t *o = out + qi * 40;
        const int16_t *zb = plane == 0 ? q->y_zbin[qi] : plane == 1 ? q->u_zbin[qi] : q->v_zbin[qi];
        const int16_t *rd = plane == 0 ? q->y_round[qi] : plane == 1 ? q->u_round[qi] : q->v_round[qi];
        const int16_t *qu = plane == 0 ? q->y_quant[qi] : plane == 1 ? q->u_quant[qi] : q->v_quant[qi];
        const int16_t *qs = plane == 0 ? q->y_quant_shift[qi] : plane == 1 ? q->u_quant_shift[qi] : q->v_quant_shift[qi];
        const int16_t *dq = plane == 0 ? d->y_dequant_QTX[qi] : plane == 1 ? d->u_dequant_QTX[qi] : d->v_dequant_QTX[qi];
        memcpy(o, zb, 16); memcpy(o + 8, rd, 16); memcpy(o + 16, qu, 16);
        memcpy(o + 24, qs, 16); memcpy(o + 32, dq, 16);
    }
    free(q); free(d);
}

int ref_get_scan(int tx_size, int tx_type, int16_t *scan, int16_t *iscan) {
    const SCAN_ORDER *so = &av1_scan_orders[tx_size][tx_type];
    const int n = av1_get_max_eob((TxSize)tx_size);
    if (!so->scan) return 0;
    memcpy(scan, so->scan, n * sizeof(int16_t));
    memcpy(iscan, so->iscan, n * sizeof(int16_t));
    return n;
}

/* qmatrix pointers exactly as av1_qm_init would bind them
 * (EbModeDecisionConfigurationProcess.c:524-554); returns the element count or 0 for flat */
int ref_get_qmatrix(int level, int plane, int tx_size, uint8_t *wt, uint8_t *iwt) {
    if (level >= NUM_QM_LEVELS - 1) return 0;
    int current = 0;
    const uint8_t *w = NULL, *iw = NULL;
    const TxSize want = av1_get_adjusted_tx_size((TxSize)tx_size);
    for (int t = 0; t < TX_SIZES_ALL; ++t) {
        const int size = tx_size_2d[t];
        if (t != (int)av1_get_adjusted_tx_size((TxSize)t)) continue;
        if (t == (int)want) {
            w = &wt_matrix_ref[level][plane >= 1][current];
            iw = &iwt_matrix_ref[level][plane >= 1][current];
        }
        current += size;
    }
    const int n = tx_size_2d[want];
    memcpy(wt, w, n); memcpy(iwt, iw, n);
    return n;
}

void ref_get_trig(int32_t *cospi /* [7][64] */, int32_t *sinpi /* [7][5] */) {
    memcpy(cospi, av1_cospi_arr_data, sizeof(av1_cospi_arr_data));
    memcpy(sinpi, av1_sinpi_arr_data, sizeof(av1_sinpi_arr_data));
}

/* 1-D kernels: kind 0 dct, 1 adst, 2 identity; inverse uses a uniform stage range */
int ref_txfm1d(int kind, int n, int inverse, const int32_t *in, int32_t *out, int cos_bit,
               int range_bits) {
    static const Txfm1dFn f[2][3][5] = {
        { { av1_fdct4_new, av1_fdct8_new, av1_fdct16_new, av1_fdct32_new, av1_fdct64_new },
          { av1_fadst4_new, av1_fadst8_new, av1_fadst16_new, NULL, NULL },
          { av1_fidentity4_c, av1_fidentity8_c, av1_fidentity16_c, av1_fidentity32_c,
            av1_fidentity64_c } },
        { { av1_idct4_new, av1_idct8_new, av1_idct16_new, av1_idct32_new, av1_idct64_new },
          { av1_iadst4_new, av1_iadst8_new, av1_iadst16_new, NULL, NULL },
          { av1_iidentity4_c, av1_iidentity8_c, av1_iidentity16_c, av1_iidentity32_c,
            av1_iidentity64_c } } };
    int l = n == 4 ? 0 : n == 8 ? 1 : n == 16 ? 2 : n == 32 ? 3 : 4;
    Txfm1dFn fn = f[inverse][kind][l];
    if (!fn) return -1;
    int8_t sr[MAX_TXFM_STAGE_NUM];
    memset(sr, range_bits, sizeof(sr));
    int32_t tmp[64];
    memcpy(tmp, in, n * sizeof(int32_t));
    fn(tmp, out, (int8_t)cos_bit, sr);
    return 0;
}

/* per-(size,type) facts used to cross-check our generated constant tables */
void ref_get_cfg(int tx_size, int tx_type, int32_t *out /* 16 ints */) {
    Txfm2DFlipCfg f, i;
    Av1TransformConfig((TxType)tx_type, (TxSize)tx_size, &f);
    av1_get_inv_txfm_cfg((TxType)tx_type, (TxSize)tx_size, &i);
    out[0] = tx_size_wide[tx_size]; out[1] = tx_size_high[tx_size];
    out[2] = f.shift[0]; out[3] = f.shift[1]; out[4] = f.shift[2];
    out[5] = f.cos_bit_col; out[6] = f.cos_bit_row;
    out[7] = i.shift[0]; out[8] = i.shift[1];
    out[9] = f.ud_flip; out[10] = f.lr_flip;
    out[11] = (int)f.txfm_type_col; out[12] = (int)f.txfm_type_row;
    out[13] = av1_get_tx_scale((TxSize)tx_size);
    out[14] = av1_get_max_eob((TxSize)tx_size);
    out[15] = (int)av1_get_adjusted_tx_size((TxSize)tx_size);
}

/* ---- the reference's residual and transform-domain distortion kernels ---------- */
#include "EbPictureOperators.h"

/* ResidualKernel through the RTCD pointer (AVX2 when bound) for 8-bit, residual_kernel16bit
 * (Codec/EbPictureOperators.c:134) for 16-bit pixels */
void ref_residual(const void *input, uint32_t input_stride, const void *pred, uint32_t pred_stride,
                  int16_t *residual, uint32_t residual_stride, int w, int h, int bd) {
    if (bd == 8)
        ResidualKernel((uint8_t *)input, input_stride, (uint8_t *)pred, pred_stride, residual,
                       residual_stride, (uint32_t)w, (uint32_t)h);
    else
        residual_kernel16bit((uint16_t *)input, input_stride, (uint16_t *)pred, pred_stride,
                             residual, residual_stride, (uint32_t)w, (uint32_t)h);
}

/* the call sequence of Codec/EbFullLoop.c:808-835 for one luma TU */
void ref_tu_distortion(int32_t *coeff, int32_t *recon_coeff, int tx_size, int eob,
                       uint64_t three_quad_energy, uint64_t dist[2], int use_avx2) {
    EbPictureBufferDesc_t a, b;
    memset(&a, 0, sizeof(a));
    memset(&b, 0, sizeof(b));
    a.buffer_y = (uint8_t *)coeff;
    b.buffer_y = (uint8_t *)recon_coeff;
    uint64_t y[DIST_CALC_TOTAL], cb[DIST_CALC_TOTAL], cr[DIST_CALC_TOTAL];
    picture_full_distortion32_bits(&a, 0, 0, &b, 0, 0, (uint32_t)tx_size_wide[tx_size],
                                   (uint32_t)tx_size_high[tx_size], 0, 0, y, cb, cr, (uint32_t)eob,
                                   0, 0, COMPONENT_LUMA, use_avx2 ? ASM_AVX2 : ASM_NON_AVX2);
    y[DIST_CALC_RESIDUAL] += three_quad_energy;
    y[DIST_CALC_PREDICTION] += three_quad_energy;
    const int32_t shift = (MAX_TX_SCALE - av1_get_tx_scale((TxSize)tx_size)) * 2;
    dist[0] = RIGHT_SIGNED_SHIFT(y[DIST_CALC_RESIDUAL], shift);
    dist[1] = RIGHT_SIGNED_SHIFT(y[DIST_CALC_PREDICTION], shift);
}

/* ---- which tx types ProductFullLoopTxSearch evaluates (Codec/EbFullLoop.c:946-982),
 * from the reference's own get_ext_tx_set_type / av1_ext_tx_used / txsize_sqr_up_map
 * (Codec/EbDefinitions.h) and its shortlist table allowed_tx_set_a (Codec/EbFullLoop.c:883) ------ */
extern uint8_t allowed_tx_set_a[TX_SIZES_ALL][TX_TYPES];
uint32_t ref_tx_search_mask(int tx_size, int is_inter, int reduced_tx_set_used, int tx_search_reduced_set) {
    const TxSize txSize = (TxSize)tx_size;
    const TxSetType tx_set_type = get_ext_tx_set_type(txSize, is_inter, reduced_tx_set_used);
    int32_t allowed_tx_mask[TX_TYPES] = { 0 };
    int32_t allowed_tx_num = 0;
    for (int32_t tx_type_index = DCT_DCT; tx_type_index < TX_TYPES; ++tx_type_index) {
        TxType tx_type = (TxType)tx_type_index;
        allowed_tx_mask[tx_type] = 1;
        const TxType ref_tx_type = ((!av1_ext_tx_used[tx_set_type][tx_type]) ||
                                    txsize_sqr_up_map[txSize] > TX_32X32) ? DCT_DCT : tx_type;
        if (tx_type != ref_tx_type) allowed_tx_mask[tx_type] = 0;
        allowed_tx_num += allowed_tx_mask[tx_type];
    }
    if (allowed_tx_num == 0) allowed_tx_mask[DCT_DCT] = 1;
    uint32_t m = 0;
    for (int32_t t = 0; t < TX_TYPES; ++t) {
        if (!allowed_tx_mask[t]) continue;
        if (tx_search_reduced_set && !allowed_tx_set_a[txSize][t]) continue;
        m |= 1u << t;
    }
    return m;
}

/* ---- av1_txb_init_levels (Codec/EbRateDistortionCost.c:125, AVX2 twin
 * ASM_AVX2/encodetxb_avx2.c:21) on a caller buffer laid out like levels_buf ---------------------- */
void av1_txb_init_levels_c(const tran_low_t *const coeff, const int32_t width, const int32_t height,
    uint8_t *const levels);
void av1_txb_init_levels_avx2(const tran_low_t *const coeff, const int32_t width, const int32_t height,
    uint8_t *const levels);
void ref_txb_init_levels(const int32_t *coeff, int width, int height, uint8_t *buf, int use_avx2) {
    uint8_t *levels = buf + TX_PAD_TOP * (width + TX_PAD_HOR);   /* set_levels */
    if (use_avx2) av1_txb_init_levels_avx2(coeff, width, height, levels);
    else av1_txb_init_levels_c(coeff, width, height, levels);
}
