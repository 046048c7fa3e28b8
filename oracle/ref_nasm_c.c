/*
 * TEST INFRASTRUCTURE ONLY.  The reference implements these 32 symbols in NASM (.asm files under
 * Source/Lib/Common/ASM_SSE2: EbPictureOperators_SSE2.asm, intrapred_sse2.asm,
 * highbd_intrapred_sse2_.asm, highbd_variance_impl_sse2.asm, x64RegisterUtil.asm); there is no
 * assembler in this image.  None of them is on the residual-coding path (forward transform /
 * quantize / inverse transform).  The kernel oracle never reaches them; the whole-encoder
 * oracle (oracle/Makefile target `encapp`) does, so they get plain C bodies
 * written from their prototypes and documented contracts.  Both arms of the bitstream comparison
 * (CPU path and GPU path) run the same bodies.
 */
#include <stddef.h>
#include <stdint.h>
#include <string.h>

/* x64RegisterUtil.asm:25 -- emms; nothing to do for SSE2+ code on x86-64 */
void RunEmms(void) {}

/* EbPictureOperators_SSE2.asm:629 -- bsr: index of the highest set bit */
uint32_t Log2f_SSE2(uint32_t x) { return x ? 31u - (uint32_t)__builtin_clz(x) : 0u; }

/* EbPictureOperators_SSE2.asm:12 (prototype ASM_SSE2/EbMcp_SSE2.h:19) */
void picture_copy_kernel_sse2(uint8_t *src, uint32_t src_stride, uint8_t *dst, uint32_t dst_stride,
                              uint32_t area_width, uint32_t area_height) {
    for (uint32_t y = 0; y < area_height; ++y)
        memcpy(dst + (size_t)y * dst_stride, src + (size_t)y * src_stride, area_width);
}

/* EbPictureOperators_SSE2.asm:359 (prototype EbMcp_SSE2.h:27) -- pavgb: (a + b + 1) >> 1 */
void picture_average_kernel_sse2(uint8_t *src0, uint32_t src0_stride, uint8_t *src1, uint32_t src1_stride,
                                 uint8_t *dst, uint32_t dst_stride, uint32_t area_width,
                                 uint32_t area_height) {
    for (uint32_t y = 0; y < area_height; ++y)
        for (uint32_t x = 0; x < area_width; ++x)
            dst[(size_t)y * dst_stride + x] =
                (uint8_t)((src0[(size_t)y * src0_stride + x] + src1[(size_t)y * src1_stride + x] + 1) >> 1);
}

/* EbPictureOperators_SSE2.asm:214-357 (prototypes ASM_SSE2/EbPictureOperators_SSE2.h:16-42) */
static void zero_out(int16_t *coeff_buffer, uint32_t coeff_stride, uint32_t coeff_origin_index, int n) {
    for (int y = 0; y < n; ++y)
        memset(coeff_buffer + coeff_origin_index + (size_t)y * coeff_stride, 0, (size_t)n * sizeof(int16_t));
}
#define ZERO(name, n)                                                                              \
    void name(int16_t *coeff_buffer, uint32_t coeff_stride, uint32_t coeff_origin_index,          \
              uint32_t area_width, uint32_t area_height) {                                         \
        (void)area_width; (void)area_height;                                                       \
        zero_out(coeff_buffer, coeff_stride, coeff_origin_index, n);                               \
    }
ZERO(zero_out_coeff4x4_sse, 4) ZERO(zero_out_coeff8x8_sse2, 8)
ZERO(zero_out_coeff16x16_sse2, 16) ZERO(zero_out_coeff32x32_sse2, 32)

/* highbd_variance_impl_sse2.asm (prototypes ASM_SSE2/highbd_variance_sse2.c:22-28) */
static uint32_t hbd_var(const uint16_t *src, int32_t src_stride, const uint16_t *ref, int32_t ref_stride,
                        uint32_t *sse, int32_t *sum, int n) {
    uint32_t s2 = 0;
    int32_t s = 0;
    for (int y = 0; y < n; ++y)
        for (int x = 0; x < n; ++x) {
            const int32_t d = (int32_t)src[y * src_stride + x] - (int32_t)ref[y * ref_stride + x];
            s += d;
            s2 += (uint32_t)(d * d);
        }
    *sse = s2;
    *sum = s;
    return 0;
}
uint32_t aom_highbd_calc8x8var_sse2(const uint16_t *src, int32_t src_stride, const uint16_t *ref,
                                    int32_t ref_stride, uint32_t *sse, int32_t *sum) {
    return hbd_var(src, src_stride, ref, ref_stride, sse, sum, 8);
}
uint32_t aom_highbd_calc16x16var_sse2(const uint16_t *src, int32_t src_stride, const uint16_t *ref,
                                      int32_t ref_stride, uint32_t *sse, int32_t *sum) {
    return hbd_var(src, src_stride, ref, ref_stride, sse, sum, 16);
}

/* intrapred_sse2.asm / highbd_intrapred_sse2_.asm: the reference's own C forms of the same
 * predictors (Source/Lib/Common/Codec/EbIntraPrediction.c) have identical prototypes */
#define PRED8(kind, size)                                                                          \
    void aom_##kind##_predictor_##size##_c(uint8_t *, ptrdiff_t, const uint8_t *, const uint8_t *); \
    void aom_##kind##_predictor_##size##_sse2(uint8_t *dst, ptrdiff_t stride, const uint8_t *above, \
                                              const uint8_t *left) {                               \
        aom_##kind##_predictor_##size##_c(dst, stride, above, left);                               \
    }
#define PRED8_ALL(kind) PRED8(kind, 4x4) PRED8(kind, 8x8) PRED8(kind, 16x16)
PRED8_ALL(dc) PRED8_ALL(dc_top) PRED8_ALL(dc_left) PRED8_ALL(dc_128) PRED8_ALL(v) PRED8_ALL(h)
#define PRED16(kind, size)                                                                         \
    void aom_highbd_##kind##_predictor_##size##_c(uint16_t *, ptrdiff_t, const uint16_t *,         \
                                                  const uint16_t *, int32_t);                      \
    void aom_highbd_##kind##_predictor_##size##_sse2(uint16_t *dst, ptrdiff_t stride,              \
                                                     const uint16_t *above, const uint16_t *left,  \
                                                     int32_t bd) {                                 \
        aom_highbd_##kind##_predictor_##size##_c(dst, stride, above, left, bd);                    \
    }
PRED16(dc, 4x4) PRED16(dc, 8x8) PRED16(v, 4x4) PRED16(v, 8x8)
