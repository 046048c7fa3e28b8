/*
 * TEST INFRASTRUCTURE ONLY.  The reference implements these 32 symbols in NASM (.asm files
 * under Source/Lib/Common/ASM_SSE2 etc.); there is no assembler in this image.  None of them is
 * on the residual-coding path (fwd transform / quantize / inverse transform), so for the kernel
 * oracle they abort loudly if ever reached.  RunEmms (== aom_clear_system_state,
 * Source/Lib/Common/Codec/EbDefinitions.h:311) is a no-op on x86-64 SSE2+ code.
 */
#include <stdio.h>
#include <stdlib.h>
#define DIE(n) void n(void) { fprintf(stderr, "oracle/_ref: NASM-only symbol " #n " reached\n"); abort(); }
void RunEmms(void) {}
DIE(Log2f_SSE2)
DIE(aom_dc_128_predictor_16x16_sse2) DIE(aom_dc_128_predictor_4x4_sse2) DIE(aom_dc_128_predictor_8x8_sse2)
DIE(aom_dc_left_predictor_16x16_sse2) DIE(aom_dc_left_predictor_4x4_sse2) DIE(aom_dc_left_predictor_8x8_sse2)
DIE(aom_dc_predictor_16x16_sse2) DIE(aom_dc_predictor_4x4_sse2) DIE(aom_dc_predictor_8x8_sse2)
DIE(aom_dc_top_predictor_16x16_sse2) DIE(aom_dc_top_predictor_4x4_sse2) DIE(aom_dc_top_predictor_8x8_sse2)
DIE(aom_h_predictor_16x16_sse2) DIE(aom_h_predictor_4x4_sse2) DIE(aom_h_predictor_8x8_sse2)
DIE(aom_highbd_calc16x16var_sse2) DIE(aom_highbd_calc8x8var_sse2)
DIE(aom_highbd_dc_predictor_4x4_sse2) DIE(aom_highbd_dc_predictor_8x8_sse2)
DIE(aom_highbd_v_predictor_4x4_sse2) DIE(aom_highbd_v_predictor_8x8_sse2)
DIE(aom_v_predictor_16x16_sse2) DIE(aom_v_predictor_4x4_sse2) DIE(aom_v_predictor_8x8_sse2)
DIE(picture_average_kernel_sse2) DIE(picture_copy_kernel_sse2)
DIE(zero_out_coeff16x16_sse2) DIE(zero_out_coeff32x32_sse2) DIE(zero_out_coeff4x4_sse) DIE(zero_out_coeff8x8_sse2)
