// Fused residual-coding kernel: forward 2-D transform -> quantize / dequantize / eob ->
// inverse 2-D transform + reconstruction, one launch per transform size, many blocks per CTA.
//
// Replaces, for a whole batch of transform blocks, the reference's per-block chain
//   av1_estimate_transform     (Source/Lib/Common/Codec/EbTransforms.c:4918)
//   av1_quantize_inv_quantize  (Source/Lib/Common/Codec/EbFullLoop.c:653)
//   av1_inv_transform_recon    (Source/Lib/Common/Codec/EbTransforms.c:8905 / :8935)
//
// Work decomposition (B200: 148 SMs x 4 SMSPs, INT32 pipes are the co-bottleneck with HBM):
//   * a CTA of 128 threads owns NB = 128 / min(W,H) blocks of one size;
//   * a thread owns one whole column (then one whole row) of a block in registers, so every
//     butterfly is straight-line IMAD/SHF code with immediate weights and no shuffles inside a
//     1-D transform; the only exchange is the 2-D transpose through padded shared memory;
//   * the row owner runs forward row -> quantize -> inverse row back to back on its registers, so
//     coefficients never round-trip through memory between the three reference calls.
// The phases are plain functions of (task index, shared buffer) so that the host emulation in
// tests/ can run the very same code without a GPU.
#pragma once
#include <stddef.h>
#include <stdint.h>
#include <type_traits>
#include <utility>
#include "txfm1d.cuh"

#if !defined(__CUDACC__)  // host emulation build (tests only): the two vector helpers used below
struct int4 { int x, y, z, w; };
static inline int4 make_int4(int x, int y, int z, int w) { return int4{ x, y, z, w }; }
#endif

namespace svt {

enum : int { STAGE_FWD = 1, STAGE_QUANT = 2, STAGE_INV = 4, STAGE_ALL = 7,
             // internal: write all W x H coefficients at stride W (what the RTCD-level
             // av1_fwd_txfm2d_WxH returns before av1_estimate_transform packs 64-point sizes)
             STAGE_FWD_FULL = 8,
             // residual = source - prediction computed while staging (ResidualKernel,
             // EbPictureOperators.c:134-191) instead of read from `residual`
             STAGE_RESIDUAL = 16 };
enum : uint32_t { META_TYPE_MASK = 0xF, META_BD10 = 1u << 4, META_PLANE_SHIFT = 5,
                  META_CLAMP16 = 1u << 8, META_USE_QM = 1u << 9,
                  // decoded from the tx type once, on the host (make_meta): 1-D kind of the column
                  // and row transforms (0 DCT, 1 ADST, 2 identity), the two flips, the scan class
                  META_VKIND_SHIFT = 10, META_HKIND_SHIFT = 12, META_UD = 1u << 14, META_LR = 1u << 15,
                  META_SCAN_SHIFT = 16 };

struct alignas(16) DevBlock {  // 32 bytes (two 128-bit loads); device copy of a public SvtGpuBlock, bucketed by size
    uint32_t res_off, res_stride;  // int16 elements
    uint32_t rec_off, rec_stride;  // pixels
    uint32_t coeff_off;            // int32 elements, multiple of 4
    uint32_t qrow;                 // row of qtables[][5][8]
    uint32_t id;                   // index in the caller's list (eob / energy slot)
    uint32_t meta;
};

struct KParams {
    const DevBlock* blocks;
    int n_blocks;
    int stages;
    const int16_t* residual;
    const void* pred;       // uint8 (bd 8) or uint16 (bd 10)
    void* recon;            // may alias pred (in place, as the reference does)
    const int32_t* coeff_in;  // input of the quantizer / inverse when STAGE_FWD is off
    int32_t* coeff_out;     // optional: transform coefficients (packed like the reference's)
    int32_t* qcoeff;
    int32_t* dqcoeff;
    uint16_t* eob;
    uint64_t* energy;       // three_quad_energy, 64-point sizes
    const int16_t* qtables;
    const int16_t* iscan;   // all iscan tables back to back
    const uint8_t* qm_wt;   // optional quantization matrices
    const uint8_t* qm_iwt;
    const void* src;        // STAGE_RESIDUAL: source pixels, geometry of pred
    unsigned long long* dist;  // optional [n_blocks][2]: tuFullDistortion, EbFullLoop.c:808-835
    uint8_t* levels;           // optional padded level maps (av1_txb_init_levels), SURVEY 8f-4
    const uint32_t* levels_off;  // [n_blocks] byte offset of block id's map
    int bd_uniform;         // 8 / 10 when every block of the plan has that bit depth, else 0
    int any_clamp16;        // some block asks for the 8-bit C quantizer variant (META_CLAMP16)
    int iscan_off[19][3];   // [tx_size][scan class] -> offset into iscan
    int qm_off[2][19];      // [plane > 0][tx_size] -> offset into qm_wt / qm_iwt
};

// Switches of a launch, as a policy type.  CfgRuntime reads them from KParams (any stage
// combination, optional outputs); CfgFull fixes them for the headline call -- all three stages,
// qcoeff and dqcoeff written, no coefficient output -- so that the ~40 uniform branches a warp
// would otherwise execute disappear from its instruction stream (16x16: 69.3 -> 64.9 us).
struct CfgRuntime {
    SVT_HD static int stg(const KParams& p) { return p.stages; }
    SVT_HD static bool has_q(const KParams& p) { return p.qcoeff != nullptr; }
    SVT_HD static bool has_dq(const KParams& p) { return p.dqcoeff != nullptr; }
    SVT_HD static bool has_cf(const KParams& p) { return p.coeff_out != nullptr; }
    SVT_HD static bool has_lv(const KParams& p) { return p.levels != nullptr; }
    SVT_HD static bool bd10(const DevBlock& d) { return d.meta & META_BD10; }
    SVT_HD static bool clamp16(const DevBlock& d) { return d.meta & META_CLAMP16; }
    static constexpr int BD_STATIC = 0;  // bit depth not known at compile time
    static constexpr bool SRC_STAGED = false;
};
// CFG = 100 * shape + bit depth (8 or 10: the pixel type is compiled in too).  Shapes:
//   0  headline: forward + quantise + inverse, qcoeff and dqcoeff written
//   1  mode decision: the same on source - prediction, distortion pair out, no
//      coefficient arrays
//   2  tx-type search candidates (8f-2): forward + quantise on source - prediction, distortion
//      pair out, no coefficient arrays, no reconstruction
template <int CFG>
struct CfgFull {
    static constexpr int BD = CFG % 100, SHAPE = CFG / 100, BD_STATIC = BD;
    // 8-bit source - prediction shapes: both pixel blocks are staged with cp.async (prediction
    // where the reconstruction phase expects it, source behind it: 2 x W x H bytes fit the
    // residual's W x H int16) and the column owner subtracts as it loads.  No separate
    // difference pass, no second fetch of the prediction.
    static constexpr bool SRC_STAGED = BD == 8 && SHAPE != 0;
    static constexpr int STAGES = SHAPE == 0 ? STAGE_ALL
                                : (SHAPE == 1 ? (STAGE_ALL | STAGE_RESIDUAL) : (STAGE_FWD | STAGE_QUANT | STAGE_RESIDUAL));
    SVT_HD static constexpr int stg(const KParams&) { return STAGES; }
    SVT_HD static constexpr bool has_q(const KParams&) { return SHAPE == 0; }
    SVT_HD static constexpr bool has_dq(const KParams&) { return SHAPE == 0; }
    SVT_HD static constexpr bool has_cf(const KParams&) { return false; }
    SVT_HD static constexpr bool has_lv(const KParams&) { return false; }
    SVT_HD static constexpr bool bd10(const DevBlock&) { return BD == 10; }
    SVT_HD static constexpr bool clamp16(const DevBlock&) { return false; }
    // the launch-side test for this configuration
    static bool matches(const KParams& p) {
        if (p.stages != STAGES || p.coeff_out || p.levels || p.bd_uniform != BD || p.any_clamp16) return false;
        return SHAPE == 0 ? (p.qcoeff && p.dqcoeff && !p.dist) : (!p.qcoeff && !p.dqcoeff && p.dist);
    }
};

// ---- per-size constants (reference Source/Lib/Common/Codec/EbTransforms.h:121-156, :268-286) ----
SVT_HD constexpr int cmin(int a, int b) { return a < b ? a : b; }
SVT_HD constexpr int cmax(int a, int b) { return a > b ? a : b; }
SVT_HD constexpr int fwd_shift(int w, int h, int i) {
    // {shift0, shift1, shift2}; positive = left shift, negative = rounding right shift
    const int mx = cmax(w, h), mn = cmin(w, h);
    if (w == 64 && h == 64) return i == 0 ? 0 : -2;
    if (w == 32 && h == 64) return i == 0 ? 0 : -2;
    if (w == 64 && h == 32) return i == 0 ? 2 : (i == 1 ? -4 : -2);
    if (w == 16 && h == 64) return i == 0 ? 0 : (i == 1 ? -2 : 0);
    if (w == 64 && h == 16) return i == 0 ? 2 : (i == 1 ? -4 : 0);
    if (i == 0) return 2;
    if (i == 2) return 0;
    if (mx == 4) return 0;
    if (mx == 8) return -1;
    if (mx == 16) return mn == 4 ? -1 : -2;
    return mn == 8 ? -2 : -4;  // mx == 32
}
SVT_HD constexpr int inv_shift0(int w, int h) {
    const int mx = cmax(w, h), mn = cmin(w, h);
    if (mx == 4) return 0;
    if (mx == 8) return mn == 4 ? 0 : 1;
    if (mx == 16) return mn == 16 ? 2 : 1;
    if (mx == 32) return (mn == 32 || mn == 8) ? 2 : 1;
    return (mn == 64 || mn == 16) ? 2 : 1;  // mx == 64
}
SVT_HD constexpr int fwd_cos_col(int w, int h) {
    constexpr int t[5][5] = { { 13, 13, 13, 0, 0 }, { 13, 13, 13, 12, 0 }, { 13, 13, 13, 12, 13 },
                              { 0, 13, 13, 12, 13 }, { 0, 0, 13, 12, 13 } };
    return t[ilog2(w) - 2][ilog2(h) - 2];
}
SVT_HD constexpr int fwd_cos_row(int w, int h) {
    constexpr int t[5][5] = { { 13, 13, 12, 0, 0 }, { 13, 13, 13, 12, 0 }, { 13, 13, 12, 13, 12 },
                              { 0, 12, 13, 12, 11 }, { 0, 0, 12, 11, 10 } };
    return t[ilog2(w) - 2][ilog2(h) - 2];
}
// vtx_tab / htx_tab (EbTransforms.h:87-98), 2 bits per tx_type: 0 DCT 1 ADST 2 FLIPADST 3 IDTX
SVT_HD constexpr uint32_t pack2(const int (&v)[16]) {
    uint32_t r = 0;
    for (int i = 0; i < 16; ++i) r |= uint32_t(v[i]) << (2 * i);
    return r;
}
SVT_HD int vtx_of(int t) {
    constexpr int v[16] = { 0, 1, 0, 1, 2, 0, 2, 1, 2, 3, 0, 3, 1, 3, 2, 3 };
    return (pack2(v) >> (2 * t)) & 3;
}
SVT_HD int htx_of(int t) {
    constexpr int v[16] = { 0, 0, 1, 1, 0, 2, 2, 2, 1, 3, 3, 0, 3, 1, 3, 2 };
    return (pack2(v) >> (2 * t)) & 3;
}
SVT_HD int kind_of(int tx1d) { return tx1d == 3 ? 2 : (tx1d != 0); }
// DevBlock::meta of a block (host side: bucket_blocks and the per-call slots)
SVT_HD uint32_t make_meta(int tx_type, int bit_depth, int plane, bool clamp16, bool use_qm) {
    const int v = vtx_of(tx_type), h = htx_of(tx_type);
    const int cls = tx_type < 10 ? 0 : ((tx_type & 1) ? 2 : 1);  // av1_scan_orders, EbTransforms.h:3349
    return uint32_t(tx_type) | (bit_depth == 10 ? META_BD10 : 0u) | (uint32_t(plane) << META_PLANE_SHIFT) |
           (clamp16 ? META_CLAMP16 : 0u) | (use_qm ? META_USE_QM : 0u) |
           (uint32_t(kind_of(v)) << META_VKIND_SHIFT) | (uint32_t(kind_of(h)) << META_HKIND_SHIFT) |
           (v == 2 ? META_UD : 0u) | (h == 2 ? META_LR : 0u) | (uint32_t(cls) << META_SCAN_SHIFT);
}
SVT_HD int vkind(const DevBlock& d) { return (d.meta >> META_VKIND_SHIFT) & 3; }
SVT_HD int hkind(const DevBlock& d) { return (d.meta >> META_HKIND_SHIFT) & 3; }

SVT_HD constexpr int tx_id(int w, int h) {  // TxSize enum, EbDefinitions.h
    if (w == h) return ilog2(w) - 2;
    constexpr int ws[14] = { 4, 8, 8, 16, 16, 32, 32, 64, 4, 16, 8, 32, 16, 64 };
    constexpr int hs[14] = { 8, 4, 16, 8, 32, 16, 64, 32, 16, 4, 32, 8, 64, 16 };
    for (int i = 0; i < 14; ++i)
        if (ws[i] == w && hs[i] == h) return 5 + i;
    return -1;
}

// ---- CTA geometry and shared-memory layout --------------------------------------------------------
//   buf   int32 [NB][BLK]      2-D transpose buffer; row pitch LD = W + 4 keeps every row 16-byte
//                              aligned (128-bit row loads/stores) and conflict-free; block pitch
//                              BLK == W (mod 32) keeps blocks that share a warp in disjoint banks
//   stage int16 [NB][H][W]     residual staged by 128-bit coalesced copies (cp.async); after the
//                              forward column pass the same bytes hold the prediction block and
//                              then the reconstruction (pixels), copied back out 128 bits at a time
#ifndef SVT_OCC16  // experiment knobs: resident 256-thread CTA equivalents for the 16- / 32- / 64-point sizes
#define SVT_OCC16 4
#endif
#ifndef SVT_OCC32
#define SVT_OCC32 4  // v15: 41.6 vs 42.4 us for 32x32 at 64 registers (no spill in the flat kernels); 3 for the qmatrix / distortion variants
#endif
#ifndef SVT_OCC64
#define SVT_OCC64 2
#endif
#ifndef SVT_OCC8  // experiment knob: resident 256-thread CTA equivalents for the 4- and 8-point sizes
#define SVT_OCC8 5  // 6 (42 registers per thread) measured slower: 8x8 51.4 vs 50.5 us, 4x4 23.1 vs 21.2
#endif
template <int W, int H>
struct Geo {
    static constexpr int TXS = tx_id(W, H);
    // threads per CTA: 256 for the three sizes whose buckets fill the GPU many times over and run
    // fastest that way (16x16 70 vs 73 us, 32x32 51 vs 52, 4x4 equal); 128 elsewhere -- the same
    // warps per SM in twice as many CTAs fills the small grids better (8x8 55 vs 57 us,
    // 16x32 14.6 vs 17.0, 32x16 14.1 vs 16.9, 8x16 13.5 vs 14.8, 64x64 25.3 vs 27.2)
#ifdef SVT_NT_ALL
    static constexpr int NT = SVT_NT_ALL;
#else
    static constexpr int NT = (W == H && (W == 16 || W == 32 || W == 4)) ? 256 : 128;
#endif
    // R: how many times a thread runs each phase (on R different blocks; the phase loops are
    // rolled).  R = 2 was measured slower (fewer resident warps), see DESIGN.md section 6.
    static constexpr int R = 1;
    static constexpr int NB = R * NT / cmin(W, H);    // blocks per CTA
    static constexpr int LD = W + 4;
    static constexpr int BLK = ((H * LD - W + 31) / 32) * 32 + W;
    static constexpr int KW = cmin(W, 32), KH = cmin(H, 32);  // kept (coded) region
    static constexpr int LOG_SCALE = (W * H > 256) + (W * H > 1024);
    static constexpr bool RECT = (ilog2(W) - ilog2(H) == 1) || (ilog2(H) - ilog2(W) == 1);
    static constexpr int NCLS = cmax(W, H) >= 32 ? 1 : 3;     // 32/64-point: DCT_DCT / IDTX only
    static constexpr int BUF_WORDS = NB * BLK;
    // staged block pitch (int16 units): one extra row so that blocks sharing a warp start in
    // different banks
    static constexpr int SP = W * H + W;
    static constexpr int STAGE_WORDS = (NB * SP + 1) / 2;
    static constexpr int SMEM_BYTES = 4 * (BUF_WORDS + STAGE_WORDS);
    // resident CTAs per SM the register allocation is capped for (__launch_bounds__): what shared
    // memory allows, but no more than 5 / 4 / 3 / 2 for 8- / 16- / 32- / 64-point sizes -- measured on B200,
    // CTAs with 64 / 80 / 128 registers per thread beat more, register-starved ones (16x16: 69 vs
    // 74 us at 5 CTAs, 8x8: 58 vs 67 us at 8; 32x32: 56 us at 3 vs 60 us at 2)
    // HEAVY: the qmatrix and distortion variants carry 64-bit arithmetic and spill at the tighter cap
    static constexpr int ctas_per_sm(bool heavy) {
        return cmin((cmax(W, H) >= 64 ? SVT_OCC64 : (cmax(W, H) >= 32 ? (heavy ? 3 : SVT_OCC32) : (cmax(W, H) >= 16 ? SVT_OCC16 : SVT_OCC8))) * 256 / NT,
                    cmax(1, (228 * 1024) / (SMEM_BYTES + 1024 + 108 * NB)));
    }
    static constexpr int CTAS_PER_SM = ctas_per_sm(false);
};

// one of {DC, AC}.  Where the flat quantizer runs in its two-multiply form (two_mul below),
// round / quant / qshift hold the pre-shifted terms 2 round, (quant + 2^16) << 15 and
// quant_shift << (16 + log_scale) instead of the table values.
struct QSet { int zbin, round, quant, qshift, dequant; };
struct QRow { QSet dc, ac; };  // zbin / round already scaled by log_scale
template <int NB>
struct CtaShared {  // small per-CTA state next to the buffers above (108 bytes per block)
    DevBlock blk[NB];
    QRow q[NB];  // this block's quantizer row, pre-scaled
    int eob[NB];
    unsigned long long energy[NB];
    unsigned long long dist[NB][2];  // sum (coeff - dq)^2, sum coeff^2 over the coded region
    int fast[NB];  // the block is inside fast_limit: its inverse transform runs without clamps
};

// read-only global load through L1 (LDG.CONSTANT on the device)
template <class T>
SVT_HD T ld_ro(const T* p) {
#if defined(__CUDA_ARCH__)
    return __ldg(p);
#else
    return *p;
#endif
}

// ---- staging copies ---------------------------------------------------------------------------------
// rows x row_bytes between a strided global block and a dense shared block, `chunk` bytes at a
// time (16, 8, 4, 2 or 1: the widest the addresses allow).  On the device the global->shared
// direction uses cp.async (LDGSTS) for chunks >= 4 bytes, so the copy is in flight while the CTA
// does other work; the host build is the plain loop.
SVT_HD int align_chunk(size_t addr_bits, int row_bytes) {
    int c = 16;
    while (c > 1 && ((addr_bits | size_t(row_bytes)) & size_t(c - 1))) c >>= 1;
    return c;
}
template <int CH>
SVT_HD void copy_chunk(void* dst, const void* src) {
    if constexpr (CH == 16) *static_cast<int4*>(dst) = *static_cast<const int4*>(src);
    else if constexpr (CH == 8) *static_cast<long long*>(dst) = *static_cast<const long long*>(src);
    else if constexpr (CH == 4) *static_cast<int*>(dst) = *static_cast<const int*>(src);
    else if constexpr (CH == 2) *static_cast<short*>(dst) = *static_cast<const short*>(src);
    else *static_cast<char*>(dst) = *static_cast<const char*>(src);
}
template <int CH>
SVT_HD void g2s_chunk(void* smem_dst, const void* gsrc) {
#if defined(__CUDA_ARCH__)
    if constexpr (CH >= 4) {
        const unsigned sa = static_cast<unsigned>(__cvta_generic_to_shared(smem_dst));
        if constexpr (CH == 16)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(sa), "l"(gsrc) : "memory");
        else
            asm volatile("cp.async.ca.shared.global [%0], [%1], %2;\n" ::"r"(sa), "l"(gsrc), "n"(CH) : "memory");
        return;
    }
#endif
    copy_chunk<CH>(smem_dst, gsrc);
}
SVT_HD void g2s_commit_wait() {
#if defined(__CUDA_ARCH__)
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;\n" ::: "memory");
#endif
}
// one block: global (strided) <-> shared (dense) by a copy team of NTHR threads.
// `flip`: global row r lands in shared row ROWS-1-r (and comes back from it).  The up/down flip of
// the FLIPADST types (EbTransforms.c:3739-3744, :8258-8264) is folded into the copies this way,
// so the transform phases never branch on it.
// Fast path: base address and stride aligned to min(16, ROW_BYTES) -- every trip count and
// index split is a compile-time constant.  Anything else (the encoder may pass any offset) goes
// through one shared out-of-line loop with the widest chunk the addresses allow.
SVT_NOINLINE void stage_slow(char* s, char* g, int rows, int row_bytes, size_t gstride, bool flip,
                             int tid, int nthr, bool in) {
    const int ch = align_chunk(size_t(g) | gstride, row_bytes);
    const int per_row = row_bytes / ch;
    for (int i = tid; i < rows * per_row; i += nthr) {
        const int r = i / per_row, k = i - r * per_row;
        char* sp = s + (flip ? rows - 1 - r : r) * row_bytes + k * ch;
        char* gp = g + size_t(r) * gstride + k * ch;
        char* dst = in ? sp : gp;
        const char* src = in ? gp : sp;
        switch (ch) {
        case 16: copy_chunk<16>(dst, src); break;
        case 8: copy_chunk<8>(dst, src); break;
        case 4: copy_chunk<4>(dst, src); break;
        case 2: copy_chunk<2>(dst, src); break;
        default: copy_chunk<1>(dst, src); break;
        }
    }
}
template <int ROWS, int ROW_BYTES, int NTHR, bool IN>
SVT_HD void stage_copy(char* s, char* g, size_t gstride, bool flip, int tid) {
    constexpr int CH = ROW_BYTES < 16 ? ROW_BYTES : 16;
    if ((size_t(g) | gstride) & size_t(CH - 1)) {
        stage_slow(s, g, ROWS, ROW_BYTES, gstride, flip, tid, NTHR, IN);
        return;
    }
    constexpr int PER_ROW = ROW_BYTES / CH, TOTAL = ROWS * PER_ROW, ITERS = (TOTAL + NTHR - 1) / NTHR;
#pragma unroll
    for (int it = 0; it < ITERS; ++it) {
        const int i = tid + it * NTHR;
        if (TOTAL % NTHR == 0 || i < TOTAL) {
            const int r = i / PER_ROW, k = i % PER_ROW;
            char* sp = s + (flip ? ROWS - 1 - r : r) * ROW_BYTES + k * CH;
            char* gp = g + size_t(r) * gstride + k * CH;
            if constexpr (IN) g2s_chunk<CH>(sp, gp);
            else copy_chunk<CH>(gp, sp);
        }
    }
}

// residual = source - prediction while staging (residual_kernel_c / residual_kernel16bit,
// EbPictureOperators.c:134-191): two plain loads per chunk, int16 differences into the staged block.
template <int ROWS, int W, int NTHR, class PIX>
SVT_HD void stage_diff(int16_t* s, const PIX* src, const PIX* prd, size_t gstride, bool flip, int tid) {
    constexpr int PB = int(sizeof(PIX));
    constexpr int CH = W * PB < 16 ? W * PB : 16;  // bytes per chunk
    constexpr int NPX = CH / PB;                   // pixels per chunk
    constexpr int PER_ROW = W / NPX, TOTAL = ROWS * PER_ROW, ITERS = (TOTAL + NTHR - 1) / NTHR;
    const bool fast = !((size_t(src) | size_t(prd) | (gstride * PB)) & size_t(CH - 1));
#pragma unroll
    for (int it = 0; it < ITERS; ++it) {
        const int i = tid + it * NTHR;
        if (TOTAL % NTHR == 0 || i < TOTAL) {
            const int r = i / PER_ROW, k = i % PER_ROW;
            int16_t* d = s + (flip ? ROWS - 1 - r : r) * W + k * NPX;
            const PIX* a = src + size_t(r) * gstride + k * NPX;
            const PIX* b = prd + size_t(r) * gstride + k * NPX;
            PIX va[NPX], vb[NPX];
            if (fast) {
                copy_chunk<CH>(va, a);
                copy_chunk<CH>(vb, b);
            } else {
#pragma unroll
                for (int j = 0; j < NPX; ++j) { va[j] = a[j]; vb[j] = b[j]; }
            }
#pragma unroll
            for (int j = 0; j < NPX; ++j) d[j] = int16_t(int(va[j]) - int(vb[j]));
        }
    }
}

// ---- 1-D dispatch with the kind as a template argument ----------------------------------------------
// Each variant contains its own loads and stores, so the three kinds never merge register state.
// POST: round_shift(., POST) of the outputs (av1_round_shift_array_c), folded into the last
// butterfly stage for DCT / ADST (txfm1d.cuh: hbp) and applied separately after the identity
template <int N, int BIT, int KIND, int POST = 0>
SVT_HD void fwd1d_k(int (&x)[N]) {
    if constexpr (KIND == 0) fdct<N, BIT, POST>(x);
    else if constexpr (KIND == 1) fadst<N, BIT, POST>(x);
    else {
        identity<N>(x);
        if constexpr (POST > 0) {
#pragma unroll
            for (int i = 0; i < N; ++i) x[i] = rshift<(POST > 0 ? POST : 1)>(x[i]);
        }
    }
}
template <int N, int KIND, class K>
SVT_HD void inv1d_k(int (&x)[N], K k) {
    if constexpr (KIND == 0) idct<N, 12>(x, k);
    else if constexpr (KIND == 1) iadst<N, 12>(x, k);
    else identity<N>(x);
}

#if defined(__CUDA_ARCH__)
template <int OFF>
__device__ __forceinline__ int lds_s16_off(unsigned base) {
    int v;
    // volatile keeps it ordered against the barriers and the cp.async waits (all of them volatile asm
    // or barrier intrinsics); no memory clobber, so ordinary loads and stores may still move past it
    asm volatile("ld.shared.s16 %0, [%1+%2];" : "=r"(v) : "r"(base), "n"(OFF));
    return v;
}
template <int STRIDE_BYTES, int N, int... R>
__device__ __forceinline__ void lds_s16_col(unsigned base, int (&x)[N], std::integer_sequence<int, R...>) {
    ((x[R] = lds_s16_off<R * STRIDE_BYTES>(base)), ...);
}
#endif
// ---- phase 1: forward column transform, staged residual -> transpose buffer ------------------------
// SRC8: `res` holds the 8-bit prediction block (W x H bytes) followed by the 8-bit source block
// instead of an int16 residual (CfgFull::SRC_STAGED)
template <int W, int H, int KIND, bool SRC8 = false>
SVT_HD void fwd_col_k(const DevBlock& d, int c, int* buf, const int16_t* res) {
    using G = Geo<W, H>;
    constexpr int S0 = fwd_shift(W, H, 0), S1 = -fwd_shift(W, H, 1);
    const bool lr = d.meta & META_LR;
    int x[H];
    if constexpr (SRC8) {
        const uint8_t* px = reinterpret_cast<const uint8_t*>(res);
#pragma unroll
        for (int r = 0; r < H; ++r) x[r] = (int(px[W * H + r * W + c]) - int(px[r * W + c])) * (1 << S0);
    } else {
#if defined(__CUDA_ARCH__) && !defined(SVT_NO_LDS_S16)
        // nvcc turns an int16 shared load into LDS.U16 + PRMT; ld.shared.s16 is one LDS.S16
        lds_s16_col<W * 2>(unsigned(__cvta_generic_to_shared(res + c)), x, std::make_integer_sequence<int, H>());
#pragma unroll
        for (int r = 0; r < H; ++r) x[r] *= (1 << S0);  // rows already flipped
#else
#pragma unroll
        for (int r = 0; r < H; ++r) x[r] = int(res[r * W + c]) * (1 << S0);  // rows already flipped
#endif
    }
    fwd1d_k<H, fwd_cos_col(W, H), KIND, (S1 > 0 ? S1 : 0)>(x);  // includes round_shift(., -shift[1])
    int* dst = buf + (lr ? W - 1 - c : c);  // flip left to right, EbTransforms.c:3752-3757
#pragma unroll
    for (int r = 0; r < H; ++r) dst[r * G::LD] = x[r];
}
template <int W, int H, bool SRC8 = false>
SVT_HD void phase_fwd_col(const DevBlock& d, int c, int* buf, const int16_t* res) {
    const int kind = vkind(d);
    if constexpr (H == 64) fwd_col_k<W, H, 0, SRC8>(d, c, buf, res);
    else if constexpr (H == 32) {
        if (kind == 2) fwd_col_k<W, H, 2, SRC8>(d, c, buf, res); else fwd_col_k<W, H, 0, SRC8>(d, c, buf, res);
    } else {
        if (kind == 0) fwd_col_k<W, H, 0, SRC8>(d, c, buf, res);
        else if (kind == 1) fwd_col_k<W, H, 1, SRC8>(d, c, buf, res);
        else fwd_col_k<W, H, 2, SRC8>(d, c, buf, res);
    }
}

// ---- quantizer for one coefficient ---------------------------------------------------------------
// Two-multiply form of the flat quantizer: for log_scale < 2 and not in the kernels that also run
// the qmatrix quantizer (which needs the table values) nor in the stand-alone slots (WIDE: their
// callers hand in arbitrary rows).  -DSVT_QUANT_WIDE switches it off (A/B builds).
SVT_HD constexpr bool two_mul(int log_scale, bool wide) {
#if defined(SVT_QUANT_WIDE)
    return false && log_scale < 2 && !wide;
#else
    return log_scale < 2 && !wide;
#endif
}
// one table row (raw zbin / round / quant / quant_shift / dequant of one position class) -> QSet
template <int LOG_SCALE, bool WIDE>
SVT_HD QSet make_qset(int zbin, int round, int quant, int qshift, int dequant) {
    constexpr int rnd = LOG_SCALE ? (1 << (LOG_SCALE - 1)) : 0;
    QSet e;
    // ROUND_POWER_OF_TWO(zbin / round, log_scale), EbFullLoop.c:246, :280
    e.zbin = (zbin + rnd) >> LOG_SCALE;
    e.round = (round + rnd) >> LOG_SCALE;
    e.quant = quant; e.qshift = qshift; e.dequant = dequant;
    if constexpr (two_mul(LOG_SCALE, WIDE)) {
        e.round = int(uint32_t(e.round) << 1);
        e.quant = int(uint32_t(quant + 65536) << 15);         // quant is an int16: 2^15 <= quant + 2^16 < 2^17
        e.qshift = int(uint32_t(qshift) << (16 + LOG_SCALE));  // needs 0 <= quant_shift < 2^(16 - log_scale)
    }
    return e;
}
template <int LOG_SCALE, bool WIDE>
SVT_HD QRow load_qrow(const int16_t* t) {
    QRow q;
    q.dc = make_qset<LOG_SCALE, WIDE>(t[0], t[8], t[16], t[24], t[32]);
    q.ac = make_qset<LOG_SCALE, WIDE>(t[1], t[9], t[17], t[25], t[33]);
    return q;
}
SVT_HD uint32_t umulhi32(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
    return __umulhi(a, b);
#else
    return uint32_t((uint64_t(a) * b) >> 32);
#endif
}
// flat (no qmatrix): aom_highbd_quantize_b*_avx2 arithmetic
// (Source/Lib/Common/ASM_AVX2/highbd_quantize_intrin_avx2.c:80-125), equal to
// highbd_quantize_b_helper_c (EbFullLoop.c:239) for the tables av1_build_quantizer produces.
// Branch-free: the level is computed for every position and zeroed below zbin.  t1max is 32767
// for the 8-bit C variant (EbFullLoop.c:86-88) and INT_MAX otherwise.  Returns |level|.
// asum: running sum of |dqcoeff| (the bound that lets the inverse transform drop its clamps, see
// fast_limit below)
// WIDE: the 64-bit form for every log_scale (the stand-alone quantizer slots, whose callers hand
// in their own table rows)
template <int LOG_SCALE, bool WIDE = false>
SVT_HD int quant_flat(int cf, const QSet& q, int t1max, int& qc, int& dqc, unsigned& asum) {
    const int sgn = (cf >> 31) | 1;
    const int ab = cf < 0 ? -cf : cf;
    int aq;
    constexpr bool kTwoMul = two_mul(LOG_SCALE, WIDE);
    if constexpr (kTwoMul) {
        // ((t1 * quant) >> 16) + t1 == (t1 * (quant + 2^16)) >> 16 == umulhi(2 t1, (quant + 2^16) << 15), and
        // (t2 * quant_shift) >> (16 - log_scale) == umulhi(t2, quant_shift << (16 + log_scale)): two
        // multiplies and no 64-bit shifts (17 -> 14 instructions per coefficient).  Exact for
        // 0 <= t1 < 2^31 and 0 <= quant_shift < 2^(16 - log_scale); invert_quant produces
        // quant_shift = 2^(16 - l) with l >= 2 (EbModeDecisionConfigurationProcess.c:322), and
        // the table upload refuses anything else.
        const uint32_t r2 = uint32_t(q.round);  // 2 round
        const uint32_t t1x2 = t1max == 0x7fffffff ? (uint32_t(ab) << 1) + r2
                                                  : uint32_t(ab + int(r2 >> 1) > t1max ? t1max : ab + int(r2 >> 1)) << 1;
        aq = int(umulhi32(umulhi32(t1x2, uint32_t(q.quant)), uint32_t(q.qshift)));
    } else {  // 64-point sizes: quant_shift << 18 can reach 2^32 (quant_shift = 2^14 at qindex 0)
        int t1 = ab + q.round;
        t1 = t1 > t1max ? t1max : t1;
        const int64_t p1 = int64_t(t1) * q.quant;
        const uint32_t t2 = uint32_t(uint64_t(p1) >> 16) + uint32_t(t1);
        const uint64_t p2 = uint64_t(t2) * uint32_t(q.qshift);
        aq = int(uint32_t(p2 >> (16 - LOG_SCALE)));
    }
    aq = ab >= q.zbin ? aq : 0;
    qc = int(unsigned(aq) * unsigned(sgn));
    if constexpr (LOG_SCALE == 0) {
        dqc = int(unsigned(qc) * unsigned(q.dequant));
        asum = unsigned(aq) * unsigned(q.dequant) + asum;  // one multiply-add on the FMA pipe
    } else {
        const int adq = int(unsigned(aq) * unsigned(q.dequant)) >> LOG_SCALE;
        dqc = int(unsigned(adq) * unsigned(sgn));
        asum += unsigned(adq);
    }
    return aq;
}
template <int LOG_SCALE, bool WIDE = false>
SVT_HD int quant_flat(int cf, const QSet& q, int t1max, int& qc, int& dqc) {
    unsigned unused = 0;
    return quant_flat<LOG_SCALE, WIDE>(cf, q, t1max, qc, dqc, unused);
}
// with quantization matrices: highbd_quantize_b_helper_c, EbFullLoop.c:239-296
template <int LOG_SCALE>
SVT_HD int quant_qm(int cf, const QSet& q, int wt, int iwt, int& qc, int& dqc, unsigned& asum) {
    const int sign = cf >> 31;
    const int ab = (cf ^ sign) - sign;
    const int cw = int(unsigned(cf) * unsigned(wt));
    qc = 0; dqc = 0;
    int aq = 0;
    if (cw >= q.zbin * 32 || cw <= -q.zbin * 32) {
        const int64_t tw = (int64_t(ab) + q.round) * wt;
        const int64_t t2 = ((tw * q.quant) >> 16) + tw;
        aq = int((t2 * q.qshift) >> (16 - LOG_SCALE + 5));
        const int dq = (q.dequant * iwt + 16) >> 5;
        const int adq = int(unsigned(aq) * unsigned(dq)) >> LOG_SCALE;
        qc = (aq ^ sign) - sign;
        dqc = (adq ^ sign) - sign;
        asum += unsigned(adq);
    }
    return aq;
}

// ---- when the inverse transform needs no clamps ----------------------------------------------------
// The reference clamps the input of the inverse row pass to bd + 8 bits, every add/sub of the row
// networks to the same range, the input of the column pass to 16 bits and every add/sub of the
// column networks to 16 bits (EbTransforms.c:8215-8291, clamp_value :5458).  None of those can
// bind for a block whose dequantised coefficients are small in L1 norm S = sum |dqcoeff|:
//   * every add/sub node of the inverse DCT / ADST networks is a combination sum a_i x_i of the
//     inputs with |a_i| <= 1 (checked numerically for every size by tests/host_emul.cpp:
//     emul_inverse_node_gain), so |node| <= sum |x_i| + E, E = accumulated half_btf rounding
//     (0.5 per rotation, fewer than 512 rotations in the largest network);
//   * a row's inputs are a subset of the block's coefficients (rect scaling only shrinks them), so
//     S + E < 2^(bd+7) covers the input clamp and every row node;
//   * a column's inputs are row outputs shifted down by inv_shift0 = s0; an identity row scales
//     by G <= 4, the other kinds by <= 1, so a column's L1 norm is at most
//     (G S + KH E) / 2^s0 + KH, which must stay below 2^15 - E.
// Blocks inside the limit run the clamp-free networks (forward-transform cost: one instruction per
// add/sub instead of two); the others run the clamped ones.  Same results by construction, and
// both are pinned to the reference by the tests (small and full-range residuals).
constexpr int kFastE = 512;
SVT_HD constexpr int id_gain(int n) { return n == 4 ? 2 : (n == 8 ? 2 : (n == 16 ? 3 : 4)); }
template <int W, int H, int KIND_ROW>
SVT_HD constexpr unsigned fast_limit(bool bd10) {
    constexpr int KH = cmin(H, 32), S0 = inv_shift0(W, H), G = KIND_ROW == 2 ? id_gain(W) : 1;
    constexpr long long col = ((((32767LL - kFastE - KH) << S0) - (long long)KH * kFastE)) / G;
    const long long row = (1LL << ((bd10 ? 10 : 8) + 7)) - 1 - kFastE;
    return unsigned(col < row ? col : row);
}
// host emulation only: the decision comes from the caller (who runs the block's rows twice), and
// the row's share of S goes back to it
struct RowAux { unsigned sum = 0; int fast = 0; };

// ---- qcoeff / dqcoeff of 8- and 16-wide rows: 256-bit stores by the row owner ------------------------
// sm_100 has 256-bit global stores (st.global.v8.b32 -> STG.E.ENL2.256): one full 32-byte sector
// per thread and instruction.  In the flat kernels the owner of an 8- or 16-wide row stores its
// q / dq row itself, eight coefficients at a time.  What it replaces: for 16-wide rows the
// hand-over through shared memory described below (8 STS.128 + 8 LDS.128 + 4 warp syncs per
// thread, a third of the 16x16 kernel's shared-memory wavefronts: 54.9 -> 53.3 us); for 8-wide
// rows two 128-bit stores per array that each wrote HALF a sector (8x8: 51.5 -> 46.0 us -- partial
// sector writes are what hurts: the same direct stores at 128 bits took the 32x32 kernel from
// 39.6 to 59.0 us).  32-wide rows keep the hand-over: eight levels live next to the 32 dequantised
// values spill in the 64-register kernels (44.1 vs 39.6 us).  Needs 32-byte aligned rows
// (checked per block; 16-byte aligned rows take two 128-bit stores).
#if !defined(SVT_NO_DIRECT256)
#define SVT_DIRECT256 1
#else
#define SVT_DIRECT256 0
#endif
#if !defined(SVT_DIRECT256_MAXKW)
#define SVT_DIRECT256_MAXKW 16  // 32-wide rows: eight levels live at once spill in the 64-register kernels (measured slower)
#endif
#if !defined(SVT_DIRECT256_MINKW)
#define SVT_DIRECT256_MINKW 8
#endif
#if !defined(SVT_DIRECT256_QM)
#define SVT_DIRECT256_QM 1  // the qmatrix kernels too (both of their quantizer loops)
#endif
SVT_HD constexpr bool direct256(bool qm, int kw) {
    return SVT_DIRECT256 && (SVT_DIRECT256_QM || !qm) && kw >= SVT_DIRECT256_MINKW && kw <= SVT_DIRECT256_MAXKW;
}
SVT_HD void store8(int32_t* g, const int* v, bool al32) {
#if defined(__CUDA_ARCH__)
    if (al32) {
        asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(g), "r"(v[0]), "r"(v[1]), "r"(v[2]),
                     "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]) : "memory");
        return;
    }
#else
    (void)al32;
#endif
    *reinterpret_cast<int4*>(g) = make_int4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<int4*>(g + 4) = make_int4(v[4], v[5], v[6], v[7]);
}
// ---- qcoeff / dqcoeff leave the SM through the thread's own transpose-buffer row --------------------
// A thread owns one row, so a direct store would have the 32 lanes of a warp write 16 bytes each
// at a pitch of one row: 32 different cache lines per instruction, 8x the L1 wavefronts of a
// coalesced store (the q/dq stores alone were ~25 % of the kernel).  Instead the quantized row is
// parked in the thread's buffer row (free between the forward read and the inverse write), and the
// LB lanes that share a block copy the block's coefficients out together, consecutive lanes
// writing consecutive 16-byte chunks; then the same for the dequantized row.  Only __syncwarp over
// the lanes of one block is needed: they always run the same code (same block, same tx type).
// The host build (emulation tests) has one task at a time and stores directly.
template <int W, int H, class C = CfgRuntime>
SVT_HD void emit_q(const KParams& p, int* row, uint32_t cbase, int c0, const int (&qv)[4], const int (&x)[W]) {
#if defined(__CUDA_ARCH__)
    if constexpr (Geo<W, H>::KW >= 16) {
        *reinterpret_cast<int4*>(row + c0) = make_int4(qv[0], qv[1], qv[2], qv[3]);
        return;
    }
#endif
    {
        // rows of <= 32 bytes: consecutive lanes already write (nearly) consecutive addresses
        if (C::has_q(p)) *reinterpret_cast<int4*>(p.qcoeff + cbase + c0) = make_int4(qv[0], qv[1], qv[2], qv[3]);
        if (C::has_dq(p))
            *reinterpret_cast<int4*>(p.dqcoeff + cbase + c0) = make_int4(x[c0], x[c0 + 1], x[c0 + 2], x[c0 + 3]);
    }
}
template <int W, int H, class C = CfgRuntime>
SVT_HD void flush_q(const KParams& p, const DevBlock& d, int r, int* row, const int (&x)[W]) {
#if defined(__CUDA_ARCH__)
    using G = Geo<W, H>;
    constexpr int KW = G::KW;
    if constexpr (KW < 16) return;
    constexpr int LB = H < 32 ? H : 32;   // lanes of a warp that share a block (its coded rows)
    constexpr int CPR = KW / 4;           // 16-byte chunks per coded row
    const int lane = threadIdx.x & 31;
    const int li = r & (LB - 1);
    const unsigned bmask = LB == 32 ? 0xffffffffu : (((1u << LB) - 1u) << (lane - li));
    const int* rows0 = row - li * G::LD;  // first row of the group
    const uint32_t gbase = d.coeff_off + uint32_t(r - li) * KW;
    // Which 16-byte chunk (row-major index into the group's rows) lane li copies in step i.
    // 32-wide rows: consecutive lanes take consecutive chunks (a quarter-warp reads one whole row:
    // conflict-free at the 36-word pitch).  16-wide rows at the 20-word pitch: two consecutive rows
    // overlap in one bank group, so a quarter-warp reading rows (a, a + 1) paid a 2-way conflict on
    // every hand-over load -- 16 extra wavefronts per 16x16 block, two thirds of the kernel's
    // shared-memory conflicts (ncu, profiles/r02_v13_*).  Rows (a, a + 4) are exactly half a bank
    // cycle apart (5 * 4 chunks = 4 mod 8), so each quarter-warp takes those; a group of four lanes
    // still writes one whole 64-byte row to global memory (full sectors).
    auto chunk_of = [&](int i) {
        if constexpr (KW == 16 && LB >= 8) {
            const int t = li & 3, g = li >> 2, m = g & 1, Q = g >> 1;
            constexpr int ROWS_PER_STEP = LB / 4;  // 2, 4 or 8 rows of the group per step
            const int row = ROWS_PER_STEP == 2 ? i + 4 * m
                          : (ROWS_PER_STEP == 4 ? 8 * (i >> 1) + 2 * (i & 1) + Q + 4 * m : 8 * i + Q + 4 * m);
            return row * CPR + t;
        } else {
            return li + LB * i;
        }
    };
    // either output may be absent (a caller that only wants eob, distortion and reconstruction)
    if (C::has_q(p)) {
        __syncwarp(bmask);
#pragma unroll
        for (int i = 0; i < CPR; ++i) {
            const int j = chunk_of(i);
            *reinterpret_cast<int4*>(p.qcoeff + gbase + j * 4) =
                *reinterpret_cast<const int4*>(rows0 + (j / CPR) * G::LD + (j % CPR) * 4);
        }
        __syncwarp(bmask);
    }
    if (C::has_dq(p)) {
#pragma unroll
        for (int c = 0; c < KW; c += 4)
            *reinterpret_cast<int4*>(row + c) = make_int4(x[c], x[c + 1], x[c + 2], x[c + 3]);
        __syncwarp(bmask);
#pragma unroll
        for (int i = 0; i < CPR; ++i) {
            const int j = chunk_of(i);
            *reinterpret_cast<int4*>(p.dqcoeff + gbase + j * 4) =
                *reinterpret_cast<const int4*>(rows0 + (j / CPR) * G::LD + (j % CPR) * 4);
        }
        __syncwarp(bmask);
    }
#endif
}

// DIST: also accumulate this row's share of sum (coeff - dq)^2 and sum coeff^2 over the coded
// region (full_distortion_kernel32_bits, EbPictureOperators.c:283-313) into dsum[0..1].
// The inverse row pass of one row (inv_txfm2d_add_c row pass, EbTransforms.c:8215-8236), from the
// dequantised row in x[] to the transposing store.  FAST: the block is inside fast_limit, so no
// clamp can bind and none is executed.
template <int W, int H, int KIND, class C, bool FAST>
SVT_HD void inv_row_store(int (&x)[W], int* row, bool bd10) {
    using G = Geo<W, H>;
    constexpr int KW = G::KW;
    const Clamp krow_c(bd10 ? 18 : 16);  // clamp_buf(bd + 8) and the row stage range: 16 / 18
#pragma unroll
    for (int c = 0; c < W; ++c) {
        if (c >= KW) x[c] = 0;  // 64-wide: re-expansion with zeros, EbTransforms.c:8299
        else {
            if constexpr (G::RECT) x[c] = mul_round12(x[c], 2896);
            if constexpr (!FAST) x[c] = krow_c(x[c]);
        }
    }
    if constexpr (FAST) inv1d_k<W, KIND>(x, NoClamp());
    else inv1d_k<W, KIND>(x, krow_c);
    constexpr int IS0 = inv_shift0(W, H);
    // clamp_buf(max(bd + 6, 16)) = 16 bits at the input of the column pass
    // (EbTransforms.c:8241) is applied here, before the transposing store.  With 8-bit
    // blocks it cannot bind after a DCT row (its last stage is a 16-bit clamped add/sub and
    // round_shift only shrinks) nor after an 8/16-point ADST row followed by a shift (the
    // sign permutation can turn -2^15 into 2^15, which the shift halves), so it is elided
    // where the bit depth is a compile-time 8.
    constexpr bool IN_RANGE = FAST || (C::BD_STATIC == 8 && (KIND == 0 || (KIND == 1 && W >= 8 && IS0 > 0)));
    const Clamp k16(16);
#pragma unroll
    for (int c = 0; c < W; c += 4) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if constexpr (IS0 > 0) x[c + j] = rshift<(IS0 > 0 ? IS0 : 1)>(x[c + j]);
            if constexpr (!IN_RANGE) x[c + j] = k16(x[c + j]);
        }
        *reinterpret_cast<int4*>(row + c) = make_int4(x[c], x[c + 1], x[c + 2], x[c + 3]);
    }
}

// ---- phase 2: forward row -> (energy, pack) -> quantize -> inverse row, all on one thread's registers
// isc: this block's iscan table (global memory).  Returns the row's eob candidate and
// discarded-region energy.
// fast_flag: the block's shared "no clamp can bind" flag (written here, read by the column pass)
template <int W, int H, int KIND, bool QM, bool DIST, class C = CfgRuntime>
SVT_HD void row_k(const KParams& p, const DevBlock& d, const QRow& q, int r, int* buf, const int16_t* isc,
                  int& eob_max, unsigned long long& energy_sum, unsigned long long (&dsum)[2],
                  int* fast_flag, RowAux& aux, int* eob_slot = nullptr) {
    static_assert(!(QM && DIST), "the distortion outputs are built for the flat quantizer only");
    using G = Geo<W, H>;
    dsum[0] = dsum[1] = 0;
    constexpr int KW = G::KW, KH = G::KH, LS = G::LOG_SCALE;
    const bool bd10 = C::bd10(d);
    int* row = buf + r * G::LD;
    int x[W];
    eob_max = 0;
    energy_sum = 0;
    unsigned asum = 0;  // this row's share of S = sum |dqcoeff| (fast_limit)
    const uint32_t cbase = d.coeff_off + r * KW;

    if (C::stg(p) & STAGE_FWD) {
#pragma unroll
        for (int c = 0; c < W; c += 4) {
            const int4 v = *reinterpret_cast<const int4*>(row + c);
            x[c] = v.x; x[c + 1] = v.y; x[c + 2] = v.z; x[c + 3] = v.w;
        }
        constexpr int S2 = -fwd_shift(W, H, 2);
        fwd1d_k<W, fwd_cos_row(W, H), KIND, (S2 > 0 ? S2 : 0)>(x);  // includes round_shift(., -shift[2])
        if constexpr (G::RECT) {
#pragma unroll
            for (int c = 0; c < W; ++c) x[c] = mul_round12(x[c], 5793);  // EbTransforms.c:3769-3776
        }
        if constexpr (KW != W || KH != H) {  // HandleTransform*_c, EbTransforms.c:4351-4731
#pragma unroll
            for (int c = 0; c < W; ++c)
                if (r >= KH || c >= KW)
                    energy_sum += (unsigned long long)(int64_t(x[c]) * int64_t(x[c]));
        }
        if (C::stg(p) & STAGE_FWD_FULL) {
#pragma unroll
            for (int c = 0; c < W; c += 4)
                *reinterpret_cast<int4*>(p.coeff_out + d.coeff_off + r * W + c) =
                    make_int4(x[c], x[c + 1], x[c + 2], x[c + 3]);
        } else if (C::has_cf(p) && r < KH) {
#pragma unroll
            for (int c = 0; c < KW; c += 4)
                *reinterpret_cast<int4*>(p.coeff_out + cbase + c) =
                    make_int4(x[c], x[c + 1], x[c + 2], x[c + 3]);
        }
    } else if (r < KH) {
#pragma unroll
        for (int c = 0; c < KW; c += 4) {
            const int4 v = *reinterpret_cast<const int4*>(p.coeff_in + cbase + c);
            x[c] = v.x; x[c + 1] = v.y; x[c + 2] = v.z; x[c + 3] = v.w;
        }
    }

    if ((C::stg(p) & STAGE_QUANT) && r < KH) {
        const QSet qa = q.ac;                      // registers, not repeated shared loads
        const QSet q0 = r == 0 ? q.dc : qa;        // position 0 of the block is DC
        const int t1max = C::clamp16(d) ? 32767 : 0x7fffffff;
        const bool use_qm = QM && (d.meta & META_USE_QM);
        // eob = 1 + the largest scan index of a non-zero level.  Along a row every AV1 scan order
        // (default diagonal, mrow, mcol; all 19 sizes) visits the columns left to right, i.e.
        // iscan[r][c] grows with c (checked for every table by tests/test_kernel_emulation.py), so
        // the row's candidate is iscan[r][last non-zero column] + 1: one table read per row
        // instead of one per coefficient.
        int last = -1;
        // this row of the padded level map (av1_txb_init_levels, EbRateDistortionCost.c:125),
        // its 4 zero bytes of horizontal padding, and this thread's share of the zero rows above
        // and below the block (2 + 4 rows + 16 bytes = 6 * LP / 4 + 4 words, dealt round robin)
        uint8_t* lvrow = nullptr;
        if (C::has_lv(p)) {
            constexpr int LP = KW + 4;  // pitch of the map
            uint8_t* base = p.levels + p.levels_off[d.id];
            lvrow = base + (2 + r) * LP;
            *reinterpret_cast<uint32_t*>(lvrow + KW) = 0u;
            constexpr int TOPW = 2 * LP / 4, PADW = TOPW + LP + 4;
            for (int k = r; k < PADW; k += KH) {
                uint8_t* q4 = k < TOPW ? base + 4 * k : base + (2 + KH) * LP + 4 * (k - TOPW);
                *reinterpret_cast<uint32_t*>(q4) = 0u;
            }
        }
        constexpr bool DIRECT = direct256(QM, KW);
        if (!use_qm) {
            if constexpr (DIRECT) {
                const bool al32 = !(((C::has_q(p) ? size_t(p.qcoeff) : 0) | (C::has_dq(p) ? size_t(p.dqcoeff) : 0) |
                                     (size_t(d.coeff_off) * 4)) & 31);
#pragma unroll
                for (int c0 = 0; c0 < KW; c0 += 8) {
                    int qv[8];
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        uint32_t lv4 = 0;
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const int c = c0 + 4 * h + j;
                            int dqc;
                            const int cf = x[c];
                            const int aq = quant_flat<LS, QM>(cf, c == 0 ? q0 : qa, t1max, qv[4 * h + j], dqc, asum);
                            if (lvrow) lv4 |= uint32_t(aq > 127 ? 127 : aq) << (8 * j);
                            if constexpr (DIST) {
                                const int dd = cf - dqc;
                                dsum[0] += (unsigned long long)((long long)dd * dd);
                                dsum[1] += (unsigned long long)((long long)cf * cf);
                            }
                            x[c] = dqc;
                            last = aq != 0 ? c : last;
                        }
                        if (lvrow) *reinterpret_cast<uint32_t*>(lvrow + c0 + 4 * h) = lv4;
                    }
                    if (C::has_q(p)) store8(p.qcoeff + cbase + c0, qv, al32);
                    if (C::has_dq(p)) store8(p.dqcoeff + cbase + c0, &x[c0], al32);
                }
            } else {
#pragma unroll
            for (int c0 = 0; c0 < KW; c0 += 4) {
                int qv[4];
                uint32_t lv4 = 0;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int c = c0 + j;
                    int dqc;
                    const int cf = x[c];
                    const int aq = quant_flat<LS, QM>(cf, c == 0 ? q0 : qa, t1max, qv[j], dqc, asum);
                    if (lvrow) lv4 |= uint32_t(aq > 127 ? 127 : aq) << (8 * j);
                    if constexpr (DIST) {
                        const int dd = cf - dqc;
                        dsum[0] += (unsigned long long)((long long)dd * dd);
                        dsum[1] += (unsigned long long)((long long)cf * cf);
                    }
                    x[c] = dqc;
                    last = aq != 0 ? c : last;
                }
                if (lvrow) *reinterpret_cast<uint32_t*>(lvrow + c0) = lv4;
                emit_q<W, H, C>(p, row, cbase, c0, qv, x);
            }
            }
        } else if constexpr (QM) {
            const int qmo = p.qm_off[((d.meta >> META_PLANE_SHIFT) & 3) > 0][G::TXS] + r * KW;
            if constexpr (DIRECT) {
                const bool al32 = !(((C::has_q(p) ? size_t(p.qcoeff) : 0) | (C::has_dq(p) ? size_t(p.dqcoeff) : 0) |
                                     (size_t(d.coeff_off) * 4)) & 31);
#pragma unroll
                for (int c0 = 0; c0 < KW; c0 += 8) {
                    int qv[8];
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        uint32_t lv4 = 0;
                        const uint32_t w4 = ld_ro(reinterpret_cast<const uint32_t*>(p.qm_wt + qmo + c0 + 4 * h));
                        const uint32_t i4 = ld_ro(reinterpret_cast<const uint32_t*>(p.qm_iwt + qmo + c0 + 4 * h));
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const int c = c0 + 4 * h + j;
                            int dqc;
                            const int aq = quant_qm<LS>(x[c], c == 0 ? q0 : qa, int((w4 >> (8 * j)) & 255),
                                                        int((i4 >> (8 * j)) & 255), qv[4 * h + j], dqc, asum);
                            if (lvrow) lv4 |= uint32_t(aq > 127 ? 127 : aq) << (8 * j);
                            x[c] = dqc;
                            last = aq != 0 ? c : last;
                        }
                        if (lvrow) *reinterpret_cast<uint32_t*>(lvrow + c0 + 4 * h) = lv4;
                    }
                    if (C::has_q(p)) store8(p.qcoeff + cbase + c0, qv, al32);
                    if (C::has_dq(p)) store8(p.dqcoeff + cbase + c0, &x[c0], al32);
                }
            } else
#pragma unroll
            for (int c0 = 0; c0 < KW; c0 += 4) {
                int qv[4];
                // four weights per load (the library keeps every matrix 16-byte aligned)
                uint32_t lv4 = 0;
                const uint32_t w4 = ld_ro(reinterpret_cast<const uint32_t*>(p.qm_wt + qmo + c0));
                const uint32_t i4 = ld_ro(reinterpret_cast<const uint32_t*>(p.qm_iwt + qmo + c0));
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int c = c0 + j;
                    int dqc;
                    const int aq = quant_qm<LS>(x[c], c == 0 ? q0 : qa, int((w4 >> (8 * j)) & 255),
                                                int((i4 >> (8 * j)) & 255), qv[j], dqc, asum);
                    if (lvrow) lv4 |= uint32_t(aq > 127 ? 127 : aq) << (8 * j);
                    x[c] = dqc;
                    last = aq != 0 ? c : last;
                }
                if (lvrow) *reinterpret_cast<uint32_t*>(lvrow + c0) = lv4;
                emit_q<W, H, C>(p, row, cbase, c0, qv, x);
            }
        }
        eob_max = last >= 0 ? int(ld_ro(isc + r * KW + last)) + 1 : 0;
#if defined(__CUDA_ARCH__)
        {   // the block's eob = the maximum over its row owners: LB consecutive lanes of one warp, so a
            // butterfly of shuffles replaces the shared-memory atomicMax (16 lanes on one address were
            // 7.7 serialised wavefronts per 16x16 block, ncu profiles/r02_v15_*); lane 0 of the group
            // stores the result, and eob_max goes back as 0 so that the caller has nothing left to do
            constexpr int LB = H < 32 ? H : 32;
            const int lane = threadIdx.x & 31, li = r & (LB - 1);
            const unsigned bmask = LB == 32 ? 0xffffffffu : (((1u << LB) - 1u) << (lane - li));
            int m = eob_max;
#pragma unroll
            for (int k = LB / 2; k > 0; k >>= 1) { const int o = __shfl_xor_sync(bmask, m, k); m = o > m ? o : m; }
            if (li == 0) *eob_slot = m;
            eob_max = 0;
        }
#endif
        if constexpr (!direct256(QM, KW)) flush_q<W, H, C>(p, d, r, row, x);
    }

    if (C::stg(p) & STAGE_INV) {  // inv_txfm2d_add_c row pass, EbTransforms.c:8215-8236
        if (r < KH) {
            // S = sum |dqcoeff| over the block: the rows of a block are LB consecutive lanes of one
            // warp (its coded rows), so a butterfly of shuffles gives every row owner the total
            bool fast = false;
            if (C::stg(p) & STAGE_QUANT) {
#if defined(__CUDA_ARCH__)
                constexpr int LB = H < 32 ? H : 32;
                const int lane = threadIdx.x & 31, li = r & (LB - 1);
                const unsigned bmask = LB == 32 ? 0xffffffffu : (((1u << LB) - 1u) << (lane - li));
                unsigned tot = asum < (1u << 24) ? asum : (1u << 24);
#pragma unroll
                for (int m = LB / 2; m > 0; m >>= 1) tot += __shfl_xor_sync(bmask, tot, m);
                fast = tot <= fast_limit<W, H, KIND>(bd10);
                if (li == 0) *fast_flag = fast;
#else
                aux.sum = asum;
                fast = aux.fast != 0;
#endif
            } else if (r == 0) {
                *fast_flag = 0;  // inverse of caller-supplied coefficients: the clamped networks
            }
            if (fast) inv_row_store<W, H, KIND, C, true>(x, row, bd10);
            else inv_row_store<W, H, KIND, C, false>(x, row, bd10);
        }
        // rows >= KH: nothing to do, the column pass takes them as zeros (inv_col_k)
    }
}
template <int W, int H, bool QM, bool DIST = false, class C = CfgRuntime>
SVT_HD void phase_row(const KParams& p, const DevBlock& d, const QRow& q, int r, int* buf,
                      const int16_t* isc, int& e, unsigned long long& en, unsigned long long (&ds)[2],
                      int* fast_flag, RowAux& aux, int* eob_slot = nullptr) {
    const int kind = hkind(d);
    if constexpr (W == 64) row_k<W, H, 0, QM, DIST, C>(p, d, q, r, buf, isc, e, en, ds, fast_flag, aux, eob_slot);
    else if constexpr (W == 32) {
        if (kind == 2) row_k<W, H, 2, QM, DIST, C>(p, d, q, r, buf, isc, e, en, ds, fast_flag, aux, eob_slot);
        else row_k<W, H, 0, QM, DIST, C>(p, d, q, r, buf, isc, e, en, ds, fast_flag, aux, eob_slot);
    } else {
        if (kind == 0) row_k<W, H, 0, QM, DIST, C>(p, d, q, r, buf, isc, e, en, ds, fast_flag, aux, eob_slot);
        else if (kind == 1) row_k<W, H, 1, QM, DIST, C>(p, d, q, r, buf, isc, e, en, ds, fast_flag, aux, eob_slot);
        else row_k<W, H, 2, QM, DIST, C>(p, d, q, r, buf, isc, e, en, ds, fast_flag, aux, eob_slot);
    }
}
// eob rule, three_quad_energy and scaling of the per-TU distortion (EbFullLoop.c:808-835;
// picture_full_distortion32_bits, EbPictureOperators.c:349-402)
template <int LOG_SCALE>
SVT_HD void finish_dist(unsigned long long res, unsigned long long pred, int eob, unsigned long long energy,
                        unsigned long long* out) {
    if (!eob) res = pred;  // full_distortion_kernel_cbf_zero32_bits
    res += energy;
    pred += energy;
    constexpr int shift = (1 - LOG_SCALE) * 2;  // (MAX_TX_SCALE - av1_get_tx_scale) * 2
    out[0] = shift < 0 ? res << (shift < 0 ? -shift : 0) : res >> (shift > 0 ? shift : 0);
    out[1] = shift < 0 ? pred << (shift < 0 ? -shift : 0) : pred >> (shift > 0 ? shift : 0);
}

// ---- phase 3: inverse column transform + reconstruction on the staged prediction block --------------
// pix: the block's prediction pixels in shared memory (dense, W per row), overwritten in place with
// the reconstruction (highbd_clip_pixel_add + check_range, EbTransforms.c:7794-7813).
template <int W, int H, int KIND, class PIX, bool FAST>
SVT_HD void inv_col_k(const DevBlock& d, int c, const int* buf, PIX* pix, int pmax, int rmax) {
    using G = Geo<W, H>;
    const bool lr = d.meta & META_LR;
    const int* src = buf + (lr ? W - 1 - c : c);
    const Clamp k16(16);  // the column stage range: 16 bits for bd 8 / 10
    int x[H];
    // already clamped to 16 bits by the row pass.  Rows >= KH (64-row sizes) hold the inverse row
    // transform of all-zero coefficient rows, i.e. zeros (EbTransforms.c:8299): they are neither
    // written by the row pass nor read here, and as compile-time zeros they fold the first stages
    // of the 64-point network away.
#pragma unroll
    for (int r = 0; r < H; ++r) x[r] = r < G::KH ? src[r * G::LD] : 0;
    if constexpr (FAST) inv1d_k<H, KIND>(x, NoClamp());  // inside fast_limit: no clamp can bind
    else inv1d_k<H, KIND>(x, k16);
    // check_range (EbTransforms.c:7794) clamps the residual to +-(2^(7+bd) - 1 + (914 << (bd-7))),
    // i.e. beyond +-34595.  It can never bind here: every 1-D inverse output is bounded by its
    // 16-bit input clamp (butterfly outputs are re-clamped to 16 bits; identity scales by <= 4;
    // ADST4 sums <= 4 terms of weight < 1), so after round_shift(., 4) |v| <= 2^13.  The clamp is
    // therefore omitted; `rmax` documents the bound.
    (void)rmax;
#pragma unroll
    for (int r = 0; r < H; ++r) {  // pixel rows were staged flipped for the FLIPADST column types
        PIX* q = pix + r * W + c;
        const int px = int(*q) + rshift<4>(x[r]);
#if defined(__CUDA_ARCH__)
        *q = PIX(__vimin_s32_relu(px, pmax));  // clip to [0, pmax] in one instruction
#else
        *q = PIX(px < 0 ? 0 : (px > pmax ? pmax : px));
#endif
    }
}
template <int W, int H, int KIND, class C, bool FAST>
SVT_HD void inv_col_bd(const DevBlock& d, int c, const int* buf, void* pix) {
    if (C::bd10(d))
        inv_col_k<W, H, KIND, uint16_t, FAST>(d, c, buf, static_cast<uint16_t*>(pix), 1023, (1 << 17) - 1 + (914 << 3));
    else
        inv_col_k<W, H, KIND, uint8_t, FAST>(d, c, buf, static_cast<uint8_t*>(pix), 255, (1 << 15) - 1 + (914 << 1));
}
// fast: the block's flag from the row pass (fast_limit); only the DCT / long ADST columns have
// clamps to drop
template <int W, int H, class C = CfgRuntime>
SVT_HD void phase_inv_col(const DevBlock& d, int c, const int* buf, void* pix, bool fast) {
    const int kind = vkind(d);
    if constexpr (H == 64) {
        if (fast) inv_col_bd<W, H, 0, C, true>(d, c, buf, pix); else inv_col_bd<W, H, 0, C, false>(d, c, buf, pix);
    } else if constexpr (H == 32) {
        if (kind == 2) inv_col_bd<W, H, 2, C, false>(d, c, buf, pix);
        else if (fast) inv_col_bd<W, H, 0, C, true>(d, c, buf, pix);
        else inv_col_bd<W, H, 0, C, false>(d, c, buf, pix);
    } else {
        if (kind == 0) {
            if (fast) inv_col_bd<W, H, 0, C, true>(d, c, buf, pix); else inv_col_bd<W, H, 0, C, false>(d, c, buf, pix);
        } else if (kind == 1) {
            if (fast && H > 4) inv_col_bd<W, H, 1, C, true>(d, c, buf, pix);
            else inv_col_bd<W, H, 1, C, false>(d, c, buf, pix);
        } else inv_col_bd<W, H, 2, C, false>(d, c, buf, pix);
    }
}

// ---- per-block staging steps shared by the kernel and the host emulation ---------------------------
SVT_HD bool ud_flip(const DevBlock& d) { return d.meta & META_UD; }
template <int W, int H, int NTHR, class C = CfgRuntime>
SVT_HD void stage_residual(const KParams& p, const DevBlock& d, int16_t* sres, int tid) {
    if constexpr (C::SRC_STAGED) {  // prediction block, then source block, both as they are
        char* s = reinterpret_cast<char*>(sres);
        stage_copy<H, W, NTHR, true>(s, static_cast<char*>(const_cast<void*>(p.pred)) + size_t(d.rec_off),
                                     size_t(d.rec_stride), ud_flip(d), tid);
        stage_copy<H, W, NTHR, true>(s + W * H, static_cast<char*>(const_cast<void*>(p.src)) + size_t(d.rec_off),
                                     size_t(d.rec_stride), ud_flip(d), tid);
        return;
    }
    if (C::stg(p) & STAGE_RESIDUAL) {
        if (C::bd10(d))
            stage_diff<H, W, NTHR>(sres, static_cast<const uint16_t*>(p.src) + d.rec_off,
                                   static_cast<const uint16_t*>(p.pred) + d.rec_off, d.rec_stride, ud_flip(d), tid);
        else
            stage_diff<H, W, NTHR>(sres, static_cast<const uint8_t*>(p.src) + d.rec_off,
                                   static_cast<const uint8_t*>(p.pred) + d.rec_off, d.rec_stride, ud_flip(d), tid);
        return;
    }
    stage_copy<H, W * 2, NTHR, true>(reinterpret_cast<char*>(sres),
                                     reinterpret_cast<char*>(const_cast<int16_t*>(p.residual + d.res_off)),
                                     size_t(d.res_stride) * 2, ud_flip(d), tid);
}
template <int W, int H, int NTHR, class C = CfgRuntime>
SVT_HD void stage_pred(const KParams& p, const DevBlock& d, void* spix, int tid) {
    char* s = static_cast<char*>(spix);
    char* g = static_cast<char*>(const_cast<void*>(p.pred));
    if (C::bd10(d))
        stage_copy<H, W * 2, NTHR, true>(s, g + size_t(d.rec_off) * 2, size_t(d.rec_stride) * 2, ud_flip(d), tid);
    else
        stage_copy<H, W, NTHR, true>(s, g + size_t(d.rec_off), size_t(d.rec_stride), ud_flip(d), tid);
}
template <int W, int H, int NTHR, class C = CfgRuntime>
SVT_HD void unstage_recon(const KParams& p, const DevBlock& d, const void* spix, int tid) {
    char* s = static_cast<char*>(const_cast<void*>(spix));
    char* g = static_cast<char*>(p.recon);
    if (C::bd10(d))
        stage_copy<H, W * 2, NTHR, false>(s, g + size_t(d.rec_off) * 2, size_t(d.rec_stride) * 2, ud_flip(d), tid);
    else
        stage_copy<H, W, NTHR, false>(s, g + size_t(d.rec_off), size_t(d.rec_stride), ud_flip(d), tid);
}
// this block's iscan table (global memory, a few hundred bytes per size: L1-resident)
template <int W, int H>
SVT_HD const int16_t* block_iscan(const KParams& p, const DevBlock& d) {
    using G = Geo<W, H>;
    const int cls = (d.meta >> META_SCAN_SHIFT) & 3;
    return p.iscan + p.iscan_off[G::TXS][G::NCLS == 1 ? 0 : cls];
}

#if defined(__CUDACC__)
// FULL: 0 = switches read at run time (CfgRuntime); otherwise the CFG code of CfgFull
template <int W, int H, bool QM, bool DIST, int FULL>
__global__ void __launch_bounds__(Geo<W, H>::NT, Geo<W, H>::ctas_per_sm(QM || DIST))
residual_kernel(const __grid_constant__ KParams p, int first_block, int count) {
    using G = Geo<W, H>;
    using C = typename std::conditional<FULL != 0, CfgFull<FULL>, CfgRuntime>::type;
    extern __shared__ __align__(16) int smem[];
    int* buf = smem;
    int16_t* stage = reinterpret_cast<int16_t*>(smem + G::BUF_WORDS);
    __shared__ CtaShared<G::NB> sh;
    const int tid = threadIdx.x;
#ifdef SVT_TRACE
    long long tr[7];
    int tri = 0;
#define TR() do { if (tid == 0) tr[tri] = clock64(); ++tri; } while (0)
#else
#define TR() do { } while (0)
#endif
    TR();
    const int base = first_block + blockIdx.x * G::NB;
    const int nb = min(G::NB, first_block + count - base);
    // Square sizes up to 32x32: a block's W column tasks, W row tasks and its copy team are the
    // same W consecutive threads of one warp, so every hand-over between phases is warp-local and
    // __syncwarp replaces the CTA-wide barriers (warps of a CTA no longer wait for each other).
    // The owner of a block's per-block state is then the first lane of that group (`mine`).
#ifdef SVT_NO_WARP_PRIVATE
    constexpr bool WP = false;
#else
    constexpr bool WP = (W == H) && (W <= 32);
#endif
    auto sync = [&] { if constexpr (WP) __syncwarp(); else __syncthreads(); };
    // per-block state of blocks own0, own0 + OSTEP, ... belongs to this thread (if `mine`)
    constexpr int OSTEP = WP ? G::NT / W : G::NT;
    const int own0 = WP ? tid / W : tid;
    const bool mine = WP ? (tid % W == 0) : true;
    if (mine) {
        for (int own = own0; own < G::NB; own += OSTEP) {
            if (own < nb) sh.blk[own] = p.blocks[base + own];
            sh.eob[own] = 0;
            sh.energy[own] = 0ull;
            if constexpr (DIST) sh.dist[own][0] = sh.dist[own][1] = 0ull;
        }
    }
    sync();
    // the quantizer rows depend on the descriptors; they are fetched while the residual rows are
    // already in flight, not before them (one global round trip less on the CTA's critical path)
    auto fetch_qrows = [&] {
        if ((C::stg(p) & STAGE_QUANT) && mine)
            for (int own = own0; own < nb; own += OSTEP)
                sh.q[own] = load_qrow<G::LOG_SCALE, QM>(p.qtables + size_t(sh.blk[own].qrow) * 40);
    };
    TR();
    // copy teams: TPB consecutive threads move one block's rows, 16 bytes per thread per step;
    // a team serves blocks tb0, tb0 + NTEAM, ... (the same blocks its threads transform)
    constexpr int TPB = cmin(W, H), NTEAM = G::NT / TPB;
    const int tb0 = tid / TPB, tl = tid - tb0 * TPB;
    if (C::stg(p) & STAGE_FWD) {
#pragma unroll 1
        for (int tb = tb0; tb < nb; tb += NTEAM) stage_residual<W, H, TPB, C>(p, sh.blk[tb], stage + tb * G::SP, tl);
        fetch_qrows();
        g2s_commit_wait();
        sync();
        TR();
#pragma unroll 1
        for (int task = tid; task < G::NB * W; task += G::NT) {
            const int b = task / W, c = task % W;
            if (b < nb) phase_fwd_col<W, H, C::SRC_STAGED>(sh.blk[b], c, buf + b * G::BLK, stage + b * G::SP);
        }
        sync();
        TR();
    } else if (C::stg(p) & STAGE_QUANT) {
        fetch_qrows();
        sync();
    }
    // the residual is consumed: the prediction block streams into the same bytes while the row
    // phase runs
    if ((C::stg(p) & STAGE_INV) && !C::SRC_STAGED) {  // (SRC_STAGED: the prediction is already there)
#pragma unroll 1
        for (int tb = tb0; tb < nb; tb += NTEAM) stage_pred<W, H, TPB, C>(p, sh.blk[tb], stage + tb * G::SP, tl);
    }
#pragma unroll 1
    for (int task = tid; task < G::NB * H; task += G::NT) {
        const int b = task / H, r = task % H;
        if (b < nb) {
            int e;
            unsigned long long en, ds[2];
            RowAux aux;
            phase_row<W, H, QM, DIST, C>(p, sh.blk[b], sh.q[b], r, buf + b * G::BLK,
                                      block_iscan<W, H>(p, sh.blk[b]), e, en, ds, &sh.fast[b], aux, &sh.eob[b]);
            if (e) atomicMax(&sh.eob[b], e);  // (device: reduced inside row_k by shuffles, e == 0 here)
            if (G::KW != W || G::KH != H) {
#if defined(__CUDA_ARCH__) && !defined(SVT_ENERGY_ATOMIC_PER_THREAD)
                // a 64-bit shared-memory atomicAdd is a compare-and-swap loop: 64 row owners on one
                // address serialise badly.  The rows of a block are LB consecutive lanes of one warp:
                // sum there by shuffles, one atomic per group.
                constexpr int LB = H < 32 ? H : 32;
                const int lane = threadIdx.x & 31, li = r & (LB - 1);
                const unsigned bmask = LB == 32 ? 0xffffffffu : (((1u << LB) - 1u) << (lane - li));
#pragma unroll
                for (int k = LB / 2; k > 0; k >>= 1) en += __shfl_xor_sync(bmask, en, k);
                if (li == 0 && en) atomicAdd(&sh.energy[b], en);
#else
                if (en) atomicAdd(&sh.energy[b], en);
#endif
            }
            if constexpr (DIST) {
#if defined(__CUDA_ARCH__) && !defined(SVT_ENERGY_ATOMIC_PER_THREAD)
                // (rows >= KH of the 64-row sizes carry zeros here; every row owner takes part)
                constexpr int LB = H < 32 ? H : 32;
                const int lane = threadIdx.x & 31, li = r & (LB - 1);
                const unsigned bmask = LB == 32 ? 0xffffffffu : (((1u << LB) - 1u) << (lane - li));
#pragma unroll
                for (int k = LB / 2; k > 0; k >>= 1) {
                    ds[0] += __shfl_xor_sync(bmask, ds[0], k);
                    ds[1] += __shfl_xor_sync(bmask, ds[1], k);
                }
                if (li == 0 && ds[1]) {
                    atomicAdd(&sh.dist[b][0], ds[0]);
                    atomicAdd(&sh.dist[b][1], ds[1]);
                }
#else
                if (ds[1]) {
                    atomicAdd(&sh.dist[b][0], ds[0]);
                    atomicAdd(&sh.dist[b][1], ds[1]);
                }
#endif
            }
        }
    }
    g2s_commit_wait();
    sync();
    TR();
    if (mine) {
        for (int own = own0; own < nb; own += OSTEP) {
            if (C::stg(p) & STAGE_QUANT) p.eob[sh.blk[own].id] = uint16_t(sh.eob[own]);
            if ((C::stg(p) & STAGE_FWD) && p.energy) p.energy[sh.blk[own].id] = sh.energy[own];
            if constexpr (DIST)
                finish_dist<G::LOG_SCALE>(sh.dist[own][0], sh.dist[own][1], sh.eob[own], sh.energy[own],
                                          p.dist + 2 * size_t(sh.blk[own].id));
        }
    }
    if (C::stg(p) & STAGE_INV) {
#pragma unroll 1
        for (int task = tid; task < G::NB * W; task += G::NT) {
            const int b = task / W, c = task % W;
            if (b < nb) phase_inv_col<W, H, C>(sh.blk[b], c, buf + b * G::BLK, stage + b * G::SP, sh.fast[b] != 0);
        }
        sync();
        TR();
#pragma unroll 1
        for (int tb = tb0; tb < nb; tb += NTEAM) unstage_recon<W, H, TPB, C>(p, sh.blk[tb], stage + tb * G::SP, tl);
        TR();
    }
#ifdef SVT_TRACE
    if (tid == 0 && nb >= 6 && p.energy)
        for (int i = 0; i < 6; ++i) p.energy[sh.blk[i].id] = (unsigned long long)(tr[i + 1] - tr[i]);
#endif
#undef TR
}

// Stand-alone quantizer behind the RTCD slots aom_[highbd_]quantize_b[_32x32|_64x64]
// (aom_dsp_rtcd.h:321-342): the caller supplies n_coeffs and its own iscan, not a tx_size.
// One CTA per call; eob = max(iscan + 1) over non-zero levels.
template <int LOG_SCALE>
__global__ void __launch_bounds__(256)
quant_kernel(const int32_t* coeff, int n, const int16_t* tab /* 5 x {DC, AC} */,
             const int16_t* iscan, int32_t* qcoeff, int32_t* dqcoeff, uint16_t* eob) {
    const QSet dc = make_qset<LOG_SCALE, true>(tab[0], tab[2], tab[4], tab[6], tab[8]);
    const QSet ac = make_qset<LOG_SCALE, true>(tab[1], tab[3], tab[5], tab[7], tab[9]);
    __shared__ int s_eob;
    if (threadIdx.x == 0) s_eob = 0;
    __syncthreads();
    int e = 0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        int qc, dqc;
        const int aq = quant_flat<LOG_SCALE, true>(coeff[i], i == 0 ? dc : ac, 0x7fffffff, qc, dqc);
        qcoeff[i] = qc;
        dqcoeff[i] = dqc;
        if (aq != 0) { const int s = int(iscan[i]) + 1; e = s > e ? s : e; }
    }
    if (e) atomicMax(&s_eob, e);
    __syncthreads();
    if (threadIdx.x == 0) *eob = uint16_t(s_eob);
}
#endif

}  // namespace svt
