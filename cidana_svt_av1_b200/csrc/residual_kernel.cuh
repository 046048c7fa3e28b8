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
#include "txfm1d.cuh"

#if !defined(__CUDACC__)  // host emulation build (tests only): the two vector helpers used below
struct int4 { int x, y, z, w; };
static inline int4 make_int4(int x, int y, int z, int w) { return int4{ x, y, z, w }; }
#endif

namespace svt {

enum : int { STAGE_FWD = 1, STAGE_QUANT = 2, STAGE_INV = 4, STAGE_ALL = 7,
             // internal: write all W x H coefficients at stride W (what the RTCD-level
             // av1_fwd_txfm2d_WxH returns before av1_estimate_transform packs 64-point sizes)
             STAGE_FWD_FULL = 8 };
enum : uint32_t { META_TYPE_MASK = 0xF, META_BD10 = 1u << 4, META_PLANE_SHIFT = 5,
                  META_CLAMP16 = 1u << 8, META_USE_QM = 1u << 9 };

struct DevBlock {  // 32 bytes; device copy of a public SvtGpuBlock, bucketed by size
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
    int iscan_off[19][3];   // [tx_size][scan class] -> offset into iscan
    int qm_off[2][19];      // [plane > 0][tx_size] -> offset into qm_wt / qm_iwt
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

SVT_HD constexpr int tx_id(int w, int h) {  // TxSize enum, EbDefinitions.h
    if (w == h) return ilog2(w) - 2;
    constexpr int ws[14] = { 4, 8, 8, 16, 16, 32, 32, 64, 4, 16, 8, 32, 16, 64 };
    constexpr int hs[14] = { 8, 4, 16, 8, 32, 16, 64, 32, 16, 4, 32, 8, 64, 16 };
    for (int i = 0; i < 14; ++i)
        if (ws[i] == w && hs[i] == h) return 5 + i;
    return -1;
}

template <int W, int H>
struct Geo {
    static constexpr int TXS = tx_id(W, H);
    static constexpr int NT = 128;                    // threads per CTA
    static constexpr int NB = NT / cmin(W, H);        // blocks per CTA
    static constexpr int LD = W + 1;                  // padded row pitch (words)
    // block pitch == W (mod 32): blocks sharing a warp land in disjoint banks
    static constexpr int BLK = ((H * LD - W + 31) / 32) * 32 + W;
    static constexpr int KW = cmin(W, 32), KH = cmin(H, 32);  // kept (coded) region
    static constexpr int LOG_SCALE = (W * H > 256) + (W * H > 1024);
    static constexpr bool RECT = (ilog2(W) - ilog2(H) == 1) || (ilog2(H) - ilog2(W) == 1);
    static constexpr int SMEM_WORDS = NB * BLK;
};

struct CtaShared {  // small per-CTA state next to the transpose buffer
    DevBlock blk[32];
    int eob[32];
    unsigned long long energy[32];
};

// ---- phase 1: forward column transforms -> shared (transposed hand-off) --------------------------
template <int W, int H>
SVT_HD void phase_fwd_col(const KParams& p, const DevBlock& d, int c, int* buf /* this block */) {
    using G = Geo<W, H>;
    constexpr int S0 = fwd_shift(W, H, 0), S1 = -fwd_shift(W, H, 1);
    const int t = d.meta & META_TYPE_MASK;
    const int vt = vtx_of(t), ht = htx_of(t);
    const bool ud = vt == 2, lr = ht == 2;
    const int16_t* src = p.residual + d.res_off + c;
    int x[H];
#pragma unroll
    for (int r = 0; r < H; ++r) {
        const int rr = ud ? H - 1 - r : r;
        x[r] = int(src[size_t(rr) * d.res_stride]) * (1 << S0);
    }
    fwd1d<H, fwd_cos_col(W, H)>(x, kind_of(vt));
    const int cc = lr ? W - 1 - c : c;
#pragma unroll
    for (int r = 0; r < H; ++r) {
        if constexpr (S1 > 0) x[r] = rshift<(S1 > 0 ? S1 : 1)>(x[r]);
        buf[r * G::LD + cc] = x[r];
    }
}

// ---- quantizer for one coefficient ---------------------------------------------------------------
struct QSet { int zbin, round, quant, qshift, dequant; };  // one of {DC, AC}
struct QRow { QSet dc, ac; };  // zbin / round already scaled by log_scale
template <int LOG_SCALE>
SVT_HD QRow load_qrow(const int16_t* t) {
    constexpr int rnd = LOG_SCALE ? (1 << (LOG_SCALE - 1)) : 0;
    QRow q;
    // ROUND_POWER_OF_TWO(zbin / round, log_scale), EbFullLoop.c:246, :280
    q.dc = QSet{ (int(t[0]) + rnd) >> LOG_SCALE, (int(t[8]) + rnd) >> LOG_SCALE, t[16], t[24], t[32] };
    q.ac = QSet{ (int(t[1]) + rnd) >> LOG_SCALE, (int(t[9]) + rnd) >> LOG_SCALE, t[17], t[25], t[33] };
    return q;
}
// flat (no qmatrix): aom_highbd_quantize_b*_avx2 arithmetic
// (Source/Lib/Common/ASM_AVX2/highbd_quantize_intrin_avx2.c:80-125), equal to
// highbd_quantize_b_helper_c (EbFullLoop.c:239) for the tables av1_build_quantizer produces.
template <int LOG_SCALE>
SVT_HD void quant_flat(int cf, const QSet& q, bool clamp16, int& qc, int& dqc) {
    const int sign = cf >> 31;
    const int ab = (cf ^ sign) - sign;
    qc = 0; dqc = 0;
    if (ab >= q.zbin) {
        int t1 = ab + q.round;
        if (clamp16) t1 = t1 > 32767 ? 32767 : t1;  // 8-bit C variant, EbFullLoop.c:86-88
        const int64_t p1 = int64_t(t1) * q.quant;
        const uint32_t t2 = uint32_t(uint64_t(p1) >> 16) + uint32_t(t1);
        const uint64_t p2 = uint64_t(t2) * uint32_t(q.qshift);
        const int aq = int(uint32_t(p2 >> (16 - LOG_SCALE)));
        const int adq = int(unsigned(aq) * unsigned(q.dequant)) >> LOG_SCALE;
        qc = (aq ^ sign) - sign;
        dqc = (adq ^ sign) - sign;
    }
}
// with quantization matrices: highbd_quantize_b_helper_c, EbFullLoop.c:239-296
template <int LOG_SCALE>
SVT_HD void quant_qm(int cf, const QSet& q, int wt, int iwt, int& qc, int& dqc) {
    const int sign = cf >> 31;
    const int ab = (cf ^ sign) - sign;
    const int cw = int(unsigned(cf) * unsigned(wt));
    qc = 0; dqc = 0;
    if (cw >= q.zbin * 32 || cw <= -q.zbin * 32) {
        const int64_t tw = (int64_t(ab) + q.round) * wt;
        const int64_t t2 = ((tw * q.quant) >> 16) + tw;
        const int aq = int((t2 * q.qshift) >> (16 - LOG_SCALE + 5));
        const int dq = (q.dequant * iwt + 16) >> 5;
        const int adq = int(unsigned(aq) * unsigned(dq)) >> LOG_SCALE;
        qc = (aq ^ sign) - sign;
        dqc = (adq ^ sign) - sign;
    }
}

// ---- phase 2: forward row -> (energy, pack) -> quantize -> inverse row -> shared ------------------
// Returns this row's eob candidate and discarded-region energy.
template <int W, int H>
SVT_HD void phase_row(const KParams& p, const DevBlock& d, int r, int* buf, int& eob_max,
                      unsigned long long& energy_sum) {
    using G = Geo<W, H>;
    constexpr int KW = G::KW, KH = G::KH, LS = G::LOG_SCALE;
    const int t = d.meta & META_TYPE_MASK;
    const int ht = htx_of(t);
    const bool bd10 = d.meta & META_BD10;
    int* row = buf + r * G::LD;
    int x[W];
    eob_max = 0;
    energy_sum = 0;
    const uint32_t cbase = d.coeff_off + r * KW;

    if (p.stages & STAGE_FWD) {
#pragma unroll
        for (int c = 0; c < W; ++c) x[c] = row[c];
        fwd1d<W, fwd_cos_row(W, H)>(x, kind_of(ht));
        constexpr int S2 = -fwd_shift(W, H, 2);
#pragma unroll
        for (int c = 0; c < W; ++c) {
            if constexpr (S2 > 0) x[c] = rshift<(S2 > 0 ? S2 : 1)>(x[c]);
            if constexpr (G::RECT) x[c] = mul_round12(x[c], 5793);  // EbTransforms.c:3769-3776
        }
        if constexpr (KW != W || KH != H) {  // HandleTransform*_c, EbTransforms.c:4351-4731
#pragma unroll
            for (int c = 0; c < W; ++c)
                if (r >= KH || c >= KW)
                    energy_sum += (unsigned long long)(int64_t(x[c]) * int64_t(x[c]));
        }
        if (p.stages & STAGE_FWD_FULL) {
#pragma unroll
            for (int c = 0; c < W; c += 4)
                *reinterpret_cast<int4*>(p.coeff_out + d.coeff_off + r * W + c) =
                    make_int4(x[c], x[c + 1], x[c + 2], x[c + 3]);
        } else if (p.coeff_out && r < KH) {
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

    if ((p.stages & STAGE_QUANT) && r < KH) {
        const QRow q = load_qrow<LS>(p.qtables + size_t(d.qrow) * 40);
        const QSet q0 = r == 0 ? q.dc : q.ac;
        const int cls = t < 10 ? 0 : ((t & 1) ? 2 : 1);  // av1_scan_orders, EbTransforms.h:3349
        const int16_t* iscan = p.iscan + p.iscan_off[G::TXS][cls] + r * KW;
        const bool clamp16 = d.meta & META_CLAMP16;
        const bool use_qm = (d.meta & META_USE_QM) && p.qm_wt != nullptr;
        const int qmo = p.qm_off[((d.meta >> META_PLANE_SHIFT) & 3) > 0][G::TXS] + r * KW;
        int e = 0;
#pragma unroll
        for (int c0 = 0; c0 < KW; c0 += 4) {
            int qv[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int c = c0 + j;
                int dqc;
                const QSet& qs = (c == 0) ? q0 : q.ac;  // position 0 of the block is DC
                if (!use_qm) quant_flat<LS>(x[c], qs, clamp16, qv[j], dqc);
                else quant_qm<LS>(x[c], qs, p.qm_wt[qmo + c], p.qm_iwt[qmo + c], qv[j], dqc);
                x[c] = dqc;
                if (qv[j] != 0) { const int s = int(iscan[c]) + 1; e = s > e ? s : e; }
            }
            *reinterpret_cast<int4*>(p.qcoeff + cbase + c0) = make_int4(qv[0], qv[1], qv[2], qv[3]);
            *reinterpret_cast<int4*>(p.dqcoeff + cbase + c0) =
                make_int4(x[c0], x[c0 + 1], x[c0 + 2], x[c0 + 3]);
        }
        eob_max = e;
    }

    if (p.stages & STAGE_INV) {  // inv_txfm2d_add_c row pass, EbTransforms.c:8215-8236
        if (r < KH) {
            const Clamp kin(bd10 ? 18 : 16);   // clamp_buf(bd + 8)
            const Clamp krow(bd10 ? 18 : 16);  // av1_gen_inv_stage_range rows: 16 / 18
#pragma unroll
            for (int c = 0; c < W; ++c) {
                if (c >= KW) x[c] = 0;  // 64-wide: re-expansion with zeros, EbTransforms.c:8299
                else {
                    if constexpr (G::RECT) x[c] = mul_round12(x[c], 2896);
                    x[c] = kin(x[c]);
                }
            }
            inv1d<W>(x, kind_of(ht), krow);
            constexpr int IS0 = inv_shift0(W, H);
#pragma unroll
            for (int c = 0; c < W; ++c) {
                if constexpr (IS0 > 0) x[c] = rshift<(IS0 > 0 ? IS0 : 1)>(x[c]);
                row[c] = x[c];
            }
        } else {
#pragma unroll
            for (int c = 0; c < W; ++c) row[c] = 0;
        }
    }
}

// ---- phase 3: inverse column transforms + reconstruction ------------------------------------------
template <int W, int H>
SVT_HD void phase_inv_col(const KParams& p, const DevBlock& d, int c, const int* buf) {
    using G = Geo<W, H>;
    const int t = d.meta & META_TYPE_MASK;
    const int vt = vtx_of(t), ht = htx_of(t);
    const bool ud = vt == 2, lr = ht == 2;
    const bool bd10 = d.meta & META_BD10;
    const int cc = lr ? W - 1 - c : c;
    const Clamp kmid(16);  // clamp_buf(max(bd + 6, 16)) == 16 for bd 8 and 10
    const Clamp kcol(16);  // av1_gen_inv_stage_range columns
    int x[H];
#pragma unroll
    for (int r = 0; r < H; ++r) x[r] = kmid(buf[r * G::LD + cc]);
    inv1d<H>(x, kind_of(vt), kcol);
    // highbd_clip_pixel_add + check_range, EbTransforms.c:7794-7813
    const int pmax = bd10 ? 1023 : 255;
    const int rmax = bd10 ? ((1 << 17) - 1 + (914 << 3)) : ((1 << 15) - 1 + (914 << 1));
    const size_t o = size_t(d.rec_off) + c;
#pragma unroll
    for (int r = 0; r < H; ++r) {
        int v = rshift<4>(x[ud ? H - 1 - r : r]);
        v = v > rmax ? rmax : (v < -rmax - 1 ? -rmax - 1 : v);
        const size_t a = o + size_t(r) * d.rec_stride;
        int px = bd10 ? int(static_cast<const uint16_t*>(p.pred)[a])
                      : int(static_cast<const uint8_t*>(p.pred)[a]);
        px += v;
        px = px < 0 ? 0 : (px > pmax ? pmax : px);
        if (bd10) static_cast<uint16_t*>(p.recon)[a] = uint16_t(px);
        else static_cast<uint8_t*>(p.recon)[a] = uint8_t(px);
    }
}

#if defined(__CUDACC__)
template <int W, int H>
__global__ void __launch_bounds__(Geo<W, H>::NT)
residual_kernel(const __grid_constant__ KParams p, int first_block, int count) {
    using G = Geo<W, H>;
    __shared__ int buf[G::SMEM_WORDS];
    __shared__ CtaShared sh;
    const int tid = threadIdx.x;
    const int base = first_block + blockIdx.x * G::NB;
    const int nb = min(G::NB, first_block + count - base);
    if (tid < G::NB) {
        if (tid < nb) sh.blk[tid] = p.blocks[base + tid];
        sh.eob[tid] = 0;
        sh.energy[tid] = 0ull;
    }
    __syncthreads();
    if (p.stages & STAGE_FWD) {
        for (int task = tid; task < G::NB * W; task += G::NT) {
            const int b = task / W, c = task % W;
            if (b < nb) phase_fwd_col<W, H>(p, sh.blk[b], c, buf + b * G::BLK);
        }
        __syncthreads();
    }
    for (int task = tid; task < G::NB * H; task += G::NT) {
        const int b = task / H, r = task % H;
        if (b < nb) {
            int e;
            unsigned long long en;
            phase_row<W, H>(p, sh.blk[b], r, buf + b * G::BLK, e, en);
            if (e) atomicMax(&sh.eob[b], e);
            if (G::KW != W || G::KH != H) { if (en) atomicAdd(&sh.energy[b], en); }
        }
    }
    __syncthreads();
    if (tid < nb) {
        if (p.stages & STAGE_QUANT) p.eob[sh.blk[tid].id] = uint16_t(sh.eob[tid]);
        if ((p.stages & STAGE_FWD) && p.energy) p.energy[sh.blk[tid].id] = sh.energy[tid];
    }
    if (p.stages & STAGE_INV) {
        for (int task = tid; task < G::NB * W; task += G::NT) {
            const int b = task / W, c = task % W;
            if (b < nb) phase_inv_col<W, H>(p, sh.blk[b], c, buf + b * G::BLK);
        }
    }
}

// Stand-alone quantizer behind the RTCD slots aom_[highbd_]quantize_b[_32x32|_64x64]
// (aom_dsp_rtcd.h:321-342): the caller supplies n_coeffs and its own iscan, not a tx_size.
// One CTA per call; eob = max(iscan + 1) over non-zero levels.
template <int LOG_SCALE>
__global__ void __launch_bounds__(256)
quant_kernel(const int32_t* coeff, int n, const int16_t* tab /* 5 x {DC, AC} */,
             const int16_t* iscan, int32_t* qcoeff, int32_t* dqcoeff, uint16_t* eob) {
    constexpr int rnd = LOG_SCALE ? (1 << (LOG_SCALE - 1)) : 0;
    const QSet dc{ (int(tab[0]) + rnd) >> LOG_SCALE, (int(tab[2]) + rnd) >> LOG_SCALE, tab[4], tab[6], tab[8] };
    const QSet ac{ (int(tab[1]) + rnd) >> LOG_SCALE, (int(tab[3]) + rnd) >> LOG_SCALE, tab[5], tab[7], tab[9] };
    __shared__ int s_eob;
    if (threadIdx.x == 0) s_eob = 0;
    __syncthreads();
    int e = 0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        int qc, dqc;
        quant_flat<LOG_SCALE>(coeff[i], i == 0 ? dc : ac, false, qc, dqc);
        qcoeff[i] = qc;
        dqcoeff[i] = dqc;
        if (qc != 0) { const int s = int(iscan[i]) + 1; e = s > e ? s : e; }
    }
    if (e) atomicMax(&s_eob, e);
    __syncthreads();
    if (threadIdx.x == 0) *eob = uint16_t(s_eob);
}
#endif

}  // namespace svt
