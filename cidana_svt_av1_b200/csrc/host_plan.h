// Host-side logic of libsvtgpu that needs no CUDA: scan-table generation, block validation and
// bucketing.  Header-only so the CPU-only tests can exercise it (tests/host_emul.cpp).
#pragma once
#include <stdint.h>
#include <algorithm>
#include <vector>
#include "../../include/svt_gpu.h"
#include "residual_kernel.cuh"

namespace svt {

static const int kTxW[19] = { 4, 8, 16, 32, 64, 4, 8, 8, 16, 16, 32, 32, 64, 4, 16, 8, 32, 16, 64 };
static const int kTxH[19] = { 4, 8, 16, 32, 64, 8, 4, 16, 8, 32, 16, 64, 32, 16, 4, 32, 8, 64, 16 };

inline int coded_w(int s) { return kTxW[s] > 32 ? 32 : kTxW[s]; }
inline int coded_h(int s) { return kTxH[s] > 32 ? 32 : kTxH[s]; }
inline int max_eob(int s) { return coded_w(s) * coded_h(s); }  // av1_get_max_eob, EbTransforms.h:3872

// is_txfm_allowed (AV1 5.11; reference test/TxfmCommon.h:172)
inline bool tx_pair_valid(int s, int t) {
    if (s < 0 || s > 18 || t < 0 || t > 15) return false;
    const int m = kTxW[s] > kTxH[s] ? kTxW[s] : kTxH[s];
    if (m == 64) return t == 0;
    if (m == 32) return t == 0 || t == 9;
    return true;
}

// ---- tx-type search: which types ProductFullLoopTxSearch evaluates -------------------
// Bit t of the result = tx type t is tried for a luma block of `tx_size`
// (Source/Lib/Common/Codec/EbFullLoop.c:946-982):
//   * the type must belong to the extended-tx set of the size: get_ext_tx_set_type +
//     av1_ext_tx_used (EbDefinitions.h:1472-1500) -- DCT only above 32x32 (square-up), DCT (+IDTX
//     for inter) at 32x32, otherwise by intra/inter, the reduced-set flag and whether the
//     square-down size is 16x16;
//   * with tx_search_reduced_set the encoder's own shortlist allowed_tx_set_a
//     (EbFullLoop.c:883-902) is applied on top;
//   * a size whose set comes out empty still tries DCT_DCT (:973-975).
inline uint32_t tx_search_mask(int tx_size, int is_inter, int reduced_tx_set_used, int tx_search_reduced_set) {
    if (tx_size < 0 || tx_size > 18) return 0;
    enum : uint32_t { DCTONLY = 0x0001, DCT_IDTX = 0x0201, DTT4_IDTX = 0x020F, DTT4_IDTX_1DDCT = 0x0E0F,
                      DTT9_IDTX_1DDCT = 0x0FFF, ALL16 = 0xFFFF };
    const int w = kTxW[tx_size], h = kTxH[tx_size];
    const int up = w > h ? w : h, down = w < h ? w : h;  // txsize_sqr_up_map / txsize_sqr_map
    uint32_t set;
    if (up > 32) set = DCTONLY;
    else if (up == 32) set = is_inter ? DCT_IDTX : DCTONLY;
    else if (reduced_tx_set_used) set = is_inter ? DCT_IDTX : DTT4_IDTX;
    else if (is_inter) set = down == 16 ? DTT9_IDTX_1DDCT : ALL16;
    else set = down == 16 ? DTT4_IDTX : DTT4_IDTX_1DDCT;
    if (tx_search_reduced_set) {
        // allowed_tx_set_a, one mask per TxSize: the four DCT/ADST combinations + IDTX, V_DCT, H_DCT
        // for the small sizes (8x8 and 8x16 add V_ADST / H_ADST, 16x8 adds V_ADST / V_FLIPADST),
        // {DCT_DCT, IDTX} for the 32-point sizes, DCT_DCT for the 64-point sizes
        static const uint16_t kSetA[19] = { 0x0E0F, 0xAE0F, 0x0E0F, 0x0201, 0x0001, 0x0E0F, 0x0E0F, 0xAE0F,
                                            0x5E0F, 0x0201, 0x0201, 0x0001, 0x0001, 0x0E0F, 0x0E0F, 0x0201,
                                            0x0201, 0x0001, 0x0001 };
        set &= kSetA[tx_size];
    }
    return set ? set : DCTONLY;
}

// One candidate per (block, tried type): a copy of the block with that tx_type, its own output
// slot (`cand` index = eob / dist slot) and, when coefficients are wanted, its own coefficient
// range.  cand must hold 16 * n entries.  Returns the candidate count.
inline int tx_search_expand(const SvtGpuBlock* blocks, int n, int is_inter, int reduced_tx_set_used,
                            int tx_search_reduced_set, int want_coeff, SvtGpuBlock* cand,
                            int32_t* cand_block, int64_t* coeff_elems) {
    int k = 0;
    int64_t co = 0;
    for (int i = 0; i < n; ++i) {
        const uint32_t m = tx_search_mask(blocks[i].tx_size, is_inter, reduced_tx_set_used, tx_search_reduced_set);
        for (int t = 0; t < 16; ++t) {
            if (!((m >> t) & 1)) continue;
            cand[k] = blocks[i];
            cand[k].tx_type = uint8_t(t);
            cand[k].coeff_off = want_coeff ? uint32_t(co) : 0;
            if (want_coeff) co += max_eob(blocks[i].tx_size);
            if (cand_block) cand_block[k] = i;
            ++k;
        }
    }
    if (coeff_elems) *coeff_elems = co;
    return k;
}

// Scan order of the coded region (kw x kh, 64-point sizes use their 32-wide counterpart), as a
// rule instead of the reference's literal tables (EbTransforms.h:331-3348, selected by
// av1_scan_orders :3349): class 1 "mrow" is raster order, class 2 "mcol" is column-major, class 0
// walks the anti-diagonals -- always top-right to bottom-left for tall blocks, always bottom-left
// to top-right for wide blocks, and alternating (odd diagonals downwards) for squares.
inline void make_scan(int s, int cls, std::vector<int16_t>& scan) {
    const int kw = coded_w(s), kh = coded_h(s);
    scan.clear();
    if (cls == 1) {
        for (int i = 0; i < kw * kh; ++i) scan.push_back(int16_t(i));
    } else if (cls == 2) {
        for (int c = 0; c < kw; ++c)
            for (int r = 0; r < kh; ++r) scan.push_back(int16_t(r * kw + c));
    } else {
        for (int d = 0; d < kw + kh - 1; ++d) {
            const bool down = kh > kw ? true : (kw > kh ? false : (d & 1));
            const int r_lo = d - (kw - 1) > 0 ? d - (kw - 1) : 0;
            const int r_hi = d < kh - 1 ? d : kh - 1;
            if (down) for (int r = r_lo; r <= r_hi; ++r) scan.push_back(int16_t(r * kw + (d - r)));
            else      for (int r = r_hi; r >= r_lo; --r) scan.push_back(int16_t(r * kw + (d - r)));
        }
    }
}
inline void make_iscan(int s, int cls, std::vector<int16_t>& iscan) {
    std::vector<int16_t> scan;
    make_scan(s, cls, scan);
    iscan.assign(scan.size(), 0);
    for (size_t i = 0; i < scan.size(); ++i) iscan[scan[i]] = int16_t(i);
}
// all 19 x 3 tables back to back (sizes that share a coded shape share a table)
inline void build_iscan_tables(std::vector<int16_t>& all, int off[19][3]) {
    all.clear();
    for (int s = 0; s < 19; ++s)
        for (int c = 0; c < 3; ++c) {
            int share = -1;
            for (int s2 = 0; s2 < s; ++s2)
                if (coded_w(s2) == coded_w(s) && coded_h(s2) == coded_h(s)) { share = s2; break; }
            if (share >= 0) { off[s][c] = off[share][c]; continue; }
            std::vector<int16_t> t;
            make_iscan(s, c, t);
            off[s][c] = int(all.size());
            all.insert(all.end(), t.begin(), t.end());
        }
}

// Validates the caller's list and buckets it by tx_size.  Inside a size bucket the caller's order is
// kept in windows of 2^sort_window_log2() blocks, and each window is stable-sorted by the 1-D kinds of the tx type: warps then
// mostly see one type (no divergence between 1-D kinds), while blocks that are neighbours in the
// frame -- and share 32-byte DRAM sectors of the pixel rows -- are still processed within a few
// CTAs of each other, so the second half of a sector is an L2 hit, not another DRAM read.
// Returns 0 or the 1-based index of the first bad block.
#include <stdlib.h>
// tuning knob SVTGPU_SORT_WINDOW_LOG2 (8..24).  Measured on B200, 1080p8, ms per step:
// 10 -> 0.2605, 12 -> 0.2574, 13 -> 0.2556, 14 -> 0.2524, 16 -> 0.2616, 20 (whole bucket) -> 0.2767:
// longer runs of one tx type keep the instruction working set small, until neighbouring blocks
// stop meeting in L2.
inline int sort_window_log2() {
    static const int v = [] {
        const char* e = getenv("SVTGPU_SORT_WINDOW_LOG2");
        const int x = e ? atoi(e) : 14;
        return x < 8 ? 8 : (x > 24 ? 24 : x);
    }();
    return v;
}
struct BucketScratch {  // reused across calls by callers that bucket a picture per submission
    std::vector<uint32_t> ord;
    std::vector<int> pos;
};
// `out` must hold n descriptors (it may be page-locked staging memory: it is written exactly once,
// in three sequential passes over `in`)
inline int bucket_blocks(const SvtGpuBlock* in, int n, DevBlock* out, int first[19], int count[19],
                         BucketScratch& sc) {
    const int kSortWindowLog2 = sort_window_log2(), kSortWindow = 1 << kSortWindowLog2;
    uint16_t valid[19];
    for (int s = 0; s < 19; ++s) {
        valid[s] = 0;
        for (int t = 0; t < 16; ++t) valid[s] |= uint16_t(tx_pair_valid(s, t)) << t;
    }
    // sort key of a tx type: its (column kind, row kind) pair -- the flips are data, not code, so
    // e.g. ADST_DCT and FLIPADST_DCT run the same instructions and sort together
    uint8_t tkey[16];
    for (int t = 0; t < 16; ++t) tkey[t] = uint8_t(kind_of(vtx_of(t)) * 3 + kind_of(htx_of(t)));
    sc.ord.resize(size_t(n));
    uint32_t* ord = sc.ord.data();  // position of the block inside its size bucket, caller order
    int hist[19] = {};
    for (int i = 0; i < n; ++i) {
        const SvtGpuBlock& b = in[i];
        if (b.tx_size >= 19 || b.tx_type >= 16 || !((valid[b.tx_size] >> b.tx_type) & 1)) return i + 1;
        if ((b.bit_depth != 8 && b.bit_depth != 10) || (b.coeff_off & 3) || b.plane > 2) return i + 1;
        ord[i] = uint32_t(hist[b.tx_size]++);
    }
    int acc = 0, key0[19], nkeys = 0;
    for (int s = 0; s < 19; ++s) {
        first[s] = acc;
        count[s] = hist[s];
        acc += hist[s];
        key0[s] = nkeys;  // keys of a size: (window, type), window-major
        nkeys += ((hist[s] + kSortWindow - 1) >> kSortWindowLog2) * 16;
    }
    sc.pos.assign(size_t(nkeys) + 1, 0);
    int* pos = sc.pos.data();
    for (int i = 0; i < n; ++i)
        ++pos[key0[in[i].tx_size] + int(ord[i] >> kSortWindowLog2) * 16 + tkey[in[i].tx_type] + 1];
    for (int k = 0; k < nkeys; ++k) pos[k + 1] += pos[k];
    for (int i = 0; i < n; ++i) {
        const SvtGpuBlock& b = in[i];
        DevBlock d;
        d.res_off = b.res_off; d.res_stride = b.res_stride;
        d.rec_off = b.rec_off; d.rec_stride = b.rec_stride;
        d.coeff_off = b.coeff_off; d.qrow = b.qrow; d.id = uint32_t(i);
        d.meta = make_meta(b.tx_type, b.bit_depth, b.plane, b.flags & SVTGPU_BLK_CLAMP16,
                           b.flags & SVTGPU_BLK_USE_QM);
        out[pos[key0[b.tx_size] + int(ord[i] >> kSortWindowLog2) * 16 + tkey[b.tx_type]]++] = d;
    }
    return 0;
}
inline int bucket_blocks(const SvtGpuBlock* in, int n, std::vector<DevBlock>& out, int first[19],
                         int count[19]) {
    BucketScratch sc;
    out.resize(size_t(n));
    return bucket_blocks(in, n, out.data(), first, count, sc);
}

}  // namespace svt
