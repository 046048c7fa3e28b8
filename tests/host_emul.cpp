// TEST INFRASTRUCTURE ONLY.  Runs the product's device phase functions
// (cidana_svt_av1_b200/csrc/residual_kernel.cuh) on the host, task by task, so that the kernel
// arithmetic can be checked against the oracle in the CPU-only test tier.  This is not a CPU
// fallback: nothing in the product links or loads it.
#include "../cidana_svt_av1_b200/csrc/host_plan.h"
#include <random>
#include <vector>
using namespace svt;

// how many blocks took the clamp-free inverse (shared by the parts; defined in part 0)
extern long long g_emul_blocks, g_emul_fast_blocks;
#if !defined(EMUL_PART) || EMUL_PART == 0
long long g_emul_blocks = 0, g_emul_fast_blocks = 0;
extern "C" void emul_fast_stats(long long* blocks, long long* fast) {
    *blocks = g_emul_blocks; *fast = g_emul_fast_blocks; g_emul_blocks = g_emul_fast_blocks = 0;
}
#endif

template <int W, int H, class C = CfgRuntime>
static void run_size(const KParams& p, int first, int count) {
    using G = Geo<W, H>;
    std::vector<int> buf(G::BLK);
    alignas(16) static thread_local int16_t stage[64 * 64];
    for (int i = first; i < first + count; ++i) {
        const DevBlock& d = p.blocks[i];
        if (p.stages & STAGE_FWD) {
            stage_residual<W, H, 1, C>(p, d, stage, 0);
            for (int c = 0; c < W; ++c) phase_fwd_col<W, H, C::SRC_STAGED>(d, c, buf.data(), stage);
        }
        if ((p.stages & STAGE_INV) && !C::SRC_STAGED) stage_pred<W, H, 1, C>(p, d, stage, 0);
        int eob = 0;
        unsigned long long energy = 0;
        QRow q = {};
        if (p.stages & STAGE_QUANT)  // the kernel instantiation that runs this batch: QM = (matrices uploaded and no dist)
            q = (!p.dist && p.qm_wt) ? load_qrow<G::LOG_SCALE, true>(p.qtables + size_t(d.qrow) * 40)
                                     : load_qrow<G::LOG_SCALE, false>(p.qtables + size_t(d.qrow) * 40);
        unsigned long long dist[2] = { 0, 0 };
        int fast_flag = 0;
        // The kernel decides per block, from S = sum |dqcoeff| (a warp-level reduction over the row
        // owners), whether the inverse transform runs without clamps (fast_limit).  Here the rows
        // run one after another, so the block is run twice: once with the clamped networks to get
        // S, and -- if S is inside the limit -- again from the saved forward-column output with
        // the clamp-free ones, which is what the kernel would have executed.
        const std::vector<int> saved = buf;
        for (int pass = 0; pass < 2; ++pass) {
            unsigned long long total = 0;
            eob = 0; energy = 0; dist[0] = dist[1] = 0;
            for (int r = 0; r < H; ++r) {
                int e; unsigned long long en, ds[2];
                RowAux aux;
                aux.fast = pass;
                const int16_t* isc_b = block_iscan<W, H>(p, d);
                if (p.dist) phase_row<W, H, false, true, C>(p, d, q, r, buf.data(), isc_b, e, en, ds, &fast_flag, aux);
                else if (p.qm_wt) phase_row<W, H, true, false, C>(p, d, q, r, buf.data(), isc_b, e, en, ds, &fast_flag, aux);
                else phase_row<W, H, false, false, C>(p, d, q, r, buf.data(), isc_b, e, en, ds, &fast_flag, aux);
                eob = e > eob ? e : eob;
                energy += en;
                dist[0] += ds[0];
                dist[1] += ds[1];
                total += aux.sum < (1u << 24) ? aux.sum : (1u << 24);
            }
            if (pass == 1) break;
            const int hk = hkind(d);
            const bool bd10 = C::bd10(d);
            const unsigned lim = hk == 0 ? fast_limit<W, H, 0>(bd10) : (hk == 1 ? fast_limit<W, H, 1>(bd10) : fast_limit<W, H, 2>(bd10));
            const bool can = (p.stages & STAGE_QUANT) && (p.stages & STAGE_INV) && total <= lim;
            if (!can) break;
            buf = saved;
            fast_flag = 1;
            ++g_emul_fast_blocks;
        }
        ++g_emul_blocks;
        if (p.dist) finish_dist<G::LOG_SCALE>(dist[0], dist[1], eob, energy, p.dist + 2 * size_t(d.id));
        if (p.stages & STAGE_QUANT) p.eob[d.id] = (uint16_t)eob;
        if ((p.stages & STAGE_FWD) && p.energy) p.energy[d.id] = energy;
        if (p.stages & STAGE_INV) {
            for (int c = 0; c < W; ++c) phase_inv_col<W, H, C>(d, c, buf.data(), stage, fast_flag != 0);
            unstage_recon<W, H, 1, C>(p, d, stage, 0);
        }
    }
}

// The compiled-in headline configuration (CfgFull<8> / <10>) when the launch qualifies for it,
// exactly as svt_gpu_launch.cu selects it; `allow_full` = 0 keeps the run-time switches.
template <int W, int H>
static void run_size_cfg(const KParams& p, int first, int count, int allow_full) {
    if (allow_full && p.dist && CfgFull<108>::matches(p)) run_size<W, H, CfgFull<108>>(p, first, count);
    else if (allow_full && p.dist && CfgFull<208>::matches(p)) run_size<W, H, CfgFull<208>>(p, first, count);
    else if (allow_full && !p.qm_wt && CfgFull<8>::matches(p)) run_size<W, H, CfgFull<8>>(p, first, count);
    else if (allow_full && CfgFull<10>::matches(p)) run_size<W, H, CfgFull<10>>(p, first, count);
    else run_size<W, H>(p, first, count);
}
// The size cases are spread over EMUL_PARTS translation units (the same file compiled with
// -DEMUL_PART=k, in parallel, by tests/emul_lib.py): the fully unrolled 32- and 64-point networks
// make a single unit take minutes to compile.  Part 0 also carries the C API below.
#ifndef EMUL_PART
#define EMUL_PART 0
#endif
#define EMUL_PARTS 6
int emul_run_bucket_part1(const KParams& p, int tx_size, int first, int count, int allow_full);
int emul_run_bucket_part2(const KParams& p, int tx_size, int first, int count, int allow_full);
int emul_run_bucket_part3(const KParams& p, int tx_size, int first, int count, int allow_full);
int emul_run_bucket_part4(const KParams& p, int tx_size, int first, int count, int allow_full);
int emul_run_bucket_part5(const KParams& p, int tx_size, int first, int count, int allow_full);
#define CASE(id, w, h) case id: run_size_cfg<w, h>(p, first, count, allow_full); return 0;
#if EMUL_PART == 1
int emul_run_bucket_part1(const KParams& p, int tx_size, int first, int count, int allow_full) {
    switch (tx_size) { CASE(4, 64, 64) }
    return -1;
}
#elif EMUL_PART == 2
int emul_run_bucket_part2(const KParams& p, int tx_size, int first, int count, int allow_full) {
    switch (tx_size) { CASE(11, 32, 64) CASE(17, 16, 64) }
    return -1;
}
#elif EMUL_PART == 3
int emul_run_bucket_part3(const KParams& p, int tx_size, int first, int count, int allow_full) {
    switch (tx_size) { CASE(12, 64, 32) CASE(18, 64, 16) }
    return -1;
}
#elif EMUL_PART == 4
int emul_run_bucket_part4(const KParams& p, int tx_size, int first, int count, int allow_full) {
    switch (tx_size) { CASE(3, 32, 32) CASE(9, 16, 32) CASE(10, 32, 16) }
    return -1;
}
#elif EMUL_PART == 5
int emul_run_bucket_part5(const KParams& p, int tx_size, int first, int count, int allow_full) {
    switch (tx_size) { CASE(15, 8, 32) CASE(16, 32, 8) CASE(2, 16, 16) CASE(7, 8, 16) CASE(8, 16, 8) }
    return -1;
}
#else
static int g_allow_full = 0;
extern "C" void emul_allow_full(int on) { g_allow_full = on; }
static int run_bucket(const KParams& p, int tx_size, int first, int count) {
    const int allow_full = g_allow_full;
    switch (tx_size) {
        CASE(0, 4, 4) CASE(1, 8, 8) CASE(5, 4, 8) CASE(6, 8, 4) CASE(13, 4, 16) CASE(14, 16, 4)
    case 4: return emul_run_bucket_part1(p, tx_size, first, count, allow_full);
    case 11: case 17: return emul_run_bucket_part2(p, tx_size, first, count, allow_full);
    case 12: case 18: return emul_run_bucket_part3(p, tx_size, first, count, allow_full);
    case 3: case 9: case 10: return emul_run_bucket_part4(p, tx_size, first, count, allow_full);
    case 15: case 16: case 2: case 7: case 8: return emul_run_bucket_part5(p, tx_size, first, count, allow_full);
    }
    return -1;
}
#endif
#undef CASE

#if EMUL_PART == 0
// Same contract as svtgpu_plan_create + svtgpu_run, with host pointers in `bufs`.
extern "C" int emul_run_batch(const SvtGpuBlock* blocks, int n, const SvtGpuBuffers* bufs,
                              const int16_t* qtables, const uint8_t* wt, const uint8_t* iwt,
                              const int32_t* qm_off, int stages) {
    std::vector<DevBlock> dev;
    int first[19], count[19];
    const int bad = bucket_blocks(blocks, n, dev, first, count);
    if (bad) return -bad;
    std::vector<int16_t> iscan;
    KParams p = {};
    build_iscan_tables(iscan, p.iscan_off);
    p.blocks = dev.data(); p.n_blocks = n; p.stages = stages;
    int n10 = 0;
    for (int i = 0; i < n; ++i) {
        n10 += blocks[i].bit_depth == 10;
        p.any_clamp16 |= (blocks[i].flags & SVTGPU_BLK_CLAMP16) != 0;
    }
    p.bd_uniform = n10 == 0 ? 8 : (n10 == n ? 10 : 0);
    p.residual = bufs->residual; p.pred = bufs->pred; p.recon = bufs->recon;
    p.coeff_in = bufs->coeff_in; p.coeff_out = bufs->coeff; p.qcoeff = bufs->qcoeff;
    p.dqcoeff = bufs->dqcoeff; p.eob = bufs->eob; p.energy = bufs->energy;
    p.src = bufs->src; p.dist = reinterpret_cast<unsigned long long*>(bufs->dist);
    p.levels = bufs->levels; p.levels_off = bufs->levels_off;
    p.qtables = qtables; p.iscan = iscan.data(); p.qm_wt = wt; p.qm_iwt = iwt;
    for (int i = 0; i < 2 * 19; ++i) (&p.qm_off[0][0])[i] = qm_off ? qm_off[i] : 0;
    for (int s = 0; s < 19; ++s)
        if (count[s]) run_bucket(p, s, first[s], count[s]);
    return 0;
}
extern "C" int emul_get_iscan(int tx_size, int cls, int16_t* out) {
    std::vector<int16_t> t;
    make_iscan(tx_size, cls, t);
    for (size_t i = 0; i < t.size(); ++i) out[i] = t[i];
    return (int)t.size();
}
extern "C" int emul_cospi(int bit, int j) { return kCos.v[bit - 10][j]; }
extern "C" unsigned emul_tx_search_mask(int s, int inter, int reduced, int shortlist) {
    return tx_search_mask(s, inter, reduced, shortlist);
}
extern "C" int emul_tx_search_expand(const SvtGpuBlock* blocks, int n, int inter, int reduced, int shortlist,
                                     int want_coeff, SvtGpuBlock* cand, int32_t* cand_block, int64_t* coeff_elems) {
    return tx_search_expand(blocks, n, inter, reduced, shortlist, want_coeff, cand, cand_block, coeff_elems);
}

// ---- the two facts fast_limit (residual_kernel.cuh) rests on ---------------------------------------
// (1) every add/sub node of the inverse DCT / ADST networks has |coefficient| <= 1 with respect to
//     every single input: run unit vectors (scaled by 2^14, so that rounding is negligible) through
//     the networks with a policy that records the largest add/sub result.
static long long g_node_max;
struct TrackNodes {
    int operator()(int v) const { return v; }
    int add(int a, int b) const { const long long r = (long long)a + b; g_node_max = r > g_node_max ? r : (-r > g_node_max ? -r : g_node_max); return int(r); }
    int sub(int a, int b) const { const long long r = (long long)a - b; g_node_max = r > g_node_max ? r : (-r > g_node_max ? -r : g_node_max); return int(r); }
};
template <int N, int KIND>
static double node_gain() {
    const int S = 1 << 14;
    long long worst = 0;
    for (int i = 0; i < N; ++i) {
        int x[N] = {};
        x[i] = S;
        g_node_max = 0;
        inv1d_k<N, KIND>(x, TrackNodes());
        worst = g_node_max > worst ? g_node_max : worst;
        for (int k = 0; k < N; ++k) { const long long a = x[k] < 0 ? -(long long)x[k] : x[k]; worst = a > worst ? a : worst; }
    }
    return double(worst) / S;
}
extern "C" double emul_inverse_node_gain(int n, int kind) {
    if (kind == 0) return n == 4 ? node_gain<4, 0>() : n == 8 ? node_gain<8, 0>() : n == 16 ? node_gain<16, 0>() : n == 32 ? node_gain<32, 0>() : node_gain<64, 0>();
    if (kind == 1) return n == 4 ? node_gain<4, 1>() : n == 8 ? node_gain<8, 1>() : node_gain<16, 1>();
    return -1;
}
// (2) with sum |x_i| + kFastE below the clamp range, the clamped and the clamp-free network agree:
//     returns the number of vectors (of `count` random ones with that L1 norm) on which they differ
template <int N, int KIND>
static int clampfree_diff(unsigned seed, int l1, int bits, int count) {
    unsigned st = seed * 2654435761u + 12345u;
    auto rnd = [&] { st = st * 1664525u + 1013904223u; return st >> 8; };
    int bad = 0;
    for (int it = 0; it < count; ++it) {
        int x[N], y[N];
        // spread l1 over a random number of positions with random signs (spiky and flat vectors)
        const int k = 1 + int(rnd() % N);
        for (int i = 0; i < N; ++i) x[i] = 0;
        int left = l1;
        for (int j = 0; j < k; ++j) {
            const int pos = int(rnd() % N);
            const int v = j == k - 1 ? left : int(rnd() % unsigned(left + 1));
            left -= v;
            const int cur = x[pos] < 0 ? -x[pos] : x[pos];
            x[pos] = (rnd() & 1) ? (cur + v) : -(cur + v);
        }
        for (int i = 0; i < N; ++i) y[i] = x[i];
        inv1d_k<N, KIND>(x, Clamp(bits));
        inv1d_k<N, KIND>(y, NoClamp());
        for (int i = 0; i < N; ++i) bad += x[i] != y[i];
    }
    return bad;
}
extern "C" int emul_clampfree_diff(int n, int kind, unsigned seed, int l1, int bits, int count) {
    if (kind == 0) return n == 4 ? clampfree_diff<4, 0>(seed, l1, bits, count) : n == 8 ? clampfree_diff<8, 0>(seed, l1, bits, count)
                        : n == 16 ? clampfree_diff<16, 0>(seed, l1, bits, count) : n == 32 ? clampfree_diff<32, 0>(seed, l1, bits, count)
                        : clampfree_diff<64, 0>(seed, l1, bits, count);
    return n == 8 ? clampfree_diff<8, 1>(seed, l1, bits, count) : clampfree_diff<16, 1>(seed, l1, bits, count);
}
extern "C" int emul_fast_e() { return kFastE; }

// The two-multiply form of the flat quantizer (quant_flat) against its 64-bit form: every row of
// `tables` (40 int16 each), DC and AC, log_scale 0 and 1, |coeff| over a dense low range, every
// power of two +-1 and `n_random` random magnitudes below 2^limit_bits, both signs, both values
// of the 16-bit clamp.  Returns the number of differing (level, dequantised value) pairs.
template <int LS>
static long long quant_forms_diff(const int16_t* tables, int n_rows, int limit_bits, int n_random, unsigned seed) {
    long long bad = 0;
    std::vector<int> mags;
    for (int v = 0; v < 6000; ++v) mags.push_back(v);
    for (int b = 12; b < limit_bits; ++b) for (int d = -2; d <= 2; ++d) mags.push_back((1 << b) + d);
    std::mt19937 rng(seed);
    for (int i = 0; i < n_random; ++i) mags.push_back(int(rng() & ((1u << limit_bits) - 1u)));
    for (int r = 0; r < n_rows; ++r) {
        const QRow q = load_qrow<LS, false>(tables + size_t(r) * 40), qw = load_qrow<LS, true>(tables + size_t(r) * 40);
        for (int k = 0; k < 2; ++k) {
            const QSet& e = k ? q.ac : q.dc;
            const QSet& ew = k ? qw.ac : qw.dc;
            for (int m : mags)
                for (int sg = -1; sg <= 1; sg += 2)
                    for (int t1max : { 32767, 0x7fffffff }) {
                        int qa, da, qb, db;
                        unsigned sa = 0, sb = 0;
                        const int la = quant_flat<LS, false>(sg * m, e, t1max, qa, da, sa);
                        const int lb = quant_flat<LS, true>(sg * m, ew, t1max, qb, db, sb);
                        bad += la != lb || qa != qb || da != db || sa != sb;
                    }
        }
    }
    return bad;
}
// quant_qm (two-multiply form below 2^21, 64-bit form above) against the reference's arithmetic
// written out in 64 bits (highbd_quantize_b_helper_c, EbFullLoop.c:239-296), raw table rows.
template <int LS>
static long long quant_qm_diff(const int16_t* tables, int n_rows, int n_random, unsigned seed) {
    long long bad = 0;
    std::vector<int> mags;
    for (int v = 0; v < 3000; ++v) mags.push_back(v);
    for (int b = 12; b < 30; ++b) for (int d = -2; d <= 2; ++d) mags.push_back((1 << b) + d);
    for (int d = -70000; d <= 70000; d += 4999) mags.push_back((1 << 21) + d);  // around the guard
    std::mt19937 rng(seed);
    for (int i = 0; i < n_random; ++i) mags.push_back(int(rng() & ((1u << 24) - 1u)));
    const int wts[] = { 1, 7, 32, 33, 64, 127, 128, 200, 255 };
    for (int r = 0; r < n_rows; r += 3) {
        const QRow q = load_qrow<LS, true>(tables + size_t(r) * 40);
        for (int k = 0; k < 2; ++k) {
            const QSet& e = k ? q.ac : q.dc;
            for (int m : mags)
                for (int sg = -1; sg <= 1; sg += 2)
                    for (int wt : wts) {
                        const int iwt = 288 - wt > 255 ? 255 : (288 - wt < 1 ? 1 : 288 - wt);
                        int qa, da;
                        unsigned sa = 0;
                        const int la = quant_qm<LS>(sg * m, e, wt, iwt, qa, da, sa);
                        // the reference, 64-bit throughout
                        const int cf = sg * m, sign = cf >> 31, ab = (cf ^ sign) - sign;
                        const int cw = int(unsigned(cf) * unsigned(wt));
                        int qb = 0, db = 0, lb = 0;
                        if (cw >= e.zbin * 32 || cw <= -e.zbin * 32) {
                            const int64_t tw = (int64_t(ab) + e.round) * wt;
                            const int64_t t2 = ((tw * e.quant) >> 16) + tw;
                            lb = int((t2 * e.qshift) >> (16 - LS + 5));
                            const int dq = (e.dequant * iwt + 16) >> 5;
                            const int adq = int(unsigned(lb) * unsigned(dq)) >> LS;
                            qb = (lb ^ sign) - sign;
                            db = (adq ^ sign) - sign;
                        }
                        bad += la != lb || qa != qb || da != db;
                    }
        }
    }
    return bad;
}
extern "C" long long emul_quant_qm_diff(const int16_t* tables, int n_rows, int log_scale, int n_random, unsigned seed) {
    return log_scale == 0 ? quant_qm_diff<0>(tables, n_rows, n_random, seed)
         : log_scale == 1 ? quant_qm_diff<1>(tables, n_rows, n_random, seed) : quant_qm_diff<2>(tables, n_rows, n_random, seed);
}
extern "C" long long emul_quant_forms_diff(const int16_t* tables, int n_rows, int log_scale, int limit_bits,
                                           int n_random, unsigned seed) {
    return log_scale == 0 ? quant_forms_diff<0>(tables, n_rows, limit_bits, n_random, seed)
                          : quant_forms_diff<1>(tables, n_rows, limit_bits, n_random, seed);
}
#endif  // EMUL_PART == 0
