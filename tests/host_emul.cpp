// TEST INFRASTRUCTURE ONLY.  Runs the product's device phase functions
// (cidana_svt_av1_b200/csrc/residual_kernel.cuh) on the host, task by task, so that the kernel
// arithmetic can be checked against the oracle in the CPU-only test tier.  This is not a CPU
// fallback: nothing in the product links or loads it.
#include "../cidana_svt_av1_b200/csrc/host_plan.h"
#include <vector>
using namespace svt;

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
        if (p.stages & STAGE_QUANT) q = load_qrow<G::LOG_SCALE>(p.qtables + size_t(d.qrow) * 40);
        unsigned long long dist[2] = { 0, 0 };
        for (int r = 0; r < H; ++r) {
            int e; unsigned long long en, ds[2];
            const int16_t* isc_b = block_iscan<W, H>(p, d);
            if (p.dist) phase_row<W, H, false, true, C>(p, d, q, r, buf.data(), isc_b, e, en, ds);
            else if (p.qm_wt) phase_row<W, H, true, false, C>(p, d, q, r, buf.data(), isc_b, e, en, ds);
            else phase_row<W, H, false, false, C>(p, d, q, r, buf.data(), isc_b, e, en, ds);
            eob = e > eob ? e : eob;
            energy += en;
            dist[0] += ds[0];
            dist[1] += ds[1];
        }
        if (p.dist) finish_dist<G::LOG_SCALE>(dist[0], dist[1], eob, energy, p.dist + 2 * size_t(d.id));
        if (p.stages & STAGE_QUANT) p.eob[d.id] = (uint16_t)eob;
        if ((p.stages & STAGE_FWD) && p.energy) p.energy[d.id] = energy;
        if (p.stages & STAGE_INV) {
            for (int c = 0; c < W; ++c) phase_inv_col<W, H, C>(d, c, buf.data(), stage);
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
static int g_allow_full = 0;
extern "C" void emul_allow_full(int on) { g_allow_full = on; }
static int run_bucket(const KParams& p, int tx_size, int first, int count) {
    switch (tx_size) {
#define CASE(id, w, h) case id: run_size_cfg<w, h>(p, first, count, g_allow_full); return 0;
        CASE(0, 4, 4) CASE(1, 8, 8) CASE(2, 16, 16) CASE(3, 32, 32) CASE(4, 64, 64)
        CASE(5, 4, 8) CASE(6, 8, 4) CASE(7, 8, 16) CASE(8, 16, 8) CASE(9, 16, 32)
        CASE(10, 32, 16) CASE(11, 32, 64) CASE(12, 64, 32) CASE(13, 4, 16) CASE(14, 16, 4)
        CASE(15, 8, 32) CASE(16, 32, 8) CASE(17, 16, 64) CASE(18, 64, 16)
#undef CASE
    }
    return -1;
}

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
