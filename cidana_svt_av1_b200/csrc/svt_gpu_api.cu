// libsvtgpu.so: C ABI (include/svt_gpu.h) over the fused residual kernel.  Level C (batch API).
// Levels A/B (RTCD slots and entry points with host pointers) live in svt_gpu_rtcd.cu.
//
// There is no CPU fallback anywhere in this library: every entry point either launches the
// sm_100a kernels or fails loudly.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <atomic>
#include <new>
#include <vector>

#include "../../include/svt_gpu.h"
#include "host_plan.h"
#include "residual_kernel.cuh"  // KParams, stage flags (no kernel is instantiated in this file)
#include "svt_gpu_internal.h"

using namespace svt;

static thread_local char g_err[512] = "";
void svtgpu_set_error(const char* fmt, const char* a, const char* b, int line) {
    snprintf(g_err, sizeof(g_err), fmt, a, b, line);
}
extern "C" const char* svtgpu_last_error(void) { return g_err; }

std::atomic<long long> g_total_launches{ 0 };  // every context of the process (levels A / B use one per thread)
extern "C" long long svtgpu_total_launch_count(void) { return g_total_launches.load(); }

// ---- context -------------------------------------------------------------------------------------
extern "C" int svtgpu_create(SvtGpuContext** out, int device) {
    if (!out) return SVTGPU_ERR_ARG;
    *out = nullptr;
    int n = 0;
    CK(cudaGetDeviceCount(&n));
    if (device < 0 || device >= n) { svtgpu_set_error("%s%s (line %d)", "no such CUDA device", "", __LINE__); return SVTGPU_ERR_ARG; }
    CK(cudaSetDevice(device));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) {
        svtgpu_set_error("%s%s (line %d)", "libsvtgpu is built for sm_100a only; device is ", prop.name, __LINE__);
        return SVTGPU_ERR_CUDA;
    }
    SvtGpuContext* c = new (std::nothrow) SvtGpuContext();
    if (!c) return SVTGPU_ERR_NOMEM;
    c->device = device;
    for (int pl = 0; pl < 2; ++pl)
        for (int s = 0; s < 19; ++s) c->qm_off[pl][s] = -1;
    const int rc = [&]() -> int {
        CK(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
        for (int i = 0; i < SvtGpuContext::kSide; ++i) {
            CK(cudaStreamCreateWithFlags(&c->side[i], cudaStreamNonBlocking));
            CK(cudaEventCreateWithFlags(&c->side_done[i], cudaEventDisableTiming));
        }
        CK(cudaEventCreateWithFlags(&c->fork, cudaEventDisableTiming));
        std::vector<int16_t> iscan;
        build_iscan_tables(iscan, c->iscan_off);
        CK(cudaMalloc(&c->d_iscan, iscan.size() * sizeof(int16_t)));
        CK(cudaMemcpy(c->d_iscan, iscan.data(), iscan.size() * sizeof(int16_t), cudaMemcpyHostToDevice));
        return SVTGPU_OK;
    }();
    if (rc != SVTGPU_OK) {  // nothing half-built survives an early error
        svtgpu_destroy(c);
        return rc;
    }
    *out = c;
    return SVTGPU_OK;
}

extern "C" void svtgpu_destroy(SvtGpuContext* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    svtgpu_host_release(c);
    if (c->stream) cudaStreamSynchronize(c->stream);
    for (int i = 0; i < SvtGpuContext::kSide; ++i) {
        if (c->side[i]) { cudaStreamSynchronize(c->side[i]); cudaStreamDestroy(c->side[i]); }
        if (c->side_done[i]) cudaEventDestroy(c->side_done[i]);
    }
    if (c->fork) cudaEventDestroy(c->fork);
    cudaFree(c->d_iscan);
    cudaFree(c->d_qtables);
    cudaFree(c->d_qm_wt);
    cudaFree(c->d_qm_iwt);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

extern "C" int svtgpu_upload_quant_tables(SvtGpuContext* c, const int16_t* tables, int n_rows) {
    if (!c || !tables || n_rows <= 0) return SVTGPU_ERR_ARG;
    // The fused kernels multiply by quant_shift << (16 + log_scale) (quant_flat): exact for
    // 0 <= quant_shift < 2^15, which covers everything invert_quant can produce
    // (1 << (16 - msb(d)) with d >= 4, EbModeDecisionConfigurationProcess.c:322-329).
    for (int r = 0; r < n_rows; ++r)
        for (int k = 0; k < 2; ++k)
            if (tables[size_t(r) * 40 + 24 + k] < 0) {
                char msg[160];
                snprintf(msg, sizeof msg, "svtgpu_upload_quant_tables: row %d has a negative quant_shift (%d); "
                         "av1_build_quantizer produces 1 << (16 - msb(q))", r, int(tables[size_t(r) * 40 + 24 + k]));
                svtgpu_set_error("%s%s (line %d)", msg, "", __LINE__);
                return SVTGPU_ERR_ARG;
            }
    CK(cudaSetDevice(c->device));
    CK(cudaStreamSynchronize(c->stream));
    cudaFree(c->d_qtables);
    c->d_qtables = nullptr;
    const size_t bytes = size_t(n_rows) * 40 * sizeof(int16_t);
    CK(cudaMalloc(&c->d_qtables, bytes));
    CK(cudaMemcpy(c->d_qtables, tables, bytes, cudaMemcpyHostToDevice));
    c->n_qrows = n_rows;
    return SVTGPU_OK;
}

extern "C" int svtgpu_upload_qmatrix(SvtGpuContext* c, const uint8_t* wt, const uint8_t* iwt,
                                     int n_bytes, const int32_t* off) {
    if (!c || !wt || !iwt || n_bytes <= 0 || !off) return SVTGPU_ERR_ARG;
    CK(cudaSetDevice(c->device));
    CK(cudaStreamSynchronize(c->stream));
    cudaFree(c->d_qm_wt);
    cudaFree(c->d_qm_iwt);
    c->d_qm_wt = c->d_qm_iwt = nullptr;
    for (int pl = 0; pl < 2; ++pl)
        for (int s = 0; s < 19; ++s) c->qm_off[pl][s] = -1;  // nothing usable until this upload succeeds
    // repack: every (plane class, size) matrix at a 16-byte aligned offset, so that the kernel
    // can fetch four weights with one 32-bit load
    std::vector<uint8_t> pw, pi;
    int noff[2][19];
    for (int pl = 0; pl < 2; ++pl)
        for (int s = 0; s < 19; ++s) {
            const int o = off[pl * 19 + s], n = max_eob(s);
            noff[pl][s] = -1;
            if (o < 0) continue;
            if (o + n > n_bytes) return SVTGPU_ERR_ARG;
            noff[pl][s] = int(pw.size());
            pw.insert(pw.end(), wt + o, wt + o + n);
            pi.insert(pi.end(), iwt + o, iwt + o + n);
            while (pw.size() & 15) { pw.push_back(0); pi.push_back(0); }
        }
    if (pw.empty()) return SVTGPU_ERR_ARG;
    cudaError_t e = cudaMalloc(&c->d_qm_wt, pw.size());
    if (e == cudaSuccess) e = cudaMalloc(&c->d_qm_iwt, pi.size());
    if (e == cudaSuccess) e = cudaMemcpy(c->d_qm_wt, pw.data(), pw.size(), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(c->d_qm_iwt, pi.data(), pi.size(), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {  // leave the context without matrices rather than with half of them
        cudaFree(c->d_qm_wt);
        cudaFree(c->d_qm_iwt);
        c->d_qm_wt = c->d_qm_iwt = nullptr;
        CK(e);
    }
    memcpy(c->qm_off, noff, sizeof(c->qm_off));
    return SVTGPU_OK;
}

// ---- plans ---------------------------------------------------------------------------------------
extern "C" int svtgpu_plan_create(SvtGpuContext* c, const SvtGpuBlock* blocks, int n, SvtGpuPlan** out) {
    if (!c || !out || n < 0 || (n > 0 && !blocks)) return SVTGPU_ERR_ARG;
    *out = nullptr;
    CK(cudaSetDevice(c->device));
    SvtGpuPlan* pl = new (std::nothrow) SvtGpuPlan();
    if (!pl) return SVTGPU_ERR_NOMEM;
    pl->ctx = c;
    pl->n = n;
    std::vector<DevBlock> dev;
    const int bad = bucket_blocks(blocks, n, dev, pl->first, pl->count);
    if (bad) {
        char idx[32];
        snprintf(idx, sizeof(idx), "%d", bad - 1);
        svtgpu_set_error("%s%s (line %d)", "invalid SvtGpuBlock at index ", idx, __LINE__);
        delete pl;
        return SVTGPU_ERR_ARG;
    }
    uint32_t max_qrow = 0;
    for (int i = 0; i < n; ++i) max_qrow = dev[i].qrow > max_qrow ? dev[i].qrow : max_qrow;
    pl->max_qrow = max_qrow;
    int n10 = 0;
    for (int i = 0; i < n; ++i) {
        pl->uses_qm |= (dev[i].meta & META_USE_QM) != 0;
        if (dev[i].meta & META_USE_QM)
            pl->qm_sizes[((dev[i].meta >> META_PLANE_SHIFT) & 3) > 0] |= 1u << blocks[dev[i].id].tx_size;
        n10 += (dev[i].meta & META_BD10) != 0;
        pl->any_clamp16 |= (dev[i].meta & META_CLAMP16) != 0;
    }
    pl->bd_uniform = n10 == 0 ? 8 : (n10 == n ? 10 : 0);
    for (int i = 0; i < n; ++i) {
        const SvtGpuBlock& b = blocks[i];
        const size_t w = size_t(kTxW[b.tx_size]), h = size_t(kTxH[b.tx_size]);
        pl->need_res = std::max(pl->need_res, size_t(b.res_off) + (h - 1) * b.res_stride + w);
        pl->need_pix = std::max(pl->need_pix, size_t(b.rec_off) + (h - 1) * b.rec_stride + w);
        pl->need_coeff = std::max(pl->need_coeff, size_t(b.coeff_off) + size_t(max_eob(b.tx_size)));
    }
    if (n > 0) {
        if (cudaMalloc(&pl->d_blocks, size_t(n) * sizeof(DevBlock)) != cudaSuccess) {
            delete pl;
            cudaGetLastError();
            svtgpu_set_error("%s%s (line %d)", "out of device memory for the plan's descriptors", "", __LINE__);
            return SVTGPU_ERR_NOMEM;
        }
        cudaError_t e = cudaMemcpyAsync(pl->d_blocks, dev.data(), size_t(n) * sizeof(DevBlock), cudaMemcpyHostToDevice, c->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
        if (e != cudaSuccess) {
            cudaFree(pl->d_blocks);
            delete pl;
            CK(e);
        }
    }
    *out = pl;
    return SVTGPU_OK;
}

extern "C" void svtgpu_plan_destroy(SvtGpuPlan* pl) {
    if (!pl) return;
    cudaSetDevice(pl->ctx->device);
    cudaStreamSynchronize(pl->ctx->stream);
    cudaFree(pl->d_blocks);
    delete pl;
}

// ---- run -----------------------------------------------------------------------------------------
extern "C" int svtgpu_run(SvtGpuContext* c, const SvtGpuPlan* pl, const SvtGpuBuffers* b, int stages,
                          void* cuda_stream) {
    if (stages & ~(SVTGPU_STAGE_ALL | SVTGPU_STAGE_RESIDUAL)) return SVTGPU_ERR_ARG;
    return svtgpu_run_internal(c, pl, b, stages, cuda_stream);
}

int svtgpu_run_internal(SvtGpuContext* c, const SvtGpuPlan* pl, const SvtGpuBuffers* b, int stages,
                        void* cuda_stream) {
    if (!c || !pl || !b || pl->ctx != c || !(stages & SVTGPU_STAGE_ALL)) return SVTGPU_ERR_ARG;
    const bool fwd = stages & STAGE_FWD, qnt = stages & STAGE_QUANT, inv = stages & STAGE_INV;
    const bool diff = stages & STAGE_RESIDUAL;
    if (diff && (!fwd || !b->src || !b->pred)) return SVTGPU_ERR_ARG;
    if (fwd && !diff && !b->residual) return SVTGPU_ERR_ARG;
    if (b->dist && (!fwd || !qnt || pl->uses_qm)) return SVTGPU_ERR_ARG;
    if (b->levels && (!qnt || !b->levels_off)) return SVTGPU_ERR_ARG;
    if (!fwd && !b->coeff_in) return SVTGPU_ERR_ARG;
    if (fwd && !qnt && !b->coeff) return SVTGPU_ERR_ARG;
    if (qnt && (!b->eob || !c->d_qtables)) return SVTGPU_ERR_ARG;
    if (qnt && !b->dist && !b->qcoeff) return SVTGPU_ERR_ARG;  // qcoeff optional only next to dist
    // dqcoeff is always optional: the inverse transform reads the dequantised row from registers
    if (qnt && pl->max_qrow >= uint32_t(c->n_qrows)) return SVTGPU_ERR_ARG;
    if (qnt && pl->uses_qm) {
        // a block flagged USE_QM must find its matrix: no silent flat quantisation, no read in
        // front of the table
        if (!c->d_qm_wt || !c->d_qm_iwt) {
            svtgpu_set_error("%s%s (line %d)", "the plan has SVTGPU_BLK_USE_QM blocks but no qmatrix was uploaded ",
                             "(svtgpu_upload_qmatrix)", __LINE__);
            return SVTGPU_ERR_ARG;
        }
        for (int pc = 0; pc < 2; ++pc)
            for (int s = 0; s < 19; ++s)
                if (((pl->qm_sizes[pc] >> s) & 1) && c->qm_off[pc][s] < 0) {
                    char what[64];
                    snprintf(what, sizeof(what), "tx_size %d, plane class %d", s, pc);
                    svtgpu_set_error("%s%s (line %d)", "SVTGPU_BLK_USE_QM block without an uploaded matrix: ", what, __LINE__);
                    return SVTGPU_ERR_ARG;
                }
    }
    if (inv && (!b->pred || !b->recon)) return SVTGPU_ERR_ARG;
    if (fwd && !qnt && inv) return SVTGPU_ERR_ARG;  // inverse needs dequantized input
    if (pl->n == 0) return SVTGPU_OK;
    CK(cudaSetDevice(c->device));
    cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : c->stream;

    KParams p;
    memset(&p, 0, sizeof(p));
    p.blocks = pl->d_blocks; p.n_blocks = pl->n; p.stages = stages; p.bd_uniform = pl->bd_uniform; p.any_clamp16 = pl->any_clamp16;
    p.residual = b->residual; p.pred = b->pred; p.recon = b->recon;
    p.coeff_in = b->coeff_in; p.coeff_out = b->coeff; p.qcoeff = b->qcoeff; p.dqcoeff = b->dqcoeff;
    p.eob = b->eob; p.energy = b->energy;
    p.src = b->src; p.dist = reinterpret_cast<unsigned long long*>(b->dist);
    p.levels = b->levels; p.levels_off = b->levels_off;
    p.qtables = c->d_qtables; p.iscan = c->d_iscan;
    p.qm_wt = pl->uses_qm ? c->d_qm_wt : nullptr; p.qm_iwt = pl->uses_qm ? c->d_qm_iwt : nullptr;
    memcpy(p.iscan_off, c->iscan_off, sizeof(p.iscan_off));
    memcpy(p.qm_off, c->qm_off, sizeof(p.qm_off));

    // One launch per transform size present.  Buckets are independent, so they are spread over
    // side streams (fork/join on events) and overlap on the device instead of draining one after
    // another; a single bucket goes straight onto the caller's stream.
    int nb = 0;
    for (int s = 0; s < 19; ++s) nb += pl->count[s] > 0;
    static const int n_side = [] {  // tuning knob SVTGPU_SIDE_STREAMS (0 = every bucket on the caller's stream)
        const char* e = getenv("SVTGPU_SIDE_STREAMS");
        const int v = e ? atoi(e) : 3;  // measured on B200 (1080p8, ms/step): 0 -> 0.352; smallest first 2 / 3 / 4 -> 0.290 / 0.287 / 0.292 (kernel v8); v11: 3 / 5 / 7 / 11 -> 0.248 / 0.256 / 0.279 / 0.270
        return v < 0 ? 0 : (v > SvtGpuContext::kSide ? SvtGpuContext::kSide : v);
    }();
    if (nb == 1 || n_side == 0) {
        for (int s = 0; s < 19; ++s)
            if (pl->count[s]) { CK(launch_bucket(s, p, pl->first[s], pl->count[s], st)); c->launches++, ++g_total_launches; }
        return SVTGPU_OK;
    }
    CK(cudaEventRecord(c->fork, st));
    int k = 0;
    // buckets sorted by work
    int order[19];
    for (int s = 0; s < 19; ++s) order[s] = s;
    for (int i = 0; i < 19; ++i)
        for (int j = i + 1; j < 19; ++j) {
            const long long wi = (long long)pl->count[order[i]] * kTxW[order[i]] * kTxH[order[i]];
            const long long wj = (long long)pl->count[order[j]] * kTxW[order[j]] * kTxH[order[j]];
            if (wj > wi) { int t = order[i]; order[i] = order[j]; order[j] = t; }
        }
    // SVTGPU_ORDER: 3 (default, see below; v12, three streams: 0.238 vs 0.248 ms/step for 1);
    // 1 = smallest first, round robin (the under-filled
    // small grids start at once and the big ones fill in behind them; measured 0.287 vs 0.306
    // ms/step for largest first on three streams); 0 = largest first; 2 = the four largest round
    // robin over the first n_side - 1 streams, everything else on the last stream (0.318)
    static const int order_mode = [] { const char* e = getenv("SVTGPU_ORDER"); return e ? atoi(e) : 3; }();
    // 3 = smallest first; the four largest queue behind each other on stream 0 (two big grids never
    // share the SMs -- each fills the GPU by itself, together they only compete for the instruction
    // cache), the rest round robin over the other streams and run in their shadow
    bool used[SvtGpuContext::kSide] = {};
    int rank_of = 0, n_present = 0;
    for (int s = 0; s < 19; ++s) n_present += pl->count[s] > 0;
    for (int ii = 0; ii < 19; ++ii) {
        const int i = (order_mode == 1 || order_mode == 3) ? 18 - ii : ii;
        const int s = order[i];
        if (!pl->count[s]) continue;
        int si = k++ % n_side;
        if (order_mode == 2 && n_side >= 2) si = rank_of < 4 ? rank_of % (n_side - 1) : n_side - 1;
        if (order_mode == 3 && n_side >= 2) si = rank_of >= n_present - 4 ? 0 : 1 + rank_of % (n_side - 1);
        ++rank_of;
        if (!used[si]) { CK(cudaStreamWaitEvent(c->side[si], c->fork, 0)); used[si] = true; }
        CK(launch_bucket(s, p, pl->first[s], pl->count[s], c->side[si]));
        c->launches++, ++g_total_launches;
    }
    for (int si = 0; si < SvtGpuContext::kSide; ++si)
        if (used[si]) {
            CK(cudaEventRecord(c->side_done[si], c->side[si]));
            CK(cudaStreamWaitEvent(st, c->side_done[si], 0));
        }
    return SVTGPU_OK;
}

extern "C" size_t svtgpu_levels_size(int tx_size) {
    if (tx_size < 0 || tx_size > 18) return 0;
    return size_t(coded_w(tx_size) + 4) * size_t(coded_h(tx_size) + 6) + 16;  // TX_PAD_HOR, TX_PAD_VER, TX_PAD_END
}

extern "C" uint32_t svtgpu_tx_search_mask(int tx_size, int is_inter, int reduced_tx_set_used,
                                          int tx_search_reduced_set) {
    return tx_search_mask(tx_size, is_inter, reduced_tx_set_used, tx_search_reduced_set);
}
extern "C" int svtgpu_tx_search_expand(const SvtGpuBlock* blocks, int n_blocks, int is_inter,
                                       int reduced_tx_set_used, int tx_search_reduced_set, int want_coeff,
                                       SvtGpuBlock* cand, int32_t* cand_block, int64_t* coeff_elems) {
    if (n_blocks < 0 || (n_blocks && (!blocks || !cand))) return SVTGPU_ERR_ARG;
    return tx_search_expand(blocks, n_blocks, is_inter, reduced_tx_set_used, tx_search_reduced_set,
                            want_coeff, cand, cand_block, coeff_elems);
}

extern "C" int svtgpu_sync(SvtGpuContext* c) {
    if (!c) return SVTGPU_ERR_ARG;
    CK(cudaSetDevice(c->device));
    CK(cudaStreamSynchronize(c->stream));
    return SVTGPU_OK;
}
extern "C" long long svtgpu_launch_count(const SvtGpuContext* c) { return c ? (long long)c->launches : 0; }

// ---- memory helpers ------------------------------------------------------------------------------
extern "C" int svtgpu_malloc(SvtGpuContext* c, void** p, size_t bytes) {
    if (!c || !p) return SVTGPU_ERR_ARG;
    CK(cudaSetDevice(c->device));
    if (cudaMalloc(p, bytes ? bytes : 1) != cudaSuccess) { cudaGetLastError(); return SVTGPU_ERR_NOMEM; }
    return SVTGPU_OK;
}
extern "C" int svtgpu_free(SvtGpuContext* c, void* p) {
    if (!c) return SVTGPU_ERR_ARG;
    CK(cudaSetDevice(c->device));
    CK(cudaFree(p));
    return SVTGPU_OK;
}
extern "C" int svtgpu_malloc_host(SvtGpuContext* c, void** p, size_t bytes) {
    if (!c || !p) return SVTGPU_ERR_ARG;
    CK(cudaSetDevice(c->device));
    if (cudaMallocHost(p, bytes ? bytes : 1) != cudaSuccess) { cudaGetLastError(); return SVTGPU_ERR_NOMEM; }
    return SVTGPU_OK;
}
extern "C" int svtgpu_free_host(SvtGpuContext* c, void* p) {
    if (!c) return SVTGPU_ERR_ARG;
    CK(cudaFreeHost(p));
    return SVTGPU_OK;
}
extern "C" int svtgpu_memcpy_h2d(SvtGpuContext* c, void* d, const void* h, size_t bytes) {
    if (!c) return SVTGPU_ERR_ARG;
    CK(cudaSetDevice(c->device));
    CK(cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return SVTGPU_OK;
}
extern "C" int svtgpu_memcpy_d2h(SvtGpuContext* c, void* h, const void* d, size_t bytes) {
    if (!c) return SVTGPU_ERR_ARG;
    CK(cudaSetDevice(c->device));
    CK(cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return SVTGPU_OK;
}
extern "C" int svtgpu_memset(SvtGpuContext* c, void* d, int v, size_t bytes) {
    if (!c) return SVTGPU_ERR_ARG;
    CK(cudaSetDevice(c->device));
    CK(cudaMemsetAsync(d, v, bytes, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return SVTGPU_OK;
}

extern "C" int svtgpu_get_iscan(int tx_size, int cls, int16_t* out) {
    if (tx_size < 0 || tx_size > 18 || cls < 0 || cls > 2 || !out) return SVTGPU_ERR_ARG;
    std::vector<int16_t> t;
    make_iscan(tx_size, cls, t);
    memcpy(out, t.data(), t.size() * sizeof(int16_t));
    return int(t.size());
}
