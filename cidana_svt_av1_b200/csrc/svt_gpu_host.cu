// Host-buffer pipeline: svtgpu_submit_host / svtgpu_wait_host (include/svt_gpu.h).
// A ring of kHostSlots lanes, each with its own stream and device buffers.  A submission does
//   sort descriptors (host) -> H2D descriptors, residual, prediction -> kernels -> D2H outputs
// on its lane's stream; lanes overlap each other, so in steady state the two copy engines and
// the SMs are all busy.  Nothing here computes on the host: a CUDA failure is returned, never
// worked around.
#include <stdio.h>
#include <string.h>
#include <new>
#include <vector>
#include "host_plan.h"
#include "svt_gpu_internal.h"

using namespace svt;

namespace {

long long g_host_allocs = 0;  // (re)allocations made by the pipeline: 0 in steady state

template <class T>
int grow(T*& p, size_t& cap, size_t need, cudaStream_t st, size_t headroom_div = 4) {
    if (need <= cap) return SVTGPU_OK;
    ++g_host_allocs;
    CK(cudaStreamSynchronize(st));
    if (p) CK(cudaFree(p));
    p = nullptr;
    cap = 0;
    const size_t want = need + need / headroom_div;
    if (cudaMalloc(reinterpret_cast<void**>(&p), want * sizeof(T)) != cudaSuccess) {
        cudaGetLastError();
        svtgpu_set_error("%s%s (line %d)", "out of device memory in the host pipeline", "", __LINE__);
        return SVTGPU_ERR_NOMEM;
    }
    cap = want;
    return SVTGPU_OK;
}

int slot_init(HostSlot& s) {
    if (s.st) return SVTGPU_OK;
    CK(cudaStreamCreateWithFlags(&s.st, cudaStreamNonBlocking));
    CK(cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming));
    return SVTGPU_OK;
}

int slot_wait(HostSlot& s) {
    if (s.busy) {
        CK(cudaEventSynchronize(s.done));
        s.busy = false;
    }
    return SVTGPU_OK;
}

}  // namespace

#define TRY(x)                        \
    do {                              \
        const int rc_ = (x);          \
        if (rc_ != SVTGPU_OK) return rc_; \
    } while (0)

// fault injection for the tests: the k-th grow() from now on fails as if the device were out of
// memory (0 = off).  Not declared in the public header.
static int g_fail_grow_in = 0;
extern "C" void svtgpu_debug_fail_host_alloc(int kth) { g_fail_grow_in = kth; }

namespace {
template <class T>
int grow_checked(T*& p, size_t& cap, size_t need, cudaStream_t st, size_t headroom_div = 4) {
    if (need > cap && g_fail_grow_in > 0 && --g_fail_grow_in == 0) {
        // behave exactly like a failed cudaMalloc inside grow(): old buffer gone, capacity 0
        CK(cudaStreamSynchronize(st));
        if (p) CK(cudaFree(p));
        p = nullptr;
        cap = 0;
        svtgpu_set_error("%s%s (line %d)", "out of device memory in the host pipeline", " (injected)", __LINE__);
        return SVTGPU_ERR_NOMEM;
    }
    return grow(p, cap, need, st, headroom_div);
}
}  // namespace

extern "C" int svtgpu_submit_host(SvtGpuContext* c, const SvtGpuHostFrame* f, int stages, int* ticket) {
    if (!c || !f || f->n_blocks < 0 || (f->n_blocks && !f->blocks && !f->plan)) return SVTGPU_ERR_ARG;
    if (!(stages & SVTGPU_STAGE_ALL) || (stages & ~(SVTGPU_STAGE_ALL | SVTGPU_STAGE_RESIDUAL))) return SVTGPU_ERR_ARG;
    if (f->bit_depth != 8 && f->bit_depth != 10) return SVTGPU_ERR_ARG;
    const bool fwd = stages & STAGE_FWD, qnt = stages & STAGE_QUANT, inv = stages & STAGE_INV;
    if (!fwd || !qnt) return SVTGPU_ERR_ARG;  // the pipeline carries the fused path (optionally without INV)
    const bool diff = stages & SVTGPU_STAGE_RESIDUAL;
    if (!f->eob || (diff ? (!f->src || !f->pred) : !f->residual)) return SVTGPU_ERR_ARG;
    if (!f->dist && !f->qcoeff) return SVTGPU_ERR_ARG;  // dqcoeff alone is never useful; either may be absent next to dist
    if (f->dqcoeff && !f->qcoeff) return SVTGPU_ERR_ARG;
    if (inv && (!f->pred || !f->recon)) return SVTGPU_ERR_ARG;
    if (f->plan && (f->plan->ctx != c || f->plan->n != f->n_blocks)) {
        svtgpu_set_error("%s%s (line %d)", "SvtGpuHostFrame.plan belongs to another context or block count", "", __LINE__);
        return SVTGPU_ERR_ARG;
    }
    CK(cudaSetDevice(c->device));
    HostSlot& s = c->host[c->host_next];
    c->host_next = (c->host_next + 1) % SvtGpuContext::kHostSlots;
    TRY(slot_init(s));
    TRY(slot_wait(s));
    const int n = f->n_blocks;
    const int my_ticket = c->host_tickets++;
    s.ticket = my_ticket;
    if (ticket) *ticket = my_ticket;
    if (n == 0) return SVTGPU_OK;

    SvtGpuPlan local;
    const SvtGpuPlan* plan = f->plan;
    if (plan) {
        // a plan made once for this picture geometry (svtgpu_plan_create): descriptors are already
        // validated, sorted and resident; nothing is re-bucketed or uploaded per submission
        if (plan->bd_uniform != f->bit_depth) return SVTGPU_ERR_ARG;
        if ((!diff && plan->need_res > f->residual_elems) || plan->need_pix > f->pixel_elems ||
            plan->need_coeff > f->coeff_elems) {
            svtgpu_set_error("%s%s (line %d)", "a block of the plan reaches beyond residual_elems / pixel_elems / coeff_elems", "", __LINE__);
            return SVTGPU_ERR_ARG;
        }
    } else {
        // descriptors: validate + bucket on the host, straight into page-locked staging
        local.ctx = c;
        local.n = n;
        if (size_t(n) > s.cap_hblocks) {
            CK(cudaStreamSynchronize(s.st));
            ++g_host_allocs;
            if (s.h_blocks) CK(cudaFreeHost(s.h_blocks));
            s.h_blocks = nullptr;
            s.cap_hblocks = 0;
            CK(cudaMallocHost(reinterpret_cast<void**>(&s.h_blocks), size_t(n) * 2 * sizeof(DevBlock)));
            s.cap_hblocks = size_t(n) * 2;
        }
        // per-block buffers are small (32 + 26 bytes per block) and the block count of a picture moves
        // with its partitioning: 2x headroom so that a lane settles after its first picture
        TRY(grow_checked(s.d_blocks, s.cap_blocks, size_t(n), s.st, 1));
        static thread_local BucketScratch scratch;
        const int bad = bucket_blocks(f->blocks, n, s.h_blocks, local.first, local.count, scratch);
        if (bad) {
            char idx[32];
            snprintf(idx, sizeof(idx), "%d", bad - 1);
            svtgpu_set_error("%s%s (line %d)", "invalid SvtGpuBlock at index ", idx, __LINE__);
            return SVTGPU_ERR_ARG;
        }
        for (int i = 0; i < n; ++i) {
            const SvtGpuBlock& b = f->blocks[i];
            if (b.bit_depth != f->bit_depth) return SVTGPU_ERR_ARG;
            // one bad descriptor must not make the kernel read or write outside this lane's buffers
            const size_t w = size_t(kTxW[b.tx_size]), h = size_t(kTxH[b.tx_size]);
            const bool bad_res = !diff && size_t(b.res_off) + (h - 1) * b.res_stride + w > f->residual_elems;
            const bool bad_pix = (inv || diff) && size_t(b.rec_off) + (h - 1) * b.rec_stride + w > f->pixel_elems;
            const bool bad_cf = f->qcoeff && size_t(b.coeff_off) + size_t(max_eob(b.tx_size)) > f->coeff_elems;
            if (bad_res || bad_pix || bad_cf) {
                char idx[32];
                snprintf(idx, sizeof(idx), "%d", i);
                svtgpu_set_error("%s%s (line %d)", "a block reaches beyond residual_elems / pixel_elems / coeff_elems: index ", idx, __LINE__);
                return SVTGPU_ERR_ARG;
            }
            local.max_qrow = b.qrow > local.max_qrow ? b.qrow : local.max_qrow;
            local.uses_qm |= (b.flags & SVTGPU_BLK_USE_QM) != 0;
            if (b.flags & SVTGPU_BLK_USE_QM) local.qm_sizes[b.plane > 0] |= 1u << b.tx_size;
            local.any_clamp16 |= (b.flags & SVTGPU_BLK_CLAMP16) != 0;
        }
        local.d_blocks = s.d_blocks;
        local.bd_uniform = f->bit_depth;  // checked per block above
        plan = &local;
    }

    const size_t px = f->bit_depth == 8 ? 1 : 2;
    if (!diff) TRY(grow_checked(s.d_res, s.cap_res, f->residual_elems, s.st));
    if (diff) {
        char* sp = static_cast<char*>(s.d_src);
        const int rc = grow_checked(sp, s.cap_src, f->pixel_elems * px, s.st);
        s.d_src = sp;
        TRY(rc);
    }
    if (inv || diff) {
        char* pix = static_cast<char*>(s.d_pix);
        const int rc = grow_checked(pix, s.cap_pix, f->pixel_elems * px, s.st);
        s.d_pix = pix;
        TRY(rc);
    }
    // every buffer carries its own capacity: a failed allocation leaves that one buffer empty
    // (pointer null, capacity 0) and the next submission simply tries again
    if (f->qcoeff) TRY(grow_checked(s.d_q, s.cap_q, f->coeff_elems, s.st));
    if (f->dqcoeff) TRY(grow_checked(s.d_dq, s.cap_dq, f->coeff_elems, s.st));
    TRY(grow_checked(s.d_eob, s.cap_eob, size_t(n), s.st, 1));
    TRY(grow_checked(s.d_energy, s.cap_energy, size_t(n), s.st, 1));
    if (f->dist) TRY(grow_checked(s.d_dist, s.cap_dist, size_t(n) * 2, s.st, 1));

    if (plan == &local)
        CK(cudaMemcpyAsync(s.d_blocks, s.h_blocks, size_t(n) * sizeof(DevBlock), cudaMemcpyHostToDevice, s.st));
    if (diff) CK(cudaMemcpyAsync(s.d_src, f->src, f->pixel_elems * px, cudaMemcpyHostToDevice, s.st));
    else CK(cudaMemcpyAsync(s.d_res, f->residual, f->residual_elems * sizeof(int16_t), cudaMemcpyHostToDevice, s.st));
    if (inv || diff) CK(cudaMemcpyAsync(s.d_pix, f->pred, f->pixel_elems * px, cudaMemcpyHostToDevice, s.st));

    SvtGpuBuffers b;
    memset(&b, 0, sizeof(b));
    b.residual = s.d_res;
    b.pred = s.d_pix;
    b.recon = s.d_pix;  // in place, like the reference
    b.qcoeff = f->qcoeff ? s.d_q : nullptr;
    b.dqcoeff = f->dqcoeff ? s.d_dq : nullptr;
    b.eob = s.d_eob;
    b.energy = s.d_energy;
    b.src = s.d_src;
    b.dist = f->dist ? s.d_dist : nullptr;
    TRY(svtgpu_run_internal(c, plan, &b, stages, s.st));

    if (f->qcoeff)
        CK(cudaMemcpyAsync(f->qcoeff, s.d_q, f->coeff_elems * sizeof(int32_t), cudaMemcpyDeviceToHost, s.st));
    if (f->dqcoeff)
        CK(cudaMemcpyAsync(f->dqcoeff, s.d_dq, f->coeff_elems * sizeof(int32_t), cudaMemcpyDeviceToHost, s.st));
    if (f->dist)
        CK(cudaMemcpyAsync(f->dist, s.d_dist, size_t(n) * 2 * sizeof(uint64_t), cudaMemcpyDeviceToHost, s.st));
    CK(cudaMemcpyAsync(f->eob, s.d_eob, size_t(n) * sizeof(uint16_t), cudaMemcpyDeviceToHost, s.st));
    if (f->energy)
        CK(cudaMemcpyAsync(f->energy, s.d_energy, size_t(n) * sizeof(uint64_t), cudaMemcpyDeviceToHost, s.st));
    if (inv) CK(cudaMemcpyAsync(f->recon, s.d_pix, f->pixel_elems * px, cudaMemcpyDeviceToHost, s.st));
    CK(cudaEventRecord(s.done, s.st));
    s.busy = true;
    return SVTGPU_OK;
}

extern "C" long long svtgpu_debug_host_allocs(void) { return g_host_allocs; }

extern "C" int svtgpu_wait_host(SvtGpuContext* c, int ticket) {
    if (!c) return SVTGPU_ERR_ARG;
    CK(cudaSetDevice(c->device));
    for (int i = 0; i < SvtGpuContext::kHostSlots; ++i) {
        HostSlot& s = c->host[i];
        if (ticket < 0 || s.ticket == ticket) TRY(slot_wait(s));
    }
    return SVTGPU_OK;
}

// called by svtgpu_destroy
void svtgpu_host_release(SvtGpuContext* c) {
    for (int i = 0; i < SvtGpuContext::kHostSlots; ++i) {
        HostSlot& s = c->host[i];
        if (!s.st) continue;
        cudaStreamSynchronize(s.st);
        cudaFree(s.d_blocks);
        cudaFreeHost(s.h_blocks);
        cudaFree(s.d_res);
        cudaFree(s.d_pix);
        cudaFree(s.d_q);
        cudaFree(s.d_dq);
        cudaFree(s.d_eob);
        cudaFree(s.d_energy);
        cudaFree(s.d_dist);
        cudaFree(s.d_src);
        cudaEventDestroy(s.done);
        cudaStreamDestroy(s.st);
        s = HostSlot();
    }
}
