// Several GPUs behind one handle, for a C caller that is ONE process (the encoder): pictures are
// independent on this path, so submission i goes to device i mod n -- whole pictures
// per device, no data-path collective, nothing shared between devices but the caller's host
// buffers.  Each device keeps its own context (tables, lanes, streams); the calls below only pick
// the context and forward.  bench.py keeps the process-per-GPU mode (torchrun) the driver asks for;
// this is the same sharding for callers that cannot fork.
#include <new>

#include "svt_gpu_internal.h"

struct SvtGpuMulti {
    int n = 0;
    SvtGpuContext* ctx[SVTGPU_MAX_DEVICES] = {};
    long long submitted = 0;                  // pictures submitted so far: the round-robin cursor
    int local_ticket[SVTGPU_MAX_DEVICES] = {};  // last ticket issued by each context
};

extern "C" int svtgpu_multi_device_of(int n_devices, long long submission_index) {
    if (n_devices <= 0 || submission_index < 0) return -1;
    return int(submission_index % n_devices);
}

extern "C" int svtgpu_create_multi(SvtGpuMulti** out, const int* device_ids, int n) {
    if (!out || !device_ids || n <= 0 || n > SVTGPU_MAX_DEVICES) {
        svtgpu_set_error("%s%s (line %d)", "svtgpu_create_multi: need 1..SVTGPU_MAX_DEVICES device ids", "", __LINE__);
        return SVTGPU_ERR_ARG;
    }
    *out = nullptr;
    SvtGpuMulti* m = new (std::nothrow) SvtGpuMulti();
    if (!m) return SVTGPU_ERR_NOMEM;
    for (int i = 0; i < n; ++i) {
        const int rc = svtgpu_create(&m->ctx[i], device_ids[i]);
        if (rc != SVTGPU_OK) {
            for (int j = 0; j < i; ++j) svtgpu_destroy(m->ctx[j]);
            delete m;
            return rc;
        }
        m->n = i + 1;
    }
    *out = m;
    return SVTGPU_OK;
}

extern "C" void svtgpu_multi_destroy(SvtGpuMulti* m) {
    if (!m) return;
    for (int i = 0; i < m->n; ++i) svtgpu_destroy(m->ctx[i]);
    delete m;
}

extern "C" int svtgpu_multi_count(const SvtGpuMulti* m) { return m ? m->n : 0; }
extern "C" SvtGpuContext* svtgpu_multi_context(SvtGpuMulti* m, int i) {
    return (m && i >= 0 && i < m->n) ? m->ctx[i] : nullptr;
}

extern "C" int svtgpu_multi_upload_quant_tables(SvtGpuMulti* m, const int16_t* tables, int n_rows) {
    if (!m) return SVTGPU_ERR_ARG;
    for (int i = 0; i < m->n; ++i) {
        const int rc = svtgpu_upload_quant_tables(m->ctx[i], tables, n_rows);
        if (rc != SVTGPU_OK) return rc;
    }
    return SVTGPU_OK;
}

extern "C" int svtgpu_multi_upload_qmatrix(SvtGpuMulti* m, const uint8_t* wt, const uint8_t* iwt, int n_bytes,
                                           const int32_t* off) {
    if (!m) return SVTGPU_ERR_ARG;
    for (int i = 0; i < m->n; ++i) {
        const int rc = svtgpu_upload_qmatrix(m->ctx[i], wt, iwt, n_bytes, off);
        if (rc != SVTGPU_OK) return rc;
    }
    return SVTGPU_OK;
}

// ticket = the global submission index; the device it ran on is svtgpu_multi_device_of(n, ticket).
// frame->plan, when given, must belong to the context of THAT device (svtgpu_multi_context).
extern "C" int svtgpu_multi_submit_host(SvtGpuMulti* m, const SvtGpuHostFrame* frame, int stages, long long* ticket) {
    if (!m || !frame) return SVTGPU_ERR_ARG;
    const long long idx = m->submitted;
    const int dev = svtgpu_multi_device_of(m->n, idx);
    int t = -1;
    const int rc = svtgpu_submit_host(m->ctx[dev], frame, stages, &t);
    if (rc != SVTGPU_OK) return rc;
    m->local_ticket[dev] = t;
    m->submitted = idx + 1;
    if (ticket) *ticket = idx;
    return SVTGPU_OK;
}

// ticket < 0: everything submitted so far, on every device.  Otherwise every submission that the
// device of `ticket` has been given up to now (a device drains its lanes in order).
extern "C" int svtgpu_multi_wait_host(SvtGpuMulti* m, long long ticket) {
    if (!m) return SVTGPU_ERR_ARG;
    if (ticket >= m->submitted) return SVTGPU_ERR_ARG;
    for (int i = 0; i < m->n; ++i) {
        if (ticket >= 0 && i != svtgpu_multi_device_of(m->n, ticket)) continue;
        const int rc = svtgpu_wait_host(m->ctx[i], -1);
        if (rc != SVTGPU_OK) return rc;
    }
    return SVTGPU_OK;
}
