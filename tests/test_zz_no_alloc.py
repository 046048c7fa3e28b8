"""svtgpu_quantize_b_noalloc (include/svt_gpu.h): the allocation-free variant of the stand-alone
quantizer, whose only working memory is a fixed file-scope __device__ array.  No counterpart in
the reference, so the checker is the existing
level-A slot svtgpu_aom_highbd_quantize_b[_32x32|_64x64] on the same inputs: bit-exact, like every
integer comparison in this suite.  Named test_zz_* so that pytest collects it last."""
import ctypes

import numpy as np
import pytest

import cidana_svt_av1_b200 as pkg
from golden_util import load_quant_tables

MAX_COEFFS = 32768  # SVTGPU_NOALLOC_MAX_COEFFS


def P(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _fn():
    L = pkg.load_library()
    f = L.svtgpu_quantize_b_noalloc
    f.restype = ctypes.c_int
    f.argtypes = [ctypes.c_void_p, ctypes.c_ssize_t, ctypes.c_int] + [ctypes.c_void_p] * 10
    return L, f


def test_rejects_sizes_beyond_the_maximum_without_touching_the_device():
    """The size check comes before any CUDA call, so it is testable on the CPU tier too."""
    L, f = _fn()
    dummy = np.zeros(8, np.int16)
    one = ctypes.c_void_p(256)  # never dereferenced: the call must fail on the size alone
    rc = f(one, MAX_COEFFS + 1, 0, P(dummy), P(dummy), P(dummy), P(dummy), P(dummy), one, one, one, one, None)
    assert rc == -1  # SVTGPU_ERR_ARG
    msg = L.svtgpu_last_error().decode()
    assert "SVTGPU_NOALLOC_MAX_COEFFS" in msg and str(MAX_COEFFS) in msg, msg
    assert f(one, 0, 0, P(dummy), P(dummy), P(dummy), P(dummy), P(dummy), one, one, one, one, None) == -1
    assert f(one, 16, 3, P(dummy), P(dummy), P(dummy), P(dummy), P(dummy), one, one, one, one, None) == -1


def _case(rng, n, amp, zero_tail):
    cf = rng.integers(-amp, amp + 1, n).astype(np.int32)
    if zero_tail:
        cf[n // 8:] = 0
    iscan = rng.permutation(n).astype(np.int16)
    return cf, iscan


@pytest.mark.gpu
def test_noalloc_quantizer_equals_the_slot_on_two_calls_in_a_row():
    import torch
    L, f = _fn()
    dev = torch.device("cuda", 0)
    slots = [L.svtgpu_aom_highbd_quantize_b, L.svtgpu_aom_highbd_quantize_b_32x32,
             L.svtgpu_aom_highbd_quantize_b_64x64]
    tabs = load_quant_tables()[(10, 0)]
    rng = np.random.default_rng(77)
    # back to back on the shared scratch array: a 4-CTA call with a large eob first, then smaller
    # ones whose eob must not inherit anything from it, then the largest size the slot accepts again
    cases = [(4096, 0, 60, 1 << 15, False), (256, 0, 200, 40, True), (1024, 1, 120, 3000, False),
             (1024, 2, 255, 9, True), (4096, 2, 20, 1 << 17, False), (16, 0, 0, 5000, False)]
    st = torch.cuda.Stream(dev)
    for rep in range(2):
        for n, ls, qi, amp, zero_tail in cases:
            cf, iscan = _case(rng, n, amp, zero_tail)
            tb = tabs[qi]
            q_s = np.ones(n, np.int32); dq_s = np.ones(n, np.int32); e_s = ctypes.c_uint16(7)
            slots[ls](P(cf), ctypes.c_ssize_t(n), 0, P(tb[0]), P(tb[1]), P(tb[2]), P(tb[3]), P(q_s), P(dq_s),
                      P(tb[4]), ctypes.byref(e_s), P(iscan), P(iscan))
            d_cf = torch.from_numpy(cf).to(dev)
            d_is = torch.from_numpy(iscan).to(dev)
            d_q = torch.full((n,), 1, dtype=torch.int32, device=dev)
            d_dq = torch.full((n,), 1, dtype=torch.int32, device=dev)
            d_e = torch.full((1,), 7, dtype=torch.int16, device=dev)
            torch.cuda.synchronize(dev)
            with torch.cuda.device(dev):
                rc = f(d_cf.data_ptr(), n, ls, P(tb[0]), P(tb[1]), P(tb[2]), P(tb[3]), P(tb[4]),
                       d_q.data_ptr(), d_dq.data_ptr(), d_e.data_ptr(), d_is.data_ptr(), st.cuda_stream)
            assert rc == 0, L.svtgpu_last_error().decode()
            st.synchronize()
            what = (rep, n, ls, qi)
            assert np.array_equal(d_q.cpu().numpy(), q_s), what
            assert np.array_equal(d_dq.cpu().numpy(), dq_s), what
            assert int(d_e.cpu().numpy().view(np.uint16)[0]) == e_s.value, what


@pytest.mark.gpu
def test_noalloc_quantizer_up_to_the_maximum_size():
    """n_coeffs = SVTGPU_NOALLOC_MAX_COEFFS (32 CTAs, every partial slot of the scratch array in
    use) against the slot run on the same data in chunks it accepts: levels are position-wise, and
    the eob is the maximum of iscan + 1 over non-zero levels."""
    import torch
    L, f = _fn()
    dev = torch.device("cuda", 0)
    tabs = load_quant_tables()[(10, 0)]
    rng = np.random.default_rng(5)
    n = MAX_COEFFS
    tb = tabs[90]
    for rep in range(2):
        cf = rng.integers(-3000, 3001, n).astype(np.int32)
        cf[rng.random(n) < 0.7] = 0
        iscan = rng.permutation(n).astype(np.int16)
        d_cf, d_is = torch.from_numpy(cf).to(dev), torch.from_numpy(iscan).to(dev)
        d_q = torch.zeros(n, dtype=torch.int32, device=dev)
        d_dq = torch.zeros(n, dtype=torch.int32, device=dev)
        d_e = torch.zeros(1, dtype=torch.int16, device=dev)
        torch.cuda.synchronize(dev)
        with torch.cuda.device(dev):
            rc = f(d_cf.data_ptr(), n, 0, P(tb[0]), P(tb[1]), P(tb[2]), P(tb[3]), P(tb[4]), d_q.data_ptr(),
                   d_dq.data_ptr(), d_e.data_ptr(), d_is.data_ptr(), None)
        assert rc == 0, L.svtgpu_last_error().decode()
        torch.cuda.synchronize(dev)
        q_g, dq_g = d_q.cpu().numpy(), d_dq.cpu().numpy()
        # the slot on 4096-coefficient chunks; chunk 0 carries the DC position, the others are
        # compared from their second coefficient on and their first against an AC-only call
        want_q = np.zeros(n, np.int32); want_dq = np.zeros(n, np.int32)
        ident = np.arange(4096, dtype=np.int16)
        for c0 in range(0, n, 4096):
            part = cf[c0:c0 + 4096].copy()
            if c0:  # position 0 of the chunk is an AC coefficient of the big call: shift by one
                part = np.concatenate([[0], part[:-1]]).astype(np.int32)
            q_s = np.zeros(4096, np.int32); dq_s = np.zeros(4096, np.int32); e_s = ctypes.c_uint16(0)
            L.svtgpu_aom_highbd_quantize_b(P(part), ctypes.c_ssize_t(4096), 0, P(tb[0]), P(tb[1]), P(tb[2]),
                                           P(tb[3]), P(q_s), P(dq_s), P(tb[4]), ctypes.byref(e_s), P(ident), P(ident))
            if c0:
                want_q[c0:c0 + 4095], want_dq[c0:c0 + 4095] = q_s[1:], dq_s[1:]
                last = np.array([0, cf[c0 + 4095]], np.int32)
                q1 = np.zeros(2, np.int32); dq1 = np.zeros(2, np.int32)
                L.svtgpu_aom_highbd_quantize_b(P(last), ctypes.c_ssize_t(2), 0, P(tb[0]), P(tb[1]), P(tb[2]),
                                               P(tb[3]), P(q1), P(dq1), P(tb[4]), ctypes.byref(e_s), P(ident), P(ident))
                want_q[c0 + 4095], want_dq[c0 + 4095] = q1[1], dq1[1]
            else:
                want_q[:4096], want_dq[:4096] = q_s, dq_s
        assert np.array_equal(q_g, want_q) and np.array_equal(dq_g, want_dq), rep
        nz = np.nonzero(want_q)[0]
        want_eob = int(iscan[nz].astype(np.int64).max()) + 1 if len(nz) else 0
        assert int(d_e.cpu().numpy().view(np.uint16)[0]) == want_eob, rep
