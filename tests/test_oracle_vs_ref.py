"""Pins the oracle (oracle/svt_oracle.c) to the reference compiled from source
(oracle/_ref/libsvtref.so): 1-D kernels, trig tables, per-size configuration, forward 2-D +
64-point packing/energy, quantizer (C 10-bit, C 8-bit clamp, AVX2), inverse + reconstruction.
Mirrors test/FwdTxfm2dAsmTest.cc, test/InvTxfm2dAsmTest.cc, test/QuantAsmTest.cc of the
reference.  Skipped where the reference build is absent (it needs /root/reference)."""
import ctypes

import numpy as np
import pytest

from oracle_lib import (P, TX_H, TX_W, aligned, aligned_copy, max_eob, oracle, ref, scan_class,
                        tx_scale, valid_pairs)
from golden_util import load_quant_tables, load_scans

pytestmark = pytest.mark.skipif(ref() is None, reason="oracle/_ref not built")


def test_trig_tables():
    R, O = ref(), oracle()
    c = np.zeros((7, 64), np.int32)
    s = np.zeros((7, 5), np.int32)
    R.ref_get_trig(P(c), P(s))
    for b in range(10, 17):
        assert np.array_equal(np.ctypeslib.as_array(O.svto_cospi(b), (64,)), c[b - 10])
        assert np.array_equal(np.ctypeslib.as_array(O.svto_sinpi(b), (5,)), s[b - 10])


def test_cfg_tables():
    R, O = ref(), oracle()
    for s, t in valid_pairs():
        a = np.zeros(16, np.int32)
        b = np.zeros(16, np.int32)
        R.ref_get_cfg(s, t, P(a))
        O.svto_get_cfg(s, t, P(b))
        assert np.array_equal(a[:11], b[:11]) and np.array_equal(a[13:15], b[13:15]), (s, t, a, b)


@pytest.mark.parametrize("kind,n", [(0, 4), (0, 8), (0, 16), (0, 32), (0, 64), (1, 4), (1, 8),
                                    (1, 16), (2, 4), (2, 8), (2, 16), (2, 32), (2, 64)])
def test_txfm1d(kind, n):
    R, O = ref(), oracle()
    rng = np.random.default_rng(100 * kind + n)
    for inv in (0, 1):
        for bit in (10, 11, 12, 13):
            for rb in ((0,) if not inv else (16, 18)):
                for amp in (1, 100, 3000, 30000, 120000):
                    for k in range(40):
                        x = rng.integers(-amp, amp + 1, n).astype(np.int32)
                        if k == 0:
                            x[:] = 0
                        a = np.zeros(64, np.int32)
                        b = np.zeros(64, np.int32)
                        R.ref_txfm1d(kind, n, inv, P(x), P(a), bit, rb if inv else 20)
                        O.svto_txfm1d(kind, n, inv, P(x), P(b), bit, rb)
                        assert np.array_equal(a, b), (kind, n, inv, bit, rb, amp)


def _rand_block(rng, w, h, bd, stride, k):
    m = (1 << bd) - 1
    res = rng.integers(-m, m + 1, (h, stride)).astype(np.int16)
    if k == 0:
        res[:] = m
    elif k == 1:
        res[:] = -m
    elif k == 2:
        res[:] = 0
    elif k == 3:  # checkerboard extremes
        res[:] = np.where((np.add.outer(np.arange(h), np.arange(stride)) & 1) == 0, m, -m)
    return res


@pytest.mark.parametrize("bd", [8, 10])
@pytest.mark.parametrize("asm", [0, 1])
def test_estimate_transform_all_pairs(bd, asm):
    """FwdTxfm2dAsmTest: every valid (size,type), C table and AVX2 table."""
    O = oracle()
    R = ref(asm)
    rng = np.random.default_rng(13596 + bd)
    for s, t in valid_pairs():
        w, h = TX_W[s], TX_H[s]
        for k in range(12):
            stride = 64 if k & 1 else w
            res = aligned_copy(_rand_block(rng, w, h, bd, stride, k))
            a = aligned(w * h, np.int32, 7)
            b = np.full(w * h, 9, np.int32)
            ea = ctypes.c_uint64(0)
            eb = ctypes.c_uint64(0)
            R.ref_estimate_transform(P(res), stride, P(a), s, t, bd, ctypes.byref(ea))
            O.svto_estimate_transform(P(res), stride, P(b), s, t, bd, ctypes.byref(eb))
            n = max_eob(s)
            assert np.array_equal(a[:n], b[:n]), (s, t, k)
            assert ea.value == eb.value, (s, t, k)


@pytest.mark.parametrize("bd", [8, 10])
def test_quantize_variants(bd):
    """QuantAsmTest: random + min/max coefficients at every qindex; C (Q2), C 8-bit clamp (Q3)
    and the AVX2 kernel the encoder binds for both depths (Q4)."""
    O, R = oracle(), ref(1)
    tabs = load_quant_tables()[(bd, 0)]
    scans = load_scans()
    rng = np.random.default_rng(bd)
    lim = 1 << (7 + bd)
    for s in (0, 1, 2, 3, 4, 5, 8, 9, 12, 13, 16, 17):
        n = max_eob(s)
        ls = tx_scale(s)
        for t in (0, 10, 11):
            scan, _ = scans[(s, scan_class(t))]
            for qi in list(range(0, 256, 5)) + [255]:
                tb = tabs[qi]
                for k in range(3):
                    cf = aligned_copy(rng.integers(-lim, lim, n).astype(np.int32))
                    if k == 1:
                        cf[:] = cf >> 6  # mostly-below-zbin case
                    if k == 2:
                        cf[:] = 0
                        cf[rng.integers(0, n, 4)] = [lim - 1, -lim, 1, -1]
                    outs = []
                    for lib, variant in ((R, 1), (R, 0), (O, None)):
                        q = aligned(n, np.int32, 3)
                        dq = aligned(n, np.int32, 5)
                        e = ctypes.c_uint16(0)
                        if lib is R:
                            R.ref_quantize(P(cf), n, P(tb[0]), P(tb[1]), P(tb[2]), P(tb[3]),
                                           P(tb[4]), P(q), P(dq), ctypes.byref(e), s, t, 10,
                                           variant, None, None)
                        else:
                            O.svto_quantize(P(cf), n, P(tb[0]), P(tb[1]), P(tb[2]), P(tb[3]),
                                            P(tb[4]), P(q), P(dq), ctypes.byref(e), P(scan), ls,
                                            None, None, 0)
                        outs.append((q, dq, e.value))
                    for a in outs[:2]:
                        assert np.array_equal(a[0], outs[2][0]), (s, t, qi, k)
                        assert np.array_equal(a[1], outs[2][1]), (s, t, qi, k)
                        assert a[2] == outs[2][2], (s, t, qi, k)
                    if bd == 8:  # the C 8-bit helper with its INT16 clamp
                        q = aligned(n, np.int32); dq = aligned(n, np.int32); e = ctypes.c_uint16()
                        q2 = np.zeros(n, np.int32); dq2 = np.zeros(n, np.int32); e2 = ctypes.c_uint16()
                        R.ref_quantize(P(cf), n, P(tb[0]), P(tb[1]), P(tb[2]), P(tb[3]), P(tb[4]),
                                       P(q), P(dq), ctypes.byref(e), s, t, 8, 2, None, None)
                        O.svto_quantize(P(cf), n, P(tb[0]), P(tb[1]), P(tb[2]), P(tb[3]), P(tb[4]),
                                        P(q2), P(dq2), ctypes.byref(e2), P(scan), ls, None, None, 1)
                        assert np.array_equal(q, q2) and np.array_equal(dq, dq2) and e.value == e2.value


def test_quantize_qmatrix():
    """Q2 with qmatrix: pinned only by highbd_quantize_b_helper_c itself."""
    O, R = oracle(), ref(1)
    tabs = load_quant_tables()[(10, 0)]
    scans = load_scans()
    rng = np.random.default_rng(5)
    for s in (0, 1, 2, 3, 4, 7, 10, 11, 14, 18):
        n = max_eob(s)
        for level in (0, 5, 7, 14):
            wt = np.zeros(1024, np.uint8)
            iwt = np.zeros(1024, np.uint8)
            assert R.ref_get_qmatrix(level, 0, s, P(wt), P(iwt)) == n
            scan, _ = scans[(s, 0)]
            for qi in (0, 40, 100, 180, 255):
                tb = tabs[qi]
                cf = aligned_copy(rng.integers(-(1 << 15), 1 << 15, n).astype(np.int32))
                q = aligned(n, np.int32); dq = aligned(n, np.int32); e = ctypes.c_uint16()
                q2 = np.zeros(n, np.int32); dq2 = np.zeros(n, np.int32); e2 = ctypes.c_uint16()
                R.ref_quantize(P(cf), n, P(tb[0]), P(tb[1]), P(tb[2]), P(tb[3]), P(tb[4]), P(q),
                               P(dq), ctypes.byref(e), s, 0, 10, 1, P(wt), P(iwt))
                O.svto_quantize(P(cf), n, P(tb[0]), P(tb[1]), P(tb[2]), P(tb[3]), P(tb[4]), P(q2),
                                P(dq2), ctypes.byref(e2), P(scan), tx_scale(s), P(wt), P(iwt), 0)
                assert np.array_equal(q, q2) and np.array_equal(dq, dq2) and e.value == e2.value


@pytest.mark.parametrize("bd", [8, 10])
@pytest.mark.parametrize("asm", [0, 1])
def test_inverse_recon_all_pairs(bd, asm):
    """InvTxfm2dAsmTest: inputs are forward transforms of random residuals (optionally
    quantized/dequantized), random prediction; C table and the AVX2/SSE4/SSSE3 table."""
    O = oracle()
    R = ref(asm)
    tabs = load_quant_tables()[(bd, 0)]
    scans = load_scans()
    rng = np.random.default_rng(77 + bd)
    for s, t in valid_pairs():
        w, h = TX_W[s], TX_H[s]
        n = max_eob(s)
        scan, _ = scans[(s, scan_class(t))]
        for k in range(8):
            res = _rand_block(rng, w, h, bd, w, k + 4)
            if k >= 4:
                res = (res // 8).astype(np.int16)
            cf = np.zeros(w * h, np.int32)
            en = ctypes.c_uint64()
            O.svto_estimate_transform(P(res), w, P(cf), s, t, bd, ctypes.byref(en))
            dq = cf[:n].copy()
            eob = n
            if k & 1:  # through the quantizer
                tb = tabs[[20, 90, 160, 230][k // 2]]
                q = np.zeros(n, np.int32); dq = np.zeros(n, np.int32); e = ctypes.c_uint16()
                O.svto_quantize(P(cf), n, P(tb[0]), P(tb[1]), P(tb[2]), P(tb[3]), P(tb[4]), P(q),
                                P(dq), ctypes.byref(e), P(scan), tx_scale(s), None, None, 0)
                eob = e.value
                if eob == 0:
                    continue
            stride = 80
            if bd == 8:
                pa = aligned_copy(rng.integers(0, 256, (h, stride)).astype(np.uint8))
                pb = pa.copy()
                R.ref_inv_transform_recon(P(aligned_copy(dq)), P(pa), stride, s, t, 8, eob)
                O.svto_inv_txfm2d_add8(P(dq), P(pb), stride, s, t)
            else:
                pa = aligned_copy(rng.integers(0, 1024, (h, stride)).astype(np.uint16))
                pb = pa.copy()
                R.ref_inv_transform_recon(P(aligned_copy(dq)), P(pa), stride, s, t, 10, eob)
                O.svto_inv_txfm2d_add(P(dq), P(pb), stride, s, t, 10)
            assert np.array_equal(pa, pb), (s, t, k, asm)


# ---- residual = source - prediction, transform-domain distortion ----------------------
@pytest.mark.parametrize("bd", [8, 10])
@pytest.mark.parametrize("asm", [0, 1])
def test_residual_kernel(bd, asm):
    """svto_residual == ResidualKernel (C / AVX2) and residual_kernel16bit on every block shape."""
    R, O = ref(asm), oracle()
    rng = np.random.default_rng(31 + bd)
    pix = np.uint8 if bd == 8 else np.uint16
    for s in range(19):
        w, h = TX_W[s], TX_H[s]
        stride = 96
        src = aligned_copy(rng.integers(0, 1 << bd, (h, stride)).astype(pix))
        prd = aligned_copy(rng.integers(0, 1 << bd, (h, stride)).astype(pix))
        if s % 3 == 0:  # extremes: full-range differences
            src[:, :w], prd[:, :w] = (1 << bd) - 1, 0
            src[::2, :w], prd[::2, :w] = 0, (1 << bd) - 1
        a = aligned((h, 80), np.int16, fill=7)
        b = aligned((h, 80), np.int16, fill=7)
        R.ref_residual(P(src), stride, P(prd), stride, P(a), 80, w, h, bd)
        O.svto_residual(P(src), stride, P(prd), stride, P(b), 80, w, h, bd)
        assert np.array_equal(a, b), (s, bd, asm)
        assert np.array_equal(a[:, :w], src[:, :w].astype(np.int32) - prd[:, :w].astype(np.int32))


@pytest.mark.parametrize("asm", [0, 1])
def test_tu_distortion(asm):
    """svto_tu_distortion == picture_full_distortion32_bits + energy + shift (EbFullLoop.c:808-835),
    C and AVX2 kernels, with and without non-zero coefficients, every transform size."""
    R, O = ref(asm), oracle()
    rng = np.random.default_rng(77)
    u64 = ctypes.c_uint64
    for s in range(19):
        n = max_eob(s)
        for amp, eob, energy in ((200, 5, 0), (1 << 17, n, 123456789), (1 << 15, 0, 0), (3, 0, 99)):
            cf = aligned_copy(rng.integers(-amp, amp + 1, n).astype(np.int32))
            dq = aligned_copy((cf + rng.integers(-50, 51, n)).astype(np.int32))
            a = np.zeros(2, np.uint64)
            b = np.zeros(2, np.uint64)
            R.ref_tu_distortion(P(cf), P(dq), s, eob, u64(energy), P(a), asm)
            O.svto_tu_distortion(P(cf), P(dq), s, eob, u64(energy), P(b))
            assert np.array_equal(a, b), (s, amp, eob, a, b)


@pytest.mark.parametrize("bd,use_qm", [(8, False), (10, False), (10, True)])
def test_pool_frame_batch_equals_the_restatement(bd, use_qm):
    """The harness's pthread pool (bench.py's CPU arm and the checker of its parity gate) over a
    mixed-size frame batch == the plain-C restatement, including the qmatrix routing; the
    reference skips the inverse when eob == 0, which leaves the prediction in place either way."""
    from cidana_svt_av1_b200 import workloads as wl
    from golden_util import load_qmatrices, load_quant_tables, load_scans
    from oracle_lib import CpuRunner, build_tables, qm_pack, run_oracle, run_ref
    batch = wl.make_frame_batch(256, 192, 2, bd, seed=91, qindex=100, laplace_b=20.0, use_qm=use_qm)
    qt = load_quant_tables()
    tables = build_tables({p: qt[(bd, p)] for p in range(3)})
    qm = qm_pack(load_qmatrices(), 6) if use_qm else None
    want = run_oracle(batch, tables, load_scans(), qm=qm, threads=4)
    for threads in (1, 5):
        got = run_ref(batch, tables, threads=threads, qm=qm)
        for k in ("qcoeff", "dqcoeff", "eob", "energy", "recon"):
            assert np.array_equal(got[k], want[k]), (k, threads)
    r = CpuRunner(batch, tables, scans=load_scans(), qm=qm)
    assert r.kind == "reference"
    assert r.run(3) > 0
    r.reset()
    r.run(3)
    for k, v in r.outputs().items():
        assert np.array_equal(v, want[k]), k
