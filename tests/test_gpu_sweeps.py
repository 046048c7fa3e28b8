"""GPU tier: the reference's own sweeps, on the device.

* QuantAsmTest (test/QuantAsmTest.cc:225-292): all 256 qindex x {random, DC/AC min-max, zero}
  through all six quantizer slots, and all 256 qindex through the fused path;
* InvTxfm2dAsmTest (test/InvTxfm2dAsmTest.cc:219-251): forward-transformed inputs truncated at
  every eob in scan order through the inverse path (batch API and the RTCD slots);
* the extreme residual patterns (all +-max, checkerboards, impulses) that guard the clamp
  elisions of the compiled-in variants, on the device;
* the clamp-free / clamped split of the inverse transform exercised on both sides of its limit.

The checker is the compiled reference (oracle/_ref, its AVX2 RTCD binding -- present on the GPU box
as a prebuilt file) wherever it exists, else the oracle restatement.  Bit-exact everywhere."""
import ctypes

import numpy as np
import pytest

import cidana_svt_av1_b200 as pkg
from cidana_svt_av1_b200 import workloads as wl
from golden_util import load_quant_tables, load_scans
from oracle_lib import P, aligned, aligned_copy, build_tables, max_eob, oracle, ref, run_oracle, run_ref, scan_class, tx_scale

pytestmark = pytest.mark.gpu
KEYS = ("qcoeff", "dqcoeff", "eob", "energy", "recon")
_ctx = {}


def tables(bd):
    qt = load_quant_tables()
    return build_tables({p: qt[(bd, p)] for p in range(3)})


def ctx(bd):
    if bd not in _ctx:
        g = pkg.SvtGpu(0)
        g.upload_quant_tables(tables(bd))
        _ctx[bd] = g
    return _ctx[bd]


def checker(batch, asm=1):
    if ref() is not None:
        return run_ref(batch, tables(batch.bd), threads=8, asm=asm)
    return run_oracle(batch, tables(batch.bd), load_scans(), threads=8)


def check(a, b, what):
    for k in KEYS:
        assert np.array_equal(a[k], b[k]), (what, k, np.nonzero(a[k] != b[k])[0][:5])


def ref_quant(cf, tb, s, t, bd):
    """The quantizer the encoder binds for this bit depth (RTCD pointer), or the oracle."""
    n = len(cf)
    q = aligned(n, np.int32, 3); dq = aligned(n, np.int32, 5); e = ctypes.c_uint16(9)
    R = ref()
    if R is not None:
        R.ref_quantize(P(cf), n, P(tb[0]), P(tb[1]), P(tb[2]), P(tb[3]), P(tb[4]), P(q), P(dq), ctypes.byref(e),
                       s, t, bd, 0, None, None)
    else:
        scan, _ = load_scans()[(s, scan_class(t))]
        oracle().svto_quantize(P(cf), n, P(tb[0]), P(tb[1]), P(tb[2]), P(tb[3]), P(tb[4]), P(q), P(dq),
                               ctypes.byref(e), P(scan), tx_scale(s), None, None, 0)
    return q, dq, e.value


@pytest.mark.parametrize("bd", [8, 10])
def test_quant_slots_every_qindex(bd):
    """All six slots: aom_quantize_b / _32x32 / _64x64 (8-bit) and the highbd three (10-bit), at
    every qindex, with QuantAsmTest's coefficient patterns."""
    L = pkg.load_library()
    pre = "svtgpu_aom_quantize_b" if bd == 8 else "svtgpu_aom_highbd_quantize_b"
    slots = {2: getattr(L, pre), 3: getattr(L, pre + "_32x32"), 4: getattr(L, pre + "_64x64")}
    tabs = load_quant_tables()[(bd, 0)]
    scans = load_scans()
    rng = np.random.default_rng(100 + bd)
    cmin, cmax = -(1 << (7 + bd)), (1 << (7 + bd)) - 1
    for s, fn in slots.items():
        n = max_eob(s)
        scan, iscan = scans[(s, 0)]
        for qi in range(256):
            tb = tabs[qi]
            pats = [rng.integers(cmin, cmax + 1, n).astype(np.int32), (rng.integers(cmin, cmax + 1, n) >> 7).astype(np.int32)]
            if qi % 25 == 0:  # input_dcac_minmax_q_n
                for dc in (cmin, cmax):
                    for ac in (cmin, cmax):
                        z = np.zeros(n, np.int32); z[0] = dc; z[1] = ac
                        pats.append(z)
            if qi == 0:
                pats.append(np.zeros(n, np.int32))  # input_zero_all
                z = np.zeros(n, np.int32); z[0] = rng.integers(cmin, cmax + 1)  # input_random_dc_only
                pats.append(z)
            for k, cf in enumerate(pats):
                cf = aligned_copy(cf)
                want = ref_quant(cf, tb, s, 0, bd)
                q = np.ones(n, np.int32); dq = np.ones(n, np.int32); e = ctypes.c_uint16(7)
                fn(P(cf), ctypes.c_ssize_t(n), 0, P(tb[0]), P(tb[1]), P(tb[2]), P(tb[3]), P(q), P(dq), P(tb[4]),
                   ctypes.byref(e), P(scan), P(iscan))
                assert np.array_equal(q, want[0]) and np.array_equal(dq, want[1]) and e.value == want[2], (s, qi, k)


@pytest.mark.parametrize("bd", [8, 10])
def test_fused_path_every_qindex(bd):
    """The fused path with qrow drawn from all 256 qindex, full-range and small residuals."""
    pairs = [(s, t) for s in (0, 1, 2, 3, 4, 7, 10, 13, 16, 18) for t in (0, 1, 6, 9, 10, 15) if wl.is_valid(s, t)]
    for amp_div in (1, 16):
        batch = wl.make_packed_batch(pairs, bd, 64, seed=40 + bd + amp_div, qindex=tuple(range(256)), amp_div=amp_div)
        assert len(np.unique(batch.blocks["qrow"])) == 256
        check(ctx(bd).run_host(batch), checker(batch), (bd, amp_div))


def _truncated_coeffs(bd, pairs, per_eob_step):
    """For every (size, type): its forward transform of a random block, truncated after the first
    eob coefficients in scan order, for eob = 1 .. max (InvTxfm2dAsmTest::run_with_eob)."""
    scans = load_scans()
    g = ctx(bd)
    base = wl.make_packed_batch(pairs, bd, 1, seed=77 + bd)
    coeff = g.run_staged(base, stages=1, want_coeff=True)["coeff"]
    blocks, chunks, eobs = [], [], []
    off = 0
    for i, (s, t) in enumerate(pairs):
        n = max_eob(s)
        c0 = int(base.blocks["coeff_off"][i])
        full = coeff[c0:c0 + n]
        scan, _ = scans[(s, scan_class(t))]
        step = 1 if n <= 256 else per_eob_step
        for eob in sorted(set(list(range(1, min(n, 64) + 1)) + list(range(1, n + 1, step)) + [n])):
            cf = np.zeros(n, np.int32)
            keep = scan[:eob].astype(np.int64)
            cf[keep] = full[keep]
            b = base.blocks[i].copy()
            b["coeff_off"] = off
            blocks.append(b)
            chunks.append(cf)
            eobs.append(eob)
            off += n
    blk = np.array(blocks, dtype=wl.BLOCK_DTYPE)
    batch = wl.Batch(blk, base.residual, base.pred, off, bd, "eob_sweep")
    return batch, np.concatenate(chunks), np.array(eobs)


@pytest.mark.parametrize("bd", [8, 10])
def test_inverse_at_every_eob(bd):
    """Blocks that reuse one prediction block each (the recon of a block is compared block by
    block against the reference run on a private copy of the prediction, with the eob passed so
    that the reference's eob-dependent shortcuts are the ones compared with)."""
    pairs = [(s, t) for s in range(19) for t in (0, 3, 9, 12) if wl.is_valid(s, t)]
    batch, cin, eobs = _truncated_coeffs(bd, pairs, per_eob_step=13)
    # every block needs its own pixels: lay the blocks out densely, one after the other
    s_arr = batch.blocks["tx_size"]
    area = wl.AREA[s_arr]
    off = np.zeros(len(s_arr), np.int64)
    np.cumsum(area[:-1], out=off[1:])
    rng = np.random.default_rng(5)
    pred = rng.integers(0, 1 << bd, int(area.sum())).astype(np.uint8 if bd == 8 else np.uint16)
    blk = batch.blocks.copy()
    blk["rec_off"] = off
    blk["res_off"] = off
    blk["rec_stride"] = np.array(wl.TX_W)[s_arr]
    blk["res_stride"] = blk["rec_stride"]
    b2 = wl.Batch(blk, np.zeros(len(pred), np.int16), pred, batch.n_coeff, bd, "eob_sweep")
    got = ctx(bd).run_staged(b2, stages=4, coeff_in=cin)["recon"]
    R = ref()
    want = pred.copy()
    if R is not None:
        cbuf = aligned_copy(cin)
        wbuf = aligned_copy(want)
        for i in range(len(blk)):
            s, t = int(blk["tx_size"][i]), int(blk["tx_type"][i])
            px = 1 if bd == 8 else 2
            R.ref_inv_transform_recon(ctypes.c_void_p(cbuf.ctypes.data + 4 * int(blk["coeff_off"][i])),
                                      ctypes.c_void_p(wbuf.ctypes.data + px * int(off[i])), int(wl.TX_W[s]), s, t, bd,
                                      int(eobs[i]))
        want = wbuf
    else:
        O = oracle()
        for i in range(len(blk)):
            s, t = int(blk["tx_size"][i]), int(blk["tx_type"][i])
            px = 1 if bd == 8 else 2
            fn = O.svto_inv_txfm2d_add8 if bd == 8 else O.svto_inv_txfm2d_add
            args = [ctypes.c_void_p(cin.ctypes.data + 4 * int(blk["coeff_off"][i])),
                    ctypes.c_void_p(want.ctypes.data + px * int(off[i])), int(wl.TX_W[s]), s, t]
            fn(*(args + ([bd] if bd != 8 else [])))
    bad = np.nonzero(got != want)[0]
    assert len(bad) == 0, (bd, len(bad), bad[:5])
    # a sample of the same blocks through the level-A slots / level-B entry points (host pointers)
    for i in list(range(0, len(blk), 97))[:60]:
        s, t = int(blk["tx_size"][i]), int(blk["tx_type"][i])
        w, h = wl.TX_W[s], wl.TX_H[s]
        cf = np.ascontiguousarray(cin[int(blk["coeff_off"][i]):int(blk["coeff_off"][i]) + max_eob(s)])
        p0 = np.ascontiguousarray(pred[int(off[i]):int(off[i]) + w * h])
        if bd == 8:
            out = pkg.av1_inv_transform_recon8bit(cf, p0.copy(), w, s, t, int(eobs[i]))
        else:
            out = pkg.av1_inv_transform_recon(cf, p0.copy(), w, s, 2, t, int(eobs[i]))
        assert np.array_equal(out, want[int(off[i]):int(off[i]) + w * h]), (s, t, int(eobs[i]))


def _pattern_batch(bd, pattern, pairs, n_per=6, qindex=(0, 40, 255)):
    batch = wl.make_packed_batch(pairs, bd, n_per, seed=3, qindex=qindex)
    m = (1 << bd) - 1
    res = batch.residual
    idx = np.arange(len(res))
    if pattern == "max":
        res[:] = m
    elif pattern == "min":
        res[:] = -m
    elif pattern == "checker":  # +-max checkerboard inside every block (blocks are dense, stride = w)
        for i in range(batch.n_blocks):
            s = int(batch.blocks["tx_size"][i]); w, h = wl.TX_W[s], wl.TX_H[s]
            o = int(batch.blocks["res_off"][i])
            yy, xx = np.mgrid[0:h, 0:w]
            res[o:o + w * h] = np.where((xx + yy) & 1, m, -m).ravel()
    elif pattern == "rows":  # alternating +-max rows
        for i in range(batch.n_blocks):
            s = int(batch.blocks["tx_size"][i]); w, h = wl.TX_W[s], wl.TX_H[s]
            o = int(batch.blocks["res_off"][i])
            yy, xx = np.mgrid[0:h, 0:w]
            res[o:o + w * h] = np.where(yy & 1, m, -m).ravel()
    elif pattern == "impulse":  # one +-max sample per block, elsewhere zero
        res[:] = 0
        for i in range(batch.n_blocks):
            s = int(batch.blocks["tx_size"][i]); n = wl.TX_W[s] * wl.TX_H[s]
            o = int(batch.blocks["res_off"][i])
            res[o + (i * 7) % n] = m if i & 1 else -m
    elif pattern == "mixed":
        res[:] = np.where(idx % 3 == 0, m, np.where(idx % 3 == 1, -m, 0))
    # bright and dark predictions so that the reconstruction clip binds on both sides
    batch.pred[:] = np.where((idx // 64) & 1, m, 0).astype(batch.pred.dtype)
    return batch


@pytest.mark.parametrize("bd", [8, 10])
@pytest.mark.parametrize("pattern", ["max", "min", "checker", "rows", "impulse", "mixed"])
def test_extreme_patterns_on_the_device(bd, pattern):
    """The compiled-in variants (CfgFull<8> / <10>) leave out the 16-bit clamp at the column input
    where it cannot bind and omit check_range; these patterns are the ones that come closest.

    8-bit: the reference's C and AVX2 / SSSE3 builds agree and both are checked.  10-bit: on these
    saturated inputs the reference's own SSE4.1 / AVX2 inverse departs from its C inverse
    (inv_txfm2d_add_c with every clamp, EbTransforms.c:8180-8291) for 16x16 DCT_ADST / DCT_FLIPADST at
    qindex 255 and for 4x16 with an identity row (V_DCT, V_ADST, V_FLIPADST) -- a divergence inside
    the reference (its InvTxfm2dAsmTest only feeds transforms of random residuals).  The GPU path
    follows the C definition, like the oracle; the SIMD build is checked to differ from it only in
    those classes."""
    batch = _pattern_batch(bd, pattern, wl.valid_pairs())
    got = ctx(bd).run_host(batch)
    check(got, checker(batch, asm=0), (bd, pattern, "C"))
    if ref() is not None:
        simd = checker(batch, asm=1)
        for k in ("qcoeff", "dqcoeff", "eob", "energy"):
            assert np.array_equal(got[k], simd[k]), (bd, pattern, k)
        bad = np.nonzero(got["recon"] != simd["recon"])[0]
        if bd == 8:
            assert len(bad) == 0
        else:
            owner = np.searchsorted(batch.blocks["rec_off"], bad, side="right") - 1
            classes = {(int(batch.blocks["tx_size"][i]), int(batch.blocks["tx_type"][i])) for i in owner}
            assert classes <= {(2, 2), (2, 5), (13, 10), (13, 12), (13, 14)}, classes


@pytest.mark.parametrize("bd", [8, 10])
def test_both_sides_of_the_clamp_free_limit(bd):
    """Residual amplitudes scaled so that sum |dqcoeff| of the blocks straddles fast_limit
    (residual_kernel.cuh): some blocks of every size run the clamp-free inverse, some the clamped
    one, inside the same launch, sometimes inside the same warp."""
    pairs = [(s, t) for s in range(19) for t in (0, 1, 5, 9, 11) if wl.is_valid(s, t)]
    for amp_div, qi in ((2, (20,)), (4, (60,)), (8, (60, 120)), (24, (120,)), (64, (200,))):
        batch = wl.make_packed_batch(pairs, bd, 24, seed=amp_div, amp_div=amp_div, qindex=qi)
        # every third block a different scale: neighbours in a warp take different paths
        sc = np.repeat(np.arange(batch.n_blocks) % 3, wl.AREA[batch.blocks["tx_size"]])
        batch.residual[:] = (batch.residual.astype(np.int32) >> sc[:len(batch.residual)]).astype(np.int16)
        check(ctx(bd).run_host(batch), checker(batch), (bd, amp_div))
