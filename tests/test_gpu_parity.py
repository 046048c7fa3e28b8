"""GPU tier (run on a B200: `pytest -m gpu`): the CUDA path, called through the C ABI of
libsvtgpu.so, against the oracle on the same seeded inputs, against the golden vectors produced by
the reference build, and -- at BASELINE.json's full sizes -- through size-independent properties.
Bit-exact everywhere (integer path)."""
import ctypes

import numpy as np
import pytest

import cidana_svt_av1_b200 as pkg
from cidana_svt_av1_b200 import workloads as wl
from golden_util import load_fused_vectors, load_qmatrices, load_quant_tables, load_scans
from oracle_lib import P, build_tables, max_eob, qm_pack, run_oracle, run_oracle_md

pytestmark = pytest.mark.gpu
KEYS = ("qcoeff", "dqcoeff", "eob", "energy", "recon")
_ctx = {}


def ctx(bd, qm=None):
    key = (bd, qm is not None)
    if key not in _ctx:
        g = pkg.SvtGpu(0)
        g.upload_quant_tables(tables(bd))
        if qm is not None:
            g.upload_qmatrix(*qm)
        _ctx[key] = g
    return _ctx[key]


def tables(bd):
    qt = load_quant_tables()
    return build_tables({p: qt[(bd, p)] for p in range(3)})


def run_gpu(batch, stages=7, qm=None, coeff_in=None, want_coeff=False):
    g = ctx(batch.bd, qm)
    if coeff_in is None and not want_coeff and stages == 7:
        return g.run_host(batch)  # the host pipeline (svtgpu_submit_host / svtgpu_wait_host)
    return g.run_staged(batch, stages=stages, coeff_in=coeff_in, want_coeff=want_coeff)


def check(a, b, what):
    for k in KEYS:
        assert np.array_equal(a[k], b[k]), (what, k, np.nonzero(a[k] != b[k])[0][:5])


@pytest.mark.parametrize("bd", [8, 10])
@pytest.mark.parametrize("strided", [False, True])
def test_fused_all_pairs(bd, strided):
    batch = wl.make_packed_batch(wl.valid_pairs(), bd, 40, seed=13596 + bd, strided=strided)
    check(run_gpu(batch), run_oracle(batch, tables(bd), load_scans(), threads=8), f"bd{bd}")


@pytest.mark.parametrize("bd", [8, 10])
def test_small_residuals_and_zero_blocks(bd):
    batch = wl.make_packed_batch(wl.valid_pairs(), bd, 17, seed=5, amp_div=64, qindex=(200, 255))
    batch.residual[: len(batch.residual) // 3] = 0
    check(run_gpu(batch), run_oracle(batch, tables(bd), load_scans(), threads=8), "small")


def test_ragged_tail_and_single_block():
    """CTA tails: block counts that do not fill the last CTA, down to one block; empty list."""
    for n_per in (1, 3, 33):
        batch = wl.make_packed_batch(wl.valid_pairs(), 8, n_per, seed=n_per)
        check(run_gpu(batch), run_oracle(batch, tables(8), load_scans(), threads=8), n_per)
    g = ctx(8)
    empty = g.plan(np.zeros(0, wl.BLOCK_DTYPE))
    assert empty.n_blocks == 0


@pytest.mark.parametrize("bd", [8, 10])
def test_coefficient_rows_aligned_to_16_bytes_only(bd):
    """coeff_off is promised to be a multiple of 4 coefficients, not 8: the 256-bit row stores of
    the 8- and 16-wide sizes must fall back to 128-bit ones (store8), block by block -- here every
    second block sits 16 bytes off a 32-byte boundary."""
    base = wl.make_packed_batch(wl.valid_pairs(), bd, 9, seed=77 + bd)
    blocks = base.blocks.copy()
    blocks["coeff_off"] += (4 * (1 + (np.arange(len(blocks)) // 2))).astype(np.uint32)  # odd multiples of 4 for half of the blocks
    assert (blocks["coeff_off"] % 8 == 4).any() and (blocks["coeff_off"] % 8 == 0).any()
    batch = wl.Batch(blocks, base.residual, base.pred, base.n_coeff + 4 * (1 + len(blocks) // 2) + 4, bd, "odd")
    got, want = run_gpu(batch), run_oracle(batch, tables(bd), load_scans(), threads=8)
    covered = np.zeros(batch.n_coeff, bool)  # the gaps between blocks belong to nobody: not compared
    for off, s in zip(blocks["coeff_off"], blocks["tx_size"]):
        covered[int(off):int(off) + max_eob(int(s))] = True
    for k in ("qcoeff", "dqcoeff"):
        assert np.array_equal(got[k][covered], want[k][covered]), k
    for k in ("eob", "energy", "recon"):
        assert np.array_equal(got[k], want[k]), k


def test_clamp16_variant_8bit():
    pairs = [(s, 0) for s in range(19)]
    batch = wl.make_packed_batch(pairs, 8, 9, seed=3, qindex=(0, 1, 2), flags=wl.BLK_CLAMP16)
    check(run_gpu(batch), run_oracle(batch, tables(8), load_scans(), clamp16=1), "clamp16")


def test_qmatrix_10bit():
    qm = qm_pack(load_qmatrices(), 6)
    pairs = [(s, t) for s, t in wl.valid_pairs() if t < 9]
    batch = wl.make_packed_batch(pairs, 10, 7, seed=9, qindex=(100,), flags=wl.BLK_USE_QM)
    check(run_gpu(batch, qm=qm), run_oracle(batch, tables(10), load_scans(), qm=qm), "qm")


def test_stage_split_equals_fused():
    batch = wl.make_packed_batch(wl.valid_pairs(), 10, 5, seed=21)
    fused = run_gpu(batch, want_coeff=True)
    f = run_gpu(batch, stages=1, want_coeff=True)
    assert np.array_equal(f["coeff"], fused["coeff"])
    q = run_gpu(batch, stages=2, coeff_in=f["coeff"])
    assert np.array_equal(q["qcoeff"], fused["qcoeff"]) and np.array_equal(q["eob"], fused["eob"])
    i = run_gpu(batch, stages=4, coeff_in=q["dqcoeff"])
    assert np.array_equal(i["recon"], fused["recon"])


def test_golden_vectors_from_reference_build():
    g = load_fused_vectors()
    for bd in (8, 10):
        pairs = wl.valid_pairs()
        batch = wl.make_packed_batch(pairs, bd, 1, seed=0)
        for i, (s, t) in enumerate(pairs):
            k = f"{bd}_{s}_{t}"
            o, n = int(batch.blocks["res_off"][i]), wl.TX_W[s] * wl.TX_H[s]
            batch.residual[o:o + n] = g["res_" + k].ravel()
            batch.pred[o:o + n] = g["pred_" + k].ravel()
            batch.blocks["qrow"][i] = int(g["meta_" + k][0])
        out = run_gpu(batch, want_coeff=True)
        for i, (s, t) in enumerate(pairs):
            k = f"{bd}_{s}_{t}"
            co, n = int(batch.blocks["coeff_off"][i]), max_eob(s)
            o, a = int(batch.blocks["res_off"][i]), wl.TX_W[s] * wl.TX_H[s]
            meta = g["meta_" + k]
            assert np.array_equal(out["coeff"][co:co + n], g["coeff_" + k]), k
            assert np.array_equal(out["qcoeff"][co:co + n], g["q_" + k]), k
            assert np.array_equal(out["dqcoeff"][co:co + n], g["dq_" + k]), k
            assert out["eob"][i] == meta[1], k
            assert int(out["energy"][i]) == int(meta[2]) + (int(meta[3]) << 32), k
            assert np.array_equal(out["recon"][o:o + a], g["rec_" + k].ravel()), k


def test_frame_batches_vs_oracle():
    b8 = wl.make_frame_batch(640, 384, 2, 8, seed=20240)
    check(run_gpu(b8), run_oracle(b8, tables(8), load_scans(), threads=8), "frame8")
    b10 = wl.make_frame_batch(512, 256, 1, 10, seed=20241, qindex=100, laplace_b=24.0)
    check(run_gpu(b10), run_oracle(b10, tables(10), load_scans(), threads=8), "frame10")
    qm = qm_pack(load_qmatrices(), 6)
    bq = wl.make_frame_batch(512, 256, 1, 10, seed=20241, qindex=100, laplace_b=24.0, use_qm=True)
    check(run_gpu(bq, qm=qm), run_oracle(bq, tables(10), load_scans(), qm=qm, threads=8), "frame_qm")


def test_host_pipeline_many_frames_reused_buffers():
    """svtgpu_submit_host / svtgpu_wait_host: more pictures than lanes, page-locked buffers reused
    across runs, results land in the caller's arrays; and the path without reconstruction."""
    g = ctx(8)
    batch = wl.make_frame_batch(320, 192, 7, 8, seed=99)
    assert len(batch.frames()) == 7
    want = run_oracle(batch, tables(8), load_scans(), threads=8)
    hb = g.host_buffers(batch, pinned=True)
    for rep in range(3):
        for v in hb.out.values():
            if v is not None:
                v[:] = 0x5A if v.dtype == np.uint8 else 12345
        check(g.run_host(batch, bufs=hb), want, f"pipeline rep {rep}")
    # pageable buffers, tickets waited one by one
    hb2 = g.host_buffers(batch, pinned=False)
    tickets = [g.submit_host(fr, hb2.frame_out(k), 7, *hb2.frame_in(k)) for k, fr in enumerate(batch.frames())]
    assert tickets == sorted(tickets) and len(set(tickets)) == 7
    for t in tickets:
        g.wait_host(t)
    g.wait_host()
    check(hb2.out, want, "pipeline pageable")
    # forward + quantiser only: reconstruction buffer untouched
    hb3 = g.host_buffers(batch, pinned=False)
    hb3.out["recon"][:] = 7
    out = g.run_host(batch, stages=3, bufs=hb3)
    for k in ("qcoeff", "dqcoeff", "eob", "energy"):
        assert np.array_equal(out[k], want[k]), k
    assert (out["recon"] == 7).all()
    # a bad block is refused, not worked around
    bad = batch.frames()[0]
    blk = bad.blocks.copy()
    blk["tx_type"][0] = 15
    blk["tx_size"][0] = 4  # 64x64 only allows DCT_DCT
    bb = wl.Batch(blk, bad.residual, bad.pred, bad.n_coeff, 8, "bad")
    with pytest.raises(pkg.api.SvtGpuError):
        g.submit_host(bb, hb2.frame_out(0))
    g.wait_host()


def test_source_minus_prediction_and_distortion():
    """SURVEY 8f-1 on the device: residual = source - prediction computed while staging
    (ResidualKernel), per-block distortion pairs (picture_full_distortion32_bits + energy + scaling),
    with and without the coefficient arrays, through the host pipeline and the batch API."""
    scans = load_scans()
    for bd in (8, 10):
        g = ctx(bd)
        for qi in ((20,), (255,)):
            batch = wl.make_packed_batch(wl.valid_pairs(), bd, 9, seed=40 + bd, qindex=qi).with_source()
            want = run_oracle_md(batch, tables(bd), scans, threads=8)
            hb = g.host_buffers(batch, pinned=False, dist=True)
            got = g.run_host(batch, stages=7 | 16, bufs=hb)
            check(got, want, f"md bd{bd} q{qi}")
            assert np.array_equal(got["dist"], want["dist"]), np.nonzero(got["dist"] != want["dist"])[0][:5]
            lean = g.run_host(batch, stages=7 | 16, bufs=g.host_buffers(batch, pinned=False, dist=True, coeffs=False))
            for k in ("eob", "energy", "recon", "dist"):
                assert np.array_equal(lean[k], want[k]), k
            assert lean["qcoeff"] is None
    # frames (strided planes, flips, mixed sizes), pinned buffers, the plain-residual route with dist
    g = ctx(8)
    frame = wl.make_frame_batch(640, 384, 3, 8, seed=11).with_source()
    want = run_oracle_md(frame, tables(8), scans, threads=8)
    got = g.run_host(frame, stages=7 | 16, bufs=g.host_buffers(frame, pinned=True, dist=True))
    check(got, want, "md frames")
    assert np.array_equal(got["dist"], want["dist"])
    got2 = g.run_host(frame, stages=7, bufs=g.host_buffers(frame, pinned=True, dist=True))
    check(got2, want, "dist with a given residual")
    assert np.array_equal(got2["dist"], want["dist"])
    # unaligned planes take the per-pixel path
    odd = wl.make_packed_batch([(2, 0), (0, 5), (7, 9)], 8, 5, seed=3).with_source()
    odd.src = np.concatenate([np.zeros(3, odd.src.dtype), odd.src])[3:]
    odd.pred = np.concatenate([np.zeros(1, odd.pred.dtype), odd.pred])[1:]
    want = run_oracle_md(odd, tables(8), scans)
    got = g.run_host(odd, stages=7 | 16, bufs=g.host_buffers(odd, pinned=False, dist=True))
    check(got, want, "md unaligned")
    assert np.array_equal(got["dist"], want["dist"])
    # qmatrix + dist is refused
    qm = qm_pack(load_qmatrices(), 6)
    gq = ctx(10, qm)
    bq = wl.make_packed_batch([(1, 0)], 10, 2, seed=1, qindex=(100,), flags=wl.BLK_USE_QM).with_source()
    with pytest.raises(pkg.api.SvtGpuError):
        gq.run_host(bq, stages=7 | 16, bufs=gq.host_buffers(bq, pinned=False, dist=True))
    gq.wait_host()


# ---- levels A and B: the reference's own signatures, host pointers ---------------------------------
def _inv_slot_call(L, s, name, dq, pix, stride, t, bd, eob):
    fn = getattr(L, "svtgpu_av1_inv_txfm2d_add_" + name)
    u8, i32 = ctypes.c_uint8, ctypes.c_int32
    if s <= 4:
        fn(P(dq), P(pix), i32(stride), u8(t), i32(bd))
    elif s in (5, 6, 13, 14):
        fn(P(dq), P(pix), i32(stride), u8(t), u8(s), i32(bd))
    else:
        fn(P(dq), P(pix), i32(stride), u8(t), u8(s), i32(eob), i32(bd))


@pytest.mark.parametrize("bd", [8, 10])
def test_rtcd_slots_and_entry_points(bd):
    from oracle_lib import oracle, scan_class, tx_scale
    O, L = oracle(), pkg.load_library()
    scans = load_scans()
    tabs = load_quant_tables()[(bd, 0)]
    rng = np.random.default_rng(bd)
    m = (1 << bd) - 1
    for s, t in wl.valid_pairs():
        w, h, n = wl.TX_W[s], wl.TX_H[s], max_eob(s)
        name = f"{w}x{h}"
        stride = 72
        res = rng.integers(-m, m + 1, (h, stride)).astype(np.int16)
        # level A forward: full w*h coefficients
        want = np.zeros(w * h, np.int32)
        O.svto_fwd_txfm2d(P(res), stride, P(want), s, t, bd)
        got = np.zeros(w * h, np.int32)
        getattr(L, "svtgpu_av1_fwd_txfm2d_" + name)(P(res), P(got), ctypes.c_uint32(stride),
                                                     ctypes.c_uint8(t), ctypes.c_uint8(bd))
        assert np.array_equal(got, want), ("fwd slot", s, t)
        # level B forward: packed + energy
        cf_o = np.zeros(w * h, np.int32)
        en = ctypes.c_uint64()
        O.svto_estimate_transform(P(res), stride, P(cf_o), s, t, bd, ctypes.byref(en))
        cf_g, en_g = pkg.av1_estimate_transform(res, stride, s, t, bit_increment=0 if bd == 8 else 2)
        assert np.array_equal(cf_g, cf_o[:n]) and en_g == en.value, ("estimate", s, t)
        # quantizer: slot (caller's iscan) and entry point (library's scan choice)
        tb = tabs[[15, 110, 240][(s + t) % 3]]
        scan, iscan = scans[(s, scan_class(t))]
        q_o = np.zeros(n, np.int32); dq_o = np.zeros(n, np.int32); e_o = ctypes.c_uint16()
        O.svto_quantize(P(cf_g), n, P(tb[0]), P(tb[1]), P(tb[2]), P(tb[3]), P(tb[4]), P(q_o), P(dq_o),
                        ctypes.byref(e_o), P(scan), tx_scale(s), None, None, 0)
        slot = ("svtgpu_aom_quantize_b" if bd == 8 else "svtgpu_aom_highbd_quantize_b") + \
               ["", "_32x32", "_64x64"][tx_scale(s)]
        q_g = np.zeros(n, np.int32); dq_g = np.zeros(n, np.int32); e_g = ctypes.c_uint16()
        getattr(L, slot)(P(cf_g), ctypes.c_ssize_t(n), 0, P(tb[0]), P(tb[1]), P(tb[2]), P(tb[3]), P(q_g),
                         P(dq_g), P(tb[4]), ctypes.byref(e_g), P(scan), P(iscan))
        assert np.array_equal(q_g, q_o) and np.array_equal(dq_g, dq_o) and e_g.value == e_o.value, ("quant slot", s, t)
        q_b, dq_b, e_b = pkg.av1_quantize_inv_quantize(cf_g, tb, s, t, bit_increment=0 if bd == 8 else 2)
        assert np.array_equal(q_b, q_o) and np.array_equal(dq_b, dq_o) and e_b == e_o.value, ("qiq", s, t)
        # inverse: highbd slot on uint16 pixels, and the entry points
        p16 = rng.integers(0, m + 1, (h, stride)).astype(np.uint16)
        want16 = p16.copy()
        O.svto_inv_txfm2d_add(P(dq_o), P(want16), stride, s, t, bd)
        got16 = p16.copy()
        _inv_slot_call(L, s, name, dq_o, got16, stride, t, bd, e_o.value)
        assert np.array_equal(got16, want16), ("inv slot", s, t)
        if bd == 8:
            p8 = p16.astype(np.uint8)
            got8 = pkg.av1_inv_transform_recon8bit(dq_o, p8.copy(), stride, s, t, e_o.value)
            assert np.array_equal(got8, want16.astype(np.uint8)), ("recon8bit", s, t)
            prm = (ctypes.c_uint8 * 24)()
            prm[0], prm[1] = t, s
            got8b = p8.copy()
            L.svtgpu_av1_inv_txfm_add(P(dq_o), P(got8b), ctypes.c_int32(stride), prm)
            assert np.array_equal(got8b, got8), ("inv_txfm_add", s, t)
        else:
            gotb = pkg.av1_inv_transform_recon(dq_o, p16.copy(), stride, s, 2, t, e_o.value)
            assert np.array_equal(gotb, want16), ("recon", s, t)


def test_quant_slot_skip_block_and_extremes():
    """QuantAsmTest's DC/AC min/max inputs and skip_block."""
    from oracle_lib import oracle
    O, L = oracle(), pkg.load_library()
    scans = load_scans()
    tabs = load_quant_tables()[(10, 0)]
    scan, iscan = scans[(2, 0)]
    n = 256
    for qi in (0, 25, 50, 100, 255):
        tb = tabs[qi]
        for val in (-(1 << 17), (1 << 17) - 1, 32767, -32768):
            cf = np.zeros(n, np.int32)
            cf[0] = val
            cf[1:] = -val
            q_o = np.zeros(n, np.int32); dq_o = np.zeros(n, np.int32); e_o = ctypes.c_uint16()
            O.svto_quantize(P(cf), n, P(tb[0]), P(tb[1]), P(tb[2]), P(tb[3]), P(tb[4]), P(q_o),
                            P(dq_o), ctypes.byref(e_o), P(scan), 0, None, None, 0)
            q_g = np.ones(n, np.int32); dq_g = np.ones(n, np.int32); e_g = ctypes.c_uint16(7)
            L.svtgpu_aom_highbd_quantize_b(P(cf), ctypes.c_ssize_t(n), 0, P(tb[0]), P(tb[1]), P(tb[2]),
                                           P(tb[3]), P(q_g), P(dq_g), P(tb[4]), ctypes.byref(e_g),
                                           P(scan), P(iscan))
            assert np.array_equal(q_g, q_o) and np.array_equal(dq_g, dq_o) and e_g.value == e_o.value
    q_g = np.ones(n, np.int32); dq_g = np.ones(n, np.int32); e_g = ctypes.c_uint16(7)
    L.svtgpu_aom_highbd_quantize_b(P(cf), ctypes.c_ssize_t(n), 1, P(tb[0]), P(tb[1]), P(tb[2]), P(tb[3]),
                                   P(q_g), P(dq_g), P(tb[4]), ctypes.byref(e_g), P(scan), P(iscan))
    assert not q_g.any() and not dq_g.any() and e_g.value == 0


# ---- BASELINE.json full sizes: size-independent properties -----------------------------------------
def _subset(batch, idx):
    sub = wl.Batch(batch.blocks[idx].copy(), batch.residual, batch.pred, 0, batch.bd, "sub")
    return wl._finish(sub.blocks, batch.residual, batch.pred, batch.bd, "sub")


@pytest.mark.parametrize("cfg", ["1080p8", "4k10qm"])
def test_full_size_properties(cfg):
    if cfg == "1080p8":
        batch = wl.make_frame_batch(1920, 1080, 16, 8, seed=20240)
        qm = None
    else:
        batch = wl.make_frame_batch(3840, 2160, 4, 10, seed=20241, qindex=100, laplace_b=24.0, use_qm=True)
        qm = qm_pack(load_qmatrices(), 6)
    out = run_gpu(batch, qm=qm, want_coeff=True)
    # (1) determinism / idempotence of the launch (inputs are not modified)
    again = run_gpu(batch, qm=qm)
    for k in KEYS:
        assert np.array_equal(out[k], again[k]), k
    # (2) EVERY block equals the checker: the compiled reference over the whole batch (its AVX2 RTCD
    # path, plus highbd_quantize_b_helper_c for the qmatrix blocks) where oracle/_ref exists on the
    # box, else the oracle restatement
    from oracle_lib import ref, run_ref
    if ref() is not None:
        ora = run_ref(batch, tables(batch.bd), threads=16, qm=qm)
    else:
        ora = run_oracle(batch, tables(batch.bd), load_scans(), qm=qm, threads=16)
    for k in KEYS:
        assert np.array_equal(out[k], ora[k]), (cfg, k, np.nonzero(out[k] != ora[k])[0][:5])
    # (3) the three RTCD-level stages chained == fused
    q = run_gpu(batch, stages=2, qm=qm, coeff_in=out["coeff"])
    assert np.array_equal(q["qcoeff"], out["qcoeff"]) and np.array_equal(q["eob"], out["eob"])
    inv = run_gpu(batch, stages=4, qm=qm, coeff_in=out["dqcoeff"])
    assert np.array_equal(inv["recon"], out["recon"])
    # (4) zero residual: nothing survives and reconstruction == prediction
    zero = wl.Batch(batch.blocks, np.zeros_like(batch.residual), batch.pred, batch.n_coeff, batch.bd, "z")
    z = run_gpu(zero, qm=qm)
    assert not z["qcoeff"].any() and not z["eob"].any() and np.array_equal(z["recon"], batch.pred)
    # (5) eob is consistent with qcoeff: eob == 0 <=> block all zero
    offs = batch.blocks["coeff_off"].astype(np.int64)
    nz = np.add.reduceat((out["qcoeff"] != 0).astype(np.int64), offs)
    assert np.array_equal(nz > 0, out["eob"] > 0)
