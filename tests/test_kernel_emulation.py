"""CPU tier: the product's device code (txfm1d.cuh / residual_kernel.cuh / host_plan.h), compiled
for the host, against the oracle and the golden fixtures.  The GPU tier (test_gpu_parity.py) runs
the same comparisons through the C ABI on a B200."""
import ctypes

import numpy as np
import pytest

from cidana_svt_av1_b200 import workloads as wl
from emul_lib import emul, run_emul
from golden_util import load_fused_vectors, load_qmatrices, load_quant_tables, load_scans
from oracle_lib import P, build_tables, max_eob, oracle, qm_pack, run_oracle, run_oracle_md

KEYS = ("qcoeff", "dqcoeff", "eob", "energy", "recon")


def _tables(bd):
    qt = load_quant_tables()
    return build_tables({p: qt[(bd, p)] for p in range(3)})


def _check(a, b, what):
    for k in KEYS:
        assert np.array_equal(a[k], b[k]), (what, k, np.nonzero(a[k] != b[k])[0][:5])


def test_cospi_matches_reference_closed_form():
    O, E = oracle(), emul()
    for bit in range(10, 14):
        ref = np.ctypeslib.as_array(O.svto_cospi(bit), (64,))
        got = np.array([E.emul_cospi(bit, j) for j in range(64)])
        assert np.array_equal(ref, got)


def test_generated_scans_match_av1_scan_orders():
    E = emul()
    scans = load_scans()
    for s in range(19):
        for c in range(3):
            n = max_eob(s)
            out = np.zeros(n, np.int16)
            assert E.emul_get_iscan(s, c, P(out)) == n
            assert np.array_equal(out, scans[(s, c)][1]), (s, c)


def test_iscan_grows_along_every_row():
    """The kernel takes a row's eob candidate from its LAST non-zero column (row_k in
    residual_kernel.cuh); that is only the maximum of iscan over the row's non-zero positions
    if iscan grows with the column index, for every size and scan class of av1_scan_orders."""
    for (s, cls), (_, iscan) in load_scans().items():
        kw, kh = min(int(wl.TX_W[s]), 32), min(int(wl.TX_H[s]), 32)
        m = iscan.astype(np.int64).reshape(kh, kw)
        assert (np.diff(m, axis=1) > 0).all(), (s, cls)


@pytest.mark.parametrize("bd", [8, 10])
@pytest.mark.parametrize("strided", [False, True])
def test_fused_all_pairs(bd, strided):
    batch = wl.make_packed_batch(wl.valid_pairs(), bd, 3, seed=13596 + bd, strided=strided)
    tables = _tables(bd)
    a = run_emul(batch, tables)
    b = run_oracle(batch, tables, load_scans())
    _check(a, b, f"bd{bd}")
    assert (b["eob"] > 0).mean() > 0.5


@pytest.mark.parametrize("bd", [8, 10])
def test_fused_small_residuals_and_zero_blocks(bd):
    batch = wl.make_packed_batch(wl.valid_pairs(), bd, 2, seed=5, amp_div=64, qindex=(200, 255))
    batch.residual[: len(batch.residual) // 3] = 0
    tables = _tables(bd)
    _check(run_emul(batch, tables), run_oracle(batch, tables, load_scans()), "small")


def test_clamp16_variant_8bit():
    pairs = [(s, 0) for s in range(19)]
    batch = wl.make_packed_batch(pairs, 8, 3, seed=3, qindex=(0, 1, 2), flags=wl.BLK_CLAMP16)
    tables = _tables(8)
    a = run_emul(batch, tables)
    b = run_oracle(batch, tables, load_scans(), clamp16=1)
    _check(a, b, "clamp16")


def test_qmatrix_10bit():
    qm = qm_pack(load_qmatrices(), 6)
    pairs = [(s, t) for s, t in wl.valid_pairs() if t < 9]
    batch = wl.make_packed_batch(pairs, 10, 2, seed=9, qindex=(100,), flags=wl.BLK_USE_QM)
    tables = _tables(10)
    a = run_emul(batch, tables, qm=qm)
    b = run_oracle(batch, tables, load_scans(), qm=qm)
    _check(a, b, "qm")
    flat = run_oracle(batch, tables, load_scans())
    assert not np.array_equal(flat["qcoeff"], b["qcoeff"])  # the matrices do something


def test_stage_split_equals_fused():
    """fwd-only -> quant-only -> inverse-only (the RTCD-level path) == fused."""
    batch = wl.make_packed_batch(wl.valid_pairs(), 10, 1, seed=21)
    tables = _tables(10)
    fused = run_emul(batch, tables, want_coeff=True)
    f = run_emul(batch, tables, stages=1, want_coeff=True)
    assert np.array_equal(f["coeff"], fused["coeff"])
    q = run_emul(batch, tables, stages=2, coeff_in=f["coeff"])
    assert np.array_equal(q["qcoeff"], fused["qcoeff"]) and np.array_equal(q["eob"], fused["eob"])
    i = run_emul(batch, tables, stages=4, coeff_in=q["dqcoeff"])
    assert np.array_equal(i["recon"], fused["recon"])


def test_golden_vectors_from_reference_build():
    """Known answers produced by the reference's AVX2 path (tests/golden/make_golden.py)."""
    g = load_fused_vectors()
    O = oracle()
    for bd in (8, 10):
        tables = _tables(bd)
        pairs = wl.valid_pairs()
        batch = wl.make_packed_batch(pairs, bd, 1, seed=0)
        for i, (s, t) in enumerate(pairs):
            k = f"{bd}_{s}_{t}"
            o, n = int(batch.blocks["res_off"][i]), wl.TX_W[s] * wl.TX_H[s]
            batch.residual[o:o + n] = g["res_" + k].ravel()
            batch.pred[o:o + n] = g["pred_" + k].ravel()
            batch.blocks["qrow"][i] = int(g["meta_" + k][0])
        for name, out in (("emul", run_emul(batch, tables, want_coeff=True)),
                          ("oracle", run_oracle(batch, tables, load_scans()))):
            for i, (s, t) in enumerate(pairs):
                k = f"{bd}_{s}_{t}"
                co, n = int(batch.blocks["coeff_off"][i]), max_eob(s)
                o, a = int(batch.blocks["res_off"][i]), wl.TX_W[s] * wl.TX_H[s]
                meta = g["meta_" + k]
                assert np.array_equal(out["qcoeff"][co:co + n], g["q_" + k]), (name, k)
                assert np.array_equal(out["dqcoeff"][co:co + n], g["dq_" + k]), (name, k)
                assert out["eob"][i] == meta[1], (name, k)
                assert int(out["energy"][i]) == int(meta[2]) + (int(meta[3]) << 32), (name, k)
                assert np.array_equal(out["recon"][o:o + a], g["rec_" + k].ravel()), (name, k)
                if name == "emul":
                    assert np.array_equal(out["coeff"][co:co + n], g["coeff_" + k]), (name, k)


@pytest.mark.parametrize("bd", [8, 10])
def test_compiled_in_headline_configuration(bd):
    """The headline call runs with its switches compiled in (CfgFull<8> / <10>): stage flags, output
    pointers, pixel type, no clamp16 -- and, for 8-bit, without the 16-bit clamp at the column
    input after DCT / long ADST rows, where it cannot bind.  Same arithmetic, checked here on
    random full-range blocks and on patterns that drive the transforms to their extremes
    (constant +-max, checkerboards, single rows / columns, alternating stripes)."""
    E = emul()
    tables, scans = _tables(bd), load_scans()
    m = (1 << bd) - 1
    batches = [wl.make_packed_batch(wl.valid_pairs(), bd, 2, seed=91 + bd, qindex=(0, 40, 255))]
    ext = wl.make_packed_batch(wl.valid_pairs(), bd, 8, seed=92 + bd, qindex=(0, 10))
    for i, b in enumerate(ext.blocks):
        w, h = int(wl.TX_W[b["tx_size"]]), int(wl.TX_H[b["tx_size"]])
        yy, xx = np.mgrid[0:h, 0:w]
        pat = [np.full((h, w), m), np.full((h, w), -m), np.where((yy + xx) & 1, m, -m), np.where(yy & 1, m, -m),
               np.where(xx & 1, -m, m), np.where(yy == 0, m, -m), np.where(xx == w - 1, -m, m),
               np.where((yy // 2 + xx // 2) & 1, m, -m)][i % 8]
        o = int(b["res_off"])
        ext.residual[o:o + w * h] = pat.astype(np.int16).ravel()
        ext.pred[int(b["rec_off"]):int(b["rec_off"]) + w * h] = (0 if i % 2 else m)
    batches.append(ext)
    for k, batch in enumerate(batches):
        want = run_oracle(batch, tables, scans)
        E.emul_allow_full(1)
        try:
            got = run_emul(batch, tables)
        finally:
            E.emul_allow_full(0)
        _check(got, want, f"compiled-in bd{bd} batch {k}")


def test_frame_batch_small():
    batch = wl.make_frame_batch(256, 128, 1, 8, seed=20240)
    tables = _tables(8)
    _check(run_emul(batch, tables), run_oracle(batch, tables, load_scans()), "frame8")
    batch = wl.make_frame_batch(192, 128, 1, 10, seed=20241, qindex=100, laplace_b=24.0)
    tables = _tables(10)
    _check(run_emul(batch, tables), run_oracle(batch, tables, load_scans()), "frame10")


def test_source_minus_prediction_and_distortion():
    """residual taken as source - prediction while staging, per-block distortion pairs,
    with and without the coefficient outputs; all sizes and types, both depths, small qindex
    (eob != 0) and large (mostly eob == 0: the cbf-zero rule), packed and frame layouts."""
    scans = load_scans()
    for bd in (8, 10):
        tables = _tables(bd)
        for qi in ((20,), (255,)):
            batch = wl.make_packed_batch(wl.valid_pairs(), bd, 2, seed=40 + bd, qindex=qi).with_source()
            want = run_oracle_md(batch, tables, scans)
            # the source route and the residual route describe the same data
            _check(want, run_oracle(batch, tables, scans), "md oracle vs plain oracle")
            got = run_emul(batch, tables, stages=7 | 16, dist=True)
            _check(got, want, f"md emul bd{bd} q{qi}")
            assert np.array_equal(got["dist"], want["dist"]), np.nonzero(got["dist"] != want["dist"])[0][:5]
            lean = run_emul(batch, tables, stages=7 | 16, dist=True, coeffs=False)
            for k in ("eob", "energy", "recon", "dist"):
                assert np.array_equal(lean[k], want[k]), k
            # the same two calls with their switches compiled in (CfgFull<108> / <208> for 8-bit:
            # source and prediction staged as pixels, the column owner subtracts)
            E = emul()
            E.emul_allow_full(1)
            try:
                fast = run_emul(batch, tables, stages=7 | 16, dist=True, coeffs=False)
                cand = run_emul(batch, tables, stages=3 | 16, dist=True, coeffs=False)
            finally:
                E.emul_allow_full(0)
            for k in ("eob", "energy", "recon", "dist"):
                assert np.array_equal(fast[k], want[k]), ("compiled-in md", k)
            for k in ("eob", "energy", "dist"):
                assert np.array_equal(cand[k], want[k]), ("compiled-in tx-search shape", k)
    frame = wl.make_frame_batch(256, 128, 1, 8, seed=7).with_source()
    want = run_oracle_md(frame, _tables(8), scans)
    got = run_emul(frame, _tables(8), stages=7 | 16, dist=True)
    _check(got, want, "md frame")
    assert np.array_equal(got["dist"], want["dist"])
    assert (want["eob"] == 0).any() and (want["eob"] != 0).any()


def test_plan_rejects_bad_blocks():
    E = emul()
    batch = wl.make_packed_batch([(4, 0)], 8, 1, seed=1)
    tables = _tables(8)
    for field, val in (("tx_type", 1), ("bd", 12), ("coeff_off", 2), ("tx_size", 19)):
        bad = wl.make_packed_batch([(4, 0)], 8, 1, seed=1)
        bad.blocks[field] = val
        with pytest.raises(AssertionError):
            run_emul(bad, tables)


def test_inverse_networks_have_unit_node_gain():
    """fast_limit (residual_kernel.cuh), fact 1: no add/sub node of an inverse DCT / ADST network
    weighs any single input by more than 1, so |node| <= sum |inputs| + rounding."""
    E = emul()
    E.emul_inverse_node_gain.restype = ctypes.c_double
    for kind, sizes in ((0, (4, 8, 16, 32, 64)), (1, (4, 8, 16))):
        for n in sizes:
            g = E.emul_inverse_node_gain(n, kind)
            assert 0.5 < g <= 1.0 + 1e-3, (kind, n, g)


def test_clamp_free_inverse_equals_clamped_inside_the_limit():
    """fast_limit, fact 2: with L1 norm + kFastE inside the stage range the clamped and the
    clamp-free networks agree on every output -- tried right at the limit, for every size and kind
    and both stage ranges; and beyond the range they do differ somewhere (the test can fail)."""
    E = emul()
    e = E.emul_fast_e()
    differs_beyond = 0
    for kind, sizes in ((0, (4, 8, 16, 32, 64)), (1, (8, 16))):
        for n in sizes:
            for bits in (16, 18):
                lim = (1 << (bits - 1)) - 1 - e
                for l1 in (lim, lim - 1, lim // 2, 7):
                    assert E.emul_clampfree_diff(n, kind, 17 * n + bits, l1, bits, 3000) == 0, (kind, n, bits, l1)
                differs_beyond += E.emul_clampfree_diff(n, kind, 5, 3 * (1 << bits), bits, 3000) > 0
    assert differs_beyond > 0


def test_two_multiply_quantizer_equals_the_64_bit_form():
    """quant_flat computes the level with two 32 x 32 -> high-word multiplies where log_scale < 2.
    Checked against the 64-bit form it replaces on every row of the reference's tables (256 qindex
    x 3 planes, both bit depths, DC and AC), a dense low range of magnitudes, every power of two
    +-2 and random magnitudes below 2^30, both signs, with and without the 16-bit clamp."""
    E = emul()
    E.emul_quant_forms_diff.restype = ctypes.c_longlong
    for bd in (8, 10):
        t = np.ascontiguousarray(_tables(bd))
        assert t.dtype == np.int16 and t.size % 40 == 0
        for ls in (0, 1):
            bad = E.emul_quant_forms_diff(t.ctypes.data_as(ctypes.c_void_p), t.size // 40, ls, 30, 3000, 7 + ls)
            assert bad == 0, (bd, ls, bad)


def test_qmatrix_quantizer_equals_the_reference_arithmetic():
    """quant_qm against highbd_quantize_b_helper_c's arithmetic written out in 64 bits, coefficient by
    coefficient: a third of the table rows, nine weights, magnitudes up to 2^30, all three
    log_scales.  (A two-multiply form guarded at 2^21 passed this test too but ran slower on B200 --
    4k10qm 174 vs 183 Gcoeff/s, two code paths per coefficient -- and was dropped.)"""
    E = emul()
    E.emul_quant_qm_diff.restype = ctypes.c_longlong
    t = np.ascontiguousarray(_tables(10))
    for ls in (0, 1, 2):
        assert E.emul_quant_qm_diff(t.ctypes.data_as(ctypes.c_void_p), t.size // 40, ls, 1500, 11 + ls) == 0, ls


def test_both_inverse_paths_are_exercised():
    """The frame workloads run (almost) entirely inside fast_limit, the full-range sweep (almost)
    entirely outside: the emulation counts which inverse a block took."""
    E = emul()
    blocks, fast = ctypes.c_longlong(), ctypes.c_longlong()
    E.emul_fast_stats(ctypes.byref(blocks), ctypes.byref(fast))
    small = wl.make_frame_batch(256, 128, 1, 8, seed=3)
    _check(run_emul(small, _tables(8)), run_oracle(small, _tables(8), load_scans()), "frame")
    E.emul_fast_stats(ctypes.byref(blocks), ctypes.byref(fast))
    assert blocks.value == small.n_blocks and fast.value > 0.9 * blocks.value, (blocks.value, fast.value)
    full = wl.make_packed_batch([(2, 0), (3, 0), (1, 5)], 8, 20, seed=4)
    _check(run_emul(full, _tables(8)), run_oracle(full, _tables(8), load_scans()), "sweep")
    E.emul_fast_stats(ctypes.byref(blocks), ctypes.byref(fast))
    assert fast.value < 0.2 * blocks.value, (blocks.value, fast.value)
