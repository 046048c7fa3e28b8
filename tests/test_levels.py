"""the padded level map the coefficient-rate estimate starts from
(av1_txb_init_levels, EbRateDistortionCost.c:125 / :450), written by the quantiser next to qcoeff.

Pins: oracle restatement == the reference's C and AVX2 functions (where the reference is compiled);
kernel code on the host and on the GPU == the oracle applied to the oracle's qcoeff, byte for byte
over the whole padded buffer (padding rows and columns included), and nothing written outside it."""
import ctypes

import numpy as np
import pytest

import cidana_svt_av1_b200 as pkg
from cidana_svt_av1_b200 import workloads as wl
from emul_lib import run_emul
from golden_util import load_qmatrices, load_quant_tables, load_scans
from oracle_lib import P, build_tables, oracle, qm_pack, ref, run_oracle

GUARD = 0xEE


def _tables(bd):
    qt = load_quant_tables()
    return build_tables({p: qt[(bd, p)] for p in range(3)})


def _size(s):
    return (min(int(wl.TX_W[s]), 32) + 4) * (min(int(wl.TX_H[s]), 32) + 6) + 16


def _offsets(batch, gap=8):
    """Dense placement with a few guard bytes between blocks (offsets stay multiples of 4)."""
    sz = np.array([_size(int(s)) + gap for s in batch.blocks["tx_size"]], np.int64)
    off = np.zeros(len(sz) + 1, np.int64)
    np.cumsum(sz, out=off[1:])
    return off.astype(np.uint32)


def _expected(batch, qcoeff, off):
    O = oracle()
    out = np.full(int(off[-1]) + 2048, GUARD, np.uint8)
    for i, b in enumerate(batch.blocks):
        s = int(b["tx_size"])
        w, h = min(int(wl.TX_W[s]), 32), min(int(wl.TX_H[s]), 32)
        q = np.ascontiguousarray(qcoeff[int(b["coeff_off"]):int(b["coeff_off"]) + w * h])
        buf = np.zeros(_size(s), np.uint8)
        O.svto_txb_init_levels(P(q), w, h, P(buf))
        out[int(off[i]):int(off[i]) + _size(s)] = buf
    return out


def test_oracle_matches_reference_c_and_avx2():
    R = ref(1)
    if R is None:
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    O = oracle()
    rng = np.random.default_rng(3)
    for (w, h) in [(4, 4), (8, 8), (16, 16), (32, 32), (4, 8), (8, 4), (16, 4), (4, 16), (32, 8), (8, 32), (16, 32), (32, 16)]:
        for amp in (2, 200, 30000, 70000):
            q = rng.integers(-amp, amp + 1, w * h).astype(np.int32)
            q[rng.random(w * h) < 0.5] = 0
            n = (w + 4) * (h + 6) + 16
            a, b, c = (np.full(n + 64, 0x55, np.uint8) for _ in range(3))
            O.svto_txb_init_levels(P(q), w, h, P(a))
            R.ref_txb_init_levels(P(q), w, h, P(b), 0)
            R.ref_txb_init_levels(P(q), w, h, P(c), 1)
            assert np.array_equal(a[:n], b[:n]), (w, h, amp, "C")
            # the AVX2 twin leaves the TX_PAD_END = 16 bytes after the bottom rows untouched; the
            # reference's own test compares the (h + TX_PAD_VER) rows only (test/EncodeTxbAsmTest.cc:97)
            # (and packs to int16 first, so it is only defined for |qcoeff| < 2^15 -- every level
            # the quantiser can produce)
            if amp < 32768:
                assert np.array_equal(a[:n - 16], c[:n - 16]), (w, h, amp, "AVX2")
                assert (c[n - 16:] == 0x55).all()
            assert (a[n:] == 0x55).all()


@pytest.mark.parametrize("bd", [8, 10])
def test_levels_from_the_kernel_code_on_the_host(bd):
    tables, scans = _tables(bd), load_scans()
    batch = wl.make_packed_batch(wl.valid_pairs(), bd, 2, seed=17 + bd, qindex=(0, 30, 200))
    off = _offsets(batch)
    want = run_oracle(batch, tables, scans)
    got = run_emul(batch, tables, levels_off=off)
    for k in ("qcoeff", "dqcoeff", "eob", "recon"):
        assert np.array_equal(got[k], want[k]), k
    exp = _expected(batch, want["qcoeff"], off)
    assert np.array_equal(got["levels"], exp), np.nonzero(got["levels"] != exp)[0][:8]


def test_levels_with_qmatrix_on_the_host():
    tables, scans = _tables(10), load_scans()
    qm = qm_pack(load_qmatrices(), 6)
    pairs = [(s, t) for (s, t) in wl.valid_pairs() if t < 9]
    batch = wl.make_packed_batch(pairs, 10, 1, seed=5, qindex=(100,), flags=wl.BLK_USE_QM)
    off = _offsets(batch)
    want = run_oracle(batch, tables, scans, qm=qm)
    got = run_emul(batch, tables, qm=qm, levels_off=off)
    assert np.array_equal(got["qcoeff"], want["qcoeff"])
    assert np.array_equal(got["levels"], _expected(batch, want["qcoeff"], off))


@pytest.mark.gpu
@pytest.mark.parametrize("bd", [8, 10])
def test_levels_on_the_gpu(bd):
    import torch
    tables, scans = _tables(bd), load_scans()
    g = pkg.SvtGpu(0)
    g.upload_quant_tables(tables)
    for batch in (wl.make_packed_batch(wl.valid_pairs(), bd, 5, seed=27 + bd, qindex=(0, 30, 200)),
                  wl.make_frame_batch(320, 192, 2, bd, seed=9)):
        off = _offsets(batch)
        want = run_oracle(batch, tables, scans, threads=8)
        exp = _expected(batch, want["qcoeff"], off)
        dev = torch.device("cuda", 0)
        st = torch.cuda.Stream(dev)
        pdt = np.uint8 if bd == 8 else np.int16
        with torch.cuda.stream(st):
            d = dict(residual=torch.from_numpy(batch.residual).to(dev),
                     pred=torch.from_numpy(batch.pred.view(pdt)).to(dev))
            d["recon"] = d["pred"].clone()
            d["qcoeff"] = torch.zeros(batch.n_coeff, dtype=torch.int32, device=dev)
            d["dqcoeff"] = torch.zeros(batch.n_coeff, dtype=torch.int32, device=dev)
            d["eob"] = torch.zeros(batch.n_blocks, dtype=torch.int16, device=dev)
            d["energy"] = torch.zeros(batch.n_blocks, dtype=torch.int64, device=dev)
            d["levels"] = torch.full((len(exp),), GUARD, dtype=torch.uint8, device=dev)
            d["levels_off"] = torch.from_numpy(off.view(np.int32).copy()).to(dev)
            plan = g.plan(batch.blocks)
            g.run(plan, pkg.api.STAGE_ALL, st.cuda_stream, **d)
        st.synchronize()
        assert np.array_equal(d["qcoeff"].cpu().numpy(), want["qcoeff"])
        assert np.array_equal(d["eob"].cpu().numpy().view(np.uint16), want["eob"])
        got = d["levels"].cpu().numpy()
        assert np.array_equal(got, exp), np.nonzero(got != exp)[0][:8]
        plan.close()
    assert pkg.api.levels_size(2) == 20 * 22 + 16 and pkg.api.levels_size(4) == 36 * 38 + 16
