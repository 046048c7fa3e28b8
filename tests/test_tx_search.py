"""batched tx-type search (ProductFullLoopTxSearch, EbFullLoop.c:927-1100).

The reference tries the allowed tx types of a block one after another; the product expands a block
list into one candidate per (block, tried type) and runs them all through one pass of the fused
path.  Pins: the set of tried types against the reference's own tables (golden fixture generated
from the reference build, and the reference itself where it is compiled); the per-candidate
{eob, distortion, qcoeff} against the oracle evaluating every candidate on its own."""
import ctypes

import numpy as np
import pytest

import cidana_svt_av1_b200 as pkg
from cidana_svt_av1_b200 import workloads as wl
from cidana_svt_av1_b200.api import tx_search_candidates, tx_search_mask
from emul_lib import emul, run_emul
from golden_util import load_quant_tables, load_scans, load_tx_search_masks
from oracle_lib import build_tables, ref, run_oracle_md

COMBOS = [(i, r, a) for i in (0, 1) for r in (0, 1) for a in (0, 1)]


def _tables(bd):
    qt = load_quant_tables()
    return build_tables({p: qt[(bd, p)] for p in range(3)})


def test_tried_types_match_the_reference_tables():
    gold = load_tx_search_masks()
    E = emul()
    E.emul_tx_search_mask.restype = ctypes.c_uint32
    for s in range(19):
        for (i, r, a) in COMBOS:
            assert E.emul_tx_search_mask(s, i, r, a) == gold[s, i, r, a], (s, i, r, a)
            assert tx_search_mask(s, i, r, a) == gold[s, i, r, a], (s, i, r, a)  # through libsvtgpu.so
            m = int(gold[s, i, r, a])
            assert m & 1  # DCT_DCT is always tried
            assert all(wl.is_valid(s, t) for t in range(16) if (m >> t) & 1)
    R = ref(1)
    if R is not None:  # the fixture still says what the reference says
        R.ref_tx_search_mask.restype = ctypes.c_uint32
        for s in range(19):
            for (i, r, a) in COMBOS:
                assert R.ref_tx_search_mask(s, i, r, a) == gold[s, i, r, a]


def _search_batch(bd, per=2, seed=77):
    # one block per size (type is ignored by the search), luma tables, a low qindex so that most
    # candidates keep coefficients and a high one where many end with eob == 0
    pairs = [(s, 0) for s in range(19)]
    return wl.make_packed_batch(pairs, bd, per, seed=seed, qindex=(30, 200)).with_source()


def _check_candidates(batch, cand, inter, reduced, shortlist):
    gold = load_tx_search_masks()
    want = []
    for i, b in enumerate(batch.blocks):
        m = int(gold[b["tx_size"], inter, reduced, shortlist])
        want += [(i, t) for t in range(16) if (m >> t) & 1]
    got = list(zip(cand.cand_block.tolist(), cand.blocks["tx_type"].tolist()))
    assert got == want
    for k, (i, t) in enumerate(got):
        for f in ("tx_size", "bd", "plane", "res_off", "res_stride", "rec_off", "rec_stride", "qrow", "flags"):
            assert cand.blocks[f][k] == batch.blocks[f][i], f


@pytest.mark.parametrize("bd", [8, 10])
def test_candidates_through_the_kernel_code_on_the_host(bd):
    E = emul()
    scans, tables = load_scans(), _tables(bd)
    batch = _search_batch(bd)
    for (inter, reduced, shortlist) in [(0, 0, 0), (1, 0, 0), (1, 0, 1), (0, 1, 0)]:
        cand = tx_search_candidates(batch, inter, reduced, shortlist, want_coeff=True,
                                    expand=E.emul_tx_search_expand)
        _check_candidates(batch, cand, inter, reduced, shortlist)
        want = run_oracle_md(cand, tables, scans, inverse=False)
        got = run_emul(cand, tables, stages=3 | 16, dist=True)
        for k in ("eob", "dist", "qcoeff", "dqcoeff", "energy"):
            assert np.array_equal(got[k], want[k]), (k, inter, reduced, shortlist)


@pytest.mark.gpu
@pytest.mark.parametrize("bd", [8, 10])
def test_batched_search_on_the_gpu(bd):
    scans, tables = load_scans(), _tables(bd)
    g = pkg.SvtGpu(0)
    g.upload_quant_tables(tables)
    batch = _search_batch(bd, per=6, seed=78)
    for (inter, reduced, shortlist) in [(0, 0, 0), (1, 0, 0), (1, 0, 1)]:
        cand = tx_search_candidates(batch, inter, reduced, shortlist, want_coeff=True)
        _check_candidates(batch, cand, inter, reduced, shortlist)
        want = run_oracle_md(cand, tables, scans, threads=8, inverse=False)
        got = g.tx_search(batch, inter, reduced, shortlist, want_coeff=True)
        assert np.array_equal(got["cand_block"], cand.cand_block)
        for k in ("eob", "dist", "qcoeff"):
            assert np.array_equal(got[k], want[k]), (k, inter, shortlist)
        lean = g.tx_search(batch, inter, reduced, shortlist)  # what mode decision needs: no coefficients
        assert np.array_equal(lean["eob"], want["eob"]) and np.array_equal(lean["dist"], want["dist"])
        # the candidate that repeats a block's own type is that block's plain result
        plain = run_oracle_md(batch, tables, scans)
        own = lean["tx_type"] == batch.blocks["tx_type"][lean["cand_block"]]
        assert np.array_equal(lean["eob"][own], plain["eob"][lean["cand_block"][own]])
    # a frame of mixed sizes given as a residual (no source plane)
    frame = wl.make_frame_batch(256, 128, 1, bd, seed=5)
    cand = tx_search_candidates(frame, 1, 0, 0, want_coeff=True)
    src_frame = frame.with_source()
    cand_src = tx_search_candidates(src_frame, 1, 0, 0, want_coeff=True)
    want = run_oracle_md(cand_src, tables, scans, threads=8, inverse=False)
    got = g.tx_search(src_frame, 1, 0, 0)
    assert np.array_equal(got["eob"], want["eob"]) and np.array_equal(got["dist"], want["dist"])
