"""The host-buffer pipeline (svtgpu_submit_host / svtgpu_wait_host) and the error paths of the batch
API: robustness items of the round-1 review.  GPU tier; the checker is the compiled reference
(oracle/_ref) where it exists on the box, else the oracle restatement."""
import ctypes

import numpy as np
import pytest

import cidana_svt_av1_b200 as pkg
from cidana_svt_av1_b200 import workloads as wl
from cidana_svt_av1_b200.api import STAGE_ALL, HostBuffers
from golden_util import load_qmatrices, load_quant_tables, load_scans
from oracle_lib import build_tables, qm_pack, ref, run_oracle, run_ref

pytestmark = pytest.mark.gpu
KEYS = ("qcoeff", "dqcoeff", "eob", "energy", "recon")


def tables(bd):
    qt = load_quant_tables()
    return build_tables({p: qt[(bd, p)] for p in range(3)})


def checker(batch, qm=None):
    if ref() is not None:
        return run_ref(batch, tables(batch.bd), threads=8, qm=qm)
    return run_oracle(batch, tables(batch.bd), load_scans(), qm=qm, threads=8)


def fresh(bd=8):
    g = pkg.SvtGpu(0)
    g.upload_quant_tables(tables(bd))
    return g


def test_plan_reuse_equals_per_submission_bucketing():
    batch = wl.make_frame_batch(512, 256, 5, 8, seed=3)
    want = checker(batch)
    g = fresh()
    plans = g.frame_plans(batch)
    for rep in range(3):  # every picture meets every lane
        bufs = HostBuffers(batch, pinned=True)
        out = g.run_host(batch, bufs=bufs, plans=plans)
        for k in KEYS:
            assert np.array_equal(out[k], want[k]), (rep, k)
    out = g.run_host(batch)  # and without plans
    for k in KEYS:
        assert np.array_equal(out[k], want[k]), k
    # a plan of another geometry is refused
    other = g.plan(batch.frames()[0].blocks[:-1])
    fr = batch.frames()[0]
    bufs = HostBuffers(batch, pinned=False)
    with pytest.raises(pkg.SvtGpuError):
        g.submit_host(fr, bufs.frame_out(0), STAGE_ALL, *bufs.frame_in(0), plan=other)
    g.wait_host()


def test_encode_pass_shape_without_dqcoeff():
    """qcoeff + eob + recon only (what Av1EncodeLoop consumes, EbCodingLoop.c:697-731)."""
    batch = wl.make_frame_batch(384, 256, 2, 10, seed=8, qindex=90, laplace_b=30.0)
    want = checker(batch)
    g = fresh(10)
    bufs = HostBuffers(batch, pinned=False)
    bufs.out["dqcoeff"] = None
    out = g.run_host(batch, bufs=bufs)
    for k in ("qcoeff", "eob", "energy", "recon"):
        assert np.array_equal(out[k], want[k]), k


def test_allocation_failure_leaves_the_lane_usable():
    """A failed allocation of the SECOND buffer of a pair (dqcoeff after qcoeff, energy after eob)
    must not leave a capacity that says 'grown' next to a null pointer."""
    L = pkg.load_library()
    batch = wl.make_frame_batch(256, 192, 1, 8, seed=4)
    want = checker(batch)
    for kth in (1, 2, 3, 4, 5, 6, 7):
        g = fresh()
        L.svtgpu_debug_fail_host_alloc(kth)
        failed = 0
        for _ in range(4):  # three lanes + one reuse
            try:
                g.run_host(batch)
            except pkg.SvtGpuError as e:
                failed += 1
                assert "out of device memory" in str(e)
        L.svtgpu_debug_fail_host_alloc(0)
        assert failed == 1, kth
        for _ in range(4):  # every lane again, including the one that failed
            out = g.run_host(batch)
            for k in KEYS:
                assert np.array_equal(out[k], want[k]), (kth, k)
        g.close()


def test_block_beyond_the_host_arrays_is_refused():
    batch = wl.make_frame_batch(256, 128, 1, 8, seed=5)
    g = fresh()
    for field, delta in (("res_off", batch.residual.size), ("rec_off", batch.pred.size), ("coeff_off", batch.n_coeff)):
        bad = wl.Batch(batch.blocks.copy(), batch.residual, batch.pred, batch.n_coeff, 8, "bad")
        bad.blocks[field][7] += delta - (delta % 4)
        with pytest.raises(pkg.SvtGpuError) as e:
            g.run_host(bad)
        assert "beyond" in str(e.value), field
    out = g.run_host(batch)  # the context is still good
    assert np.array_equal(out["eob"], checker(batch)["eob"])


def test_qmatrix_blocks_without_an_upload_are_an_error():
    """A USE_QM block must never be quantised flat in silence."""
    pairs = [(s, 0) for s in range(19)]
    batch = wl.make_packed_batch(pairs, 10, 3, seed=9, qindex=(100,), flags=wl.BLK_USE_QM)
    g = fresh(10)
    with pytest.raises(pkg.SvtGpuError) as e:
        g.run_host(batch)
    assert "qmatrix" in str(e.value)
    with pytest.raises(pkg.SvtGpuError):
        g.run_staged(batch)
    # matrices for a subset of the sizes only: the sizes without one are named
    wt, iwt, off = qm_pack(load_qmatrices(), 6)
    off = off.copy()
    off[0, 2] = -1
    g.upload_qmatrix(wt, iwt, off)
    with pytest.raises(pkg.SvtGpuError) as e:
        g.run_host(batch)
    assert "tx_size 2" in str(e.value)
    g.upload_qmatrix(*qm_pack(load_qmatrices(), 6))
    qm = qm_pack(load_qmatrices(), 6)
    out = g.run_host(batch)
    want = checker(batch, qm=qm)
    for k in KEYS:
        assert np.array_equal(out[k], want[k]), k


def test_negative_quant_shift_is_refused_at_upload():
    """The fused kernels multiply by quant_shift << (16 + log_scale); the upload refuses the one
    thing that breaks it (a negative quant_shift, which av1_build_quantizer never produces)."""
    t = tables(8).copy().reshape(-1, 40)
    g = pkg.SvtGpu(0)
    g.upload_quant_tables(t)
    bad = t.copy()
    bad[5, 25] = -3
    with pytest.raises(pkg.SvtGpuError) as e:
        g.upload_quant_tables(bad)
    assert "quant_shift" in str(e.value)
    batch = wl.make_frame_batch(128, 128, 1, 8, seed=2)
    assert np.array_equal(g.run_host(batch)["eob"], checker(batch)["eob"])  # the earlier tables still serve


def test_multi_device_handle_round_robins_pictures():
    """svtgpu_create_multi through the C ABI: every visible device (the same device twice on a
    one-GPU box, which exercises the same bookkeeping), pictures round-robin, outputs == checker."""
    import torch
    n_dev = torch.cuda.device_count()
    ids = list(range(n_dev)) if n_dev >= 2 else [0, 0]
    batch = wl.make_frame_batch(384, 256, 7, 8, seed=12)
    want = checker(batch)
    m = pkg.SvtGpuMulti(ids)
    m.upload_quant_tables(tables(8))
    for rep in range(2):
        out = m.run_host(batch)
        for k in KEYS:
            assert np.array_equal(out[k], want[k]), (rep, k)
    m.close()
    with pytest.raises(pkg.SvtGpuError):
        pkg.SvtGpuMulti([0, 99])
