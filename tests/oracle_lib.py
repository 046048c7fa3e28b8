"""ctypes bindings for the CHECKERS (test infrastructure only):

* ``oracle()``  -> oracle/libsvt_oracle.so, our plain-C restatement (always buildable: gcc only)
* ``ref()``     -> oracle/_ref/libsvtref.so, the reference itself compiled from /root/reference
                   (present in the build container and, as a prebuilt file, on the GPU box;
                   ``None`` when it was never built)

Nothing in the product package imports this module.
"""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ODIR = os.path.join(ROOT, "oracle")

TX_W = [4, 8, 16, 32, 64, 4, 8, 8, 16, 16, 32, 32, 64, 4, 16, 8, 32, 16, 64]
TX_H = [4, 8, 16, 32, 64, 8, 4, 16, 8, 32, 16, 64, 32, 16, 4, 32, 8, 64, 16]
TX_NAMES = ["4X4", "8X8", "16X16", "32X32", "64X64", "4X8", "8X4", "8X16", "16X8", "16X32",
            "32X16", "32X64", "64X32", "4X16", "16X4", "8X32", "32X8", "16X64", "64X16"]


def P(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


class SvtoBlock(ctypes.Structure):
    _fields_ = [("tx_size", ctypes.c_uint8), ("tx_type", ctypes.c_uint8), ("bd", ctypes.c_uint8),
                ("plane", ctypes.c_uint8), ("res_off", ctypes.c_uint32),
                ("res_stride", ctypes.c_uint32), ("rec_off", ctypes.c_uint32),
                ("rec_stride", ctypes.c_uint32), ("coeff_off", ctypes.c_uint32),
                ("qrow", ctypes.c_uint32), ("flags", ctypes.c_uint32)]


BLOCK_DTYPE = np.dtype([("tx_size", "u1"), ("tx_type", "u1"), ("bd", "u1"), ("plane", "u1"),
                        ("res_off", "<u4"), ("res_stride", "<u4"), ("rec_off", "<u4"),
                        ("rec_stride", "<u4"), ("coeff_off", "<u4"), ("qrow", "<u4"), ("flags", "<u4")])
assert BLOCK_DTYPE.itemsize == ctypes.sizeof(SvtoBlock) == 32

_oracle = None
_ref = None


def oracle():
    global _oracle
    if _oracle is None:
        so = os.path.join(ODIR, "libsvt_oracle.so")
        src = os.path.join(ODIR, "svt_oracle.c")
        if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
            subprocess.check_call(["make", "-s", "-C", ODIR, "oracle"])
        _oracle = ctypes.CDLL(so)
        _oracle.svto_cospi.restype = ctypes.POINTER(ctypes.c_int32)
        _oracle.svto_sinpi.restype = ctypes.POINTER(ctypes.c_int32)
    return _oracle


def ref(asm=1):
    """The compiled reference, or None.  asm=1: the AVX2 RTCD table (what the encoder runs)."""
    global _ref
    if _ref is None:
        so = os.path.join(ODIR, "_ref", "libsvtref.so")
        if not os.path.exists(so):
            return None
        _ref = ctypes.CDLL(so)
    _ref.ref_init(int(asm))
    return _ref


def valid_pairs():
    o = oracle()
    return [(s, t) for s in range(19) for t in range(16) if o.svto_is_valid(s, t)]


def max_eob(s):
    return min(TX_W[s], 32) * min(TX_H[s], 32)


def tx_scale(s):
    pels = TX_W[s] * TX_H[s]
    return int(pels > 256) + int(pels > 1024)


def scan_class(t):
    return 0 if t < 10 else (2 if t & 1 else 1)


def aligned(shape, dtype, fill=0, align=64):
    """numpy array whose data pointer is `align`-byte aligned (the reference's AVX2 kernels use
    aligned 256-bit stores on their coefficient buffers)."""
    dtype = np.dtype(dtype)
    n = int(np.prod(shape))
    raw = np.zeros(n * dtype.itemsize + align, np.uint8)
    off = (-raw.ctypes.data) % align
    a = raw[off:off + n * dtype.itemsize].view(dtype).reshape(shape)
    a[...] = fill
    return a


def aligned_copy(x, align=64):
    a = aligned(x.shape, x.dtype, align=align)
    a[...] = x
    return a


# ---- whole-batch runners (the checker side of the parity tests and of bench.py) -----------------
from golden_util import build_tables, qm_pack  # noqa: E402,F401  (fixture helpers)


def _slices(n, threads):
    step = (n + threads - 1) // threads
    return [(i, min(n, i + step)) for i in range(0, n, step)]


def run_oracle(batch, tables, scans, clamp16=0, qm=None, threads=1, first=0, last=None):
    """oracle/svt_oracle.c over batch.blocks[first:last] -> dict of outputs (full-size arrays)."""
    import threading
    O = oracle()
    n = batch.n_blocks
    last = n if last is None else last
    q = np.zeros(batch.n_coeff, np.int32)
    dq = np.zeros(batch.n_coeff, np.int32)
    eob = np.zeros(n, np.uint16)
    energy = np.zeros(n, np.uint64)
    recon = batch.pred.copy()
    ptrs = (ctypes.c_void_p * (19 * 3))()
    keep = []
    for s in range(19):
        for c in range(3):
            a = np.ascontiguousarray(scans[(s, c)][0])
            keep.append(a)
            ptrs[s * 3 + c] = a.ctypes.data
    wt = iwt = off = None
    if qm is not None:
        wt, iwt, off = qm
    blocks = np.ascontiguousarray(batch.blocks)

    def work(a, b):
        O.svto_fused_blocks(P(blocks), a, b, P(batch.residual), P(recon), P(tables), P(q), P(dq),
                            P(eob), P(energy), ptrs, int(clamp16), P(wt), P(iwt), P(off))
    sl = [(first + a, first + b) for a, b in _slices(last - first, threads)]
    if threads == 1:
        work(*sl[0])
    else:
        th = [threading.Thread(target=work, args=x) for x in sl]
        [t.start() for t in th]
        [t.join() for t in th]
    return dict(qcoeff=q, dqcoeff=dq, eob=eob, energy=energy, recon=recon)


def run_oracle_md(batch, tables, scans, clamp16=0, threads=1, inverse=True):
    """svto_fused_blocks_md: residual from batch.src - batch.pred, plus the distortion pairs.
    inverse=False: svto_tx_search_blocks (candidates of a tx-type search: no reconstruction)."""
    import threading
    O = oracle()
    n = batch.n_blocks
    q = np.zeros(batch.n_coeff, np.int32)
    dq = np.zeros(batch.n_coeff, np.int32)
    eob = np.zeros(n, np.uint16)
    energy = np.zeros(n, np.uint64)
    dist = np.zeros((n, 2), np.uint64)
    recon = batch.pred.copy()
    ptrs = (ctypes.c_void_p * (19 * 3))()
    keep = []
    for s in range(19):
        for c in range(3):
            a = np.ascontiguousarray(scans[(s, c)][0])
            keep.append(a)
            ptrs[s * 3 + c] = a.ctypes.data
    blocks = np.ascontiguousarray(batch.blocks)

    def work(a, b):
        fn = O.svto_fused_blocks_md if inverse else O.svto_tx_search_blocks
        fn(P(blocks), a, b, P(batch.src), P(recon), P(tables), P(q), P(dq), P(eob),
           P(energy), P(dist), ptrs, int(clamp16))
    sl = _slices(n, threads)
    th = [threading.Thread(target=work, args=x) for x in sl]
    [t.start() for t in th]
    [t.join() for t in th]
    return dict(qcoeff=q, dqcoeff=dq, eob=eob, energy=energy, recon=recon, dist=dist)


def _pool(R, threads):
    R.ref_pool_start.restype = ctypes.c_int
    R.ref_pool_run.restype = ctypes.c_double
    return R.ref_pool_start(int(threads))


def run_ref(batch, tables, threads=1, first=0, last=None, asm=1, qm=None):
    """The compiled reference (AVX2 RTCD table) over the same arrays, on the harness's persistent
    worker pool.  qm = (wt, iwt, off): 10-bit blocks flagged USE_QM use highbd_quantize_b_helper_c
    with the matrices (what the reference's facade does when both pointers are set).  No clamp16."""
    R = ref(asm)
    n = batch.n_blocks
    last = n if last is None else last
    q = aligned(batch.n_coeff, np.int32)
    dq = aligned(batch.n_coeff, np.int32)
    eob = np.zeros(n, np.uint16)
    energy = np.zeros(n, np.uint64)
    recon = aligned_copy(batch.pred)
    residual = aligned_copy(batch.residual)
    blocks = np.ascontiguousarray(batch.blocks[first:last])
    wt, iwt, off = (np.ascontiguousarray(x) for x in qm) if qm is not None else (None, None, None)
    _pool(R, threads)
    R.ref_pool_run(P(blocks), len(blocks), P(residual), P(recon), P(tables), P(q), P(dq), P(eob[first:last]),
                   P(energy[first:last]), P(wt), P(iwt), P(off))
    return dict(qcoeff=q, dqcoeff=dq, eob=eob, energy=energy, recon=recon)


class CpuRunner:
    """Pre-allocated, repeatable CPU pass over a batch for timing (bench.py cpu_baseline and
    --impl reference).  kind 'reference' = oracle/_ref: the reference's AVX2 RTCD path (plus its
    highbd_quantize_b_helper_c for qmatrix blocks) on the harness's persistent pthread pool, which
    also takes the time; 'port' = oracle/svt_oracle.c on Python threads (only when oracle/_ref
    was never built).  recon is updated in place pass after pass (same work)."""

    def __init__(self, batch, tables, scans=None, qm=None, kind=None):
        if kind is None:
            kind = "reference" if ref() is not None else "port"
        self.kind, self.batch, self.qm = kind, batch, qm
        self.tables = np.ascontiguousarray(tables)
        n = batch.n_blocks
        self.q = aligned(batch.n_coeff, np.int32)
        self.dq = aligned(batch.n_coeff, np.int32)
        self.eob = np.zeros(n, np.uint16)
        self.energy = np.zeros(n, np.uint64)
        self.recon = aligned_copy(batch.pred)
        self.residual = aligned_copy(batch.residual)
        self.blocks = np.ascontiguousarray(batch.blocks)
        self.qm_arrays = tuple(np.ascontiguousarray(x) for x in qm) if qm is not None else (None, None, None)
        if kind == "port":
            self.ptrs = (ctypes.c_void_p * (19 * 3))()
            self.keep = []
            for s in range(19):
                for c in range(3):
                    a = np.ascontiguousarray(scans[(s, c)][0])
                    self.keep.append(a)
                    self.ptrs[s * 3 + c] = a.ctypes.data

    def reset(self):
        """Prediction back in `recon`: the next pass is a fresh one (its outputs are the reference
        results for the batch, what the parity gate of bench.py compares with)."""
        self.recon[...] = self.batch.pred

    def outputs(self):
        return dict(qcoeff=self.q, dqcoeff=self.dq, eob=self.eob, energy=self.energy, recon=self.recon)

    def _work(self, a, b):
        wt, iwt, off = self.qm_arrays
        oracle().svto_fused_blocks(P(self.blocks), a, b, P(self.residual), P(self.recon),
                                   P(self.tables), P(self.q), P(self.dq), P(self.eob), P(self.energy),
                                   self.ptrs, 0, P(wt), P(iwt), P(off))

    def run(self, threads, n_blocks=None):
        """One pass over the first n_blocks blocks (default: all); returns seconds."""
        import threading
        import time
        n = len(self.blocks) if n_blocks is None else int(n_blocks)
        if self.kind == "reference":
            R = ref(1)
            _pool(R, threads)
            wt, iwt, off = self.qm_arrays
            return float(R.ref_pool_run(P(self.blocks), n, P(self.residual), P(self.recon), P(self.tables),
                                        P(self.q), P(self.dq), P(self.eob), P(self.energy), P(wt), P(iwt), P(off)))
        # balance by samples, not block count
        area = np.array(TX_W)[self.blocks["tx_size"][:n]] * np.array(TX_H)[self.blocks["tx_size"][:n]]
        cum = np.cumsum(area)
        cuts = [0] + [int(np.searchsorted(cum, cum[-1] * (i + 1) / threads)) + 1 for i in range(threads - 1)]
        cuts.append(n)
        cuts = sorted(set(min(c, n) for c in cuts))
        th = [threading.Thread(target=self._work, args=(cuts[i], cuts[i + 1])) for i in range(len(cuts) - 1)]
        t0 = time.perf_counter()
        [t.start() for t in th]
        [t.join() for t in th]
        return time.perf_counter() - t0
