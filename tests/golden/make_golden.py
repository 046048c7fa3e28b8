#!/usr/bin/env python
"""Regenerates tests/golden/*.npz from the reference compiled from source
(oracle/_ref/libsvtref.so; `make -C oracle ref` needs /root/reference).  Run from the repo root:

    python tests/golden/make_golden.py

Inputs come from numpy's PCG64 with fixed seeds (we never rely on reproducing the reference
tests' std::mt19937 distributions); outputs are whatever the reference's AVX2 RTCD table
produces for them, i.e. exactly what SvtAv1EncApp executes on an AVX2 host.
"""
import ctypes
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle_lib import P, TX_H, TX_W, aligned, aligned_copy, max_eob, ref, valid_pairs  # noqa: E402


def main():
    R = ref(1)
    assert R is not None, "build oracle/_ref first: make -C oracle ref"
    only = sys.argv[1] if len(sys.argv) > 1 else None
    if only == "tx_search":  # add this fixture without touching the others
        R.ref_tx_search_mask.restype = ctypes.c_uint32
        masks = np.array([[[[R.ref_tx_search_mask(s, i, r, a) for a in (0, 1)] for r in (0, 1)] for i in (0, 1)]
                          for s in range(19)], np.uint32)
        np.savez_compressed(os.path.join(HERE, "tx_search_masks.npz"), masks=masks)
        return
    # --- quantizer tables
    qt = {}
    for bd in (8, 10):
        for pl in range(3):
            t = np.zeros((256, 5, 8), np.int16)
            R.ref_build_quant_tables(bd, pl, P(t))
            qt[f"bd{bd}_p{pl}"] = t
    np.savez_compressed(os.path.join(HERE, "quant_tables.npz"), **qt)
    # --- scans
    sc = {}
    for s in range(19):
        for c, t in enumerate((0, 10, 11)):
            a = np.zeros(1024, np.int16)
            b = np.zeros(1024, np.int16)
            n = R.ref_get_scan(s, t, P(a), P(b))
            sc[f"scan_{s}_{c}"] = a[:n].copy()
            sc[f"iscan_{s}_{c}"] = b[:n].copy()
    np.savez_compressed(os.path.join(HERE, "scans.npz"), **sc)
    # --- qmatrices (levels used by tests and the 4K config: aom_get_qmlevel(100,5,9) = 6)
    qm = {}
    for lv in (0, 6, 14):
        for pl in (0, 1):
            for s in range(19):
                wt = np.zeros(1024, np.uint8)
                iwt = np.zeros(1024, np.uint8)
                n = R.ref_get_qmatrix(lv, pl, s, P(wt), P(iwt))
                qm[f"wt_{lv}_{pl}_{s}"] = wt[:n].copy()
                qm[f"iwt_{lv}_{pl}_{s}"] = iwt[:n].copy()
    np.savez_compressed(os.path.join(HERE, "qmatrix.npz"), **qm)
    # --- fused known-answer vectors: one block per valid (size,type) per bit depth
    rng = np.random.default_rng(20240)
    fv = {}
    for bd in (8, 10):
        m = (1 << bd) - 1
        tabs = qt[f"bd{bd}_p0"]
        for s, t in valid_pairs():
            w, h = TX_W[s], TX_H[s]
            n = max_eob(s)
            amp = [m, m // 4, m // 16][(s + t) % 3]
            res = aligned_copy(rng.integers(-amp, amp + 1, (h, w)).astype(np.int16))
            pred = aligned_copy(rng.integers(0, m + 1, (h, w)).astype(np.uint8 if bd == 8 else np.uint16))
            qi = [30, 120, 200][(s * 16 + t) % 3]
            tb = tabs[qi]
            cf = aligned(w * h, np.int32)
            en = ctypes.c_uint64()
            R.ref_estimate_transform(P(res), w, P(cf), s, t, bd, ctypes.byref(en))
            q = aligned(n, np.int32)
            dq = aligned(n, np.int32)
            e = ctypes.c_uint16()
            R.ref_quantize(P(cf), n, P(tb[0]), P(tb[1]), P(tb[2]), P(tb[3]), P(tb[4]), P(q), P(dq),
                           ctypes.byref(e), s, t, bd, 0, None, None)
            rec = aligned_copy(pred)
            if e.value:
                R.ref_inv_transform_recon(P(aligned_copy(dq)), P(rec), w, s, t, bd, e.value)
            k = f"{bd}_{s}_{t}"
            fv["res_" + k] = res
            fv["pred_" + k] = pred
            fv["coeff_" + k] = cf[:n].copy()
            fv["q_" + k] = q
            fv["dq_" + k] = dq
            fv["rec_" + k] = rec
            fv["meta_" + k] = np.array([qi, e.value, en.value & 0xFFFFFFFF, en.value >> 32], np.int64)
    np.savez_compressed(os.path.join(HERE, "fused_vectors.npz"), **fv)
    # --- tx types tried by ProductFullLoopTxSearch, [size][inter][reduced set][shortlist]
    R.ref_tx_search_mask.restype = ctypes.c_uint32
    masks = np.array([[[[R.ref_tx_search_mask(s, i, r, a) for a in (0, 1)] for r in (0, 1)] for i in (0, 1)]
                      for s in range(19)], np.uint32)
    np.savez_compressed(os.path.join(HERE, "tx_search_masks.npz"), masks=masks)
    for f in ("quant_tables.npz", "scans.npz", "qmatrix.npz", "fused_vectors.npz", "tx_search_masks.npz"):
        print(f, os.path.getsize(os.path.join(HERE, f)))


if __name__ == "__main__":
    main()
