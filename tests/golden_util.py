"""Loaders for the committed fixtures under tests/golden/ (generated from the reference build by
tests/golden/make_golden.py).  Test infrastructure only."""
import os

import numpy as np

GDIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
_cache = {}


def _npz(name):
    if name not in _cache:
        _cache[name] = np.load(os.path.join(GDIR, name))
    return _cache[name]


def load_quant_tables():
    """{(bd, plane): int16[256][5][8]} = zbin, round, quant, quant_shift, dequant rows of
    av1_build_quantizer (all delta_q = 0)."""
    z = _npz("quant_tables.npz")
    return {(bd, pl): z[f"bd{bd}_p{pl}"] for bd in (8, 10) for pl in (0, 1, 2)}


def load_scans():
    """{(tx_size, scan_class): (scan int16[n], iscan int16[n])} from av1_scan_orders."""
    z = _npz("scans.npz")
    return {(s, c): (z[f"scan_{s}_{c}"], z[f"iscan_{s}_{c}"]) for s in range(19) for c in range(3)}


def load_qmatrices():
    """{(level, plane>=1, tx_size): (wt uint8[n], iwt uint8[n])} for the levels in the fixture."""
    z = _npz("qmatrix.npz")
    out = {}
    for k in z.files:
        if k.startswith("wt_"):
            _, lv, pl, s = k.split("_")
            out[(int(lv), int(pl), int(s))] = (z[k], z["i" + k])
    return out


def load_tx_search_masks():
    """uint32[19][2][2][2]: bit t = ProductFullLoopTxSearch tries tx type t, indexed
    [tx_size][is_inter][reduced_tx_set_used][tx_search_reduced_set] (from the reference build)."""
    return _npz("tx_search_masks.npz")["masks"]


def load_fused_vectors():
    return _npz("fused_vectors.npz")


def build_tables(qt):
    """[3*256][5][8] int16: rows plane*256 + qindex, from the golden quantizer tables of one bit
    depth ({plane: [256][5][8]})."""
    return np.ascontiguousarray(np.concatenate([qt[0], qt[1], qt[2]], axis=0))


def qm_pack(qms, level):
    """Concatenate the qmatrices of one level into (wt, iwt, off[2][19])."""
    wt, iwt, off = [], [], np.zeros((2, 19), np.int32)
    pos = 0
    for pl in (0, 1):
        for s in range(19):
            w, iw = qms[(level, pl, s)]
            off[pl, s] = pos
            wt.append(w)
            iwt.append(iw)
            pos += len(w)
    return np.concatenate(wt), np.concatenate(iwt), off
