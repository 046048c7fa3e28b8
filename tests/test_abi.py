"""CPU tier: libsvtgpu.so loads, exports every symbol include/svt_gpu.h declares, and fails loudly
(no fallback) when there is no GPU."""
import ctypes
import os
import re

import pytest

import cidana_svt_av1_b200 as pkg
from cidana_svt_av1_b200 import build as svt_build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    h = open(os.path.join(ROOT, "include", "svt_gpu.h")).read()
    h = re.sub(r"/\*.*?\*/", "", h, flags=re.S)
    names = set(re.findall(r"\b(svtgpu_[a-z0-9_]+)\s*\(", h))
    # macro-generated declarations: NAME_DECL(4x4) ... -> svtgpu_av1_{fwd_txfm2d,inv_txfm2d_add}_4x4
    for macro, prefix in (("SVTGPU_FWD_DECL", "svtgpu_av1_fwd_txfm2d_"),
                          ("SVTGPU_INV_SQ_DECL", "svtgpu_av1_inv_txfm2d_add_"),
                          ("SVTGPU_INV_RECT_DECL", "svtgpu_av1_inv_txfm2d_add_"),
                          ("SVTGPU_INV_4N_DECL", "svtgpu_av1_inv_txfm2d_add_")):
        for size in re.findall(macro + r"\((\d+x\d+)\)", h):
            names.add(prefix + size)
    names.update(re.findall(r"SVTGPU_QUANT_DECL\((svtgpu_[a-z0-9_]+)\)", h))
    return sorted(n for n in names if not n.endswith("_"))


def test_library_is_built_in_tree():
    svt_build.build_library()
    assert os.path.exists(svt_build.LIB)


def test_every_declared_symbol_is_exported():
    L = pkg.load_library()
    syms = declared_symbols()
    assert len(syms) >= 19 + 6 + 19 + 1 + 4 + 15
    missing = [s for s in syms if not hasattr(L, s)]
    assert not missing, missing


def test_rtcd_table_lists_all_46_slots_minus_none():
    L = pkg.load_library()

    class Entry(ctypes.Structure):
        _fields_ = [("slot", ctypes.c_char_p), ("fn", ctypes.c_void_p)]
    p = ctypes.POINTER(Entry)()
    n = L.svtgpu_rtcd_table(ctypes.byref(p))
    slots = [p[i].slot.decode() for i in range(n)]
    assert n == 45 and len(set(slots)) == 45  # 19 fwd + 6 quant + 19 inv + av1_inv_txfm_add
    for i in range(n):
        assert p[i].fn == ctypes.cast(getattr(L, "svtgpu_" + slots[i]), ctypes.c_void_p).value


def test_generated_iscan_via_abi():
    import numpy as np
    from golden_util import load_scans
    L = pkg.load_library()
    scans = load_scans()
    for s in range(19):
        for c in range(3):
            ref = scans[(s, c)][1]
            out = np.zeros(len(ref), np.int16)
            assert L.svtgpu_get_iscan(s, c, out.ctypes.data_as(ctypes.c_void_p)) == len(ref)
            assert np.array_equal(out, ref)


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(pkg.SvtGpuError):
        pkg.SvtGpu(0)


def test_product_never_touches_the_oracle():
    """The product package and its C++ sources must not reference oracle/ in any way."""
    bad = []
    pdir = os.path.join(ROOT, "cidana_svt_av1_b200")
    for dp, _, files in os.walk(pdir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f)).read()
                if re.search(r"svt_oracle|libsvtref|oracle_lib|svto_|ref_harness", txt):
                    bad.append(f)
    assert not bad, bad
