"""Host-side mirror of the reference interface over libsvtgpu.so (ctypes; see include/svt_gpu.h).

Level C (batch): ``SvtGpu`` -- context, table upload, plans, device-resident runs on torch CUDA
tensors, and ``run_host`` (host numpy in / host numpy out, copies included: the e2e call).
Level B: ``av1_estimate_transform``, ``av1_quantize_inv_quantize``, ``av1_inv_transform_recon``,
``av1_inv_transform_recon8bit`` -- same names and argument meaning as the reference's
Source/Lib/Common/Codec/EbTransforms.h:3896-3987, numpy host arrays, synchronous.
"""
import ctypes
import os

import numpy as np

from .workloads import BLOCK_DTYPE, TX_H, TX_W, coded_count

HERE = os.path.dirname(os.path.abspath(__file__))
# SVTGPU_LIB: another build of the same library (A/B experiments, tools/build_variant.sh)
LIB_PATH = os.environ.get("SVTGPU_LIB") or os.path.join(HERE, "libsvtgpu.so")

STAGE_FWD, STAGE_QUANT, STAGE_INV, STAGE_ALL = 1, 2, 4, 7
STAGE_RESIDUAL = 16  # residual = src - pred on the fly


class SvtGpuError(RuntimeError):
    pass


class _Buffers(ctypes.Structure):  # SvtGpuBuffers
    _fields_ = [(n, ctypes.c_void_p) for n in ("residual", "pred", "recon", "coeff_in", "coeff",
                                               "qcoeff", "dqcoeff", "eob", "energy", "src", "dist",
                                               "levels", "levels_off")]


class _HostFrame(ctypes.Structure):  # SvtGpuHostFrame
    _fields_ = [("blocks", ctypes.c_void_p), ("n_blocks", ctypes.c_int), ("bit_depth", ctypes.c_int),
                ("residual", ctypes.c_void_p), ("residual_elems", ctypes.c_size_t),
                ("pred", ctypes.c_void_p), ("recon", ctypes.c_void_p), ("pixel_elems", ctypes.c_size_t),
                ("qcoeff", ctypes.c_void_p), ("dqcoeff", ctypes.c_void_p), ("coeff_elems", ctypes.c_size_t),
                ("eob", ctypes.c_void_p), ("energy", ctypes.c_void_p), ("src", ctypes.c_void_p),
                ("dist", ctypes.c_void_p), ("plan", ctypes.c_void_p)]


_lib = None


def load_library():
    """Loads libsvtgpu.so.  Raises if it was never built: there is no fallback path."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise SvtGpuError(f"{LIB_PATH} is missing: run `python -m cidana_svt_av1_b200.build` "
                              "(or __graft_entry__.build()); this package has no CPU fallback")
        L = ctypes.CDLL(LIB_PATH)
        L.svtgpu_last_error.restype = ctypes.c_char_p
        L.svtgpu_launch_count.restype = ctypes.c_longlong
        L.svtgpu_launch_count.argtypes = [ctypes.c_void_p]
        L.svtgpu_run.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                 ctypes.c_void_p]
        L.svtgpu_plan_create.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                         ctypes.c_void_p]
        L.svtgpu_plan_destroy.argtypes = [ctypes.c_void_p]
        L.svtgpu_destroy.argtypes = [ctypes.c_void_p]
        L.svtgpu_sync.argtypes = [ctypes.c_void_p]
        L.svtgpu_upload_quant_tables.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
        L.svtgpu_upload_qmatrix.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                            ctypes.c_int, ctypes.c_void_p]
        L.svtgpu_submit_host.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
        L.svtgpu_wait_host.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.svtgpu_create_multi.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
        L.svtgpu_multi_destroy.argtypes = [ctypes.c_void_p]
        L.svtgpu_multi_upload_quant_tables.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
        L.svtgpu_multi_submit_host.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
        L.svtgpu_multi_wait_host.argtypes = [ctypes.c_void_p, ctypes.c_longlong]
        L.svtgpu_multi_device_of.argtypes = [ctypes.c_int, ctypes.c_longlong]
        L.svtgpu_multi_context.restype = ctypes.c_void_p
        L.svtgpu_multi_context.argtypes = [ctypes.c_void_p, ctypes.c_int]
        _lib = L
    return _lib


def lib():
    return load_library()


def _P(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


def _check(rc, what):
    if rc != 0:
        raise SvtGpuError(f"{what} failed ({rc}): {lib().svtgpu_last_error().decode()}")


class Plan:
    def __init__(self, owner, handle, n_blocks):
        self.owner, self.handle, self.n_blocks = owner, handle, n_blocks

    def close(self):
        if self.handle:
            lib().svtgpu_plan_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class SvtGpu:
    """One context per GPU (one process per GPU in multi-GPU runs)."""

    def __init__(self, device=0):
        L = load_library()
        h = ctypes.c_void_p()
        _check(L.svtgpu_create(ctypes.byref(h), int(device)), "svtgpu_create")
        self.handle = h
        self.device = int(device)

    def close(self):
        if self.handle:
            lib().svtgpu_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def launches(self):
        return int(lib().svtgpu_launch_count(self.handle))

    def upload_quant_tables(self, tables):
        t = np.ascontiguousarray(tables, np.int16).reshape(-1, 5, 8)
        _check(lib().svtgpu_upload_quant_tables(self.handle, _P(t), t.shape[0]), "upload_quant_tables")

    def upload_qmatrix(self, wt, iwt, off):
        wt = np.ascontiguousarray(wt, np.uint8)
        iwt = np.ascontiguousarray(iwt, np.uint8)
        off = np.ascontiguousarray(off, np.int32).reshape(2, 19)
        _check(lib().svtgpu_upload_qmatrix(self.handle, _P(wt), _P(iwt), wt.size, _P(off)),
               "upload_qmatrix")

    def plan(self, blocks):
        b = np.ascontiguousarray(blocks, BLOCK_DTYPE)
        h = ctypes.c_void_p()
        _check(lib().svtgpu_plan_create(self.handle, _P(b), len(b), ctypes.byref(h)), "plan_create")
        return Plan(self, h, len(b))

    def run(self, plan, stages=STAGE_ALL, stream=None, **bufs):
        """bufs: torch CUDA tensors (or raw device addresses) named as in SvtGpuBuffers."""
        def ptr(x):
            if x is None:
                return None
            return int(x) if isinstance(x, int) else x.data_ptr()
        b = _Buffers(*[ptr(bufs.get(n)) for n, _ in _Buffers._fields_])
        _check(lib().svtgpu_run(self.handle, plan.handle, ctypes.byref(b), int(stages), stream),
               "svtgpu_run")

    def sync(self):
        _check(lib().svtgpu_sync(self.handle), "svtgpu_sync")

    # ---- the end-to-end call: host buffers in, host buffers out -----------------------------------
    def host_buffers(self, batch, pinned=True, dist=False, coeffs=True):
        """Output (and page-locked input) buffers a caller of the host pipeline owns: allocate
        once, reuse for every run over batches of this geometry."""
        return HostBuffers(batch, pinned, dist, coeffs)

    def submit_host(self, frame, out, stages=STAGE_ALL, residual=None, pred=None, src=None, plan=None):
        """Queue one picture (a Batch whose offsets are relative to its own arrays) on the
        pipeline; `out` is a dict of host arrays (qcoeff, dqcoeff, eob, energy, recon and
        optionally dist) sized for this picture; qcoeff / dqcoeff may be None next to dist.
        With STAGE_RESIDUAL the residual is source - prediction (`src`, default frame.src).
        Returns the ticket.  Asynchronous: see svtgpu_submit_host."""
        residual = frame.residual if residual is None else residual
        pred = frame.pred if pred is None else pred
        if stages & STAGE_RESIDUAL:
            src = frame.src if src is None else src

        def opt(name):
            v = out.get(name)
            return _P(v).value if v is not None else None
        hf = _HostFrame(_P(frame.blocks).value, frame.n_blocks, int(frame.bd),
                        _P(residual).value if residual is not None else None,
                        residual.size if residual is not None else 0, _P(pred).value,
                        _P(out["recon"]).value, pred.size, opt("qcoeff"), opt("dqcoeff"), frame.n_coeff,
                        _P(out["eob"]).value, opt("energy"), _P(src).value if src is not None else None,
                        opt("dist"), plan.handle if plan is not None else None)
        t = ctypes.c_int(-1)
        _check(lib().svtgpu_submit_host(self.handle, ctypes.byref(hf), int(stages), ctypes.byref(t)),
               "svtgpu_submit_host")
        return t.value

    def wait_host(self, ticket=-1):
        _check(lib().svtgpu_wait_host(self.handle, int(ticket)), "svtgpu_wait_host")

    def run_host(self, batch, stages=STAGE_ALL, bufs=None, plans=None):
        """Host arrays in, host arrays out, every copy included: each picture of the batch goes
        through svtgpu_submit_host (upload, kernels and download of consecutive pictures overlap),
        then one wait.  Returns a dict of numpy arrays like the oracle runners do.
        plans: one Plan per picture (frame_plans), for callers whose geometry repeats."""
        bufs = bufs or HostBuffers(batch, pinned=False)
        for k, fr in enumerate(batch.frames()):
            res, pred, src = bufs.frame_in(k)
            self.submit_host(fr, bufs.frame_out(k), stages, res, pred, src, plans[k] if plans else None)
        self.wait_host()
        return bufs.out

    def frame_plans(self, batch):
        """One reusable plan per picture of the batch (offsets relative to the picture)."""
        return [self.plan(fr.blocks) for fr in batch.frames()]

    def tx_search(self, batch, is_inter=0, reduced_tx_set_used=0, tx_search_reduced_set=0,
                  want_coeff=False):
        """SURVEY 8f-2, ProductFullLoopTxSearch (EbFullLoop.c:927-1100) for every block of `batch`
        at once: each block is expanded into one candidate per tx type the reference would try
        (svtgpu_tx_search_expand) and all candidates go through one run of forward -> quantise
        with the distortion outputs.  `batch.src` given: residual = src - pred on the fly.
        Returns dict(cand_block, tx_type, eob, dist[, qcoeff, coeff_off]) per candidate; a
        candidate with eob == 0 and tx_type != DCT_DCT is one the reference skips (:1030)."""
        cand = tx_search_candidates(batch, is_inter, reduced_tx_set_used, tx_search_reduced_set, want_coeff)
        stages = STAGE_FWD | STAGE_QUANT | (STAGE_RESIDUAL if batch.src is not None else 0)
        bufs = HostBuffers(cand, pinned=False, dist=True, coeffs=want_coeff)
        out = self.run_host(cand, stages, bufs)
        res = dict(cand_block=cand.cand_block, tx_type=cand.blocks["tx_type"].copy(), eob=out["eob"],
                   dist=out["dist"])
        if want_coeff:
            res.update(qcoeff=out["qcoeff"], coeff_off=cand.blocks["coeff_off"].copy())
        return res

    def run_staged(self, batch, stages=STAGE_ALL, want_coeff=False, coeff_in=None, plan=None):
        """Any stage combination through the batch API (svtgpu_plan_create + svtgpu_run) on
        torch-allocated device buffers, synchronous: upload, run, download everything.  Used
        where the stages are split (forward only, quantiser on given coefficients, ...)."""
        import torch
        dev = torch.device("cuda", self.device)
        own_plan = plan is None
        if own_plan:
            plan = self.plan(batch.blocks)
        # a real (non-legacy) stream: handle 0 would mean "the context's own stream" to the library
        if getattr(self, "_stream", None) is None:
            self._stream = torch.cuda.Stream(dev)
        st = self._stream
        with torch.cuda.device(dev), torch.cuda.stream(st):
            res = torch.from_numpy(batch.residual).to(dev, non_blocking=True)
            pred = torch.from_numpy(batch.pred.view(np.uint8 if batch.bd == 8 else np.int16)).to(dev, non_blocking=True)
            recon = pred.clone()
            n = batch.n_blocks
            q = torch.zeros(batch.n_coeff, dtype=torch.int32, device=dev)
            dq = torch.zeros(batch.n_coeff, dtype=torch.int32, device=dev)
            cf = torch.zeros(batch.n_coeff, dtype=torch.int32, device=dev) if want_coeff else None
            cin = torch.from_numpy(coeff_in).to(dev) if coeff_in is not None else None
            eob = torch.zeros(n, dtype=torch.int16, device=dev)
            energy = torch.zeros(n, dtype=torch.int64, device=dev)
            self.run(plan, stages, st.cuda_stream, residual=res, pred=pred, recon=recon, coeff_in=cin,
                     coeff=cf, qcoeff=q, dqcoeff=dq, eob=eob, energy=energy)
            out = dict(qcoeff=q.cpu().numpy(), dqcoeff=dq.cpu().numpy(),
                       eob=eob.cpu().numpy().view(np.uint16), energy=energy.cpu().numpy().view(np.uint64),
                       recon=recon.cpu().numpy().view(batch.pred.dtype),
                       coeff=cf.cpu().numpy() if want_coeff else None)
        st.synchronize()
        if own_plan:
            plan.close()
        return out


class SvtGpuMulti:
    """Several GPUs from one process (svtgpu_create_multi): pictures round-robin over the devices."""

    def __init__(self, device_ids):
        L = load_library()
        ids = (ctypes.c_int * len(device_ids))(*[int(d) for d in device_ids])
        h = ctypes.c_void_p()
        _check(L.svtgpu_create_multi(ctypes.byref(h), ids, len(device_ids)), "svtgpu_create_multi")
        self.handle, self.n = h, len(device_ids)

    def close(self):
        if self.handle:
            lib().svtgpu_multi_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def upload_quant_tables(self, tables):
        t = np.ascontiguousarray(tables, np.int16).reshape(-1, 5, 8)
        _check(lib().svtgpu_multi_upload_quant_tables(self.handle, _P(t), t.shape[0]), "multi_upload_quant_tables")

    def run_host(self, batch, stages=STAGE_ALL, bufs=None):
        """Every picture of the batch through svtgpu_multi_submit_host, then one wait."""
        L = lib()
        bufs = bufs or HostBuffers(batch, pinned=False)
        for k, fr in enumerate(batch.frames()):
            res, pred, src = bufs.frame_in(k)
            out = bufs.frame_out(k)

            def opt(name):
                v = out.get(name)
                return _P(v).value if v is not None else None
            hf = _HostFrame(_P(fr.blocks).value, fr.n_blocks, int(fr.bd), _P(res).value, res.size, _P(pred).value,
                            _P(out["recon"]).value, pred.size, opt("qcoeff"), opt("dqcoeff"), fr.n_coeff,
                            _P(out["eob"]).value, opt("energy"), None, opt("dist"), None)
            t = ctypes.c_longlong(-1)
            _check(L.svtgpu_multi_submit_host(self.handle, ctypes.byref(hf), int(stages), ctypes.byref(t)),
                   "svtgpu_multi_submit_host")
            assert L.svtgpu_multi_device_of(self.n, ctypes.c_longlong(t.value)) == t.value % self.n
        _check(L.svtgpu_multi_wait_host(self.handle, ctypes.c_longlong(-1)), "svtgpu_multi_wait_host")
        return bufs.out


def levels_size(tx_size):
    """Bytes of one block's padded level map (av1_txb_init_levels layout)."""
    L = lib()
    L.svtgpu_levels_size.restype = ctypes.c_size_t
    return int(L.svtgpu_levels_size(int(tx_size)))


def tx_search_mask(tx_size, is_inter=0, reduced_tx_set_used=0, tx_search_reduced_set=0):
    """Bit t set = ProductFullLoopTxSearch evaluates tx type t for this size (host only)."""
    L = lib()
    L.svtgpu_tx_search_mask.restype = ctypes.c_uint32
    return int(L.svtgpu_tx_search_mask(int(tx_size), int(is_inter), int(reduced_tx_set_used),
                                       int(tx_search_reduced_set)))


def tx_search_candidates(batch, is_inter=0, reduced_tx_set_used=0, tx_search_reduced_set=0,
                         want_coeff=False, expand=None):
    """The candidate batch of a tx-type search: same arrays as `batch`, one block per (block, tried
    type), `cand_block[k]` = the block candidate k belongs to.  `expand` defaults to the
    library's svtgpu_tx_search_expand (tests pass the host-emulation twin)."""
    from .workloads import Batch
    expand = expand or lib().svtgpu_tx_search_expand
    blocks = np.ascontiguousarray(batch.blocks)
    n = len(blocks)
    cand = np.zeros(16 * max(n, 1), BLOCK_DTYPE)
    owner = np.zeros(16 * max(n, 1), np.int32)
    elems = ctypes.c_int64(0)
    k = expand(_P(blocks), n, int(is_inter), int(reduced_tx_set_used), int(tx_search_reduced_set),
               int(bool(want_coeff)), _P(cand), _P(owner), ctypes.byref(elems))
    if k < 0:
        raise SvtGpuError(f"svtgpu_tx_search_expand failed ({k})")
    c = Batch(cand[:k].copy(), batch.residual, batch.pred, max(int(elems.value), 4), batch.bd,
              batch.name + "_txsearch")
    c.src = batch.src
    c.cand_block = owner[:k].copy()
    return c


class HostBuffers:
    """Host-side arrays of one batch geometry: outputs always, and page-locked copies of the
    inputs when `pinned` (what a caller that wants full PCIe speed would hold).
    dist=True adds the per-block distortion pairs; coeffs=False leaves out qcoeff / dqcoeff
    (allowed only with dist: the mode-decision shape of the call)."""

    def __init__(self, batch, pinned=True, dist=False, coeffs=True):
        self.batch = batch
        self.pinned = pinned
        n, nc = batch.n_blocks, batch.n_coeff
        pdt = batch.pred.dtype
        self._keep = []

        def alloc(count, dtype):
            if not pinned:
                return np.zeros(count, dtype)
            import torch
            t = torch.zeros(max(count, 1) * np.dtype(dtype).itemsize, dtype=torch.uint8).pin_memory()
            self._keep.append(t)
            return t.numpy().view(dtype)[:count]

        def held(a):
            if a is None or not pinned:
                return a
            c = alloc(a.size, a.dtype)
            c[:] = a
            return c

        self.out = dict(qcoeff=alloc(nc, np.int32) if coeffs else None,
                        dqcoeff=alloc(nc, np.int32) if coeffs else None, eob=alloc(n, np.uint16),
                        energy=alloc(n, np.uint64), recon=alloc(batch.pred.size, pdt), coeff=None,
                        dist=alloc(2 * n, np.uint64).reshape(n, 2) if dist else None)
        self.residual, self.pred, self.src = held(batch.residual), held(batch.pred), held(batch.src)
        self._frames = batch.frames()

    def _whole(self, f):
        return len(self._frames) == 1 and f is self.batch

    def frame_in(self, k):
        f = self._frames[k]
        if self._whole(f):
            return self.residual, self.pred, self.src
        e0, e1 = f.elem0, f.elem0 + f.pred.size
        return self.residual[e0:e1], self.pred[e0:e1], (self.src[e0:e1] if self.src is not None else None)

    def frame_out(self, k):
        f = self._frames[k]
        if self._whole(f):
            return self.out
        o = self.out
        b0, c0, e0 = f.block0, f.coeff0, f.elem0

        def cut(name, a, b):
            return o[name][a:b] if o.get(name) is not None else None
        return dict(qcoeff=cut("qcoeff", c0, c0 + f.n_coeff), dqcoeff=cut("dqcoeff", c0, c0 + f.n_coeff),
                    eob=o["eob"][b0:b0 + f.n_blocks], energy=o["energy"][b0:b0 + f.n_blocks],
                    recon=o["recon"][e0:e0 + f.pred.size], dist=cut("dist", b0, b0 + f.n_blocks))


# ---- Level B mirrors: the reference's entry points on host numpy arrays ----------------------------
def av1_estimate_transform(residual_buffer, residual_stride, transform_size, transform_type,
                           bit_increment=0):
    """av1_estimate_transform (EbTransforms.c:4918).  residual_buffer: int16 array holding the
    block at `residual_stride`.  Returns (coeff int32[max_eob], three_quad_energy)."""
    L = lib()
    res = np.ascontiguousarray(residual_buffer, np.int16)
    coeff = np.zeros(TX_W[transform_size] * TX_H[transform_size], np.int32)
    energy = ctypes.c_uint64(0)
    rc = L.svtgpu_av1_estimate_transform(_P(res), ctypes.c_uint32(residual_stride), _P(coeff),
                                         ctypes.c_uint32(0), ctypes.c_uint8(transform_size),
                                         ctypes.byref(energy), None, ctypes.c_uint32(bit_increment),
                                         ctypes.c_uint8(transform_type), 0, 0, 0)
    _check(rc, "av1_estimate_transform")
    return coeff[:coded_count(transform_size)].copy(), int(energy.value)


def av1_quantize_inv_quantize(coeff, table_row, txsize, tx_type, bit_increment=0, clamp16=False):
    """av1_quantize_inv_quantize (EbFullLoop.c:653) with the quantizer table row
    (int16[5][8]: zbin, round, quant, quant_shift, dequant) that the reference would pick from the
    PictureControlSet.  Returns (quant_coeff, recon_coeff, eob)."""
    L = lib()
    cf = np.ascontiguousarray(coeff, np.int32)
    tb = np.ascontiguousarray(table_row, np.int16).reshape(5, 8)
    n = coded_count(txsize)
    q = np.zeros(n, np.int32)
    dq = np.zeros(n, np.int32)
    eob = ctypes.c_uint16(0)
    cnt = ctypes.c_uint32(0)
    L.svtgpu_av1_quantize_inv_quantize(_P(tb[0]), _P(tb[1]), _P(tb[2]), _P(tb[3]), _P(tb[4]), _P(cf),
                                       _P(q), _P(dq), ctypes.c_uint8(txsize), ctypes.byref(eob),
                                       ctypes.byref(cnt), ctypes.c_uint32(bit_increment),
                                       ctypes.c_uint8(tx_type), int(bool(clamp16)))
    return q, dq, int(eob.value)


def av1_inv_transform_recon8bit(coeff_buffer, recon_buffer, recon_stride, txsize, transform_type,
                                eob=0):
    """av1_inv_transform_recon8bit (EbTransforms.c:8935): recon_buffer (uint8, prediction in) is
    updated in place."""
    L = lib()
    cf = np.ascontiguousarray(coeff_buffer, np.int32)
    assert recon_buffer.dtype == np.uint8 and recon_buffer.flags.c_contiguous
    _check(L.svtgpu_av1_inv_transform_recon8bit(_P(cf), _P(recon_buffer), ctypes.c_uint32(recon_stride),
                                                ctypes.c_uint8(txsize), ctypes.c_uint8(transform_type),
                                                0, ctypes.c_uint32(eob)), "av1_inv_transform_recon8bit")
    return recon_buffer


def av1_inv_transform_recon(coeff_buffer, recon_buffer, recon_stride, txsize, bit_increment,
                            transform_type, eob=0):
    """av1_inv_transform_recon (EbTransforms.c:8905): 16-bit pixels; the C symbol takes the
    reference's CONVERT_TO_BYTEPTR form of the pointer, which this wrapper produces."""
    L = lib()
    cf = np.ascontiguousarray(coeff_buffer, np.int32)
    assert recon_buffer.dtype == np.uint16 and recon_buffer.flags.c_contiguous
    byteptr = ctypes.c_void_p(recon_buffer.ctypes.data >> 1)
    _check(L.svtgpu_av1_inv_transform_recon(_P(cf), byteptr, ctypes.c_uint32(recon_stride),
                                            ctypes.c_uint8(txsize), ctypes.c_uint32(bit_increment),
                                            ctypes.c_uint8(transform_type), 0, ctypes.c_uint32(eob)),
           "av1_inv_transform_recon")
    return recon_buffer
