"""Synthetic residual-coding workloads: the block lists and pixel/residual planes
that feed the GPU path, the oracle and the reference alike.

Everything here is host-side numpy; no oracle and no CUDA dependency.
"""
import numpy as np

TX_W = [4, 8, 16, 32, 64, 4, 8, 8, 16, 16, 32, 32, 64, 4, 16, 8, 32, 16, 64]
TX_H = [4, 8, 16, 32, 64, 8, 4, 16, 8, 32, 16, 64, 32, 16, 4, 32, 8, 64, 16]
TX_NAMES = ["4X4", "8X8", "16X16", "32X32", "64X64", "4X8", "8X4", "8X16", "16X8", "16X32",
            "32X16", "32X64", "64X32", "4X16", "16X4", "8X32", "32X8", "16X64", "64X16"]
TX_ID = {(w, h): i for i, (w, h) in enumerate(zip(TX_W, TX_H))}

# layout of SvtGpuBlock (include/svt_gpu.h), 32 bytes
BLOCK_DTYPE = np.dtype([("tx_size", "u1"), ("tx_type", "u1"), ("bd", "u1"), ("plane", "u1"),
                        ("res_off", "<u4"), ("res_stride", "<u4"), ("rec_off", "<u4"),
                        ("rec_stride", "<u4"), ("coeff_off", "<u4"), ("qrow", "<u4"),
                        ("flags", "<u4")])
assert BLOCK_DTYPE.itemsize == 32

BLK_CLAMP16 = 1
BLK_USE_QM = 2


def coded_count(s):
    """av1_get_max_eob: coefficients kept per block (64-point sizes keep the 32-wide part)."""
    return min(TX_W[s], 32) * min(TX_H[s], 32)


def is_valid(s, t):
    m = max(TX_W[s], TX_H[s])
    if m == 64:
        return t == 0
    if m == 32:
        return t in (0, 9)
    return True


def valid_pairs():
    return [(s, t) for s in range(19) for t in range(16) if is_valid(s, t)]


CODED = np.array([coded_count(s) for s in range(19)], np.int64)
AREA = np.array([TX_W[s] * TX_H[s] for s in range(19)], np.int64)


class Batch:
    """blocks + the arrays they index.  residual int16, pred uint8/uint16 (same geometry as
    recon), n_coeff = total int32 slots of coeff/qcoeff/dqcoeff."""

    def __init__(self, blocks, residual, pred, n_coeff, bd, name):
        self.blocks = blocks
        self.residual = residual
        self.pred = pred
        self.n_coeff = int(n_coeff)
        self.bd = bd
        self.name = name
        self.frame_bounds = None  # [(block0, block1, elem0, elem1)] when the batch is made of frames
        self._frames = None
        self.src = None           # source pixels (same geometry as pred) once with_source() made them

    @property
    def n_blocks(self):
        return len(self.blocks)

    @property
    def n_samples(self):
        return int(AREA[self.blocks["tx_size"]].sum())

    def frames(self):
        """The batch as independent pictures: one Batch per frame whose arrays are views into this
        batch's arrays and whose block offsets are relative to the frame (what a per-picture
        caller holds).  A batch without frame structure is a single picture."""
        if self._frames is None:
            if not self.frame_bounds:
                self._frames = [self]
            else:
                out = []
                for (b0, b1, e0, e1) in self.frame_bounds:
                    blk = self.blocks[b0:b1].copy()
                    c0 = int(blk["coeff_off"][0]) if len(blk) else 0
                    blk["res_off"] -= e0
                    blk["rec_off"] -= e0
                    blk["coeff_off"] -= c0
                    f = Batch(blk, self.residual[e0:e1], self.pred[e0:e1],
                              int(CODED[blk["tx_size"]].sum()), self.bd, f"{self.name}[{len(out)}]")
                    f.block0, f.coeff0, f.elem0 = b0, c0, e0
                    if self.src is not None:
                        f.src = self.src[e0:e1]
                    out.append(f)
                self._frames = out
        return self._frames

    def with_source(self):
        """The same batch as mode decision sees it: source pixels instead of a
        residual.  src = clip(pred + residual); `residual` is recomputed as src - pred so that the
        two ways of feeding the path describe the same data."""
        mval = (1 << self.bd) - 1
        src = np.clip(self.pred.astype(np.int32) + self.residual, 0, mval).astype(self.pred.dtype)
        b = Batch(self.blocks, (src.astype(np.int32) - self.pred).astype(np.int16), self.pred, self.n_coeff,
                  self.bd, self.name + "_src")
        b.src = src
        b.frame_bounds = self.frame_bounds
        return b

    def algorithmic_bytes(self):
        """residual 2 B + prediction + q/dq 8 B on the coded part + recon, plus
        32 B descriptor + 2 B eob (+ 8 B energy for 64-point sizes) per block."""
        px = 1 if self.bd == 8 else 2
        s = self.blocks["tx_size"]
        area = AREA[s].sum()
        coded = CODED[s].sum()
        n64 = int((np.maximum(np.array(TX_W)[s], np.array(TX_H)[s]) == 64).sum())
        return int(area * (2 + 2 * px) + coded * 8 + len(s) * 34 + n64 * 8)


def _finish(blocks, residual, pred, bd, name):
    coded = CODED[blocks["tx_size"]]
    off = np.zeros(len(blocks), np.int64)
    np.cumsum(coded[:-1], out=off[1:])
    blocks["coeff_off"] = off
    return Batch(blocks, residual, pred, int(coded.sum()), bd, name)


def make_packed_batch(pairs, bd, n_per, seed, strided=False, qindex=(0, 60, 120, 180, 255),
                      amp_div=1, flags=0):
    """Sweep workload: n_per blocks of every (size,type) in `pairs`, each block stored
    densely (stride = w) or at stride 64 (`strided`), residual uniform in +-(2^bd - 1)/amp_div,
    prediction uniform in [0, 2^bd).  qrow = qindex (luma tables)."""
    rng = np.random.default_rng(seed)
    m = (1 << bd) - 1
    n = len(pairs) * n_per
    blocks = np.zeros(n, BLOCK_DTYPE)
    s = np.repeat(np.array([p[0] for p in pairs], np.uint8), n_per)
    t = np.repeat(np.array([p[1] for p in pairs], np.uint8), n_per)
    blocks["tx_size"], blocks["tx_type"], blocks["bd"] = s, t, bd
    w = np.array(TX_W)[s]
    h = np.array(TX_H)[s]
    stride = np.where(strided, 64, w).astype(np.int64) if strided else w.astype(np.int64)
    span = stride * h
    off = np.zeros(n, np.int64)
    np.cumsum(span[:-1], out=off[1:])
    total = int(span.sum())
    blocks["res_off"] = off
    blocks["rec_off"] = off
    blocks["res_stride"] = stride
    blocks["rec_stride"] = stride
    blocks["qrow"] = np.asarray(qindex)[rng.integers(0, len(qindex), n)]
    blocks["flags"] = flags
    a = max(1, m // amp_div)
    residual = rng.integers(-a, a + 1, total).astype(np.int16)
    pred = rng.integers(0, m + 1, total).astype(np.uint8 if bd == 8 else np.uint16)
    return _finish(blocks, residual, pred, bd, f"sweep_bd{bd}")


# ---- frame batches -------------------------------------------------------
# SB layouts: (probability, [(tx_size name, count, luma sub-grid)])
def _grid(w, h, sb=64):
    return [(x, y) for y in range(0, sb, h) for x in range(0, sb, w)]


def _layouts():
    L = []
    L.append((0.10, [(64, 64, _grid(64, 64))]))
    L.append((0.25, [(32, 32, _grid(32, 32))]))
    L.append((0.30, [(16, 16, _grid(16, 16))]))
    L.append((0.20, [(8, 8, _grid(8, 8))]))
    # 4 x 32x16 in the top half, 4 x 16x32 in the bottom half
    L.append((0.10, [(32, 16, [(x, y) for y in (0, 16) for x in (0, 32)]),
                     (16, 32, [(x, 32) for x in (0, 16, 32, 48)])]))
    # 16 x 16x8 in the top half, 16 x 8x16 in the bottom half
    L.append((0.05, [(16, 8, [(x, y) for y in range(0, 32, 8) for x in range(0, 64, 16)]),
                     (8, 16, [(x, y) for y in (32, 48) for x in range(0, 64, 8)])]))
    return L


def _one_frame(rng, pw, ph, bd, qindex, laplace_b, use_qm, uniform):
    """Blocks (offsets relative to the frame), residual and prediction planes of one frame."""
    sbx, sby = pw // 64, ph // 64
    n_sb = sbx * sby
    lay = _layouts()
    probs = np.array([l[0] for l in lay])
    choice = rng.choice(len(lay), size=n_sb, p=probs / probs.sum())
    sb_idx = np.arange(n_sb)
    sb_x = (sb_idx % sbx) * 64
    sb_y = (sb_idx // sbx) * 64
    y_size = pw * ph
    c_w, c_h = pw // 2, ph // 2
    c_size = c_w * c_h
    frame_size = y_size + 2 * c_size
    parts = []
    for li, (_, groups) in enumerate(lay):
        sel = np.nonzero(choice == li)[0]
        if len(sel) == 0:
            continue
        for (w, h, pos) in groups:
            px = np.array([p[0] for p in pos])
            py = np.array([p[1] for p in pos])
            for plane in range(3):
                if plane == 0:
                    bw, bh, stride, base = w, h, pw, 0
                    x = (sb_x[sel][:, None] + px[None, :]).ravel()
                    y = (sb_y[sel][:, None] + py[None, :]).ravel()
                else:
                    bw, bh, stride, base = w // 2, h // 2, c_w, y_size + (plane - 1) * c_size
                    x = (sb_x[sel][:, None] // 2 + px[None, :] // 2).ravel()
                    y = (sb_y[sel][:, None] // 2 + py[None, :] // 2).ravel()
                b = np.zeros(len(x), BLOCK_DTYPE)
                b["tx_size"] = TX_ID[(bw, bh)]
                b["plane"] = plane
                b["res_off"] = base + y * stride + x
                b["res_stride"] = stride
                b["rec_stride"] = stride
                parts.append(b)
    blocks = np.concatenate(parts)
    # keep raster-ish locality: order by plane offset
    blocks = blocks[np.argsort(blocks["res_off"], kind="stable")]
    n = len(blocks)
    s = blocks["tx_size"]
    m = np.maximum(np.array(TX_W)[s], np.array(TX_H)[s])
    u = rng.random(n)
    other = rng.integers(1, 16, n)
    t = np.zeros(n, np.uint8)
    small = m <= 16
    t[small & (u >= 0.6)] = other[small & (u >= 0.6)]
    t[(m == 32) & (u >= 0.9)] = 9
    blocks["tx_type"] = t
    blocks["bd"] = bd
    blocks["qrow"] = blocks["plane"].astype(np.uint32) * 256 + qindex
    if use_qm:  # 2-D transform types only (EbFullLoop.c:532-539)
        blocks["flags"] = np.where(t < 9, BLK_USE_QM, 0)
    mval = (1 << bd) - 1
    if uniform:
        residual = rng.integers(-mval, mval + 1, frame_size).astype(np.int16)
    else:
        residual = np.clip(np.rint(rng.laplace(0.0, laplace_b, frame_size)), -mval, mval).astype(np.int16)
    pred = rng.integers(0, mval + 1, frame_size).astype(np.uint8 if bd == 8 else np.uint16)
    return blocks, residual, pred


def make_frame_batch(width, height, n_frames, bd, seed, qindex=120, laplace_b=6.0, use_qm=False,
                     uniform=False, name=None, first_frame=0):
    """Mixed-size frame batch: every 64x64 superblock draws one of six luma partitions; both
    chroma planes (4:2:0) use the half-size counterpart.  Residual/pred/recon are frame planes
    (Y, U, V per frame, back to back) with their natural strides; coefficients are packed block
    after block.  qrow = plane * 256 + qindex.

    Frame g of the job (g = first_frame .. first_frame + n_frames - 1) is a pure function of
    (seed, g), so the shards the ranks of a multi-GPU run build are slices of one global job
    (see sharding.frame_slice)."""
    pw = (width + 63) // 64 * 64
    ph = (height + 63) // 64 * 64
    frame_size = pw * ph * 3 // 2
    blocks, residual, pred = [], [], []
    for k in range(n_frames):
        rng = np.random.default_rng([seed, first_frame + k])
        b, r, p = _one_frame(rng, pw, ph, bd, qindex, laplace_b, use_qm, uniform)
        b["res_off"] += k * frame_size
        b["rec_off"] = b["res_off"]
        blocks.append(b)
        residual.append(r)
        pred.append(p)
    batch = _finish(np.concatenate(blocks), np.concatenate(residual), np.concatenate(pred), bd,
                    name or f"{width}x{height}x{n_frames}f_bd{bd}{'_qm' if use_qm else ''}")
    ends = np.cumsum([len(b) for b in blocks])
    batch.frame_bounds = [(int(e - len(b)), int(e), k * frame_size, (k + 1) * frame_size)
                          for k, (b, e) in enumerate(zip(blocks, ends))]
    return batch
