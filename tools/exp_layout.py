"""Experiment: how much of the 16x16 launch time is the frame-plane access pattern?
Times the 16x16 bucket of the 1080p8 batch (a) as is, (b) with every block DCT_DCT, (c) with the
same blocks stored densely (block after block, stride = W), (d) dense + all DCT_DCT."""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench
import cidana_svt_av1_b200 as pkg
from cidana_svt_av1_b200 import workloads as wl

size = int(sys.argv[1]) if len(sys.argv) > 1 else 2
batch = bench.make_batch("1080p8", 0)
tables = bench.load_tables(8)
g = pkg.SvtGpu(0); g.upload_quant_tables(tables)
dev = torch.device("cuda", 0)
st = torch.cuda.Stream(dev)

def run(blocks, residual, pred, label):
    with torch.cuda.stream(st):
        d = dict(residual=torch.from_numpy(residual).to(dev), pred=torch.from_numpy(pred).to(dev))
        d["recon"] = torch.empty_like(d["pred"])
        d["qcoeff"] = torch.empty(batch.n_coeff, dtype=torch.int32, device=dev)
        d["dqcoeff"] = torch.empty(batch.n_coeff, dtype=torch.int32, device=dev)
        d["eob"] = torch.empty(len(blocks), dtype=torch.int16, device=dev)
        d["energy"] = torch.empty(len(blocks), dtype=torch.int64, device=dev)
    pl = g.plan(blocks)
    for _ in range(3):
        g.run(pl, 7, st.cuda_stream, **d)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    st.synchronize(); e0.record(st)
    for _ in range(20):
        g.run(pl, 7, st.cuda_stream, **d)
    e1.record(st); st.synchronize()
    print(f"{label:34s} {e0.elapsed_time(e1) / 20 * 1e3:8.2f} us  ({len(blocks)} blocks)", flush=True)

blk = batch.blocks[batch.blocks["tx_size"] == size].copy()
w, h = wl.TX_W[size], wl.TX_H[size]
run(blk, batch.residual, batch.pred, "frame layout, mixed types")
b2 = blk.copy(); b2["tx_type"] = 0
run(b2, batch.residual, batch.pred, "frame layout, all DCT_DCT")
# dense copies of the same blocks
n = len(blk)
res = np.zeros(n * w * h, np.int16); prd = np.zeros(n * w * h, batch.pred.dtype)
rows = np.arange(h)[None, :, None]; cols = np.arange(w)[None, None, :]
src = blk["res_off"].astype(np.int64)[:, None, None] + rows * blk["res_stride"].astype(np.int64)[:, None, None] + cols
res[:] = batch.residual[src].ravel(); prd[:] = batch.pred[src].ravel()
b3 = blk.copy(); b3["res_off"] = np.arange(n, dtype=np.int64) * w * h; b3["rec_off"] = b3["res_off"]
b3["res_stride"] = w; b3["rec_stride"] = w
run(b3, res, prd, "dense layout, mixed types")
b4 = b3.copy(); b4["tx_type"] = 0
run(b4, res, prd, "dense layout, all DCT_DCT")
