"""Experiment (library built with -DSVT_TRACE): per-phase clock64 deltas of thread 0 of every CTA,
returned through the energy buffer of the first six blocks of each CTA tile."""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench
import cidana_svt_av1_b200 as pkg
from cidana_svt_av1_b200 import workloads as wl
size = int(sys.argv[1]) if len(sys.argv) > 1 else 2
nb = 256 // min(wl.TX_W[size], wl.TX_H[size])
batch = bench.make_batch("1080p8", 0)
g = pkg.SvtGpu(0); g.upload_quant_tables(bench.load_tables(8))
dev = torch.device("cuda", 0); st = torch.cuda.Stream(dev)
blk = batch.blocks[batch.blocks["tx_size"] == size].copy()
with torch.cuda.stream(st):
    d = dict(residual=torch.from_numpy(batch.residual).to(dev), pred=torch.from_numpy(batch.pred).to(dev))
    d["recon"] = torch.empty_like(d["pred"])
    d["qcoeff"] = torch.empty(batch.n_coeff, dtype=torch.int32, device=dev)
    d["dqcoeff"] = torch.empty(batch.n_coeff, dtype=torch.int32, device=dev)
    d["eob"] = torch.empty(len(blk), dtype=torch.int16, device=dev)
    d["energy"] = torch.zeros(len(blk), dtype=torch.int64, device=dev)
pl = g.plan(blk)
for _ in range(3):
    g.run(pl, 7, st.cuda_stream, **d)
st.synchronize()
en = d["energy"].cpu().numpy()
# the plan sorts blocks; ids map back to input order, so collect per tile through the sort order
order = np.lexsort((np.arange(len(blk)) // 4096 * 0 + blk["tx_type"], np.arange(len(blk)) // 4096))
tiles = len(blk) // nb
t = en[order][: tiles * nb].reshape(tiles, nb)[:, :6].astype(np.float64)
names = ["desc+qrow load", "residual wait", "fwd col phase", "row phase + pred wait", "inv col phase", "recon store issue"]
tot = t.sum(1)
print(f"size {wl.TX_W[size]}x{wl.TX_H[size]}: {tiles} CTAs, mean CTA lifetime {tot.mean():.0f} clk = {tot.mean()/1965:.2f} us")
for i, n in enumerate(names):
    print(f"  {n:24s} mean {t[:, i].mean():8.0f} clk  ({t[:, i].mean() / tot.mean() * 100:4.1f} %)   median {np.median(t[:, i]):8.0f}")
