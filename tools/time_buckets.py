#!/usr/bin/env python
"""Timing-only helper for kernel experiments whose outputs are deliberately wrong (a store left out,
a sector padded): the step and every size bucket alone, through the same DeviceJob as bench.py, with
NO parity gate -- so nothing printed here is a result, only a cost estimate.
    SVTGPU_LIB=tools/variants/libsvtgpu_X.so python tools/time_buckets.py [workload]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
import numpy as np  # noqa: E402
import torch  # noqa: E402
import cidana_svt_av1_b200 as pkg  # noqa: E402
from cidana_svt_av1_b200 import workloads as wl  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "1080p8"
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
batch = bench.make_batch(name, 0, 1)
if os.environ.get("TB_FRAC"):  # the first 1/k of the block list (first frames): buckets small enough to stay in L2 across repetitions
    k = int(os.environ["TB_FRAC"])
    batch = wl.Batch(batch.blocks[:len(batch.blocks) // k].copy(), batch.residual, batch.pred, batch.n_coeff, batch.bd, "frac")
if os.environ.get("TB_ALL_DCT"):  # every block DCT_DCT: what a kind-specialised launch would see
    batch.blocks["tx_type"] = 0
tables = bench.load_tables(batch.bd)
g = pkg.SvtGpu(0)
g.upload_quant_tables(tables)
st = torch.cuda.Stream(dev)
job = bench.DeviceJob(pkg, g, batch, tables, None, dev, st)
for _ in range(5):
    job.step()
st.synchronize()
ms, _ = job.timed(20, lambda: st.synchronize())
out = {"lib": os.environ.get("SVTGPU_LIB", "in-tree"), "step_ms": round(ms / 20, 5)}
sizes = batch.blocks["tx_size"]
for s in sorted(np.unique(sizes), key=lambda s: -int(wl.AREA[sizes[sizes == s]].sum())):
    t, _b = job.time_bucket(int(s))
    out[f"{wl.TX_W[s]}x{wl.TX_H[s]}"] = round(t * 1e3, 2)
print(json.dumps(out))
