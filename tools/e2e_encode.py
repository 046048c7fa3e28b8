#!/usr/bin/env python
"""tools/e2e_encode.py [--w 1920 --h 1080 --frames 10 --qp 30 --level AB] -> one JSON line.

BASELINE configs 1 / 5: the unmodified reference application (oracle/_ref/SvtAv1EncApp) encodes a
synthetic clip on the host cores, then again with libsvtgpu.so installed by the LD_PRELOAD shim
(integration/svt_gpu_preload.c); reports both wall times, the sizes and whether the two .ivf files
are byte-identical.  One transform block per GPU call (levels A / B), so the GPU run is
latency-bound by construction; this is the parity run, not the throughput benchmark."""
import argparse
import json
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import test_e2e_bitstream as e2e  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--w", type=int, default=1920)
ap.add_argument("--h", type=int, default=1080)
ap.add_argument("--frames", type=int, default=10)
ap.add_argument("--qp", type=int, default=30)
ap.add_argument("--level", default="ABS")
ap.add_argument("--bd", type=int, default=8)
a = ap.parse_args()
with tempfile.TemporaryDirectory() as d:
    clip, ref, gpu = (os.path.join(d, n) for n in ("clip.yuv", "ref.ivf", "gpu.ivf"))
    e2e.make_clip(clip, a.w, a.h, a.frames, bd=a.bd)
    t_ref, _ = e2e.encode(clip, ref, a.w, a.h, a.frames, a.qp, timeout=3600, bd=a.bd)
    t_gpu, log = e2e.encode(clip, gpu, a.w, a.h, a.frames, a.qp, gpu_level=a.level, timeout=3600, bd=a.bd)
    same = open(ref, "rb").read() == open(gpu, "rb").read()
    print(json.dumps({"clip": f"{a.w}x{a.h} 4:2:0 {a.bd}-bit, {a.frames} frames, synthetic", "enc_mode": 8, "qp": a.qp,
                      "level": a.level, "reference_s": round(t_ref, 2), "gpu_path_s": round(t_gpu, 2),
                      "ivf_bytes": os.path.getsize(ref), "ivf_md5": e2e.md5(ref), "byte_identical": same,
                      "preload_log": [l for l in log.splitlines() if "svtgpu" in l]}))
    sys.exit(0 if same else 1)
