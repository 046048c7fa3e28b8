#!/usr/bin/env python
"""bench.py -- fused fwd-txfm + quant + inv-recon throughput (Mcoeff/s) and HBM roofline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload 1080p8|4k10qm|4k10]
    python bench.py --impl reference ...     # the reference's CPU path on the host cores

One "step" = one pass of the hot path over one synthetic frame batch.  `value` is
device-resident throughput (CUDA events on the launching stream, max over ranks); `e2e` is the
same metric through the host-buffer call (H2D of residual/prediction/descriptors and D2H of every
output inside the timed region).  Multi-GPU: one process per GPU, frames sharded (each rank owns
its own frame batch: weak scaling), no collective on the data path.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from cidana_svt_av1_b200 import sharding  # noqa: E402
from cidana_svt_av1_b200 import workloads as wl  # noqa: E402

METRIC = "fwd-txfm+quant+inv-recon throughput"
UNIT = "Mcoeff/s"

WORKLOADS = {
    # name: (maker kwargs, description)
    "1080p8": (dict(width=1920, height=1080, n_frames=16, bd=8, seed=20240, qindex=120,
                    laplace_b=6.0), "1080p 8-bit frame batch, 16 frames, mixed TX sizes"),
    "1080p8u": (dict(width=1920, height=1080, n_frames=16, bd=8, seed=20240, qindex=120, uniform=True),
                "1080p 8-bit frame batch, 16 frames, residuals uniform over the full range (data independence)"),
    "4k10qm": (dict(width=3840, height=2160, n_frames=4, bd=10, seed=20241, qindex=100,
                    laplace_b=24.0, use_qm=True), "4K 10-bit frame batch with qmatrix, 4 frames"),
    "4k10": (dict(width=3840, height=2160, n_frames=4, bd=10, seed=20241, qindex=100,
                  laplace_b=24.0), "4K 10-bit frame batch, 4 frames"),
    "tiny": (dict(width=256, height=128, n_frames=1, bd=8, seed=1), "smoke-size batch"),
}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


def make_batch(name, rank, world=1):
    """Rank `rank`'s shard of the weak-scaling job: world x n_frames frames, frame-major slices."""
    kw = dict(WORKLOADS[name][0])
    return sharding.rank_batch(wl.make_frame_batch, kw, kw["n_frames"], rank, world)


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False
        self.proc = None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "100"], stdout=subprocess.PIPE, text=True)
            for line in self.proc.stdout:
                self.rows.append([x.strip() for x in line.split(",")])
                if self.stop_flag:
                    break
        except Exception:
            pass

    def finish(self):
        self.stop_flag = True
        if self.proc:
            self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i] == "Active"})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def load_tables(bd):
    from golden_util import build_tables, load_quant_tables
    qt = load_quant_tables()
    return build_tables({p: qt[(bd, p)] for p in range(3)})


def load_qm():
    from golden_util import load_qmatrices, qm_pack
    return qm_pack(load_qmatrices(), 6)  # aom_get_qmlevel(100, 5, 9)


# ---- the reference arm / cpu_baseline --------------------------------------------------------------
def make_cpu_runner(batch, tables, qm):
    from golden_util import load_scans
    from oracle_lib import CpuRunner
    return CpuRunner(batch, tables, scans=load_scans(), qm=qm)


def cpu_baseline(batch, tables, qm, budget_s=12.0):
    """The reference compiled from source (oracle/_ref, kind 'reference') when present and
    applicable, else the oracle port, on all host cores; best pass within a bounded time."""
    cores = os.cpu_count() or 1
    runner = make_cpu_runner(batch, tables, qm)
    best, passes, t_start = None, 0, time.perf_counter()
    while passes < 2 or (time.perf_counter() - t_start < budget_s and passes < 200):
        dt = runner.run(cores)
        best = dt if best is None else min(best, dt)
        passes += 1
    return {"value": batch.n_samples / best / 1e6, "unit": UNIT, "cores": cores, "kind": runner.kind,
            "sample": f"{passes} passes over the full {batch.name} batch ({batch.n_samples} samples), "
                      f"best pass, {cores} threads"}


def reference_arm(args, rank, world):
    """--impl reference: the reference's own CPU implementation of the path (oracle/_ref when it was
    compiled, else the oracle port) on all host threads; each step = one pass over the workload."""
    if rank != 0:
        return 0
    batch = make_batch(args.workload, 0)
    tables = load_tables(batch.bd)
    qm = load_qm() if "qm" in args.workload else None
    cores = os.cpu_count() or 1
    runner = make_cpu_runner(batch, tables, qm)
    for _ in range(args.warmup):
        runner.run(cores)
    t = sum(runner.run(cores) for _ in range(args.steps))
    val = batch.n_samples * args.steps / t / 1e6
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32",
            "data": "synthetic", "config": {"workload": WORKLOADS[args.workload][1], "name": batch.name,
                                            "samples_per_step": batch.n_samples, "blocks": batch.n_blocks},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": runner.kind,
                             "sample": f"each step = one pass over the full batch, {cores} threads"},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)
    return 0


def mode_arm(args, local_rank):
    """--mode md / txsearch (N = 1): the SURVEY 8f rows next to the headline path, device-resident.
    md       = what mode decision consumes: source + prediction in, eob + distortion pair +
               reconstruction out, no coefficient arrays (8f-1);
    txsearch = every block expanded into one candidate per tx type ProductFullLoopTxSearch tries
               (inter, full set), forward + quantise + distortion per candidate (8f-2)."""
    import torch
    import cidana_svt_av1_b200 as pkg
    from cidana_svt_av1_b200.api import STAGE_ALL, STAGE_FWD, STAGE_QUANT, STAGE_RESIDUAL, tx_search_candidates

    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    batch = make_batch(args.workload, 0).with_source()
    g = pkg.SvtGpu(local_rank)
    g.upload_quant_tables(load_tables(batch.bd))
    px = 1 if batch.bd == 8 else 2
    if args.mode == "md":
        work, stages = batch, STAGE_ALL | STAGE_RESIDUAL
        nbytes = batch.n_samples * 3 * px + batch.n_blocks * (32 + 2 + 8 + 16)
        what = "source + prediction in; eob, distortion pair, reconstruction out (no coefficient arrays)"
    else:
        work = tx_search_candidates(batch, 1, 0, 0, want_coeff=False)
        stages = STAGE_FWD | STAGE_QUANT | STAGE_RESIDUAL
        nbytes = batch.n_samples * 2 * px + work.n_blocks * (32 + 2 + 8 + 16)
        what = (f"{work.n_blocks} candidates of {batch.n_blocks} blocks (inter, full tx set); forward + quantise + "
                "distortion per candidate, source and prediction read once per block from HBM")
    plan = g.plan(work.blocks)
    st = torch.cuda.Stream(dev)
    pdt = np.uint8 if batch.bd == 8 else np.int16
    with torch.cuda.stream(st):
        d = dict(src=torch.from_numpy(batch.src.view(pdt)).to(dev), pred=torch.from_numpy(batch.pred.view(pdt)).to(dev))
        d["recon"] = torch.empty_like(d["pred"])
        d["eob"] = torch.empty(work.n_blocks, dtype=torch.int16, device=dev)
        d["energy"] = torch.empty(work.n_blocks, dtype=torch.int64, device=dev)
        d["dist"] = torch.empty(2 * work.n_blocks, dtype=torch.int64, device=dev)
    st.synchronize()
    for _ in range(max(args.warmup, 3)):
        g.run(plan, stages, st.cuda_stream, **d)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    st.synchronize()
    l0 = g.launches
    e0.record(st)
    for _ in range(args.steps):
        g.run(plan, stages, st.cuda_stream, **d)
    e1.record(st)
    st.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    hbm, which = peaks()
    emit({"metric": f"{METRIC} ({args.mode} mode)", "value": work.n_samples / ms / 1e3, "unit": UNIT,
                      "n_gpus": 1, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms,
                      "higher_is_better": True, "dtype": "int32", "data": "synthetic",
                      "config": {"workload": WORKLOADS[args.workload][1], "mode": args.mode, "what": what,
                                 "samples_per_step": work.n_samples, "blocks_per_step": work.n_blocks},
                      "roofline": {"bound": "hbm", "achieved": nbytes / ms / 1e6, "peak": hbm, "peak_source": which,
                                   "unit": "GB/s", "frac": nbytes / ms / 1e6 / hbm, "algorithmic_bytes_per_step": nbytes},
                      "gpu_launches": g.launches - l0})
    return 0


def gpu_arm(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    import cidana_svt_av1_b200 as pkg

    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    batch = make_batch(args.workload, rank, world)
    tables = load_tables(batch.bd)
    qm = load_qm() if "qm" in args.workload else None
    g = pkg.SvtGpu(local_rank)
    g.upload_quant_tables(tables)
    if qm is not None:
        g.upload_qmatrix(*qm)
    plan = g.plan(batch.blocks)
    st = torch.cuda.Stream(dev)
    pdt = np.uint8 if batch.bd == 8 else np.int16
    with torch.cuda.stream(st):
        h_res = torch.from_numpy(batch.residual).pin_memory()
        h_pred = torch.from_numpy(batch.pred.view(pdt)).pin_memory()
        d = dict(residual=h_res.to(dev), pred=h_pred.to(dev))
        d["recon"] = torch.empty_like(d["pred"])
        d["qcoeff"] = torch.empty(batch.n_coeff, dtype=torch.int32, device=dev)
        d["dqcoeff"] = torch.empty(batch.n_coeff, dtype=torch.int32, device=dev)
        d["eob"] = torch.empty(batch.n_blocks, dtype=torch.int16, device=dev)
        d["energy"] = torch.empty(batch.n_blocks, dtype=torch.int64, device=dev)
    st.synchronize()

    def step():
        g.run(plan, pkg.api.STAGE_ALL, st.cuda_stream, **d)

    def barrier():
        st.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
        time.sleep(0.3)
    l0 = g.launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(st)
    for _ in range(args.steps):
        step()
    e1.record(st)
    barrier()
    ms = e0.elapsed_time(e1)
    launches = g.launches - l0
    if sampler:
        # keep the GPU busy a little longer so the 100 ms sampler sees clocks under load
        t_end = time.time() + 1.0
        while time.time() < t_end:
            step()
        st.synchronize()
    clocks = sampler.finish() if sampler else None
    ms, total_samples = sharding.reduce_step(ms, batch.n_samples, world, dev)
    value = sharding.throughput_mcoeff(total_samples, args.steps, ms)

    if args.kernel_only:
        if rank == 0:
            emit({"metric": METRIC, "value": value, "unit": UNIT, "ms_per_step": ms / args.steps,
                  "gpu_launches": launches, "kernel_only": True})
        return 0

    # ---- roofline of the dominant kernel, timed alone on the launching stream --------------------
    hbm, which = peaks()
    sizes = batch.blocks["tx_size"]
    by_size = {int(s): int((wl.AREA[sizes[sizes == s]]).sum()) for s in np.unique(sizes)}
    dom = max(by_size, key=by_size.get)

    def time_bucket(tx_size, reps=20):
        """One size bucket alone on the launching stream: (ms per launch, algorithmic bytes)."""
        sub_blocks = batch.blocks[sizes == tx_size]
        sub = wl.Batch(sub_blocks, batch.residual, batch.pred, batch.n_coeff, batch.bd, "bucket")
        pl = g.plan(sub_blocks)
        for _ in range(3):
            g.run(pl, pkg.api.STAGE_ALL, st.cuda_stream, **d)
        k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        st.synchronize()
        k0.record(st)
        for _ in range(reps):
            g.run(pl, pkg.api.STAGE_ALL, st.cuda_stream, **d)
        k1.record(st)
        st.synchronize()
        pl.close()
        return k0.elapsed_time(k1) / reps, sub.algorithmic_bytes()

    k_ms, k_bytes = time_bucket(dom)
    achieved = k_bytes / (k_ms * 1e-3) / 1e9
    per_kernel = None
    if args.per_kernel and rank == 0:
        per_kernel = {}
        for s_ in sorted(by_size, key=by_size.get, reverse=True):
            t_, b_ = time_bucket(s_)
            per_kernel[f"{wl.TX_W[s_]}x{wl.TX_H[s_]}"] = {"us": round(t_ * 1e3, 2), "GB/s": round(b_ / t_ / 1e6, 1),
                                                        "frac": round(b_ / t_ / 1e6 / hbm, 3)}
        print("per-kernel (each bucket alone):", json.dumps(per_kernel), file=sys.stderr, flush=True)
    roof = {"bound": "hbm", "kernel": f"residual_kernel<{wl.TX_W[dom]},{wl.TX_H[dom]}>",
            "achieved": achieved, "peak": hbm, "peak_source": which, "unit": "GB/s", "frac": achieved / hbm,
            "traffic": None, "launch_ms": k_ms, "algorithmic_bytes_per_launch": k_bytes,
            "step_achieved": batch.algorithmic_bytes() / (ms / args.steps * 1e-3) / 1e9,
            "step_frac": batch.algorithmic_bytes() / (ms / args.steps * 1e-3) / 1e9 / hbm}
    tf = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tf):
        try:
            roof["traffic"] = json.load(open(tf)).get(f"{args.workload}:{roof['kernel']}")
        except Exception:
            pass

    # ---- e2e: host buffers in, host buffers out, through the public host-pipeline call -----------
    # Every step: descriptor sort + upload, H2D of residual and prediction from page-locked memory,
    # the kernels, D2H of qcoeff, dqcoeff, eob, energy and recon into page-locked memory; pictures
    # of the batch are pipelined over the library's lanes (svtgpu_submit_host / svtgpu_wait_host).
    hb = g.host_buffers(batch, pinned=True)
    h2d = batch.residual.nbytes + batch.pred.nbytes + batch.n_blocks * 32
    d2h = sum(v.nbytes for v in hb.out.values() if v is not None)

    def e2e_step():
        g.run_host(batch, bufs=hb)

    e2e_steps = max(2, min(args.steps, 10))
    for _ in range(3):  # three passes: every picture of the batch has met every lane of the pipeline
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    e2e_each = []
    L = pkg.api.lib()
    L.svtgpu_debug_host_allocs.restype = ctypes.c_longlong
    for _ in range(e2e_steps):
        t1 = time.perf_counter()
        e2e_step()
        e2e_each.append((round((time.perf_counter() - t1) * 1e3, 2), int(L.svtgpu_debug_host_allocs())))
    barrier()
    if rank == 0:
        print("e2e ms per step:", e2e_each, file=sys.stderr, flush=True)
    e2e_s = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_val = total_samples * e2e_steps / float(e2e_s.item()) / 1e6

    if rank == 0:
        cpu = None
        if world == 1 and not args.no_cpu:
            cpu = cpu_baseline(batch, tables, qm)
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
                "config": {"workload": WORKLOADS[args.workload][1], "name": batch.name,
                           "samples_per_step_per_gpu": batch.n_samples, "blocks_per_gpu": batch.n_blocks,
                           "bytes_per_step_per_gpu": batch.algorithmic_bytes(),
                           "l2": "inputs+outputs per step (>= 550 MB) exceed the 126 MB L2; no flush needed",
                           "parallelism": f"frames sharded over {world} GPU(s), no collective"},
                "roofline": roof, "cpu_baseline": cpu, "clocks": clocks, "gpu_launches": launches,
                "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
                        "d2h_bytes_per_step": int(d2h), "steps": e2e_steps}}
        emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


_REAL_STDOUT = None


def quiet_stdout():
    """Libraries (NCCL's version banner, for one) write to fd 1; the contract is ONE JSON line on
    stdout.  Everything but emit() goes to stderr from here on."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="1080p8", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--per-kernel", action="store_true", help="time every size bucket alone (stderr)")
    ap.add_argument("--mode", default="fused", choices=["fused", "md", "txsearch"],
                    help="fused = the headline path; md / txsearch = the SURVEY 8f rows (N = 1, extra lines)")
    ap.add_argument("--kernel-only", action="store_true",
                    help="timed steps only (for ncu): no roofline sub-run, e2e or cpu_baseline")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    quiet_stdout()
    if args.impl == "reference":
        return reference_arm(args, rank, world)
    if args.mode != "fused":
        return mode_arm(args, local_rank) if rank == 0 else 0
    return gpu_arm(args, rank, world, local_rank)


if __name__ == "__main__":
    sys.exit(main())
