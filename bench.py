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


def config_for(workload, batch, world):
    """The `config` object of a JSON line: one function for both arms, so that the driver sees the
    same workload under the same keys."""
    return {"workload": WORKLOADS[workload][1], "name": batch.name,
            "samples_per_step_per_gpu": batch.n_samples, "blocks_per_gpu": batch.n_blocks,
            "bytes_per_step_per_gpu": batch.algorithmic_bytes(),
            "l2": "inputs+outputs per step (>= 550 MB) exceed the 126 MB L2; no flush needed",
            "parallelism": f"frames sharded over {world} GPU(s), no collective"}


def cpu_measure(runner, cores, budget_s, min_passes=3, max_passes=200):
    """Passes of the persistent-pool reference over the whole batch within a time budget:
    (mean seconds, best seconds, passes).  One untimed pass first (page faults, pool start)."""
    runner.run(cores)
    times, t_start = [], time.perf_counter()
    while len(times) < min_passes or (time.perf_counter() - t_start < budget_s and len(times) < max_passes):
        times.append(runner.run(cores))
    return float(np.mean(times)), float(np.min(times)), len(times)


def cpu_baseline(batch, tables, qm, budget_s=8.0, runner=None):
    """The reference compiled from source (oracle/_ref, kind 'reference': its AVX2 RTCD path, and
    highbd_quantize_b_helper_c where a qmatrix applies) on a persistent pthread pool over all host
    cores; `value` is the MEAN pass (what --impl reference reports), `best` the best pass, and
    `one_thread` the same code on one core over the first sixteenth of the block list."""
    cores = os.cpu_count() or 1
    runner = runner or make_cpu_runner(batch, tables, qm)
    mean, best, passes = cpu_measure(runner, cores, budget_s)
    n1 = max(1, batch.n_blocks // 16)
    s1 = int(wl.AREA[batch.blocks["tx_size"][:n1]].sum())
    t1 = min(runner.run(1, n1) for _ in range(3))
    return {"value": batch.n_samples / mean / 1e6, "best": batch.n_samples / best / 1e6,
            "one_thread": s1 / t1 / 1e6, "unit": UNIT, "cores": cores, "kind": runner.kind,
            "sample": f"{passes} passes over the full {batch.name} batch ({batch.n_samples} samples) on a "
                      f"persistent {cores}-thread pool: mean (value) and best pass; one_thread = the first "
                      f"{n1} blocks ({s1} samples) on one core, best of 3"}


def reference_arm(args, rank, world):
    """--impl reference: the reference's own CPU implementation of the path (oracle/_ref when it was
    compiled, else the oracle port) on all host threads; each step = one pass over the workload."""
    if rank != 0:
        return 0
    batch = make_batch(args.workload, 0, args.gpus)
    tables = load_tables(batch.bd)
    qm = load_qm() if "qm" in args.workload else None
    cores = os.cpu_count() or 1
    runner = make_cpu_runner(batch, tables, qm)
    for _ in range(max(args.warmup, 1)):
        runner.run(cores)
    each = [runner.run(cores) for _ in range(args.steps)]
    t = sum(each)
    val = batch.n_samples * args.steps / t / 1e6
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32",
            "data": "synthetic", "config": config_for(args.workload, batch, args.gpus),
            "cpu_baseline": {"value": val, "best": batch.n_samples / min(each) / 1e6, "unit": UNIT, "cores": cores,
                             "kind": runner.kind,
                             "sample": f"each step = one pass over the full batch on a persistent {cores}-thread pool"},
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


PARITY_KEYS = ("qcoeff", "dqcoeff", "eob", "energy", "recon")


class DeviceJob:
    """One workload resident on this rank's GPU: plan, device buffers, the step, and the checks."""

    def __init__(self, pkg, g, batch, tables, qm, dev, st):
        import torch
        self.torch, self.pkg, self.g, self.batch, self.tables, self.qm, self.dev, self.st = \
            torch, pkg, g, batch, tables, qm, dev, st
        self.plan = g.plan(batch.blocks)
        pdt = np.uint8 if batch.bd == 8 else np.int16
        with torch.cuda.stream(st):
            d = dict(residual=torch.from_numpy(batch.residual).to(dev),
                     pred=torch.from_numpy(batch.pred.view(pdt)).to(dev))
            d["recon"] = torch.empty_like(d["pred"])
            d["qcoeff"] = torch.empty(batch.n_coeff, dtype=torch.int32, device=dev)
            d["dqcoeff"] = torch.empty(batch.n_coeff, dtype=torch.int32, device=dev)
            d["eob"] = torch.empty(batch.n_blocks, dtype=torch.int16, device=dev)
            d["energy"] = torch.empty(batch.n_blocks, dtype=torch.int64, device=dev)
        st.synchronize()
        self.d = d
        self._runner = None

    def step(self, plan=None):
        self.g.run(plan or self.plan, self.pkg.api.STAGE_ALL, self.st.cuda_stream, **self.d)

    def timed(self, steps, barrier):
        """`steps` steps between two CUDA events on the launching stream -> (ms, launches)."""
        torch = self.torch
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = self.g.launches
        barrier()
        e0.record(self.st)
        for _ in range(steps):
            self.step()
        e1.record(self.st)
        barrier()
        return e0.elapsed_time(e1), self.g.launches - l0

    def outputs(self):
        d, b = self.d, self.batch
        self.st.synchronize()
        return dict(qcoeff=d["qcoeff"].cpu().numpy(), dqcoeff=d["dqcoeff"].cpu().numpy(),
                    eob=d["eob"].cpu().numpy().view(np.uint16), energy=d["energy"].cpu().numpy().view(np.uint64),
                    recon=d["recon"].cpu().numpy().view(b.pred.dtype))

    def runner(self):
        if self._runner is None:
            self._runner = make_cpu_runner(self.batch, self.tables, self.qm)
        return self._runner

    def parity(self, got=None):
        """what was timed must be byte-identical to the CPU run.  The outputs
        the timed steps left on the device (or `got`, host arrays of the e2e call) against ONE fresh
        pass of the compiled reference over the same batch, every block, every output."""
        r = self.runner()
        r.reset()
        r.run(os.cpu_count() or 1)
        want, got = r.outputs(), got or self.outputs()
        bad = [k for k in PARITY_KEYS if not np.array_equal(got[k], want[k])]
        return {"checked_blocks": self.batch.n_blocks, "vs": r.kind, "ok": not bad, "mismatch": bad}

    def time_bucket(self, tx_size, reps=20):
        """One size bucket alone on the launching stream: (ms per launch, algorithmic bytes)."""
        torch, batch = self.torch, self.batch
        sizes = batch.blocks["tx_size"]
        sub_blocks = batch.blocks[sizes == tx_size]
        sub = wl.Batch(sub_blocks, batch.residual, batch.pred, batch.n_coeff, batch.bd, "bucket")
        pl = self.g.plan(sub_blocks)
        for _ in range(3):
            self.step(pl)
        k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.st.synchronize()
        k0.record(self.st)
        for _ in range(reps):
            self.step(pl)
        k1.record(self.st)
        self.st.synchronize()
        pl.close()
        return k0.elapsed_time(k1) / reps, sub.algorithmic_bytes()

    def close(self):
        self.plan.close()
        self.d = None


def all_ok(flag, world, dev):
    """True only if every rank's flag is true."""
    import torch
    import torch.distributed as dist
    t = torch.tensor([1.0 if flag else 0.0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return bool(t.item() > 0.5)


def int32_record(workload, kernel):
    """The secondary (integer-pipe) roofline from the kept ncu capture of this kernel version
    (profiles/int32.json: ALU / FMA pipe and issue-slot utilisation, thread instructions per
    sample); static, labelled with the capture it came from."""
    f = os.path.join(ROOT, "profiles", "int32.json")
    try:
        return json.load(open(f)).get(f"{workload}:{kernel}")
    except Exception:
        return None


def extra_workload(pkg, local_rank, rank, world, dev, st, name, steps, barrier):
    """A secondary workload (BASELINE configs 4 / 4'): device-resident steps, whole-step roofline,
    parity gate.  Own context: its tables and qmatrix differ from the headline's."""
    batch = make_batch(name, rank, world)
    tables = load_tables(batch.bd)
    qm = load_qm() if "qm" in name else None
    g = pkg.SvtGpu(local_rank)
    g.upload_quant_tables(tables)
    if qm is not None:
        g.upload_qmatrix(*qm)
    job = DeviceJob(pkg, g, batch, tables, qm, dev, st)
    for _ in range(3):
        job.step()
    ms, launches = job.timed(steps, barrier)
    par = job.parity()
    ok = all_ok(par["ok"], world, dev)
    ms, total = sharding.reduce_step(ms, batch.n_samples, world, dev)
    hbm, _ = peaks()
    gbs = batch.algorithmic_bytes() / (ms / steps * 1e-3) / 1e9
    rec = {"workload": WORKLOADS[name][1], "name": batch.name,
           "value": sharding.throughput_mcoeff(total, steps, ms) if ok else None, "unit": UNIT,
           "ms_per_step": ms / steps, "steps": steps, "samples_per_step_per_gpu": batch.n_samples,
           "roofline": {"bound": "hbm", "scope": "whole step", "achieved": gbs, "peak": hbm, "unit": "GB/s",
                        "frac": gbs / hbm, "algorithmic_bytes_per_step": batch.algorithmic_bytes()},
           "parity": dict(par, ok=ok, ranks=world), "gpu_launches": launches}
    if rank == 0 and world == 1:
        c = cpu_baseline(batch, tables, qm, budget_s=3.0, runner=job.runner())
        rec["cpu_baseline"] = c
    job.close()
    g.close()
    return rec


def sweep(pkg, g8, local_rank, dev, st):
    """BASELINE config 2 on one GPU: for every TX size and bit depth, all its legal tx types x 4096
    blocks (1024 for the 32- and 64-point sizes), uniform full-range residuals, one
    launch: throughput, HBM fraction and parity against the compiled reference per (size, bd)."""
    torch = __import__("torch")
    out = {}
    hbm, _ = peaks()
    for bd in (8, 10):
        tables = load_tables(bd)
        g = g8 if bd == 8 else pkg.SvtGpu(local_rank)
        if bd != 8:
            g.upload_quant_tables(tables)
        for s_ in range(19):
            pairs = [(s_, t) for t in range(16) if wl.is_valid(s_, t)]
            n_per = 1024 if max(wl.TX_W[s_], wl.TX_H[s_]) >= 32 else 4096
            batch = wl.make_packed_batch(pairs, bd, n_per, seed=13596 + 19 * bd + s_)
            job = DeviceJob(pkg, g, batch, tables, None, dev, st)
            for _ in range(3):
                job.step()
            reps = 10
            k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            st.synchronize()
            k0.record(st)
            for _ in range(reps):
                job.step()
            k1.record(st)
            st.synchronize()
            ms = k0.elapsed_time(k1) / reps
            par = job.parity()
            gbs = batch.algorithmic_bytes() / ms / 1e6
            out[f"{wl.TX_W[s_]}x{wl.TX_H[s_]}_bd{bd}"] = {
                "Mcoeff/s": round(batch.n_samples / ms / 1e3, 1) if par["ok"] else None, "us": round(ms * 1e3, 2),
                "frac": round(gbs / hbm, 3), "blocks": batch.n_blocks, "tx_types": len(pairs), "parity_ok": par["ok"]}
            job.close()
        if bd != 8:
            g.close()
    return out


def copy_ceiling(torch, dev, batch, hb, steps, barrier, world, total_samples):
    """What the PCIe link and the host memory allow for the headline e2e call: per picture the same
    H2D (residual, prediction) and D2H (qcoeff, dqcoeff, eob, energy, recon) byte counts, from / to
    the same page-locked buffers, on one upload and one download stream, no kernels."""
    import torch.distributed as dist
    up, down = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    frames = batch.frames()
    pairs = []
    for k, fr in enumerate(frames):
        res, pred, _ = hb.frame_in(k)
        out = hb.frame_out(k)
        h_in = [torch.from_numpy(res.view(np.uint8)), torch.from_numpy(pred.view(np.uint8))]
        h_out = [torch.from_numpy(v.view(np.uint8).reshape(-1)) for v in out.values() if v is not None]
        pairs.append((h_in, h_out))
    n_in, n_out = len(pairs[0][0]), len(pairs[0][1])
    max_in = [max(hi[j].numel() for hi, _ in pairs) for j in range(n_in)]
    max_out = [max(ho[j].numel() for _, ho in pairs) for j in range(n_out)]
    d_in = [[torch.empty(n, dtype=torch.uint8, device=dev) for n in max_in] for _ in range(3)]
    d_out = [[torch.empty(n, dtype=torch.uint8, device=dev) for n in max_out] for _ in range(3)]

    def one():
        for k, (hi, ho) in enumerate(pairs):
            lane = k % 3
            with torch.cuda.stream(up):
                for dt, ht in zip(d_in[lane], hi):
                    dt[:ht.numel()].copy_(ht, non_blocking=True)
            with torch.cuda.stream(down):
                for dt, ht in zip(d_out[lane], ho):
                    ht.copy_(dt[:ht.numel()], non_blocking=True)
        up.synchronize()
        down.synchronize()
    ok = all(t.is_pinned() for hi, ho in pairs for t in hi + ho)
    for _ in range(2):
        one()
    barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        one()
    barrier()
    wall = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(wall, op=dist.ReduceOp.MAX)
    sec = float(wall.item()) / steps
    h2d = sum(t.numel() for hi, _ in pairs for t in hi)
    d2h = sum(t.numel() for _, ho in pairs for t in ho)
    return {"value": total_samples / sec / 1e6, "unit": UNIT, "ms_per_step": sec * 1e3,
            "d2h_GBps_per_gpu": d2h / sec / 1e9, "h2d_GBps_per_gpu": h2d / sec / 1e9, "pinned": ok,
            "what": "copies only, same bytes and buffers as e2e, one upload + one download stream per GPU"}


def place_near_gpu(local_rank):
    """Run this rank on the CPUs of its GPU's NUMA node (sysfs local_cpulist of the PCI device), so
    that the page-locked buffers it allocates afterwards are node-local (first touch) and its
    copies do not cross the socket interconnect.  Host plumbing only; returns what it did."""
    info = {"numa_node": None, "cpus": None, "bound": False}
    try:
        bus = subprocess.run(["nvidia-smi", "--query-gpu=pci.bus_id", "--format=csv,noheader", "-i", str(local_rank)],
                             capture_output=True, text=True, timeout=20).stdout.strip().lower()
        if bus.count(":") == 2 and len(bus.split(":")[0]) == 8:
            bus = bus[4:]  # 00000000:1b:00.0 -> 0000:1b:00.0
        base = f"/sys/bus/pci/devices/{bus}"
        info["pci"] = bus
        info["numa_node"] = int(open(base + "/numa_node").read())
        cpulist = open(base + "/local_cpulist").read().strip()
        cpus = set()
        for part in cpulist.split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        allowed = os.sched_getaffinity(0)
        use = cpus & allowed
        info["cpus"] = cpulist
        if use and use != allowed:
            os.sched_setaffinity(0, use)
            info["bound"] = True
        info["allowed_cpus"] = len(allowed)
    except Exception as e:  # placement is an optimisation, never a requirement
        info["error"] = str(e)[:120]
    return info


def gpu_arm(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    import cidana_svt_av1_b200 as pkg

    placement = place_near_gpu(local_rank)
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
    st = torch.cuda.Stream(dev)
    job = DeviceJob(pkg, g, batch, tables, qm, dev, st)

    def barrier():
        st.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        job.step()
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
        time.sleep(0.3)
    ms, launches = job.timed(args.steps, barrier)
    sustained = None
    if sampler:
        # a seconds-long run of the same step: the clock sampler (100 ms period) sees the GPU under
        # load, and the figure shows whether the short timed region above flatters the kernel
        n_sus = max(args.steps, int(1.5 / max(ms / args.steps * 1e-3, 1e-6)))
        ms_sus, _ = job.timed(n_sus, lambda: st.synchronize())
        sustained = {"steps": n_sus, "ms_per_step": ms_sus / n_sus,
                     "value": batch.n_samples * n_sus / (ms_sus * 1e-3) / 1e6}
    clocks = sampler.finish() if sampler else None
    # parity gate: the outputs the timed steps left on the device, every block, every rank
    par = job.parity()
    parity_ok = all_ok(par["ok"], world, dev)
    if not par["ok"]:
        print(f"rank {rank}: PARITY MISMATCH vs {par['vs']} in {par['mismatch']}", file=sys.stderr, flush=True)
    ms, total_samples = sharding.reduce_step(ms, batch.n_samples, world, dev)
    value = sharding.throughput_mcoeff(total_samples, args.steps, ms)
    if not parity_ok:
        if rank == 0:
            emit({"metric": METRIC, "value": None, "unit": UNIT, "n_gpus": world,
                  "parity": dict(par, ok=False, ranks=world),
                  "error": "GPU outputs differ from the reference: no throughput is reported (BASELINE.md section 3)"})
        if world > 1:
            dist.destroy_process_group()
        return 3

    if args.kernel_only:
        if rank == 0:
            emit({"metric": METRIC, "value": value, "unit": UNIT, "ms_per_step": ms / args.steps,
                  "gpu_launches": launches, "kernel_only": True, "parity": dict(par, ranks=world)})
        return 0

    # ---- roofline of the dominant kernel, timed alone on the launching stream --------------------
    hbm, which = peaks()
    sizes = batch.blocks["tx_size"]
    by_size = {int(s): int((wl.AREA[sizes[sizes == s]]).sum()) for s in np.unique(sizes)}
    dom = max(by_size, key=by_size.get)
    k_ms, k_bytes = job.time_bucket(dom)
    achieved = k_bytes / (k_ms * 1e-3) / 1e9
    per_kernel = None
    if args.per_kernel and rank == 0:
        per_kernel = {}
        for s_ in sorted(by_size, key=by_size.get, reverse=True):
            t_, b_ = job.time_bucket(s_)
            per_kernel[f"{wl.TX_W[s_]}x{wl.TX_H[s_]}"] = {"us": round(t_ * 1e3, 2), "GB/s": round(b_ / t_ / 1e6, 1),
                                                        "frac": round(b_ / t_ / 1e6 / hbm, 3)}
        print("per-kernel (each bucket alone):", json.dumps(per_kernel), file=sys.stderr, flush=True)
    kname = f"residual_kernel<{wl.TX_W[dom]},{wl.TX_H[dom]}>"
    roof = {"bound": "hbm", "kernel": kname,
            "achieved": achieved, "peak": hbm, "peak_source": which, "unit": "GB/s", "frac": achieved / hbm,
            "traffic": None, "traffic_source": None, "launch_ms": k_ms, "algorithmic_bytes_per_launch": k_bytes,
            "step_achieved": batch.algorithmic_bytes() / (ms / args.steps * 1e-3) / 1e9,
            "step_frac": batch.algorithmic_bytes() / (ms / args.steps * 1e-3) / 1e9 / hbm,
            "int32": int32_record(args.workload, kname)}
    tf = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tf):
        try:
            tj = json.load(open(tf))
            roof["traffic"] = tj.get(f"{args.workload}:{kname}")
            roof["traffic_source"] = tj.get("_source")
        except Exception:
            pass

    # ---- e2e: host buffers in, host buffers out, through the public host-pipeline call -----------
    # Every step: H2D of residual and prediction from page-locked memory, the kernels, D2H of the
    # outputs into page-locked memory; pictures of the batch are pipelined over the library's lanes
    # (svtgpu_submit_host / svtgpu_wait_host).  The geometry repeats from step to step, so the
    # per-picture plans are made once (svtgpu_plan_create) and passed with every submission.
    L = pkg.api.lib()
    L.svtgpu_debug_host_allocs.restype = ctypes.c_longlong
    e2e_steps = max(2, min(args.steps, 10))
    plans = g.frame_plans(batch)

    def e2e_measure(work, hb, stages, plans_):
        def one():
            g.run_host(work, stages=stages, bufs=hb, plans=plans_)
        for _ in range(3):  # three passes: every picture of the batch has met every lane of the pipeline
            one()
        barrier()
        t0, each = time.perf_counter(), []
        for _ in range(e2e_steps):
            t1 = time.perf_counter()
            one()
            each.append((round((time.perf_counter() - t1) * 1e3, 2), int(L.svtgpu_debug_host_allocs())))
        barrier()
        wall = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(wall, op=dist.ReduceOp.MAX)
        return total_samples * e2e_steps / float(wall.item()) / 1e6, each

    def nbytes(*arrays):
        return int(sum(a.nbytes for a in arrays if a is not None))

    # (1) the headline: every output of the path (qcoeff, dqcoeff, eob, energy, recon)
    hb = g.host_buffers(batch, pinned=True)
    h2d = nbytes(batch.residual, batch.pred)
    d2h = nbytes(*hb.out.values())
    e2e_val, e2e_each = e2e_measure(batch, hb, pkg.api.STAGE_ALL, plans)
    if rank == 0:
        print("e2e ms per step:", e2e_each, file=sys.stderr, flush=True)
    # the host arrays the e2e call filled go through the same gate
    e2e_par = job.parity({k: hb.out[k] for k in PARITY_KEYS})
    e2e_ok = all_ok(e2e_par["ok"], world, dev)
    want = job.runner().outputs()  # the reference pass the gate just made

    # (2) what the encode pass consumes (EbCodingLoop.c:697-731 feeds recon_coeff only to the inverse
    # transform, which has already run here): qcoeff + eob + recon come back, dqcoeff stays on the GPU
    hb2 = g.host_buffers(batch, pinned=True)
    hb2.out["dqcoeff"] = None
    enc_val, _ = e2e_measure(batch, hb2, pkg.api.STAGE_ALL, plans)
    enc_ok = all_ok(all(np.array_equal(hb2.out[k], want[k]) for k in ("qcoeff", "eob", "energy", "recon")), world, dev)
    enc = {"value": enc_val if enc_ok else None, "unit": UNIT, "what": "qcoeff + eob + energy + recon out (no dqcoeff)",
           "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": nbytes(*hb2.out.values()), "parity_ok": enc_ok}

    # (3) what mode decision consumes: source + prediction in; eob, distortion pair and
    # reconstruction out, no coefficient arrays
    md = None
    if qm is None:
        src_batch = batch.with_source()
        hb3 = g.host_buffers(src_batch, pinned=True, dist=True, coeffs=False)
        md_stages = pkg.api.STAGE_ALL | pkg.api.STAGE_RESIDUAL
        md_val, _ = e2e_measure(src_batch, hb3, md_stages, plans)
        # residual = src - pred is what the reference pass transforms; its distortion pair is pinned
        # by the tests (tests/test_gpu_parity.py: run_oracle_md), here eob and recon are compared
        r3 = make_cpu_runner(src_batch, tables, qm)
        r3.run(os.cpu_count() or 1)
        w3 = r3.outputs()
        md_ok = all_ok(all(np.array_equal(hb3.out[k], w3[k]) for k in ("eob", "recon")), world, dev)
        md = {"value": md_val if md_ok else None, "unit": UNIT,
              "what": "source + prediction in; eob + distortion pair + energy + recon out (no coefficient arrays)",
              "h2d_bytes_per_step": nbytes(src_batch.src, src_batch.pred), "d2h_bytes_per_step": nbytes(*hb3.out.values()),
              "parity_ok": md_ok}
        del hb3, r3

    # (4) the ceiling of (1): the same bytes through the same page-locked buffers, copies only
    ceiling = copy_ceiling(torch, dev, batch, hb, e2e_steps, barrier, world, total_samples)
    for pl_ in plans:
        pl_.close()

    extra = {}
    if not args.no_extra:
        x_steps = max(3, min(args.steps, 10))
        for name in ("4k10qm", "4k10"):
            if name != args.workload:
                extra[name] = extra_workload(pkg, local_rank, rank, world, dev, st, name, x_steps, barrier)
        if world == 1 and batch.bd == 8 and qm is None:
            extra["sweep"] = sweep(pkg, g, local_rank, dev, st)

    if rank == 0:
        cpu = None
        if world == 1 and not args.no_cpu:
            cpu = cpu_baseline(batch, tables, qm, runner=job.runner())
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
                "config": config_for(args.workload, batch, world),
                "parity": dict(par, ok=True, ranks=world),
                "roofline": roof, "cpu_baseline": cpu, "clocks": clocks, "sustained": sustained,
                "gpu_launches": launches,
                "e2e": {"value": e2e_val if e2e_ok else None, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
                        "d2h_bytes_per_step": int(d2h), "steps": e2e_steps,
                        "parity": dict(e2e_par, ok=e2e_ok, ranks=world), "copy_only_ceiling": ceiling},
                "e2e_encode_pass": enc, "e2e_md": md, "placement": placement,
                "extra": extra}
        emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0 if e2e_ok else 3


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
    ap.add_argument("--no-extra", action="store_true", help="skip the secondary workloads (4k10qm, 4k10, sweep)")
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
