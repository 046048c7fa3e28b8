"""Multi-GPU host logic on CPU: frame-major sharding and the timed-region reductions, with two
real processes over gloo (one process per GPU, no collective on the data path)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from cidana_svt_av1_b200 import sharding  # noqa: E402
from cidana_svt_av1_b200 import workloads as wl  # noqa: E402

KW = dict(width=128, height=128, n_frames=2, bd=8, seed=77)


def test_frame_slices_partition_the_job():
    for total in (0, 1, 7, 16, 64, 1001):
        for world in (1, 2, 3, 4, 8):
            seen = []
            for r in range(world):
                sl = sharding.frame_slice(total, r, world)
                seen += list(sl)
                assert len(sl) in (total // world, total // world + 1)
            assert seen == list(range(total))  # contiguous, ordered, each frame exactly once
    with pytest.raises(ValueError):
        sharding.frame_slice(8, 8, 8)


def test_rank_batches_are_slices_of_one_global_job():
    world = 2
    whole = wl.make_frame_batch(**dict(KW, n_frames=KW["n_frames"] * world))
    parts = [sharding.rank_batch(wl.make_frame_batch, KW, KW["n_frames"], r, world) for r in range(world)]
    assert sum(p.n_blocks for p in parts) == whole.n_blocks
    assert sum(p.n_samples for p in parts) == whole.n_samples
    np.testing.assert_array_equal(np.concatenate([p.residual for p in parts]), whole.residual)
    np.testing.assert_array_equal(np.concatenate([p.pred for p in parts]), whole.pred)
    for f in ("tx_size", "tx_type", "plane", "qrow", "res_stride"):
        np.testing.assert_array_equal(np.concatenate([p.blocks[f] for p in parts]), whole.blocks[f])
    # offsets restart in every shard: each rank addresses only its own planes
    shift = len(parts[0].residual)
    np.testing.assert_array_equal(
        np.concatenate([parts[0].blocks["res_off"], parts[1].blocks["res_off"] + shift]),
        whole.blocks["res_off"])


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    from golden_util import build_tables, load_quant_tables, load_scans
    from oracle_lib import run_oracle

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    assert sharding.env_rank() == (rank, world, rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        batch = sharding.rank_batch(wl.make_frame_batch, KW, KW["n_frames"], rank, world)
        qt = load_quant_tables()
        tables = build_tables({p: qt[(8, p)] for p in range(3)})
        out = run_oracle(batch, tables, scans=load_scans())  # the shard's work, done by the CPU checker
        digest = int(np.bitwise_xor.reduce(out["qcoeff"].astype(np.int64) * 2654435761 % (1 << 31)))
        # pretend rank r took (r + 1) * 10 ms: the reported time must be the slowest rank's
        ms, total = sharding.reduce_step(10.0 * (rank + 1), batch.n_samples, world, torch.device("cpu"))
        dist.barrier()
        q.put((rank, ms, total, batch.n_samples, digest, int(out["eob"].astype(np.int64).sum())))
    finally:
        dist.destroy_process_group()


def test_two_ranks_over_gloo():
    import torch.multiprocessing as mp
    from golden_util import build_tables, load_quant_tables, load_scans
    from oracle_lib import run_oracle

    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = sorted(q.get(timeout=240) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # both ranks agree on the reduced numbers: max time, total samples
    assert [g[1] for g in got] == [20.0, 20.0]
    assert got[0][2] == got[1][2] == got[0][3] + got[1][3]
    assert sharding.throughput_mcoeff(got[0][2], 4, 20.0) == pytest.approx(got[0][2] * 4 / 0.020 / 1e6)
    # the union of the shards is the single-process job: same eob total, same coefficient digest
    whole = wl.make_frame_batch(**dict(KW, n_frames=KW["n_frames"] * world))
    qt = load_quant_tables()
    out = run_oracle(whole, build_tables({p: qt[(8, p)] for p in range(3)}), scans=load_scans())
    assert int(out["eob"].astype(np.int64).sum()) == got[0][5] + got[1][5]
    n0 = sharding.rank_batch(wl.make_frame_batch, KW, KW["n_frames"], 0, world).n_coeff
    d0 = int(np.bitwise_xor.reduce(out["qcoeff"][:n0].astype(np.int64) * 2654435761 % (1 << 31)))
    d1 = int(np.bitwise_xor.reduce(out["qcoeff"][n0:].astype(np.int64) * 2654435761 % (1 << 31)))
    assert (d0, d1) == (got[0][4], got[1][4])


def test_multi_device_partitioning_matches_the_rank_sharding():
    """svtgpu_create_multi round-robins pictures over its devices (host-only rule, no GPU needed):
    every picture goes to exactly one device, loads differ by at most one picture, and with
    world = 2 device d gets the pictures a two-rank run would give rank d under a round-robin
    frame split."""
    import ctypes
    import cidana_svt_av1_b200 as pkg
    L = pkg.load_library()
    for world in (1, 2, 3, 8):
        owner = [L.svtgpu_multi_device_of(world, ctypes.c_longlong(i)) for i in range(37)]
        assert all(0 <= d < world for d in owner)
        loads = [owner.count(d) for d in range(world)]
        assert max(loads) - min(loads) <= 1
        assert owner == [i % world for i in range(37)]
    assert L.svtgpu_multi_device_of(0, ctypes.c_longlong(1)) == -1
    assert L.svtgpu_multi_device_of(2, ctypes.c_longlong(-1)) == -1
