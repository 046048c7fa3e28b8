"""Multi-GPU sharding of the residual path: one process per GPU, whole frames per
rank, no collective on the data path.  torch.distributed carries only the barrier and the two
scalar reductions of the timed region (max time, total samples)."""
import os


def frame_slice(total_frames, rank, world):
    """Contiguous, frame-major slice of the job's frames owned by `rank`; the remainder goes to
    the low ranks.  Returns range(first, last)."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    base, extra = divmod(total_frames, world)
    first = rank * base + min(rank, extra)
    return range(first, first + base + (1 if rank < extra else 0))


def rank_batch(maker, kwargs, frames_per_gpu, rank, world):
    """The shard of a weak-scaling job of world * frames_per_gpu frames that `rank` owns."""
    sl = frame_slice(frames_per_gpu * world, rank, world)
    kw = dict(kwargs)
    kw["n_frames"] = len(sl)
    kw["first_frame"] = sl.start
    return maker(**kw)


def env_rank():
    """(rank, world, local_rank) from the torchrun environment; (0, 1, 0) outside it."""
    return (int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)),
            int(os.environ.get("LOCAL_RANK", 0)))


def reduce_step(ms, samples, world, device):
    """Max over ranks of the timed region, sum over ranks of the samples processed."""
    import torch
    import torch.distributed as dist
    t = torch.tensor([ms], dtype=torch.float64, device=device)
    n = torch.tensor([samples], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(n, op=dist.ReduceOp.SUM)
    return float(t.item()), float(n.item())


def throughput_mcoeff(total_samples, steps, ms):
    return total_samples * steps / (ms * 1e-3) / 1e6
