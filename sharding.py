"""Frame-batch sharding helpers shared by bench.py and the gloo test: stereo frames are independent units, so
rank r owns frames [lo, hi) and there is no data-path collective; only the device time is max-reduced."""


def frame_shard(total_frames, rank, world):
    base, rem = divmod(total_frames, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def aggregate_throughput(frames_local, steps, ms_local, world, dist=None):
    """Whole-job frames/s: all ranks' frames x steps over the slowest rank's device time."""
    import torch
    frames = torch.tensor([float(frames_local)], device=ms_local.device)
    ms = ms_local.clone()
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(frames, op=dist.ReduceOp.SUM)
    return float(frames.item()) * steps / (float(ms.item()) / 1e3)
