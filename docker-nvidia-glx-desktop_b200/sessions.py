"""Session -> GPU partitioning and whole-job aggregation for the multi-GPU driver.

The reference runs one desktop session per container per GPU (/root/reference xgl.yml:10,120;
README.md:24,182): sessions share no state, so the multi-GPU path has no collective on the data
path (SURVEY.md §8 e). What the ranks do exchange is bookkeeping: a barrier around the timed
region and the max of the per-rank elapsed times, through whatever torch.distributed backend the
launcher initialised (nccl on GPUs, gloo in the CPU tests).
"""
from __future__ import annotations


def partition_sessions(n_sessions: int, n_gpus: int) -> list[list[int]]:
    """session s runs on GPU s % n_gpus (config C5: 64 sessions -> 8 per GPU on 8 GPUs)."""
    if n_sessions < 0 or n_gpus < 1:
        raise ValueError("n_sessions >= 0 and n_gpus >= 1 required")
    return [[s for s in range(n_sessions) if s % n_gpus == g] for g in range(n_gpus)]


def job_throughput(units_local: int, seconds_local: float, dist=None, device=None) -> tuple[float, float, int]:
    """Whole-job rate = units processed by all ranks / the slowest rank's time.
    Returns (rate, max_seconds, total_units). dist=None means a single process."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return units_local / seconds_local, seconds_local, units_local
    import torch
    t = torch.tensor([seconds_local], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    u = torch.tensor([units_local], dtype=torch.int64, device=device)
    dist.all_reduce(u, op=dist.ReduceOp.SUM)
    return int(u.item()) / float(t.item()), float(t.item()), int(u.item())
