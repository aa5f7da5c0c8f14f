"""Sharding of a batch of independent worlds over the GPUs of one box.

Worlds never couple, so there is no exchange step and no data-path collective: rank r owns the
contiguous block of worlds [r * per_gpu, (r + 1) * per_gpu) (weak scaling) and steps it on its
own device.  torch.distributed (gloo: NCCL is not used anywhere) serves only the start barrier and the
max-over-ranks of the device time, from which the whole-job throughput is computed.
"""


def world_block(rank, worlds_per_gpu):
    """First world index and count of the block rank `rank` owns."""
    return rank * worlds_per_gpu, worlds_per_gpu


def strong_block(rank, n_ranks, total_worlds):
    """First world index and count of rank `rank` when `total_worlds` worlds are split over the ranks
    (strong scaling: BASELINE.json configs[1] read literally, 4096 worlds across the GPUs of the box)."""
    per = total_worlds // n_ranks
    extra = total_worlds % n_ranks
    first = rank * per + min(rank, extra)
    return first, per + (1 if rank < extra else 0)


def max_over_ranks(value, dist=None, device="cpu"):
    """Maximum of a python float over all ranks (identity without a process group)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return float(value)
    import torch
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def job_throughput(worlds_per_gpu, n_ranks, steps, slowest_rank_seconds):
    """World-steps/s of the whole job: all ranks' work over the slowest rank's time."""
    return worlds_per_gpu * n_ranks * steps / slowest_rank_seconds
