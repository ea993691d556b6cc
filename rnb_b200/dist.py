"""Multi-GPU plumbing for the replica layout: one process per GPU, no data-path collective.

RnB shards this path by independent units (videos / clips): every GPU runs a whole-model replica
and pulls its own work (reference ``benchmark.py:237-271``, ``config/r2p1d-aggressive.json``).
``torch.distributed`` is used only to line the ranks up (barrier) and to combine their timings and
counts; the backend is NCCL on GPUs and gloo in the CPU tests.
"""
import torch
import torch.distributed as dist


def shard_range(num_items, rank, world_size):
    """Contiguous share of ``num_items`` independent work items for ``rank``.

    Remainders go to the first ranks, the same rule RnB uses for segments (runner.py:149-151).
    """
    q, r = divmod(num_items, world_size)
    return q * rank + min(rank, r), q * (rank + 1) + min(rank + 1, r)


def barrier():
    if dist.is_available() and dist.is_initialized():
        dist.barrier()


def _reduce(value, op, device):
    if not (dist.is_available() and dist.is_initialized()):
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=op)
    return float(t.item())


def max_over_ranks(value, device='cpu'):
    """A multi-GPU time is the slowest rank's time."""
    return _reduce(value, dist.ReduceOp.MAX, device)


def sum_over_ranks(value, device='cpu'):
    """Whole-job work = sum of what every rank processed."""
    return _reduce(value, dist.ReduceOp.SUM, device)


def whole_job_throughput(units_this_rank, ms_this_rank, device='cpu'):
    """units of all ranks / max-over-ranks time, in units per second."""
    total = sum_over_ranks(units_this_rank, device)
    ms = max_over_ranks(ms_this_rank, device)
    return total / (ms * 1e-3), ms
