"""Multi-GPU plumbing (one process per GPU, torch.distributed; NCCL on the B200 box, gloo in CPU tests).

The path shards by OBSERVED ENTRY: rank r holds a contiguous block of the COO list (so the per-row
entry lists of the ranks, concatenated in rank order, are the global lists in the reference's append
order).  Contributions of different entries to a row are summed (SSVI_TF_d.py:47-56,
SSVI_TF_interface.py:581-591, SSVI_TF_robust.py:135-137), so one exchange step per mode sweep suffices:
an all-reduce of the per-row accumulator records [di | Di | si_sum | count].  Stage 4 then runs
redundantly on every rank from identical inputs, which keeps the replicated posterior bit-identical
across ranks without a second collective.

Noise slots are a function of an entry's GLOBAL position inside its row's minibatch, so an N-GPU sweep
consumes exactly the draws the 1-GPU sweep does: row_slot_offsets() gives every rank the number of
entries lower ranks hold for each row.
"""
import numpy as np
import torch
import torch.distributed as dist


def shard_range(nnz, rank, world):
    """Contiguous block [lo, hi) of the COO list owned by `rank` (sizes differ by at most one)."""
    base, rem = divmod(int(nnz), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def row_slot_offsets(local_counts, group=None):
    """local_counts: int64 tensor [n_rows] = entries of each row held by this rank (in this sweep).
    Returns (offset, total): offset[r] = entries of row r on lower ranks, total[r] = over all ranks."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    gathered = [torch.empty_like(local_counts) for _ in range(world)]
    dist.all_gather(gathered, local_counts.contiguous(), group=group)
    stack = torch.stack(gathered, dim=0)
    total = stack.sum(dim=0)
    offset = stack[:rank].sum(dim=0) if rank > 0 else torch.zeros_like(local_counts)
    return offset, total


def allreduce_accumulators(acc, group=None):
    """The exchange step: sum the per-row records over ranks, in place."""
    dist.all_reduce(acc, op=dist.ReduceOp.SUM, group=group)
    return acc


def max_over_ranks(x, group=None):
    t = x if torch.is_tensor(x) else torch.tensor([float(x)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return t
