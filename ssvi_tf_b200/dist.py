"""Multi-GPU plumbing (one process per GPU, torch.distributed; NCCL on the B200 box, gloo in CPU tests).

The path shards by OBSERVED ENTRY: rank r holds a contiguous block of the COO list (so the per-row
entry lists of the ranks, concatenated in rank order, are the global lists in the reference's append
order).  Contributions of different entries to a row are summed (SSVI_TF_d.py:47-56,
SSVI_TF_interface.py:581-591, SSVI_TF_robust.py:135-137), so one exchange step per mode sweep suffices:
an all-reduce of the per-row accumulator records [di | Di | si_sum | count].  Stage 4 then runs either
redundantly on every rank from identical inputs (bit-identical replicas without a second collective), or --
for full sweeps of SSVI_TF_d, where stage 4 of a 10k-row mode is a tenth of an 8-GPU step -- on each rank's
block of rows (owned_rows) followed by an all-gather of the updated rows of every state array
(allgather_rows), which also leaves the replicas bit-identical.

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


def owned_rows(n_rows, group=None):
    """Block [lo, hi) of a mode's rows whose stage 4 this rank runs, or None when the rows do not divide evenly
    (the caller then updates every row redundantly)."""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    if n_rows % world:
        return None
    share = n_rows // world
    return rank * share, (rank + 1) * share


def allgather_rows(tensors, lo, hi, group=None):
    """Every tensor is [n_rows, ...] with this rank's block [lo, hi) freshly updated: gather all blocks in place."""
    for t in tensors:
        assert t.is_contiguous()
        dist.all_gather_into_tensor(t.view(-1), t[lo:hi].reshape(-1).clone(), group=group)
