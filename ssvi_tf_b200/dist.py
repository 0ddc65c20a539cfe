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


def split_batch(counts, batch_size):
    """counts: int64 array [world, n_rows] = entries of each row held by each rank.  Returns [world, n_rows]: how many
    entries of every row each rank puts into the row's minibatch so that the GLOBAL minibatch has
    min(len, batch_size) entries (SSVI_TF_interface.py:170-174), each rank's share proportional to its share of the
    row's entries (floor + largest remainder, ties to the lower rank)."""
    counts = np.asarray(counts, dtype=np.int64)
    total = counts.sum(axis=0)
    B = int(batch_size)
    safe = np.maximum(total, 1)
    quota = counts * B // safe
    frac = counts * B - quota * safe
    rem = B - quota.sum(axis=0)
    order = np.argsort(-frac, axis=0, kind="stable")               # ranks by decreasing remainder
    pos = np.empty_like(order)
    np.put_along_axis(pos, order, np.arange(counts.shape[0])[:, None].repeat(counts.shape[1], 1), axis=0)
    quota = quota + (pos < rem[None, :])
    return np.where(total[None, :] > B, quota, counts)


def local_batch_sizes(local_counts, batch_size, group=None):
    """This rank's row of split_batch() (one all-gather of the per-row counts)."""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    t = torch.as_tensor(np.asarray(local_counts, dtype=np.int64))
    if dist.get_backend(group) == "nccl":
        t = t.cuda()
    gathered = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(gathered, t.contiguous(), group=group)
    stack = torch.stack(gathered, dim=0).cpu().numpy()
    return split_batch(stack, batch_size)[rank]


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


# ---------------------------------------------------------------------------------------------------------------------
# In-run parity self-check of the multi-GPU path (bench.py --gpus N runs it before timing; tests/test_gpu_multi.py
# asserts the same things): one small fixed problem, its entries sharded over the ranks (contiguous COO blocks), two
# Jacobi sweeps of every mode through BOTH stage-4 variants (replicated stage 4 / row-block stage 4 + all-gather),
# against the same problem on ONE rank without a process group.
# ---------------------------------------------------------------------------------------------------------------------
class _ShardTensor(object):
    def __init__(self, datatype, dims, coords, vals):
        self.datatype, self.dims = datatype, dims
        self.train_entries, self.train_vals = coords, vals
        self.test_entries, self.test_vals = coords[:1], vals[:1]


def selfcheck_problem(seed=9, dims=(16, 8, 8), D=5, nnz=420):
    rng = np.random.default_rng(seed)
    dims = list(dims)
    flat = rng.choice(int(np.prod(dims)), size=nnz, replace=False)
    coords = np.stack(np.unravel_index(flat, dims), axis=1).astype(np.int64)
    vals = rng.choice([-1.0, 1.0], size=len(coords))
    return dims, D, coords, vals


def _selfcheck_run(coords, vals, dims, D, device, group, shard_rows, sweeps=2):
    import ssvi_tf_b200 as pkg
    np.random.seed(4)
    m = pkg.SSVI_TF_d(_ShardTensor("binary", dims, coords, vals), D, k1=64, k2=8, batch_size=8, randstart=True,
                      device=device, noise_dtype=torch.float64, noise_seed=2, noise_slots=4, process_group=group)
    m.shard_row_update = shard_rows
    for _ in range(sweeps):
        for dim in range(m.order):
            m.sweep(dim)
            m.update_hyper_parameter(dim)
    m.engine.check_flags()
    out = [p.clone() for p in m.engine.params] + [m.engine.pSigma.clone()]
    m.engine.close()
    return out


def selfcheck(group, device, rtol=1e-9):
    """Returns "ok" or raises AssertionError.  Checks, on every rank: sharded == single-rank (rtol 1e-9: the all-reduce
    changes the summation order), the two stage-4 variants bitwise equal, and all replicas bitwise equal."""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    dims, D, coords, vals = selfcheck_problem()
    lo, hi = shard_range(len(vals), rank, world)
    got = {sr: _selfcheck_run(coords[lo:hi], vals[lo:hi], dims, D, device, group, sr) for sr in (False, True)}
    want = _selfcheck_run(coords, vals, dims, D, device, None, False)
    for sr in (False, True):
        for a, b in zip(got[sr], want):
            err = float(((a - b).abs() / (b.abs() + 1e-3)).max())
            assert err <= rtol, "multi-GPU parity: sharded sweep differs from the one-rank sweep (%.3g, row blocks %s)" % (err, sr)
    for a, b in zip(got[False], got[True]):
        assert torch.equal(a, b), "multi-GPU parity: replicated and row-block stage 4 differ"
    for a in got[True]:                                   # replicas bitwise equal: max == min over ranks
        hi_, lo_ = a.clone(), a.clone()
        dist.all_reduce(hi_, op=dist.ReduceOp.MAX, group=group)
        dist.all_reduce(lo_, op=dist.ReduceOp.MIN, group=group)
        assert torch.equal(hi_, lo_), "multi-GPU parity: replicas differ"
    return "ok"
