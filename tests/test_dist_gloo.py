"""CPU, world_size 2, gloo: the multi-GPU host logic (ssvi_tf_b200/dist.py) -- contiguous entry shards,
global noise-slot offsets, all-reduce of the per-row accumulator records -- with the oracle standing in
for the CUDA kernels.  The sharded result must equal the single-process result (linearity in entries)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import ssvi_oracle as so
from tests.helpers import random_problem, random_state


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _problem():
    rng = np.random.default_rng(21)
    dims, D, k1, k2 = [5, 4, 3], 4, 8, 4
    cfg = so.make_config("robust", "bernoulli", D, k1=k1, k2=k2, batch_size=10 ** 9)
    coords, vals = random_problem(rng, "bernoulli", dims, 41)
    st = random_state(rng, cfg, dims)
    pool = so.make_tape(rng, cfg, 3, 64)
    return cfg, dims, coords, vals, st, pool


def _acc_records(cfg, dims, coords, vals, st, pool, dim, slot_offset):
    """Oracle accumulator records [di | Di | si_sum | count] of every row of `dim` over the given entries."""
    D = cfg["D"]
    rowptr, perm = so.build_index(coords, dims)
    out = np.zeros((dims[dim], D + D * D + 2))
    for row in range(dims[dim]):
        ids = perm[dim][rowptr[dim][row]:rowptr[dim][row + 1]]
        if len(ids) == 0:
            continue
        o = int(slot_offset[row])
        tape = so.Tape(pool.z_other[o:o + len(ids)], pool.z_own[o:o + len(ids)], pool.r_w[o:o + len(ids)],
                       pool.z_f[o:o + len(ids)])
        di, Di, si, _, B = so.row_accumulate(st, cfg, dim, row, coords[ids], vals[ids], tape)
        out[row] = np.concatenate([di, Di.ravel(), [si, B]])
    return out


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from ssvi_tf_b200 import dist as sdist
    cfg, dims, coords, vals, st, pool = _problem()
    lo, hi = sdist.shard_range(len(vals), rank, world)
    res = {}
    for dim in range(3):
        local = np.bincount(coords[lo:hi, dim], minlength=dims[dim]).astype(np.int64)
        off, tot = sdist.row_slot_offsets(torch.as_tensor(local))
        acc = torch.as_tensor(_acc_records(cfg, dims, coords[lo:hi], vals[lo:hi], st, pool, dim, off.numpy()))
        sdist.allreduce_accumulators(acc)
        res[dim] = (acc.numpy(), tot.numpy(), off.numpy())
    t = sdist.max_over_ranks(float(rank + 1))
    res["max"] = float(t.item())
    # stage 4 on row blocks: every rank updates its block of the (replicated) state, then the blocks are gathered
    state = [torch.zeros((6, 5), dtype=torch.float64), torch.zeros((6, 2), dtype=torch.float64)]
    own = sdist.owned_rows(6)
    res["own"], res["own_uneven"] = own, sdist.owned_rows(7)
    for k, a in enumerate(state):
        a[own[0]:own[1]] = 100.0 * (k + 1) + torch.arange(own[0], own[1], dtype=torch.float64)[:, None]
    sdist.allgather_rows(state, own[0], own[1])
    res["gathered"] = [a.numpy().copy() for a in state]
    q.put((rank, res))
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_accumulators_equal_single_process():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    cfg, dims, coords, vals, st, pool = _problem()
    for dim in range(3):
        want = _acc_records(cfg, dims, coords, vals, st, pool, dim, np.zeros(dims[dim], dtype=np.int64))
        for r in range(world):
            acc, tot, off = got[r][dim]
            np.testing.assert_allclose(acc, want, rtol=1e-12, atol=1e-13)
            np.testing.assert_array_equal(acc[:, -1].astype(np.int64), tot)          # counts: bit-exact
            np.testing.assert_array_equal(tot, np.bincount(coords[:, dim], minlength=dims[dim]))
        np.testing.assert_array_equal(got[0][dim][2], 0)                              # rank 0 starts every row
        np.testing.assert_array_equal(got[0][dim][0], got[1][dim][0])                 # replicas stay identical
    assert got[0]["max"] == got[1]["max"] == 2.0
    assert got[0]["own"] == (0, 3) and got[1]["own"] == (3, 6) and got[0]["own_uneven"] is None
    for r in range(world):
        for k, a in enumerate(got[r]["gathered"]):
            np.testing.assert_array_equal(a, 100.0 * (k + 1) + np.arange(6.0)[:, None] * np.ones_like(a))


def test_shard_range_partitions():
    from ssvi_tf_b200.dist import shard_range
    for nnz in (0, 1, 7, 64, 1001):
        for world in (1, 2, 3, 8):
            spans = [shard_range(nnz, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == nnz
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def test_split_batch_keeps_the_global_minibatch_at_batch_size():
    """Sharded per-row updates: the ranks' shares of a row's minibatch add up to min(len, batch_size)
    (SSVI_TF_interface.py:170-174) and follow the ranks' shares of the row's entries."""
    from ssvi_tf_b200.dist import split_batch
    rng = np.random.default_rng(3)
    counts = rng.integers(0, 400, size=(4, 50))
    counts[:, 0] = 0
    counts[:, 1] = [3, 0, 1, 2]
    q = split_batch(counts, 128)
    total = counts.sum(axis=0)
    np.testing.assert_array_equal(q.sum(axis=0), np.minimum(total, 128))
    assert (q <= counts).all() and (q >= 0).all()
    big = total > 128
    ideal = counts[:, big] * 128.0 / total[big]
    assert np.abs(q[:, big] - ideal).max() < 1.0 + 1e-9


def _batch_worker(rank, world, port, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from ssvi_tf_b200 import dist as sdist
    local = np.array([[300, 10, 0, 5], [100, 200, 0, 200]][rank], dtype=np.int64)
    q.put((rank, sdist.local_batch_sizes(local, 128, None)))
    dist.barrier()
    dist.destroy_process_group()


def test_local_batch_sizes_over_gloo():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_batch_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    np.testing.assert_array_equal(got[0] + got[1], [128, 128, 0, 128])
    np.testing.assert_array_equal(got[0], [96, 6, 0, 3])
