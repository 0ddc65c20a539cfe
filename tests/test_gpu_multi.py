"""GPU, world size 2 over NCCL (skipped on a one-GPU box): an entry-sharded full sweep equals the one-GPU sweep of the
concatenated shards, with stage 4 either replicated or run on row blocks + all-gather (bit-identical replicas)."""
import os
import socket

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


class _T(object):
    def __init__(self, datatype, dims, coords, vals):
        self.datatype, self.dims = datatype, dims
        self.train_entries, self.train_vals = coords, vals
        self.test_entries, self.test_vals = coords[:1], vals[:1]


def _problem():
    rng = np.random.default_rng(9)
    dims, D = [8, 6, 4], 5
    total = int(np.prod(dims))
    flat = rng.choice(total, size=150, replace=False)
    coords = np.stack(np.unravel_index(flat, dims), axis=1).astype(np.int64)
    vals = rng.choice([-1.0, 1.0], size=len(coords))
    return dims, D, coords, vals


def _run(model, sweeps=2):
    for _ in range(sweeps):
        for dim in range(model.order):
            model.sweep(dim)
            model.update_hyper_parameter(dim)
    return [p.cpu().numpy().copy() for p in model.engine.params], model.pSigma.copy()


def _worker(rank, world, port, q):
    import torch.distributed as dist
    import ssvi_tf_b200 as pkg
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda:%d" % rank))
    from ssvi_tf_b200.dist import shard_range
    dims, D, coords, vals = _problem()
    lo, hi = shard_range(len(vals), rank, world)
    out = {}
    for shard_rows in (False, True):
        np.random.seed(4)
        m = pkg.SSVI_TF_d(_T("binary", dims, coords[lo:hi], vals[lo:hi]), D, k1=64, k2=8, batch_size=8, randstart=True,
                          device="cuda:%d" % rank, noise_dtype=torch.float64, noise_seed=2, noise_slots=4,
                          process_group=dist.group.WORLD)
        m.shard_row_update = shard_rows
        out[shard_rows] = _run(m)
        m.engine.check_flags()
    q.put((rank, out))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_gpu_sweeps_match_one_gpu_and_row_block_stage4():
    import torch.multiprocessing as mp
    import ssvi_tf_b200 as pkg
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=300) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    dims, D, coords, vals = _problem()
    np.random.seed(4)
    m = pkg.SSVI_TF_d(_T("binary", dims, coords, vals), D, k1=64, k2=8, batch_size=8, randstart=True, device="cuda:0",
                      noise_dtype=torch.float64, noise_seed=2, noise_slots=4)
    want_params, want_sigma = _run(m)
    for r in range(world):
        for mode in (False, True):
            params, sigma = got[r][mode]
            for a, b in zip(params, want_params):
                np.testing.assert_allclose(a, b, rtol=1e-9, atol=1e-12)
            np.testing.assert_allclose(sigma, want_sigma, rtol=1e-9)
        for a, b in zip(got[r][False][0], got[r][True][0]):      # same sums, same stage 4: bitwise
            np.testing.assert_array_equal(a, b)
    for mode in (False, True):
        for a, b in zip(got[0][mode][0], got[1][mode][0]):        # replicas identical
            np.testing.assert_array_equal(a, b)


def _selfcheck_worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda:%d" % rank))
    from ssvi_tf_b200 import dist as sdist
    q.put((rank, sdist.selfcheck(dist.group.WORLD, "cuda:%d" % rank)))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_in_run_selfcheck_of_bench_passes_on_every_rank():
    """ssvi_tf_b200.dist.selfcheck is what `bench.py --gpus N` runs before timing (multi_gpu_parity in its JSON line)."""
    import torch.multiprocessing as mp
    world = min(torch.cuda.device_count(), 8)
    world = 1 << (world.bit_length() - 1)
    port = _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_selfcheck_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=300) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert got == {r: "ok" for r in range(world)}
