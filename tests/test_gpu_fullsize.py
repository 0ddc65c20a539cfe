"""GPU: parity at BASELINE.json's FULL sizes through size-independent properties (the oracle finishes small cases only).

Config 2 (binary 1000^3 at 1 % density, 8e6 train entries, D = 20, SSVI_TF_d, full covariance) exactly as bench.py
builds it, and config 4's shape (500x500x500x100, D = 50, SSVI_TF_simple, diagonal) at a reduced nnz:

  * index: every mode's (rowptr, perm) is the STABLE sort of the COO list by row -- the reference's append order
    (Tensor/Tensor.py:152-169, 251-258) -- bit-exact, and the touched-row counts of a sweep equal the row lengths;
  * the slot-major sweep kernels equal the direct kernel (which the small-size tests pin to the oracle and to the
    reference's fixtures) on randomly chosen rows of the full-size problem;
  * chunking invariance: the row accumulators do not depend on how a row's entries are cut into chunks;
  * additivity over entry shards (what the multi-GPU all-reduce relies on): accumulators of two halves of the COO list,
    the second half reading the noise slots that continue the first half's, sum to the accumulators of the whole.
"""
import os
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

RTOL = 1e-9          # same arithmetic, different summation order


def _close(a, b, what):
    a, b = a.double().cpu().numpy(), b.double().cpu().numpy()
    scale = max(np.abs(b).max(), 1e-300)
    np.testing.assert_allclose(a, b, rtol=RTOL, atol=1e-11 * scale, err_msg=what)


def _model(cfg_id, scale=1.0, engine_kwargs=None):
    import bench
    c = bench.CONFIGS[cfg_id]
    data = bench.DeviceTensorData(c, "cuda", 319 + int(cfg_id[1]), scale)
    m = bench.build_model(c, data, "cuda", None, 16, 128, engine_kwargs=engine_kwargs)
    m.fresh_noise()
    return c, data, m


def _check_index(eng):
    for d, n in enumerate(eng.dims):
        col = eng.coords[:, d].long()
        counts = torch.bincount(col, minlength=n).cpu().numpy()
        rp = np.asarray(eng.rowptr_h[d])
        np.testing.assert_array_equal(rp[1:] - rp[:-1], counts)              # touched-row counts, bit-exact
        perm = eng.perm[d]
        rows_sorted = col[perm]
        dr, dp = rows_sorted[1:] - rows_sorted[:-1], perm[1:] - perm[:-1]
        assert bool(((dr > 0) | ((dr == 0) & (dp > 0))).all()), "mode %d: not the stable sort by row" % d
        assert int(torch.unique(perm).numel()) == eng.nnz                    # a permutation of the entry ids


def test_config2_full_size_properties():
    c, data, m = _model("c2")
    eng = m.engine
    assert eng.nnz == 8_000_000 and eng.sweep_supported
    _check_index(eng)
    rng = np.random.default_rng(2)
    for dim in (0, 2):
        n = eng.dims[dim]
        plan = eng.plan_sweep(dim)
        acc = eng.accumulate(plan).clone()
        eng.check_flags()
        lens = np.diff(np.asarray(eng.rowptr_h[dim]))
        np.testing.assert_array_equal(acc[:, -1].cpu().numpy(), lens.astype(np.float64))
        assert bool(torch.isfinite(acc).all())
        # sweep kernels == direct kernel on a sample of rows (same slots: entry b of a row reads slot b mod P)
        rows = np.sort(rng.choice(n, 12, replace=False))
        direct = eng.accumulate(eng.plan_full(dim, rows)).clone()
        _close(acc[torch.as_tensor(rows, device=acc.device)], direct, "c2 mode %d: sweep vs direct rows" % dim)
        # chunking invariance
        eng.sweep_chunk_entries = 16
        eng._sweep_plans.clear()
        acc16 = eng.accumulate(eng.plan_sweep(dim)).clone()
        _close(acc16, acc, "c2 mode %d: 16-entry chunks vs 64" % dim)
        eng.sweep_chunk_entries = 64
        eng._sweep_plans.clear()
    # ---- one FULL-LENGTH row of every mode against the CPU oracle (the reference's own arithmetic, pinned by
    #      tests/golden): ~8000 entries, entry b of the row reading pool slot b mod P, accumulator record compared -------
    from oracle import ssvi_oracle as so
    from ssvi_tf_b200.engine import noise_offsets
    D, k1, P = c["D"], c["k1"], int(eng.noise.slots)
    off = noise_offsets(D, k1, False)                                       # [c, k] -> position inside a noise block
    z_other = eng.noise.z_other.double().cpu().numpy()[:, :, off]           # [P, N-1, D, k1]
    z_own = eng.noise.z_own.double().cpu().numpy()[:, off]                  # [P, D, k1]
    z_other, z_own = np.swapaxes(z_other, -1, -2), np.swapaxes(z_own, -1, -2)    # numpy-natural (.., k1, D)
    ocfg = so.make_config("d", "bernoulli", D, k1=k1, k2=c["k2"], batch_size=10 ** 9)
    ost = dict(params=[p.cpu().numpy() for p in eng.params])
    coords_h, vals_h = eng.coords.cpu().numpy().astype(np.int64), eng.vals.cpu().numpy()
    for dim in range(eng.N):
        acc = eng.accumulate(eng.plan_sweep(dim)).clone()
        rp = np.asarray(eng.rowptr_h[dim])
        row = int(np.argmax(rp[1:] - rp[:-1]))                              # the mode's longest row
        ids = eng.perm[dim][int(rp[row]):int(rp[row + 1])].cpu().numpy()
        sl = np.arange(len(ids)) % P
        tape = so.Tape(z_other[sl], z_own[sl])
        di, Di, si, scale, B = so.row_accumulate(ost, ocfg, dim, row, coords_h[ids], vals_h[ids], tape)
        got = acc[row].cpu().numpy()
        assert B == len(ids) == int(got[-1]) and scale == 1.0
        np.testing.assert_allclose(got[:D], di, rtol=1e-5, atol=1e-5 * np.abs(di).max(), err_msg="c2 mode %d di" % dim)
        np.testing.assert_allclose(got[D:D + D * D].reshape(D, D), Di, rtol=1e-5, atol=1e-5 * np.abs(Di).max(),
                                   err_msg="c2 mode %d Di" % dim)
    # additivity over two entry shards (mode 1)
    from ssvi_tf_b200.engine import Engine
    dim = 1
    whole = eng.accumulate(eng.plan_sweep(dim)).clone()
    half = eng.nnz // 2
    parts, first_counts = [], None
    for lo, hi in ((0, half), (half, eng.nnz)):
        sub = Engine(eng.cfg, eng.dims, eng.coords[lo:hi], eng.vals[lo:hi], device="cuda", noise_dtype=eng.noise_dtype)
        for d in range(eng.N):
            sub.params[d].copy_(eng.params[d])
        sub.refresh_sampler()
        sub.set_noise(eng.noise)
        off = None if first_counts is None else first_counts          # global position of the shard's first entry
        parts.append(sub.accumulate(sub.plan_sweep(dim, off)).clone())
        sub.check_flags()
        if first_counts is None:
            first_counts = np.diff(np.asarray(sub.rowptr_h[dim]))
        sub.close()
    _close(parts[0] + parts[1], whole, "c2 mode 1: two shards vs whole")
    np.testing.assert_array_equal((parts[0][:, -1] + parts[1][:, -1]).cpu().numpy(), whole[:, -1].cpu().numpy())
    m.engine.close()


def test_config4_shape_sweep_kernel_matches_generic_rows():
    """Config 4's kernel instance (4-way, D = 50, k1 = 64, rank compiled in) on the full index space at 2 % of the nnz:
    the entry-blocked sweep kernel against the generic CUDA-core kernel on a sample of rows of every mode."""
    from ssvi_tf_b200.engine import Engine
    c, data, m = _model("c4", scale=0.02)
    eng = m.engine
    assert eng.sweep_supported
    _check_index(eng)
    gen = Engine(eng.cfg, eng.dims, eng.coords, eng.vals, device="cuda", noise_dtype=eng.noise_dtype, kernel_path=1)
    for d in range(eng.N):
        gen.params[d].copy_(eng.params[d])
    gen.refresh_sampler()
    gen.w_state.copy_(eng.w_state)
    gen.set_noise(eng.noise)
    assert not gen.sweep_supported
    rng = np.random.default_rng(4)
    for dim in range(eng.N):
        acc = eng.accumulate(eng.plan_sweep(dim)).clone()
        eng.check_flags()
        lens = np.diff(np.asarray(eng.rowptr_h[dim]))
        np.testing.assert_array_equal(acc[:, -1].cpu().numpy(), lens.astype(np.float64))
        rows = np.sort(rng.choice(eng.dims[dim], 4, replace=False))
        want = gen.accumulate(gen.plan_full(dim, rows)).clone()
        _close(acc[torch.as_tensor(rows, device=acc.device)], want, "c4 mode %d: sweep vs generic rows" % dim)
    gen.close()
    eng.close()


@pytest.mark.parametrize("cfg_id,scale,kwargs", [
    ("c3", 0.01, None),                                   # robust + poisson, D = 30: the three-pass sample-table sweep
    ("c5", 0.01, dict(sweep_tile_blocks=2)),              # D = 32 (four row tiles), L2-tiled launch order
])
def test_config3_config5_shapes_sweep_matches_direct_rows(cfg_id, scale, kwargs):
    """Configs 3 and 5 on their full index spaces at 1 % of the nnz: the sweep path (three-pass for SSVI_TF_robust,
    L2-tiled launch order for the 10k^3 table) against the direct DMMA kernel on a sample of rows of every mode."""
    c, data, m = _model(cfg_id, scale=scale, engine_kwargs=kwargs)
    eng = m.engine
    assert eng.sweep_supported
    _check_index(eng)
    rng = np.random.default_rng(35)
    for dim in range(eng.N):
        plan = eng.plan_sweep(dim)
        if kwargs:
            assert plan.tile_blocks == kwargs["sweep_tile_blocks"]
        acc = eng.accumulate(plan).clone()
        eng.check_flags()
        lens = np.diff(np.asarray(eng.rowptr_h[dim]))
        np.testing.assert_array_equal(acc[:, -1].cpu().numpy(), lens.astype(np.float64))
        rows = np.sort(rng.choice(eng.dims[dim], 6, replace=False))
        direct = eng.accumulate(eng.plan_full(dim, rows)).clone()
        eng.check_flags()
        _close(acc[torch.as_tensor(rows, device=acc.device)], direct, "%s mode %d: sweep vs direct rows" % (cfg_id, dim))
    eng.close()
