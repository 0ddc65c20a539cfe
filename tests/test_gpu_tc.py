"""GPU: the fp32 ARITHMETIC option of the sweep path (SSVI_TABLE_F32_TC, csrc/accumulate_tc.cuh): V, f and the row sums in
fp32, the contraction on the tcgen05 tensor cores (fp16 hi / lo split, fp32 accumulation in tensor memory).  Bar:
BASELINE.json's fp32 tolerance, 1e-4 relative on the updated means / covariances / optimiser state against the oracle;
counts bit-exact.  The accumulator records are also compared with the all-fp64 sweep to show the actual deviation."""
import numpy as np
import pytest
import torch

from oracle import ssvi_oracle as so
from tests.helpers import (assert_state_close, engine_dump_state, engine_load_state, random_problem, random_state)

pytestmark = pytest.mark.gpu


def _run(cfg, dims, coords, vals, st0, pool, dim, W, table_dtype, blocks=0, **kw):
    from ssvi_tf_b200.engine import Engine, NoisePool
    eng = Engine(cfg, dims, coords, vals, device="cuda", noise_dtype=torch.float64, sweep_chunk_entries=W,
                 table_budget_bytes=1, table_dtype=table_dtype, sweep_tile_blocks=blocks, sweep_tile_min_entries=0.0, **kw)
    eng.set_noise(NoisePool.from_tape(pool, eng.device, torch.float64, diag=False))
    engine_load_state(eng, st0)
    plan = eng.plan_sweep(dim)
    acc = eng.accumulate(plan).cpu().numpy().copy()
    eng.row_update(plan, eng.accumulate(plan))
    eng.check_flags()
    state = engine_dump_state(eng, st0)
    eng.close()
    return acc, state, plan


@pytest.mark.parametrize("dims,D,lik,P,W", [
    ([6, 5, 4], 20, "bernoulli", 4, 64),      # config 2's kernel instance (NT = 3)
    ([5, 4, 4], 32, "normal", 5, 3),          # config 5's (NT = 4), segments of 3 entries (ragged quads)
    ([4, 3, 3, 3], 10, "normal", 3, 64),      # config 1's rank, three other modes
    ([40, 3, 3], 12, "poisson", 1, 200),      # chunks longer than 128 entries: the 32-quad batches roll over
    ([7, 9], 7, "bernoulli", 2, 4),           # one other mode, one row group
    ([3, 30, 30], 32, "bernoulli", 2, 1000),  # rows of ~360 entries in one chunk: many quads through both buffers
    # ranks 33..64: two warps per entry (accumulate_tc_wide_kernel)
    ([6, 5, 4], 50, "bernoulli", 3, 64),      # the reference's real-data default rank (NT = 7: the upper warp has 3 groups)
    ([5, 4, 4], 64, "normal", 2, 3),          # NT = 8, segments of 3 entries (odd pairs)
    ([4, 3, 3, 3], 33, "normal", 3, 64),      # NT = 5 (one row group in the upper half), three other modes
    ([40, 3, 3], 40, "poisson", 1, 200),      # NT = 5, chunks longer than 64 entries: the 32-pair batches roll over
    ([7, 9], 48, "bernoulli", 2, 5),          # NT = 6, one other mode
])
def test_tensor_core_sweep_is_within_the_fp32_tolerance(dims, D, lik, P, W, capsys):
    rng = np.random.default_rng(D * 131 + P)
    N = len(dims)
    cfg = so.make_config("d", lik, D, diag=False, k1=64, k2=4, batch_size=10 ** 9)
    coords, vals = random_problem(rng, lik, dims, int(0.8 * np.prod(dims)))
    st0 = random_state(rng, cfg, dims)
    rowptr, perm = so.build_index(coords, dims)
    pool = so.make_tape(rng, cfg, N, P)
    worst, fp64_bar = 0.0, True
    for dim in range(N):
        n = dims[dim]
        acc64, _, plan64 = _run(cfg, dims, coords, vals, st0, pool, dim, W, torch.float64)
        acc_tc, state_tc, plan_tc = _run(cfg, dims, coords, vals, st0, pool, dim, W, "f32_tc")
        # a float table, which also carries the swept mode's own samples U(row, slot)
        assert plan_tc.table_bytes * 2 * (sum(dims) - n) == plan64.table_bytes * sum(dims)
        dev = np.abs(acc_tc - acc64).max() / np.abs(acc64).max()
        worst = max(worst, dev)
        assert dev < 2e-5, dev
        np.testing.assert_array_equal(acc_tc[:, -1], acc64[:, -1])                # counts: bit-exact
        entries, tapes = [], []
        for row in range(n):
            ids = perm[dim][rowptr[dim][row]:rowptr[dim][row + 1]]
            sl = np.arange(len(ids)) % P
            tapes.append(so.Tape(pool.z_other[sl], pool.z_own[sl], pool.r_w[sl], pool.z_f[sl]))
            entries.append((coords[ids], vals[ids]))
        want = so.copy_state(st0)
        so.rows_step(want, cfg, dim, np.arange(n), entries, tapes)
        # fp32 bar: 1e-4 relative; elements below 1 % of an array's largest are held to 1e-6 of that scale (an fp32 product
        # of three samples cannot do better on a near-zero covariance element)
        assert_state_close(state_tc, want, cfg, 1e-4, what="tensor-core sweep, dims=%s mode %d" % (dims, dim), atol_frac=1e-2)
        try:                                             # informational: does the update also meet the fp64 bar?
            assert_state_close(state_tc, want, cfg, 1e-5)
        except AssertionError:
            fp64_bar = False
    with capsys.disabled():
        print("\n  tcgen05 sweep vs fp64 sweep, largest accumulator deviation / max |acc|: %.2e (dims %s, D %d); updated "
              "rows within 1e-5 of the oracle as well: %s" % (worst, dims, D, fp64_bar))


def test_tensor_core_sweep_under_l2_tiling_and_on_an_f32_pool():
    """The launch-order variants the bench uses at config 5 (L2 tiles, float noise pool) change nothing for this kernel:
    same entry set, accumulators equal to the untiled run up to summation order."""
    from ssvi_tf_b200.engine import Engine, NoisePool
    rng = np.random.default_rng(7)
    dims, D, P = [7, 9, 8], 32, 3
    cfg = so.make_config("d", "normal", D, diag=False, k1=64, k2=4, batch_size=10 ** 9)
    coords, vals = random_problem(rng, "normal", dims, int(0.7 * np.prod(dims)))
    st0 = random_state(rng, cfg, dims)
    pool = so.make_tape(rng, cfg, 3, P)
    for dim in range(3):
        a1, _, _ = _run(cfg, dims, coords, vals, st0, pool, dim, 64, "f32_tc")
        a2, _, _ = _run(cfg, dims, coords, vals, st0, pool, dim, 64, "f32_tc", blocks=2)
        np.testing.assert_allclose(a2, a1, rtol=0, atol=2e-6 * np.abs(a1).max())
        np.testing.assert_array_equal(a2[:, -1], a1[:, -1])


def test_tensor_core_sweep_tuning_variants_give_the_same_records():
    """SSVI_TUNE_TC_RING (records through a cp.async.bulk ring in shared memory) changes how the table records travel, not
    what is computed: bitwise the same accumulators."""
    rng = np.random.default_rng(5)
    for dims, D, lik, P, W in (([5, 4, 4], 32, "normal", 5, 3), ([40, 3, 3], 12, "poisson", 1, 200),
                               ([4, 3, 3, 3], 10, "normal", 3, 64), ([7, 9], 20, "bernoulli", 2, 64)):
        cfg = so.make_config("d", lik, D, diag=False, k1=64, k2=4, batch_size=10 ** 9)
        coords, vals = random_problem(rng, lik, dims, int(0.8 * np.prod(dims)))
        st0 = random_state(rng, cfg, dims)
        pool = so.make_tape(rng, cfg, len(dims), P)
        for dim in range(len(dims)):
            a0, _, _ = _run(cfg, dims, coords, vals, st0, pool, dim, W, "f32_tc")
            for tuning in (16, 32):
                a1, _, _ = _run(cfg, dims, coords, vals, st0, pool, dim, W, "f32_tc", tuning=tuning)
                np.testing.assert_array_equal(a1, a0, err_msg="tuning %d, dims %s mode %d" % (tuning, dims, dim))


def test_tensor_core_sweep_flags_an_fp16_overflow():
    """An operand outside the fp16 window must not pass silently: the chunk record goes non-finite and check_flags raises."""
    from ssvi_tf_b200.engine import Engine, NoisePool
    rng = np.random.default_rng(11)
    dims, D, P = [5, 4, 4], 16, 2
    cfg = so.make_config("d", "normal", D, diag=False, k1=64, k2=4, batch_size=10 ** 9)
    coords, vals = random_problem(rng, "normal", dims, 60)
    st0 = random_state(rng, cfg, dims)
    pool = so.make_tape(rng, cfg, 3, P)
    eng = Engine(cfg, dims, coords, vals, device="cuda", noise_dtype=torch.float64, table_dtype="f32_tc", tc_scale_log2=14)
    eng.set_noise(NoisePool.from_tape(pool, eng.device, torch.float64, diag=False))
    engine_load_state(eng, st0)
    for d in range(3):                                   # samples of magnitude ~30: 2^14 * 30 * 30 is far beyond 65504
        eng.params[d][:, 0, :] *= 30.0
    eng.refresh_sampler()
    eng.accumulate(eng.plan_sweep(0))
    with pytest.raises(AssertionError, match="fp16 range"):
        eng.check_flags()
    eng.close()


def test_nation_full_sweeps_with_the_tensor_core_option(tmp_path):
    """End to end on the reference's own dataset (Baseline/Data/nation.txt, the run of real_test.py:58-84 at rank 20,
    k1 = 64): full-row sweeps of every mode through the sample-table path, once all-fp64 and once with the fp32 / tcgen05
    option, same split and same variates per seed.  The option must land on the fp64 run's test metrics (1 % per seed:
    only the arithmetic differs) and must not be worse than the UNMODIFIED reference's (1 % + three standard errors)."""
    import gzip
    import os
    import shutil
    import ssvi_tf_b200 as pkg
    e2e = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden_e2e")
    z = np.load(os.path.join(e2e, "nation_d_full_metrics.npz"))
    path = str(tmp_path / "nation.txt")
    with gzip.open(os.path.join(e2e, "nation.txt.gz"), "rb") as f, open(path, "wb") as g:
        shutil.copyfileobj(f, g)
    seeds = [int(s) for s in z["seeds"]][:6]
    sweeps = 12                                          # ~ the reference's 300 one-row-per-mode iterations / mean(dims)
    res = {}
    for table in (torch.float64, "f32_tc"):
        out = []
        for seed in seeds:
            np.random.seed(seed)
            t = pkg.binary_tensor()
            t.read_from_file(path, 0.8, 0, 0.2)
            m = pkg.SSVI_TF_d(tensor=t, rank=int(z["rank"]), mean_update="S", cov_update="N", eta=1.0, cov_eta=1.0,
                              k1=int(z["k1"]), k2=int(z["k2"]), diag=False, noise_dtype=torch.float32, noise_seed=seed,
                              noise_slots=128, engine_kwargs={"table_dtype": table})
            assert m.engine.sweep_supported and m.engine.table_dtype == table
            m.factorize(report=10 ** 9, max_iteration=sweeps - 1, detailed_report=False, rows_per_iteration="all")
            out.append(m.evaluate_test_error())
        res[table] = np.array(out)
    r64, rtc = res[torch.float64], res["f32_tc"]
    np.testing.assert_allclose(rtc[:, 0], r64[:, 0], rtol=1e-2)                   # test RMSE, seed by seed
    assert np.abs(rtc[:, 1] - r64[:, 1]).max() <= 0.01                            # test error rate: 1 % absolute
    ref = z["ref_test_rsme"][:len(seeds)]
    d = rtc[:, 0] - ref
    assert d.mean() <= 0.01 * ref.mean() + 3.0 * d.std(ddof=1) / np.sqrt(len(d)), (rtc[:, 0].tolist(), ref.tolist())


@pytest.mark.parametrize("cfg_id,scale,kwargs", [
    ("c2", 1.0, {}),                                      # config 2 at FULL size: 8e6 train entries, rank 20 (NT = 3)
    ("c5", 0.01, dict(sweep_tile_blocks=2)),              # config 5's 10k^3 index space, rank 32, L2-tiled launch order
    ("c2", 0.25, dict(rank=50)),                          # the reference's real-data default rank on config 2's tensor
])
def test_tensor_core_sweep_at_baseline_sizes_matches_the_fp64_direct_kernel(cfg_id, scale, kwargs):
    """At BASELINE.json's sizes the oracle cannot follow; the tensor-core sweep is compared with the fp64 direct-evaluation
    kernel (pinned to the oracle and the reference's fixtures at small sizes) on a sample of rows of every mode, with the
    bench's float noise pool: within 2e-5 of the largest accumulator element, counts bit-exact over ALL rows."""
    import bench
    kwargs = dict(kwargs)
    rank = kwargs.pop("rank", None)
    c = bench.CONFIGS[cfg_id]
    if rank:
        c = dict(c, D=rank)
    data = bench.DeviceTensorData(c, "cuda", 319 + int(cfg_id[1]), scale)
    m = bench.build_model(c, data, "cuda", None, 16, 128, engine_kwargs=dict(kwargs, table_dtype="f32_tc"))
    m.fresh_noise()
    eng = m.engine
    assert eng.sweep_supported and eng.table_dtype == "f32_tc"
    rng = np.random.default_rng(36)
    for dim in range(eng.N):
        acc = eng.accumulate(eng.plan_sweep(dim)).clone()
        eng.check_flags()
        lens = np.diff(np.asarray(eng.rowptr_h[dim]))
        np.testing.assert_array_equal(acc[:, -1].cpu().numpy(), lens.astype(np.float64))
        rows = np.sort(rng.choice(eng.dims[dim], 6, replace=False))
        direct = eng.accumulate(eng.plan_full(dim, rows)).clone().cpu().numpy()
        eng.check_flags()
        got = acc[torch.as_tensor(rows, device=acc.device)].cpu().numpy()
        np.testing.assert_allclose(got, direct, rtol=0, atol=2e-5 * np.abs(direct).max(),
                                   err_msg="%s mode %d: tensor-core sweep vs direct rows" % (cfg_id, dim))
    eng.close()


@pytest.mark.parametrize("model,lik,dims,D,P,W", [
    ("simple", "normal", [6, 5, 4, 3], 50, 3, 64),      # config 4's shape: 4-way, rank 50
    ("simple", "normal", [9, 7, 5], 13, 2, 5),          # odd rank (a padded component), short segments, ragged groups of four
    ("simple", "normal", [40, 3, 3], 64, 1, 200),       # chunks longer than 128 entries
    ("simple", "normal", [7, 9], 8, 4, 64),             # one other mode
    ("d", "bernoulli", [6, 5, 4, 3], 50, 3, 64),        # SSVI_TF_d: weights uniform in k (accumulate_diag_d32_kernel)
    ("d", "poisson", [9, 7, 5], 13, 2, 5),
    ("d", "normal", [40, 3, 3], 64, 1, 200),
])
def test_fp32_option_diagonal_kernels(model, lik, dims, D, P, W, capsys):
    """The fp32 arithmetic option on the diagonal-covariance sweep (csrc/accumulate_diag32.cuh): SSVI_TF_simple + normal
    likelihood (BASELINE config 4's kernel instance, closed-form weights) and SSVI_TF_d with every likelihood.  Chunk
    records against the fp64 kernels', updated state (incl. the w_sigma recurrence) against the oracle at the fp32 bar,
    counts bit-exact."""
    from ssvi_tf_b200.engine import Engine, NoisePool
    rng = np.random.default_rng(D * 17 + P)
    N = len(dims)
    cfg = so.make_config(model, lik, D, diag=True, k1=64, k2=8, batch_size=10 ** 9)
    coords, vals = random_problem(rng, lik, dims, int(0.8 * np.prod(dims)))
    st0 = random_state(rng, cfg, dims)
    rowptr, perm = so.build_index(coords, dims)
    pool = so.make_tape(rng, cfg, N, P)
    worst = 0.0
    for dim in range(N):
        n = dims[dim]
        accs, states = [], []
        for table in (torch.float64, "f32_tc"):
            eng = Engine(cfg, dims, coords, vals, device="cuda", noise_dtype=torch.float64, sweep_chunk_entries=W,
                         table_dtype=table)
            eng.set_noise(NoisePool.from_tape(pool, eng.device, torch.float64, diag=True))
            engine_load_state(eng, st0)
            plan = eng.plan_sweep(dim)
            accs.append(eng.accumulate(plan).cpu().numpy().copy())
            eng.step(plan)
            eng.check_flags()
            states.append(engine_dump_state(eng, st0))
            eng.close()
        dev = np.abs(accs[1] - accs[0]).max() / np.abs(accs[0]).max()
        worst = max(worst, dev)
        assert dev < 2e-5, dev
        np.testing.assert_array_equal(accs[1][:, -1], accs[0][:, -1])
        entries, tapes = [], []
        for row in range(n):
            ids = perm[dim][rowptr[dim][row]:rowptr[dim][row + 1]]
            sl = np.arange(len(ids)) % P
            tapes.append(so.Tape(pool.z_other[sl], pool.z_own[sl], pool.r_w[sl], pool.z_f[sl]))
            entries.append((coords[ids], vals[ids]))
        want = so.copy_state(st0)
        so.rows_step(want, cfg, dim, np.arange(n), entries, tapes)
        assert_state_close(states[1], want, cfg, 1e-4, what="fp32 diagonal kernel, dims=%s mode %d" % (dims, dim), atol_frac=1e-2)
    with capsys.disabled():
        print("\n  fp32 diagonal kernel (%s, %s) vs fp64, largest record deviation / max |record|: %.2e (dims %s, D %d)"
              % (model, lik, worst, dims, D))


@pytest.mark.parametrize("table", [torch.float64, "f32_tc"])
@pytest.mark.parametrize("D", [20, 50])
def test_sample_table_covers_the_row_window_only(table, D):
    """A shard cut along index-space blocks touches a sub-range of every mode (ssvi_set_row_windows): the sweep path samples
    its table for those rows only.  Entries confined to rows [2, 7) x [3, 9) x [1, 6) of a 9 x 11 x 8 tensor: smaller table,
    and the same accumulators as the direct kernel (which uses no table) on every row."""
    from ssvi_tf_b200.engine import Engine, NoisePool
    rng = np.random.default_rng(3 + D)
    dims, lo, sub, P = [9, 11, 8], [2, 3, 1], [5, 6, 5], 3
    cfg = so.make_config("d", "bernoulli", D, diag=False, k1=64, k2=4, batch_size=10 ** 9)
    coords, vals = random_problem(rng, "bernoulli", sub, int(0.8 * np.prod(sub)))
    coords = coords + np.asarray(lo)[None, :]
    st0 = random_state(rng, cfg, dims)
    pool = so.make_tape(rng, cfg, 3, P)
    eng = Engine(cfg, dims, coords, vals, device="cuda", noise_dtype=torch.float64, table_dtype=table, table_budget_bytes=1)
    assert [w for w in eng.row_window] == [(l, l + n - 1) for l, n in zip(lo, sub)]
    eng.set_noise(NoisePool.from_tape(pool, eng.device, torch.float64, diag=False))
    engine_load_state(eng, st0)
    rec = 8 * ((D + 7) // 8) * 64 * (4 if table == "f32_tc" else 8)               # bytes of one table record
    for dim in range(3):
        plan = eng.plan_sweep(dim)
        rows_in_table = sum(sub) - (0 if table == "f32_tc" else sub[dim])         # the tensor-core table holds U as well
        assert plan.table_bytes == rows_in_table * rec * plan.groups[0][1]
        acc = eng.accumulate(plan).cpu().numpy().copy()
        direct = eng.accumulate(eng.plan_full(dim)).cpu().numpy().copy()
        eng.check_flags()
        tol = 1e-9 if table != "f32_tc" else 2e-5
        np.testing.assert_allclose(acc, direct, rtol=0, atol=tol * np.abs(direct).max())
    eng.close()
