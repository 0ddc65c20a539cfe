"""GPU: the CUDA path (through the C ABI) against the golden fixtures (reference outputs) and against
the oracle on seeded inputs.  Parity bar (BASELINE.json): fp64, 1e-5 relative on the updated means and
covariances; row ids / counts bit-exact."""
import numpy as np
import pytest
import torch

from oracle import ssvi_oracle as so
from tests.helpers import (assert_state_close, engine_dump_state, engine_load_state, golden_cases, load_golden,
                           random_problem, random_state)

pytestmark = pytest.mark.gpu
RTOL = 1e-5


def _engine(cfg, dims, coords, vals, noise_dtype=torch.float64, chunk_entries=4):
    from ssvi_tf_b200.engine import Engine
    return Engine(cfg, dims, coords, vals, device="cuda", noise_dtype=noise_dtype, chunk_entries=chunk_entries)


@pytest.mark.parametrize("noise_dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("name", golden_cases())
def test_cuda_matches_reference_golden(name, noise_dtype):
    from ssvi_tf_b200.engine import NoisePool
    case = load_golden(name)
    cfg, dims = case["cfg"], case["dims"]
    eng = _engine(cfg, dims, case["coords"], case["vals"], noise_dtype)
    for k, step in enumerate(case["steps"]):
        engine_load_state(eng, case["state0"])
        dim, row, tape = step["dim"], step["row"], step["tape"]
        eng.set_noise(NoisePool.from_tape(tape, eng.device, noise_dtype, diag=cfg["diag"]))
        sub = None if tape.sub_idx is None else {row: tape.sub_idx}
        plan = eng.plan_minibatch(dim, [row], cfg["batch_size"], sub_idx=sub)
        # touched rows / counts are integers: exact
        assert plan.n_rows == 1 and int(plan.batch_sizes[0]) == tape.z_own.shape[0]
        eng.step(plan, fixed_covariance=case["state0"]["fixed_covariance"])
        eng.check_flags()
        got = engine_dump_state(eng, case["state0"])
        assert_state_close(got, step["after"], cfg, RTOL, what="%s step %d" % (name, k))
    eng.close()


@pytest.mark.parametrize("model,lik,diag", [("d", "normal", False), ("simple", "bernoulli", False),
                                            ("robust", "poisson", False), ("simple", "normal", True),
                                            ("robust", "bernoulli", True), ("d", "poisson", True)])
def test_cuda_batched_rows_match_oracle(model, lik, diag):
    """A whole mode (all rows, ragged lengths incl. empty rows) in ONE launch == the oracle's
    rows_step (the equivalence class of SURVEY.md 8b)."""
    from ssvi_tf_b200.engine import NoisePool
    rng = np.random.default_rng(11)
    dims, D, k1, k2, B = [9, 7, 6], 7, 16, 8, 8
    cfg = so.make_config(model, lik, D, diag=diag, k1=k1, k2=k2, batch_size=B)
    coords, vals = random_problem(rng, lik, dims, 150)
    coords = coords[coords[:, 0] != 3]                 # leave row 3 of mode 0 empty
    vals = vals[:len(coords)]
    st0 = random_state(rng, cfg, dims)
    rowptr, perm = so.build_index(coords, dims)
    eng = _engine(cfg, dims, coords, vals, chunk_entries=3)
    pool = so.make_tape(rng, cfg, 3, B)                # B slots: entry b of every row uses slot b
    eng.set_noise(NoisePool.from_tape(pool, eng.device, torch.float64, diag=diag))
    for dim in range(3):
        engine_load_state(eng, st0)
        rows = np.arange(dims[dim])
        sub, entries, tapes = {}, [], []
        for row in rows:
            ids = perm[dim][rowptr[dim][row]:rowptr[dim][row + 1]]
            n = len(ids)
            t = so.Tape(pool.z_other[:min(n, B)], pool.z_own[:min(n, B)], pool.r_w[:min(n, B)], pool.z_f[:min(n, B)])
            if n > B:
                t.sub_idx = rng.choice(n, B, replace=False)
                sub[int(row)] = t.sub_idx
            entries.append((coords[ids], vals[ids]))
            tapes.append(t)
        plan = eng.plan_minibatch(dim, rows, B, sub_idx=sub)
        eng.step(plan)
        eng.check_flags()
        want = so.copy_state(st0)
        touched = so.rows_step(want, cfg, dim, rows, entries, tapes)
        got = engine_dump_state(eng, st0)
        assert_state_close(got, want, cfg, RTOL, what="mode %d" % dim)
        # bit-exact bookkeeping: which rows were touched, and how many entries each consumed
        acc = eng._scratch["acc"][:plan.n_rows * eng.acc_len].view(plan.n_rows, eng.acc_len).cpu().numpy()
        counts = acc[:, -1].astype(np.int64)
        assert [int(r) for r in rows[counts > 0]] == touched
        want_counts = np.minimum(np.diff(rowptr[dim]), B)
        np.testing.assert_array_equal(counts, want_counts)
        if dim == 0:
            assert counts[3] == 0
            np.testing.assert_array_equal(got["params"][0][3], st0["params"][0][3])
        # hyper-parameter + stop-condition reductions
        eng.hyper_reduce(dim)
        so.update_hyper_parameter(want, cfg, dim)
        np.testing.assert_allclose(eng.pSigma.cpu().numpy()[dim], want["pSigma"][dim], rtol=RTOL)
        ho = eng.hyper_out[dim].cpu().numpy()
        np.testing.assert_allclose(ho[1:], want["norm_changes"][dim].max(axis=0), rtol=RTOL)
    eng.close()


def test_full_row_sweep_chunking_is_invariant():
    """Splitting rows into chunks of different sizes changes only the summation order."""
    from ssvi_tf_b200.engine import NoisePool
    rng = np.random.default_rng(3)
    dims, D = [5, 40, 30], 6
    cfg = so.make_config("d", "normal", D, k1=32, k2=4, batch_size=10 ** 9)
    coords, vals = random_problem(rng, "normal", dims, 900)
    st0 = random_state(rng, cfg, dims)
    outs = []
    for ce in (1, 7, 64):
        eng = _engine(cfg, dims, coords, vals, chunk_entries=ce)
        eng.set_noise(NoisePool.random(16, 3, D, 32, 4, False, eng.device, torch.float64, seed=1, diag=False))
        engine_load_state(eng, st0)
        acc = eng.accumulate(eng.plan_full(0)).cpu().numpy().copy()
        outs.append(acc)
        eng.close()
    np.testing.assert_allclose(outs[1], outs[0], rtol=1e-11, atol=1e-12)
    np.testing.assert_allclose(outs[2], outs[0], rtol=1e-11, atol=1e-12)
    np.testing.assert_array_equal(outs[0][:, -1], np.bincount(coords[:, 0], minlength=5))


def test_predict_mean_matches_oracle():
    rng = np.random.default_rng(4)
    dims, D = [6, 5, 4, 3], 9
    cfg = so.make_config("d", "normal", D, diag=True, k1=8, k2=4)
    coords, vals = random_problem(rng, "normal", dims, 100)
    st0 = random_state(rng, cfg, dims)
    eng = _engine(cfg, dims, coords, vals)
    engine_load_state(eng, st0)
    got = eng.predict_mean(eng.coords).cpu().numpy()
    np.testing.assert_allclose(got, so.predict_mean(st0, coords), rtol=1e-12)
    eng.close()


@pytest.mark.parametrize("D,k1,model,lik", [(12, 64, "d", "bernoulli"), (20, 64, "d", "bernoulli"),
                                            (20, 24, "robust", "normal"), (27, 40, "simple", "poisson"),
                                            (32, 64, "d", "normal"), (30, 64, "robust", "poisson")])
def test_mma_path_matches_oracle_and_generic(D, k1, model, lik):
    """The FP64 tensor-pipe kernel (every n-tile count) == the generic CUDA-core kernel == the oracle."""
    from ssvi_tf_b200.engine import Engine, NoisePool
    rng = np.random.default_rng(D * 100 + k1)
    dims, B, k2 = [6, 5, 4], 6, 8
    cfg = so.make_config(model, lik, D, diag=False, k1=k1, k2=k2, batch_size=B)
    coords, vals = random_problem(rng, lik, dims, 70)
    st0 = random_state(rng, cfg, dims)
    rowptr, perm = so.build_index(coords, dims)
    pool = so.make_tape(rng, cfg, 3, B)
    results = []
    for path in (0, 1):
        eng = Engine(cfg, dims, coords, vals, device="cuda", noise_dtype=torch.float32, chunk_entries=4,
                     kernel_path=path)
        eng.set_noise(NoisePool.from_tape(pool, eng.device, torch.float32, diag=False))
        engine_load_state(eng, st0)
        dim = 1
        rows = np.arange(dims[dim])
        sub = {}
        entries, tapes = [], []
        for row in rows:
            ids = perm[dim][rowptr[dim][row]:rowptr[dim][row + 1]]
            n = min(len(ids), B)
            t = so.Tape(pool.z_other[:n], pool.z_own[:n], pool.r_w[:n], pool.z_f[:n])
            if len(ids) > B:
                t.sub_idx = np.arange(len(ids))[::-1][:B].copy()
                sub[int(row)] = t.sub_idx
            entries.append((coords[ids], vals[ids]))
            tapes.append(t)
        plan = eng.plan_minibatch(dim, rows, B, sub_idx=sub)
        acc = eng.accumulate(plan).cpu().numpy().copy()
        eng.row_update(plan, eng.accumulate(plan))
        eng.check_flags()
        results.append((acc, engine_dump_state(eng, st0)))
        eng.close()
    want = so.copy_state(st0)
    so.rows_step(want, cfg, dim, rows, entries, tapes)
    for acc, got in results:
        assert_state_close(got, want, cfg, RTOL, what="D=%d" % D)
    scale = np.abs(results[1][0]).max()
    np.testing.assert_allclose(results[0][0], results[1][0], rtol=1e-9, atol=1e-12 * scale)
    # Di block of the accumulator record is exactly symmetric
    Di = results[0][0][:, D:D + D * D].reshape(-1, D, D)
    np.testing.assert_array_equal(Di, np.transpose(Di, (0, 2, 1)))


@pytest.mark.parametrize("dims,D,k1,lik,P,W,offs,model", [
    ([6, 5, 4], 20, 64, "bernoulli", 4, 64, False, "d"),     # BASELINE config 2's kernel instance (NT=3, two other modes)
    ([6, 5, 4], 12, 64, "normal", 3, 2, True, "d"),          # segments of 2 entries, shifted slots (multi-GPU shard)
    ([5, 4, 4], 32, 64, "normal", 5, 3, False, "d"),         # NT=4, D % 8 == 0 (config 5): row sums by DADD
    ([5, 4, 4], 30, 64, "poisson", 4, 64, True, "d"),
    ([5, 4, 4], 16, 64, "bernoulli", 2, 5, False, "d"),      # D % 8 == 0, NT=2
    ([7, 9], 7, 64, "bernoulli", 2, 4, False, "d"),          # matrix: one other mode
    ([4, 3, 3, 3], 10, 64, "normal", 3, 64, False, "d"),     # 4-way: three other modes (config 1's NT=2)
    ([40, 3, 3], 20, 64, "normal", 1, 200, False, "d"),      # one slot, chunks longer than 128 entries
    # 32 < D <= 64 (accumulate_wide.cuh: the contraction split over the CTA's warps); the reference's real-data rank is 50
    ([6, 5, 4], 50, 64, "bernoulli", 3, 64, False, "d"),     # NT = 7 (odd: warp 0 owns one tile row, the others two)
    ([5, 4, 4], 64, 64, "normal", 2, 3, True, "d"),          # NT = 8, every lane a live row, segments of 3, shifted slots
    ([7, 9], 40, 64, "poisson", 2, 5, False, "d"),           # NT = 5, one other mode
    ([4, 3, 3, 3], 33, 64, "normal", 2, 64, False, "d"),     # NT = 5 with 7 padding rows, three other modes
    ([40, 3, 3], 48, 64, "normal", 1, 200, False, "d"),      # NT = 6, chunks longer than 32 entries (second index batch)
    ([5, 4, 4], 30, 64, "poisson", 4, 64, False, "robust"),  # BASELINE config 3's kernel instances (three-pass path)
    ([6, 5, 4], 20, 64, "bernoulli", 3, 3, True, "robust"),
    ([6, 5, 4], 12, 64, "normal", 2, 64, False, "simple"),
    ([4, 3, 3, 3], 9, 64, "poisson", 3, 5, True, "simple"),
    ([7, 9], 32, 64, "normal", 2, 64, False, "robust"),
    ([5, 4, 4], 20, 64, "poisson", 3, 64, False, "robust+quad"),   # SSVI_TF_robust(quadrature=True), every kernel path
    ([6, 5, 4], 12, 64, "bernoulli", 2, 3, True, "robust+quad"),
])
def test_sweep_table_path_matches_direct_paths_and_oracle(dims, D, k1, lik, P, W, offs, model):
    """Full sweep through the slot-major sample-table kernels == the direct DMMA kernel == the generic CUDA-core
    kernel (same accumulators to summation order) == the oracle's rows_step on the same pooled noise."""
    from ssvi_tf_b200.engine import Engine, NoisePool
    rng = np.random.default_rng(D * 1000 + k1 + P)
    N = len(dims)
    quad = 30 if model.endswith("+quad") else 0
    cfg = so.make_config(model.split("+")[0], lik, D, diag=False, k1=k1, k2=4, batch_size=10 ** 9, quadrature=quad)
    nnz = int(0.8 * np.prod(dims))
    coords, vals = random_problem(rng, lik, dims, nnz)
    st0 = random_state(rng, cfg, dims)
    rowptr, perm = so.build_index(coords, dims)
    pool = so.make_tape(rng, cfg, N, P)
    for dim in range(N):
        n = dims[dim]
        off = rng.integers(0, 50, size=n) if offs else None
        accs, states = [], []
        for path, sweep in ((0, True), (0, False), (1, False)):
            eng = Engine(cfg, dims, coords, vals, device="cuda", noise_dtype=torch.float64, chunk_entries=3,
                         kernel_path=path, sweep_chunk_entries=W, table_budget_bytes=1)   # 1 slot per call
            eng.set_noise(NoisePool.from_tape(pool, eng.device, torch.float64, diag=False))
            engine_load_state(eng, st0)
            assert eng.sweep_supported == (path == 0)
            plan = eng.plan_sweep(dim, off) if sweep else eng.plan_full(dim, None, off)
            if sweep:
                assert len(plan.groups) == P
                np.testing.assert_array_equal(np.sort(plan.entry_ids.cpu().numpy()), np.sort(perm[dim]))
            acc = eng.accumulate(plan).cpu().numpy().copy()
            if sweep and model == "d":
                # the SSVI_TUNE_* kernel variants (include/ssvi_b200.h) must give the same records: L2 prefetch, the
                # linear-weight kernel of the normal likelihood, deeper register prefetch (rank 25..32, order 3)
                for tuning in (1, 2, 3, 4, 8, 2 | 4, 2 | 8):
                    eng.set_tuning(tuning)
                    alt = eng.accumulate(plan).cpu().numpy()
                    np.testing.assert_allclose(alt, acc, rtol=1e-9, atol=1e-12 * np.abs(acc).max(),
                                               err_msg="tuning %d" % tuning)
                    np.testing.assert_array_equal(alt[:, -1], acc[:, -1])
                eng.set_tuning(0)
                eng.check_flags()
            eng.row_update(plan, eng.accumulate(plan))
            eng.check_flags()
            accs.append(acc)
            states.append(engine_dump_state(eng, st0))
            eng.close()
        scale = np.abs(accs[2]).max()
        np.testing.assert_allclose(accs[0], accs[2], rtol=1e-9, atol=1e-12 * scale)
        np.testing.assert_allclose(accs[1], accs[2], rtol=1e-9, atol=1e-12 * scale)
        np.testing.assert_array_equal(accs[0][:, -1], np.diff(rowptr[dim]))        # counts: bit-exact
        Di = accs[0][:, D:D + D * D].reshape(-1, D, D)
        np.testing.assert_array_equal(Di, np.transpose(Di, (0, 2, 1)))
        # oracle: entry b of a row reads pool slot (b + offset) mod P
        entries, tapes = [], []
        for row in range(n):
            ids = perm[dim][rowptr[dim][row]:rowptr[dim][row + 1]]
            sl = (np.arange(len(ids)) + (0 if off is None else int(off[row]))) % P
            tapes.append(so.Tape(pool.z_other[sl], pool.z_own[sl], pool.r_w[sl], pool.z_f[sl]))
            entries.append((coords[ids], vals[ids]))
        want = so.copy_state(st0)
        so.rows_step(want, cfg, dim, np.arange(n), entries, tapes)
        for got in states:
            assert_state_close(got, want, cfg, RTOL, what="dims=%s mode %d" % (dims, dim))


@pytest.mark.parametrize("dims,D,lik,P,W", [
    ([6, 5, 4], 20, "bernoulli", 4, 64),      # config 2's kernel instance (NT = 3, ones row)
    ([5, 4, 4], 32, "normal", 5, 3),          # config 5's (NT = 4), segments of 3 entries
    ([4, 3, 3, 3], 10, "normal", 3, 64),      # config 1's rank, three other modes
    ([40, 3, 3], 12, "poisson", 1, 200),      # chunks longer than 128 entries
    ([7, 9], 7, "bernoulli", 2, 4),           # one other mode, odd rank
])
def test_sweep_float_table_is_within_the_fp32_tolerance(dims, D, lik, P, W, capsys):
    """SSVI_TABLE_F32 (the contract's fp32 option): the sample table of the sweep path stored as floats, arithmetic in
    fp64.  Updated means / covariances / optimiser state must agree with the oracle within 1e-4 (BASELINE.json's fp32
    bar); the accumulator records are compared with the double table's to show the actual deviation (~1e-7)."""
    from ssvi_tf_b200.engine import Engine, NoisePool
    rng = np.random.default_rng(D * 77 + P)
    N = len(dims)
    cfg = so.make_config("d", lik, D, diag=False, k1=64, k2=4, batch_size=10 ** 9)
    coords, vals = random_problem(rng, lik, dims, int(0.8 * np.prod(dims)))
    st0 = random_state(rng, cfg, dims)
    rowptr, perm = so.build_index(coords, dims)
    pool = so.make_tape(rng, cfg, N, P)
    worst = 0.0
    for dim in range(N):
        n = dims[dim]
        accs, states = [], []
        for tdt in (torch.float64, torch.float32):
            eng = Engine(cfg, dims, coords, vals, device="cuda", noise_dtype=torch.float64, sweep_chunk_entries=W,
                         table_budget_bytes=1, table_dtype=tdt)
            eng.set_noise(NoisePool.from_tape(pool, eng.device, torch.float64, diag=False))
            engine_load_state(eng, st0)
            plan = eng.plan_sweep(dim)
            if tdt == torch.float32:
                assert plan.table_bytes * 2 == table_bytes64          # half the table
            else:
                table_bytes64 = plan.table_bytes
            accs.append(eng.accumulate(plan).cpu().numpy().copy())
            eng.row_update(plan, eng.accumulate(plan))
            eng.check_flags()
            states.append(engine_dump_state(eng, st0))
            eng.close()
        scale = np.abs(accs[0]).max()
        dev = np.abs(accs[1] - accs[0]).max() / scale
        worst = max(worst, dev)
        assert dev < 2e-6, dev
        np.testing.assert_array_equal(accs[1][:, -1], accs[0][:, -1])              # counts: bit-exact
        entries, tapes = [], []
        for row in range(n):
            ids = perm[dim][rowptr[dim][row]:rowptr[dim][row + 1]]
            sl = np.arange(len(ids)) % P
            tapes.append(so.Tape(pool.z_other[sl], pool.z_own[sl], pool.r_w[sl], pool.z_f[sl]))
            entries.append((coords[ids], vals[ids]))
        want = so.copy_state(st0)
        so.rows_step(want, cfg, dim, np.arange(n), entries, tapes)
        assert_state_close(states[1], want, cfg, 1e-4, what="float table, dims=%s mode %d" % (dims, dim))
    with capsys.disabled():
        print("\n  float table vs double table, largest accumulator deviation / max |acc|: %.2e (dims %s, D %d)" % (worst, dims, D))


@pytest.mark.parametrize("dims,D,lik,P,W,nb,offs,model", [
    ([7, 9, 8], 32, "normal", 3, 64, 2, False, "d"),         # config 5's kernel instance, 2 x 2 tiles per slot
    ([7, 9, 8], 20, "bernoulli", 2, 2, 3, True, "d"),        # 3 x 3 tiles, segments of 2, shifted slots
    ([5, 6, 4, 3], 12, "normal", 2, 64, 2, False, "d"),      # 4-way: 2 x 2 x 2 tiles
    ([7, 9, 8], 30, "poisson", 2, 4, 2, False, "robust"),    # three-pass path over tiled chunks
])
def test_sweep_l2_tiled_launch_order_matches_untiled(dims, D, lik, P, W, nb, offs, model):
    """L2-tiled launch order (one chunk per (slot, tile of other-mode blocks, row[, segment])): the same entry set,
    the same per-row accumulators as the untiled sweep up to summation order, counts bit-exact."""
    from ssvi_tf_b200.engine import Engine, NoisePool
    rng = np.random.default_rng(D * 77 + P + nb)
    N = len(dims)
    cfg = so.make_config(model, lik, D, diag=False, k1=64, k2=4, batch_size=10 ** 9)
    coords, vals = random_problem(rng, lik, dims, int(0.7 * np.prod(dims)))
    st0 = random_state(rng, cfg, dims)
    rowptr, perm = so.build_index(coords, dims)
    pool = so.make_tape(rng, cfg, N, P)
    for dim in range(N):
        off = rng.integers(0, 50, size=dims[dim]) if offs else None
        accs, states = [], []
        for blocks in (1, nb):
            eng = Engine(cfg, dims, coords, vals, device="cuda", noise_dtype=torch.float64, chunk_entries=3,
                         sweep_chunk_entries=W, table_budget_bytes=1, sweep_tile_blocks=blocks,
                         record_budget_bytes=(1 if blocks == 1 else 40 * 8 * (D + D * D + 2)))
            eng.set_noise(NoisePool.from_tape(pool, eng.device, torch.float64, diag=False))
            engine_load_state(eng, st0)
            plan = eng.plan_sweep(dim, off)
            assert plan.tile_blocks == blocks and len(plan.groups) == P
            np.testing.assert_array_equal(np.sort(plan.entry_ids.cpu().numpy()), np.sort(perm[dim]))
            if blocks > 1:                              # every chunk's entries lie in ONE tile, slot and row
                ids = plan.entry_ids.cpu().numpy()
                cs, cl = plan.chunk_start.cpu().numpy(), plan.chunk_len.cpu().numpy()
                cr = plan.chunk_row.cpu().numpy()
                assert cl.sum() == len(ids) and cl.min() >= 1
                for c in range(plan.n_chunks):
                    e = coords[ids[cs[c]:cs[c] + cl[c]]]
                    assert np.all(e[:, dim] == cr[c])
                    for d in range(N):
                        if d != dim:
                            lo, hi = eng.row_window[d]
                            blk = (e[:, d] - lo) // -(-(hi - lo + 1) // blocks)
                            assert np.all(blk == blk[0])
                co = plan.chunk_out.cpu().numpy()
                for bt in np.unique(plan.batch_of):     # record slots: a permutation inside every batch's buffer
                    got = np.concatenate([co[c0:c1] for k, (_, _, c0, c1) in enumerate(plan.groups)
                                          if plan.batch_of[k] == bt])
                    assert sorted(got.tolist()) == list(range(len(got))) and len(got) <= plan.max_batch_records
            acc = eng.accumulate(plan).cpu().numpy().copy()
            eng.row_update(plan, eng.accumulate(plan))
            eng.check_flags()
            accs.append(acc)
            states.append(engine_dump_state(eng, st0))
            eng.close()
        scale = np.abs(accs[0]).max()
        np.testing.assert_allclose(accs[1], accs[0], rtol=1e-9, atol=1e-12 * scale)
        np.testing.assert_array_equal(accs[1][:, -1], np.diff(rowptr[dim]))
        assert_state_close(states[1], states[0], cfg, 1e-9, what="tiled dims=%s mode %d" % (dims, dim))


@pytest.mark.parametrize("dims,D,k1,model,lik,P,W,offs", [
    ([6, 5, 4], 50, 64, "simple", "normal", 4, 64, False),       # BASELINE config 4's kernel instance (3-way slice)
    ([4, 3, 3, 3], 50, 64, "simple", "normal", 3, 64, False),    # config 4: 4-way, three other modes
    ([6, 5, 4], 7, 16, "d", "bernoulli", 3, 1, True),            # odd D, one k-tile, segments of 4, shifted slots
    ([6, 5, 4], 33, 32, "d", "poisson", 2, 64, False),
    ([5, 4, 4], 20, 64, "robust", "poisson", 4, 2, True),
    ([5, 4, 4], 64, 48, "robust", "bernoulli", 2, 64, False),    # D = 64: every lane busy
    ([7, 9], 12, 64, "simple", "poisson", 3, 64, False),         # matrix
    ([5, 4, 4], 20, 64, "robust+quad", "normal", 3, 64, False),  # SSVI_TF_robust(quadrature=True), diagonal
    ([6, 5, 4], 9, 32, "simple", "normal", 2, 3, True),          # closed-form (LIN) instance: odd D, two k-tiles
    ([6, 5, 4], 64, 48, "simple", "normal", 3, 64, True),        # LIN: three k-tiles, every lane busy
    ([2, 40, 30], 10, 16, "simple", "normal", 1, 2000, False),   # LIN: ~960-entry chunks (a warp walks > 32 entries)
    ([3, 33, 29], 6, 32, "simple", "normal", 2, 2000, True),     # LIN: long chunks, odd per-warp counts (lone last entry)
    ([2, 40, 30], 10, 16, "d", "bernoulli", 1, 2000, False),     # one-entry kernel on the same long chunks
])
def test_sweep_diag_path_matches_generic_and_oracle(dims, D, k1, model, lik, P, W, offs):
    """Diagonal covariance, full sweep through the slot-major shared-noise kernel == the generic kernel == the oracle."""
    from ssvi_tf_b200.engine import Engine, NoisePool
    rng = np.random.default_rng(D * 1000 + k1 + P)
    N = len(dims)
    cfg = so.make_config(model.split("+")[0], lik, D, diag=True, k1=k1, k2=6, batch_size=10 ** 9,
                         quadrature=30 if model.endswith("+quad") else 0)
    nnz = int(0.8 * np.prod(dims))
    coords, vals = random_problem(rng, lik, dims, nnz)
    st0 = random_state(rng, cfg, dims)
    rowptr, perm = so.build_index(coords, dims)
    pool = so.make_tape(rng, cfg, N, P)
    for dim in range(N):
        n = dims[dim]
        off = rng.integers(0, 50, size=n) if offs else None
        accs, states = [], []
        for path, sweep in ((0, True), (1, False)):
            eng = Engine(cfg, dims, coords, vals, device="cuda", noise_dtype=torch.float64, chunk_entries=3,
                         kernel_path=path, sweep_chunk_entries=W)
            eng.set_noise(NoisePool.from_tape(pool, eng.device, torch.float64, diag=True))
            engine_load_state(eng, st0)
            assert eng.sweep_supported == (path == 0)
            plan = eng.plan_sweep(dim, off) if sweep else eng.plan_full(dim, None, off)
            acc = eng.accumulate(plan).cpu().numpy().copy()
            eng.row_update(plan, eng.accumulate(plan))
            eng.check_flags()
            accs.append(acc)
            states.append(engine_dump_state(eng, st0))
            eng.close()
        scale = np.abs(accs[1]).max()
        np.testing.assert_allclose(accs[0], accs[1], rtol=1e-9, atol=1e-12 * scale)
        np.testing.assert_array_equal(accs[0][:, -1], np.diff(rowptr[dim]))
        entries, tapes = [], []
        for row in range(n):
            ids = perm[dim][rowptr[dim][row]:rowptr[dim][row + 1]]
            sl = (np.arange(len(ids)) + (0 if off is None else int(off[row]))) % P
            tapes.append(so.Tape(pool.z_other[sl], pool.z_own[sl], pool.r_w[sl], pool.z_f[sl]))
            entries.append((coords[ids], vals[ids]))
        want = so.copy_state(st0)
        so.rows_step(want, cfg, dim, np.arange(n), entries, tapes)
        for got in states:
            assert_state_close(got, want, cfg, RTOL, what="dims=%s mode %d" % (dims, dim))


def test_subsampled_sweep_draws_reference_minibatches_on_device():
    """plan_subsampled_sweep: every row keeps min(len, B) DISTINCT entries of its own list (bit-exact ids / counts),
    fresh picks per call, and the sweep equals the oracle's rows_step fed the same picks (SSVI_TF_interface.py:170-174)."""
    from ssvi_tf_b200.engine import NoisePool
    rng = np.random.default_rng(23)
    dims, D, k1, k2, B = [6, 9, 5], 5, 16, 4, 8
    cfg = so.make_config("d", "bernoulli", D, k1=k1, k2=k2, batch_size=B)
    coords, vals = random_problem(rng, "bernoulli", dims, 200)
    coords = coords[coords[:, 1] != 2]                 # an empty row in mode 1
    vals = vals[:len(coords)]
    st0 = random_state(rng, cfg, dims)
    rowptr, perm = so.build_index(coords, dims)
    eng = _engine(cfg, dims, coords, vals, chunk_entries=3)
    pool = so.make_tape(rng, cfg, 3, B)
    eng.set_noise(NoisePool.from_tape(pool, eng.device, torch.float64, diag=False))
    gen = torch.Generator(device=eng.device)
    gen.manual_seed(5)
    for dim in range(3):
        engine_load_state(eng, st0)
        plan = eng.plan_subsampled_sweep(dim, B, gen)
        again = eng.plan_subsampled_sweep(dim, B, gen)
        ids = plan.entry_ids.cpu().numpy()
        lens = np.diff(rowptr[dim])
        want_counts = np.minimum(lens, B)
        np.testing.assert_array_equal(plan.batch_sizes, want_counts)
        assert len(ids) == want_counts.sum()
        assert not np.array_equal(ids, again.entry_ids.cpu().numpy())          # a fresh draw every sweep
        rows = np.arange(dims[dim])
        entries, tapes, pos = [], [], 0
        for row in rows:
            own = perm[dim][rowptr[dim][row]:rowptr[dim][row + 1]]
            seg = ids[pos:pos + want_counts[row]]
            pos += want_counts[row]
            assert len(set(seg.tolist())) == len(seg) and set(seg.tolist()) <= set(own.tolist())
            if lens[row] <= B:
                np.testing.assert_array_equal(seg, own)       # kept whole, in append order
            n = len(seg)
            t = so.Tape(pool.z_other[:n], pool.z_own[:n], pool.r_w[:n], pool.z_f[:n])
            lookup = {int(e): k for k, e in enumerate(own)}
            if lens[row] > B:
                t.sub_idx = np.array([lookup[int(e)] for e in seg])
                entries.append((coords[own], vals[own]))
            else:
                entries.append((coords[own], vals[own]))
            tapes.append(t)
        eng.step(plan)
        eng.check_flags()
        want = so.copy_state(st0)
        so.rows_step(want, cfg, dim, rows, entries, tapes)
        assert_state_close(engine_dump_state(eng, st0), want, cfg, RTOL, what="subsampled mode %d" % dim)
    eng.close()


def test_device_minibatch_draw_is_a_uniform_subset():
    """ssvi_draw_minibatches: every position of a row's list is kept with probability B / len (2000 independent rows of 40
    entries, B = 8: binomial(2000, 0.2) per position, five standard deviations), the picks of a row are distinct, and
    the same seed reproduces the draw."""
    rng = np.random.default_rng(8)
    dims, D, B = [500, 8, 5], 4, 8
    cfg = so.make_config("d", "normal", D, k1=16, k2=4, batch_size=B)
    grid = np.stack(np.meshgrid(*[np.arange(n) for n in dims], indexing="ij"), axis=-1).reshape(-1, 3)
    coords = grid[rng.permutation(len(grid))]
    vals = rng.standard_normal(len(coords))
    eng = _engine(cfg, dims, coords, vals, chunk_entries=4)
    rowptr, perm = so.build_index(coords, dims)
    counts = np.zeros(40, dtype=np.int64)
    first = None
    for call in range(4):
        gen = torch.Generator(device=eng.device)
        gen.manual_seed(100 + call)
        ids = eng.plan_subsampled_sweep(0, B, gen).entry_ids.cpu().numpy().reshape(dims[0], B)
        if call == 0:
            first = ids.copy()
        for row in range(dims[0]):
            own = perm[0][rowptr[0][row]:rowptr[0][row + 1]]           # ascending entry ids (stable sort)
            pos = np.searchsorted(own, ids[row])
            assert np.array_equal(own[pos], ids[row]) and len(set(pos.tolist())) == B
            counts[pos] += 1
    n = 4 * dims[0]
    sd = np.sqrt(n * 0.2 * 0.8)
    assert np.all(np.abs(counts - 0.2 * n) < 5 * sd), counts
    eng._sub_plans.clear()                                               # same seed, same call number: same draw
    gen = torch.Generator(device=eng.device)
    gen.manual_seed(100)
    again = eng.plan_subsampled_sweep(0, B, gen).entry_ids.cpu().numpy().reshape(dims[0], B)
    np.testing.assert_array_equal(again, first)
    eng.close()
