"""GPU: ssvi_accumulate_rows_static -- stages 1-3 + chunk reduction with the chunk-record scratch in the library's
file-scope static device array (no workspace from the caller, no allocation inside the call) -- against the existing
entry points (ssvi_accumulate + ssvi_reduce_chunks with a caller buffer), through the C ABI, on two calls in a row.
The static array keeps the previous call's contents, so the second call runs over the first call's records.
Collected last on purpose (the file name)."""
import ctypes

import numpy as np
import pytest
import torch

from oracle import ssvi_oracle as so
from tests.helpers import engine_load_state, random_problem, random_state

pytestmark = pytest.mark.gpu


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _static(eng, plan, out):
    eng._call(eng.lib.ssvi_accumulate_rows_static, plan.dim, plan.n_chunks, _ptr(plan.chunk_row), _ptr(plan.chunk_start),
              _ptr(plan.chunk_len), _ptr(plan.chunk_slot0), _ptr(plan.entry_ids), _ptr(plan.rows), _ptr(eng.w_state),
              plan.n_rows, _ptr(plan.row_chunk_ptr), _ptr(out), eng._stream())


@pytest.mark.parametrize("model,lik,diag,D,k1", [("d", "bernoulli", False, 20, 64), ("simple", "normal", True, 7, 16)])
def test_static_scratch_entry_point_matches_accumulate_plus_reduce(model, lik, diag, D, k1):
    from ssvi_tf_b200 import _lib
    from ssvi_tf_b200.engine import Engine, NoisePool
    rng = np.random.default_rng(D + k1)
    dims = [9, 7, 6]
    cfg = so.make_config(model, lik, D, diag=diag, k1=k1, k2=4, batch_size=10 ** 9)
    coords, vals = random_problem(rng, lik, dims, int(0.7 * np.prod(dims)))
    eng = Engine(cfg, dims, coords, vals, device="cuda", noise_dtype=torch.float64, chunk_entries=3)
    engine_load_state(eng, random_state(rng, cfg, dims))
    eng.set_noise(NoisePool.from_tape(so.make_tape(rng, cfg, 3, 8), eng.device, torch.float64, diag=diag))
    assert eng.lib.ssvi_static_scratch_bytes() == 8 << 20
    # two calls in a row over DIFFERENT plans: the second one finds the first one's records in the static array
    for dim, rows in ((0, None), (1, [5, 0, 3])):
        plan = eng.plan_full(dim, rows)
        assert plan.n_chunks * eng.acc_len * 8 <= eng.lib.ssvi_static_scratch_bytes()
        want = eng.accumulate(plan).clone()                      # existing entry points, caller-owned scratch
        got = torch.full_like(want, float("nan"))
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        _static(eng, plan, got)
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        eng.check_flags()
        scale = float(want.abs().max())
        np.testing.assert_allclose(got.cpu().numpy(), want.cpu().numpy(), rtol=1e-9, atol=1e-12 * scale)
        np.testing.assert_array_equal(got[:, -1].cpu().numpy(), want[:, -1].cpu().numpy())    # counts: bit-exact
    # beyond the maximum: a clear error, nothing launched
    n_too_many = eng.lib.ssvi_static_scratch_bytes() // (eng.acc_len * 8) + 1
    plan = eng.plan_full(0)
    rc = eng.lib.ssvi_accumulate_rows_static(eng.h, 0, n_too_many, _ptr(plan.chunk_row), _ptr(plan.chunk_start),
                                             _ptr(plan.chunk_len), _ptr(plan.chunk_slot0), _ptr(plan.entry_ids),
                                             _ptr(plan.rows), _ptr(eng.w_state), plan.n_rows, _ptr(plan.row_chunk_ptr),
                                             _ptr(got), eng._stream())
    assert rc == -1                                                             # SSVI_E_INVALID
    assert b"static scratch holds" in eng.lib.ssvi_last_error(eng.h)
    eng.close()
