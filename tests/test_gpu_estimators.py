"""GPU: the report-path estimators through ssvi_estimate_entries against the REFERENCE's
own outputs (tests/golden_est/*.npz, oracle/gen_estimator_golden.py) and against the oracle on other shapes."""
import glob
import json
import os

import numpy as np
import pytest
import torch

from oracle import ssvi_oracle as so
from oracle.golden_io import _unpack_state
from tests.helpers import engine_load_state, random_problem, random_state

pytestmark = pytest.mark.gpu
RTOL = 1e-5
EST_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden_est")
CASES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(EST_DIR, "*.npz")))


def _pool(z):
    f64 = lambda a: a.astype(np.float64)                        # noqa: E731
    return so.Tape(f64(z["z_other"]), f64(z["z_own"]), f64(z["r_w"]), f64(z["z_f"]))


def test_fixtures_present():
    assert len(CASES) >= 6


@pytest.mark.parametrize("noise_dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("name", CASES)
def test_estimators_match_reference_outputs(name, noise_dtype):
    from ssvi_tf_b200.engine import Engine, NoisePool
    z = np.load(os.path.join(EST_DIR, name + ".npz"))
    cfg = json.loads(str(z["cfg"]))
    dims = [int(x) for x in z["dims"]]
    state = _unpack_state("s0", z, len(dims))
    entries, ys = z["entries"], z["ys"]
    eng = Engine(cfg, dims, entries, ys, device="cuda", noise_dtype=noise_dtype)
    engine_load_state(eng, state)
    eng.set_noise(NoisePool.from_tape(_pool(z), eng.device, noise_dtype, diag=cfg["diag"]))
    e_dev = torch.as_tensor(entries.astype(np.int32)).cuda()
    y_dev = torch.as_tensor(ys).cuda()
    got_obs = eng.estimate_entries(so.EST_EXPECTED_OBS, e_dev).cpu().numpy()
    np.testing.assert_allclose(got_obs, z["ref_exp_obs"], rtol=RTOL)
    what = so.EST_ROBUST_VLB if cfg["model"] == "robust" else so.EST_LOG_PDF
    got_vlb = eng.estimate_entries(what, e_dev, y_dev).cpu().numpy()
    np.testing.assert_allclose(got_vlb, z["ref_vlb_term"], rtol=RTOL, atol=1e-9)
    np.testing.assert_allclose(got_vlb.sum(), float(z["ref_vlb_sum"]), rtol=RTOL)
    got_nll = float(-np.log(eng.estimate_entries(so.EST_PDF, e_dev, y_dev).cpu().numpy()).sum())
    np.testing.assert_allclose(got_nll, float(z["ref_nll"]), rtol=RTOL)
    eng.check_flags()
    eng.close()


@pytest.mark.parametrize("model,lik,diag,dims,D,k1,k2", [
    ("d", "bernoulli", False, [7, 6, 5], 20, 64, 8),        # BASELINE config 2's shapes
    ("robust", "poisson", False, [6, 5, 4], 30, 64, 16),    # config 3
    ("simple", "normal", True, [5, 4, 3, 3], 50, 64, 16),   # config 4 (4-way, diagonal)
    ("d", "normal", False, [9, 8], 32, 40, 4),              # matrix, k1 not a multiple of 32
])
def test_estimators_match_oracle(model, lik, diag, dims, D, k1, k2):
    from ssvi_tf_b200.engine import Engine, NoisePool
    rng = np.random.default_rng(D + k1)
    cfg = so.make_config(model, lik, D, diag=diag, k1=k1, k2=k2, batch_size=16)
    coords, vals = random_problem(rng, lik, dims, 40)
    state = random_state(rng, cfg, dims)
    for d in range(len(dims)):                                   # keep |f| moderate at large D
        state["params"][d][:, 0, :] *= 0.5
    pool = so.make_tape(rng, cfg, len(dims), 7)                  # 7 slots: wraps around, shifted by slot0
    eng = Engine(cfg, dims, coords, vals, device="cuda", noise_dtype=torch.float64)
    engine_load_state(eng, state)
    eng.set_noise(NoisePool.from_tape(pool, eng.device, torch.float64, diag=diag))
    e_dev = torch.as_tensor(coords.astype(np.int32)).cuda()
    y_dev = torch.as_tensor(vals).cuda()
    whats = [so.EST_EXPECTED_OBS, so.EST_LOG_PDF, so.EST_PDF] + ([so.EST_ROBUST_VLB] if model != "d" else [])
    for what in whats:
        got = eng.estimate_entries(what, e_dev, y_dev, slot0=3).cpu().numpy()
        want = so.estimate_entries(state, cfg, what, coords, vals, pool, slot0=3)
        np.testing.assert_allclose(got, want, rtol=RTOL, atol=1e-12, err_msg="what=%d" % what)
    eng.close()


def test_model_level_estimators_and_detailed_report(capsys):
    """estimate_vlb / estimate_KL_divergence / estimate_negative_log_likelihood on the model object equal the oracle
    fed the model's own pool; the detailed report prints every column of the reference's (:764-830)."""
    import ssvi_tf_b200 as pkg
    rng = np.random.default_rng(5)
    dims, D = [6, 5, 4], 4
    coords, vals = random_problem(rng, "bernoulli", dims, 80)

    class T(object):
        datatype = "binary"
    t = T()
    t.dims = dims
    t.train_entries, t.train_vals = coords[:60], vals[:60]
    t.test_entries, t.test_vals = coords[60:], vals[60:]
    np.random.seed(1)
    m = pkg.SSVI_TF_simple(t, D, k1=16, k2=8, batch_size=8, noise_dtype=torch.float64, noise_seed=3)
    m.factorize(report=5, max_iteration=10, detailed_report=True)
    out = capsys.readouterr().out
    assert "test_nll" in out and "E-term" in out and "nan" not in out.split("KL")[-1]
    cfg = so.make_config("simple", "bernoulli", D, k1=16, k2=8, batch_size=8)
    nz = m.engine.noise
    P = nz.slots
    state = dict(params=[p.cpu().numpy() for p in m.engine.params], pSigma=m.pSigma, pmu=np.ones(D),
                 w_sigma=float(m.w_sigma))
    blk = lambda a: a.cpu().numpy().astype(np.float64)          # noqa: E731
    from ssvi_tf_b200.engine import noise_offsets
    zo = np.zeros((P, 2, 16, D))
    zw = np.zeros((P, 16, D))
    offs = noise_offsets(D, 16, False)
    zo_raw, zw_raw = blk(nz.z_other), blk(nz.z_own)
    for k in range(16):
        for c in range(D):
            zo[:, :, k, c] = zo_raw[:, :, offs[c, k]]
            zw[:, k, c] = zw_raw[:, offs[c, k]]
    pool = so.Tape(zo, zw, blk(nz.r_w), blk(nz.z_f))
    e_term, kl = m.estimate_vlb(t.train_entries, t.train_vals)
    o_term, o_kl = so.estimate_vlb(state, cfg, t.train_entries, t.train_vals, pool)
    np.testing.assert_allclose([e_term, kl], [o_term, o_kl], rtol=RTOL)
    np.testing.assert_allclose(m.estimate_negative_log_likelihood(t.test_entries, t.test_vals),
                               so.estimate_negative_log_likelihood(state, cfg, t.test_entries, t.test_vals, pool),
                               rtol=RTOL)
    np.testing.assert_allclose(m.estimate_expectation_term_vlb(t.train_entries[0], t.train_vals[0]),
                               so.estimate_entries(state, cfg, so.EST_LOG_PDF, t.train_entries[:1], t.train_vals[:1],
                                                   pool)[0], rtol=RTOL)
