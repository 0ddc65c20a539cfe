"""Generate tests/golden_est/*.npz: the reference's report-path estimators under a noise tape.

TEST INFRASTRUCTURE ONLY.  Run in the build container (needs /root/reference):
    python -m oracle.gen_estimator_golden
For every case: a small tensor, a randomised posterior state and a pool of P noise slots; the UNMODIFIED
reference's estimate_expected_observation_via_sampling (SSVI/SSVI_TF_interface.py:1168-1189),
estimate_expectation_term_vlb (:889-908), estimate_negative_log_likelihood (:939-964) and estimate_vlb /
estimate_KL_divergence (:874-883, :911-936) run entry by entry with numpy.random patched to read the pool
(entry e reads slot e mod P: the N multivariate_normal calls take z_own, z_other[:, 0], z_other[:, 1], ...;
rayleigh takes r_w; normal takes z_f), with the transform x = m + L z of this build's noise contract.
oracle/ssvi_oracle.py:estimate_entries is asserted against the outputs to 1e-9 and inputs + outputs are written
for tests/test_oracle_golden.py (CPU) and tests/test_gpu_estimators.py (CUDA path).
The reference hard-codes k = 20 draws in the VLB term; the cases use k1 = k2 = predict_num_samples = 20 so one
pool serves the three estimators.
"""
import contextlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import ref_shim, ssvi_oracle as so          # noqa: E402
from oracle.gen_golden import DATATYPE, random_problem, random_state    # noqa: E402
from oracle.golden_io import _pack_state                # noqa: E402

OUT_DIR = os.path.join(ROOT, "tests", "golden_est")
K = 20

CASES = [
    # name, model, lik, diag, dims, D, n_entries, P
    ("d_normal_full", "d", "normal", False, [5, 4, 3], 4, 12, 5),
    ("d_bernoulli_diag", "d", "bernoulli", True, [5, 4, 3], 4, 12, 5),
    ("d_poisson_full", "d", "poisson", False, [5, 4, 3], 4, 12, 5),
    ("simple_normal_full", "simple", "normal", False, [5, 4, 3], 4, 12, 5),
    ("simple_poisson_diag", "simple", "poisson", True, [4, 3, 3, 2], 3, 10, 4),
    ("robust_poisson_full", "robust", "poisson", False, [5, 4, 3], 4, 12, 5),
    ("robust_bernoulli_diag", "robust", "bernoulli", True, [5, 4, 3], 4, 12, 16),
]


@contextlib.contextmanager
def patched_rng(pool, N):
    """numpy.random -> the pool, in the call order of the three estimators (per entry: N multivariate_normal
    draws of size k, then for the noise models rayleigh((k,)) and normal(size=(k, k)))."""
    saved = (np.random.multivariate_normal, np.random.normal, np.random.rayleigh)
    P = pool.z_own.shape[0]
    st = {"mvn": 0}

    def multivariate_normal(mean, cov, size=None):
        e, d = divmod(st["mvn"], N)
        st["mvn"] += 1
        z = pool.z_own[e % P] if d == 0 else pool.z_other[e % P, d - 1]
        assert z.shape[0] == size
        return np.asarray(mean) + z @ np.linalg.cholesky(np.asarray(cov, dtype=float)).T

    def rayleigh(scale=1.0, size=None):
        e = (st["mvn"] - 1) // N
        r = pool.r_w[e % P]
        assert r.shape == (tuple(size) if isinstance(size, tuple) else (size,))
        return scale * r

    def normal(loc=0.0, scale=1.0, size=None):
        e = (st["mvn"] - 1) // N
        z = pool.z_f[e % P]
        assert z.shape == tuple(size)
        return loc + scale * z

    np.random.multivariate_normal, np.random.normal, np.random.rayleigh = multivariate_normal, normal, rayleigh
    try:
        yield
    finally:
        np.random.multivariate_normal, np.random.normal, np.random.rayleigh = saved


def run_case(name, model, lik, diag, dims, D, n_entries, P, seed):
    rng = np.random.default_rng(seed)
    N = len(dims)
    cfg = so.make_config(model, lik, D, diag=diag, k1=K, k2=K, batch_size=8)
    coords, vals = random_problem(rng, lik, dims, n_entries)
    state = random_state(rng, cfg, dims)
    pool = so.make_tape(rng, cfg, N, P)
    obj = ref_shim.make_reference_model(model, DATATYPE[lik], dims, coords, vals, D, diag, K, K, 8)
    ref_shim.load_state(obj, state)
    obj.predict_num_samples = K
    ys = vals.astype(np.int64) if lik == "poisson" else vals
    with patched_rng(pool, N):
        exp_obs = np.array([obj.estimate_expected_observation_via_sampling(coords[e]) for e in range(n_entries)])
    with patched_rng(pool, N):
        vlb_term = np.array([obj.estimate_expectation_term_vlb(coords[e], ys[e]) for e in range(n_entries)])
    with patched_rng(pool, N):
        nll = float(obj.estimate_negative_log_likelihood(coords, ys))
    with patched_rng(pool, N):
        vlb_sum, kl = obj.estimate_vlb(coords, ys)
    # the restatement against the reference
    np.testing.assert_allclose(so.estimate_entries(state, cfg, so.EST_EXPECTED_OBS, coords, None, pool), exp_obs,
                               rtol=1e-9, err_msg=name)
    what = so.EST_ROBUST_VLB if model == "robust" else so.EST_LOG_PDF
    np.testing.assert_allclose(so.estimate_entries(state, cfg, what, coords, vals, pool), vlb_term,
                               rtol=1e-9, atol=1e-12, err_msg=name)
    np.testing.assert_allclose(so.estimate_negative_log_likelihood(state, cfg, coords, vals, pool), nll, rtol=1e-9,
                               err_msg=name)
    o_vlb, o_kl = so.estimate_vlb(state, cfg, coords, vals, pool)
    np.testing.assert_allclose([o_vlb, o_kl], [float(vlb_sum), float(kl)], rtol=1e-9, err_msg=name)
    out = dict(cfg=np.array(json.dumps(cfg)), dims=np.asarray(dims, dtype=np.int64), entries=coords, ys=vals,
               z_other=pool.z_other.astype(np.float32), z_own=pool.z_own.astype(np.float32),
               r_w=pool.r_w.astype(np.float32), z_f=pool.z_f.astype(np.float32),
               ref_exp_obs=exp_obs, ref_vlb_term=vlb_term, ref_nll=np.array(nll), ref_vlb_sum=np.array(float(vlb_sum)),
               ref_kl=np.array(float(kl)))
    _pack_state("s0", state, out)
    path = os.path.join(OUT_DIR, name + ".npz")
    np.savez_compressed(path, **out)
    return path


def main():
    assert ref_shim.available(), "needs /root/reference"
    os.makedirs(OUT_DIR, exist_ok=True)
    for i, c in enumerate(CASES):
        path = run_case(*c, seed=700 + i)
        print("wrote %-32s %6.1f KB" % (os.path.basename(path), os.path.getsize(path) / 1024.0))


if __name__ == "__main__":
    main()
