"""Generate tests/golden/*.npz by running the UNMODIFIED reference under the noise tape.

TEST INFRASTRUCTURE ONLY.  Run in the build container (needs /root/reference):
    python -m oracle.gen_golden
For every case it (1) builds a small COO tensor and a randomised posterior state,
(2) calls the reference's update_natural_param_batch(dim, row) through oracle/ref_shim.py,
(3) runs oracle/ssvi_oracle.py:row_step on the same inputs and asserts agreement to 1e-9,
(4) writes inputs + reference outputs to tests/golden/<case>.npz.
The fixtures are what pins the oracle (and through it the CUDA path) to the reference.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import ref_shim, ssvi_oracle as so          # noqa: E402
from oracle.golden_io import save_case                  # noqa: E402

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")
DATATYPE = {"normal": "real", "bernoulli": "binary", "poisson": "count"}


def random_problem(rng, lik, dims, nnz):
    total = int(np.prod(dims))
    flat = rng.choice(total, size=nnz, replace=False)
    coords = np.stack(np.unravel_index(flat, dims), axis=1).astype(np.int64)
    if lik == "normal":
        vals = rng.normal(0.5, 1.0, size=nnz)
    elif lik == "bernoulli":
        vals = rng.choice([-1.0, 1.0], size=nnz)
    else:
        vals = rng.poisson(2.0, size=nnz).astype(np.float64)
    return coords, vals


def random_state(rng, cfg, dims):
    """A generic (non-initial) state so that every term of the update is exercised."""
    D = cfg["D"]
    st = so.new_state(cfg, dims, rng=rng, randstart=False)
    for d, n in enumerate(dims):
        p = st["params"][d]
        p[:, 0, :] = rng.normal(0.4, 0.35, size=(n, D))
        for i in range(n):
            if cfg["diag"]:
                p[i, 1, :] = rng.uniform(0.05, 0.4, size=D)
            else:
                A = rng.normal(size=(D, D)) * 0.2
                p[i, 1:, :] = A @ A.T + np.diag(rng.uniform(0.05, 0.3, size=D))
        st["ada_mean"][d][...] = rng.uniform(0.5, 3.0, size=(D, n))
        st["ada_cov"][d][...] = rng.uniform(0.0, 2.0, size=(D, n))
    st["pSigma"][...] = rng.uniform(0.5, 2.0, size=len(dims))
    st["time_step"] = [int(t) for t in rng.integers(0, 4, size=len(dims))]
    st["w_sigma"] = float(rng.uniform(0.6, 1.2))
    st["w_ada_grad"] = float(rng.uniform(0.0, 2.0))
    return st


def run_case(name, model, lik, diag, dims, nnz, D, k1, k2, batch_size, seed,
             mean_update="S", cov_update="N", cov_eta=1.0, n_rows=2, fixed_covariance=False, quadrature=False):
    rng = np.random.default_rng(seed)
    cfg = so.make_config(model, lik, D, diag=diag, k1=k1, k2=k2, batch_size=batch_size,
                         mean_update=mean_update, cov_update=cov_update, cov_eta=cov_eta,
                         quadrature=30 if quadrature else 0)       # the reference fixes degree 30
    coords, vals = random_problem(rng, lik, dims, nnz)
    rowptr, perm = so.build_index(coords, dims)
    state0 = random_state(rng, cfg, dims)
    state0["fixed_covariance"] = fixed_covariance
    obj = ref_shim.make_reference_model(model, DATATYPE[lik], dims, coords, vals, D, diag, k1, k2,
                                        batch_size, mean_update=mean_update, cov_update=cov_update,
                                        cov_eta=cov_eta, quadrature=quadrature)
    steps = []
    N = len(dims)
    for s in range(n_rows):
        dim = int(rng.integers(0, N))
        # prefer a populated row; the longest row of the mode on the first step (subsampling path)
        lens = np.diff(rowptr[dim])
        row = int(np.argmax(lens)) if s == 0 else int(rng.choice(np.nonzero(lens > 0)[0]))
        ids = perm[dim][rowptr[dim][row]:rowptr[dim][row + 1]]
        tape_o = so.make_tape(rng, cfg, N, len(ids), n_obs=len(ids))
        tape_r = ref_shim.NoiseTape(tape_o.z_other, tape_o.z_own, tape_o.r_w, tape_o.z_f, tape_o.sub_idx)

        ref_shim.load_state(obj, state0)
        # bit-exact index parity: the reference's per-row list == our CSR slice, in order
        ref_obs = obj.tensor.find_observed_ui(dim, row)
        assert len(ref_obs) == len(ids)
        for r in range(len(ids)):
            assert list(ref_obs[r, 0]) == list(coords[ids[r]]) and ref_obs[r, 1] == vals[ids[r]]
        calls = ref_shim.reference_row_step(obj, dim, row, tape_r)
        ref_after = ref_shim.dump_state(obj, state0)

        ora = so.copy_state(state0)
        so.row_step(ora, cfg, dim, row, coords[ids], vals[ids], tape_o)
        for d in range(N):
            for key in ("params", "ada_mean", "norm_changes", "grad_norms") + (("ada_cov",) if diag else ()):
                np.testing.assert_allclose(ora[key][d], ref_after[key][d], rtol=1e-9, atol=1e-11,
                                           err_msg="%s %s[%d]" % (name, key, d))
        if cfg["noise_added"]:
            np.testing.assert_allclose(ora["w_sigma"], ref_after["w_sigma"], rtol=1e-9, err_msg=name)
            np.testing.assert_allclose(ora["w_ada_grad"], ref_after["w_ada_grad"], rtol=1e-9, err_msg=name)
        steps.append(dict(dim=dim, row=row, tape=tape_o, after=ref_after, calls=calls))
    path = os.path.join(GOLDEN_DIR, name + ".npz")
    save_case(path, cfg, dims, coords, vals, state0, steps)
    return path


def case_list():
    cases = []
    seed = 319          # the reference's own seed (test_scripts/test.py:14)
    for model in ("d", "simple", "robust"):
        for lik in ("normal", "bernoulli", "poisson"):
            for diag in (False, True):
                seed += 1
                name = "%s_%s_%s" % (model, lik, "diag" if diag else "full")
                cases.append(dict(name=name, model=model, lik=lik, diag=diag, dims=[6, 5, 4], nnz=60,
                                  D=5, k1=8, k2=4, batch_size=6, seed=seed, n_rows=3))
    # update-scheme variants (mean_update N, cov_update S = quirk Q8), 4-way tensor, fixed covariance
    cases.append(dict(name="d_normal_full_meanN", model="d", lik="normal", diag=False, dims=[6, 5, 4],
                      nnz=60, D=5, k1=8, k2=4, batch_size=6, seed=401, mean_update="N", n_rows=2))
    cases.append(dict(name="d_normal_diag_meanN", model="d", lik="normal", diag=True, dims=[6, 5, 4],
                      nnz=60, D=5, k1=8, k2=4, batch_size=6, seed=402, mean_update="N", n_rows=2))
    cases.append(dict(name="d_normal_full_covS", model="d", lik="normal", diag=False, dims=[6, 5, 4],
                      nnz=60, D=5, k1=8, k2=4, batch_size=6, seed=403, cov_update="S", n_rows=2))
    cases.append(dict(name="d_normal_diag_covS", model="d", lik="normal", diag=True, dims=[6, 5, 4],
                      nnz=60, D=5, k1=8, k2=4, batch_size=6, seed=404, cov_update="S", n_rows=2))
    cases.append(dict(name="simple_normal_diag_4way", model="simple", lik="normal", diag=True,
                      dims=[5, 4, 4, 3], nnz=90, D=6, k1=8, k2=8, batch_size=8, seed=405, n_rows=3))
    cases.append(dict(name="d_bernoulli_full_fixedcov", model="d", lik="bernoulli", diag=False,
                      dims=[6, 5, 4], nnz=60, D=5, k1=8, k2=4, batch_size=6, seed=406, n_rows=2,
                      fixed_covariance=True))
    cases.append(dict(name="robust_poisson_full_coveta", model="robust", lik="poisson", diag=False,
                      dims=[6, 5, 4], nnz=60, D=5, k1=8, k2=8, batch_size=6, seed=407, cov_eta=0.001,
                      n_rows=2))
    cases.append(dict(name="simple_bernoulli_full_nosub", model="simple", lik="bernoulli", diag=False,
                      dims=[6, 5, 4], nnz=40, D=5, k1=8, k2=4, batch_size=64, seed=409, n_rows=3))
    # SSVI_TF_robust(quadrature=True): Gauss-Hermite rule in place of the k2 draws (SSVI_TF_robust.py:175-214)
    for lik, diag, sd in (("normal", False, 410), ("bernoulli", True, 411), ("poisson", False, 412)):
        cases.append(dict(name="robust_%s_%s_quadrature" % (lik, "diag" if diag else "full"), model="robust",
                          lik=lik, diag=diag, dims=[6, 5, 4], nnz=60, D=5, k1=8, k2=4, batch_size=6, seed=sd,
                          n_rows=2, quadrature=True))
    # BASELINE.json config 1 at full per-row size: 50^3, D=10, k1=64, k2=128, B=128, SSVI_TF_d full
    cases.append(dict(name="config1_d_normal_full", model="d", lik="normal", diag=False, dims=[50, 50, 50],
                      nnz=12000, D=10, k1=64, k2=128, batch_size=128, seed=408, n_rows=1))
    # the kernel instances of BASELINE.json configs 2-5 (rank, k1, k2, model, likelihood, covariance kind, order) on small
    # index spaces, one subsampled row step each (B = 32 keeps the noise tape of a fixture near 1-2 MB)
    cases.append(dict(name="config2_shape_d_bernoulli_full", model="d", lik="bernoulli", diag=False, dims=[12, 10, 8],
                      nnz=600, D=20, k1=64, k2=64, batch_size=32, seed=420, n_rows=1))
    cases.append(dict(name="config3_shape_robust_poisson_full", model="robust", lik="poisson", diag=False,
                      dims=[12, 10, 8], nnz=600, D=30, k1=64, k2=64, batch_size=32, seed=421, cov_eta=0.001, n_rows=1))
    cases.append(dict(name="config4_shape_simple_normal_diag_4way", model="simple", lik="normal", diag=True,
                      dims=[6, 5, 4, 4], nnz=350, D=50, k1=64, k2=64, batch_size=32, seed=422, n_rows=1))
    cases.append(dict(name="config5_shape_d_normal_full", model="d", lik="normal", diag=False, dims=[12, 10, 8],
                      nnz=600, D=32, k1=64, k2=64, batch_size=32, seed=423, n_rows=1))
    return cases


def main():
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    only = sys.argv[1:]          # optional name filters: regenerate only the cases containing one of them
    for c in case_list():
        if only and not any(o in c["name"] for o in only):
            continue
        path = run_case(**c)
        print("wrote %-40s %7.1f KB" % (os.path.basename(path), os.path.getsize(path) / 1024.0))


if __name__ == "__main__":
    main()
