"""Generate tests/golden_e2e/*.npz: END-TO-END metrics of the UNMODIFIED reference's factorize().

TEST INFRASTRUCTURE ONLY.  Run in the build container (needs /root/reference):
    python -m oracle.gen_e2e_golden
For every case it builds a small synthetic tensor (train / test split), runs the reference's own
SSVI_TF_{d,simple,robust}.factorize (SSVI/SSVI_TF_interface.py:101-143; reports disabled) under the three
compatibility shims of oracle/ref_shim.py with NumPy's real RNG for several seeds, and records the reference's
evaluate_test_error() (:997-1030) per seed.  tests/test_gpu_e2e.py runs the B200 build's factorize on the same data
with the same hyper-parameters and checks BASELINE.json's criterion: end-to-end test RMSE / error rate within 1 % of
the reference (plus the reference's own seed-to-seed spread, since both sides are stochastic).
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import ref_shim                              # noqa: E402

OUT_DIR = os.path.join(ROOT, "tests", "golden_e2e")

CASES = [
    # name, datatype, model, diag, dims, D, k1, k2, batch, iterations, seeds
    # every row fits in one minibatch (no subsampling) and the loop runs ~70 sweeps, so the seed-to-seed spread of the
    # final metric is a few per cent instead of tens of per cent
    ("real_d_full", "real", "d", False, [12, 10, 8], 3, 32, 8, 128, 800, 8),
    ("binary_d_full", "binary", "d", False, [12, 10, 8], 3, 32, 8, 128, 800, 8),
    ("real_simple_diag", "real", "simple", True, [12, 10, 8], 3, 32, 8, 128, 800, 8),
    ("count_robust_full", "count", "robust", False, [12, 10, 8], 3, 16, 8, 128, 400, 8),
    # rows of mode 0 hold ~770 train entries (the reference subsamples 128 of them per update, :170-174): the problem on
    # which the bench's schedule -- full rows, a 32-slot pool, ~24 entries per slot -- is compared with the reference
    ("real_d_longrows", "real", "d", False, [6, 40, 40], 3, 32, 8, 128, 800, 16),
]


def synth(datatype, dims, D, rng):
    """Ground-truth CP factors and observations through the reference's data links
    (Tensor/binary_tensor.py:23-26, count_tensor.py:27-33, real_tensor.py)."""
    N = len(dims)
    mus = {"real": [0.8, 1.0, 0.6], "binary": [-0.9, 1.0, 0.8], "count": [0.0, 0.75, 1.5]}[datatype][:N]
    U = [mu + np.sqrt(0.3) * rng.standard_normal((n, D)) for mu, n in zip(mus, dims)]
    total = int(np.prod(dims))
    nnz = int(0.6 * total)
    flat = rng.choice(total, size=nnz, replace=False)
    coords = np.stack(np.unravel_index(flat, dims), axis=1).astype(np.int64)
    f = np.ones((nnz, D))
    for d in range(N):
        f *= U[d][coords[:, d]]
    f = f.sum(axis=1)
    if datatype == "real":
        vals = f + 0.05 * rng.standard_normal(nnz)
    elif datatype == "binary":
        vals = np.where(rng.random(nnz) < 1.0 / (1.0 + np.exp(-f)), 1.0, -1.0)
    else:
        vals = rng.poisson(np.maximum(np.log1p(np.exp(f)), 1e-8)).astype(np.float64)
        vals = np.minimum(vals, 30.0)
    n_train = int(0.8 * nnz)
    return coords[:n_train], vals[:n_train], coords[n_train:], vals[n_train:]


def run_reference(case, data, seed):
    name, datatype, model, diag, dims, D, k1, k2, batch, iters, _ = case
    tr_c, tr_v, te_c, te_v = data
    ref = ref_shim.import_reference()
    vals = tr_v.astype(np.int64) if datatype == "count" else tr_v
    tensor = ref_shim.make_list_tensor(datatype, dims, tr_c, vals,
                                       max_count=int(tr_v.max()) if datatype == "count" else 0,
                                       min_count=int(tr_v.min()) if datatype == "count" else 0)
    tensor.test_entries, tensor.test_vals = te_c, te_v
    cls = {"d": ref["SSVI_TF_d"], "simple": ref["SSVI_TF_simple"], "robust": ref["SSVI_TF_robust"]}[model]
    np.random.seed(seed)
    mean0 = np.ones(D) if datatype == "real" else np.zeros(D)
    obj = cls(tensor, D, diag=diag, mean0=mean0, cov0=np.eye(D) * 0.1, k1=k1, k2=k2, batch_size=batch,
              randstart=True)
    devnull = open(os.devnull, "w")
    saved = sys.stdout
    sys.stdout = devnull
    try:
        obj.factorize(report=10 ** 9, max_iteration=iters, fixed_covariance=False, to_report=[],
                      detailed_report=False)
        rsme, err = obj.evaluate_test_error()
    finally:
        sys.stdout = saved
        devnull.close()
    return float(rsme), float(err)


def _job(args):
    ci, seed = args
    case = CASES[ci]
    data = synth(case[1], case[4], case[5], np.random.default_rng(sum(map(ord, case[0]))))
    return ci, seed, run_reference(case, data, seed)


def main():
    assert ref_shim.available(), "needs /root/reference"
    import multiprocessing as mp
    os.makedirs(OUT_DIR, exist_ok=True)
    only = sys.argv[1:]                                   # optional: names of the cases to (re)generate
    jobs = [(ci, 319 + s) for ci, c in enumerate(CASES) for s in range(c[10]) if not only or c[0] in only]
    t0 = time.time()
    with mp.get_context("spawn").Pool(min(8, os.cpu_count() or 1)) as pool:
        out = pool.map(_job, jobs, chunksize=1)
    for ci, case in enumerate(CASES):
        name, datatype, model, diag, dims, D, k1, k2, batch, iters, seeds = case
        if only and name not in only:
            continue
        data = synth(datatype, dims, D, np.random.default_rng(sum(map(ord, name))))
        res = np.array([r for c, s, r in sorted(out) if c == ci])
        print("%-20s rsme %s  err %s" % (name, np.round(res[:, 0], 4), np.round(res[:, 1], 4)))
        np.savez_compressed(os.path.join(OUT_DIR, name + ".npz"), datatype=datatype, model=model, diag=diag,
                            dims=np.array(dims), D=D, k1=k1, k2=k2, batch_size=batch, iterations=iters,
                            train_entries=data[0], train_vals=data[1], test_entries=data[2], test_vals=data[3],
                            ref_rsme=res[:, 0], ref_err=res[:, 1])
    print("%.0f s" % (time.time() - t0))


if __name__ == "__main__":
    main()
