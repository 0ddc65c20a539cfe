#!/usr/bin/env python
"""Final test RMSE of the B200 factorize over many seeds for one tests/golden_e2e case. usage: e2e_seeds.py case n"""
import contextlib, io, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ssvi_tf_b200 as pkg
from tests.test_gpu_e2e import _Split
z = np.load(os.path.join(ROOT, "tests", "golden_e2e", sys.argv[1] + ".npz")); t = _Split(z)
D = int(z["D"]); cls = {"d": pkg.SSVI_TF_d, "simple": pkg.SSVI_TF_simple, "robust": pkg.SSVI_TF_robust}[str(z["model"])]
out = []
for seed in range(100, 100 + int(sys.argv[2])):
    np.random.seed(seed)
    m = cls(t, D, diag=bool(z["diag"]), mean0=np.ones(D) if t.datatype == "real" else np.zeros(D), cov0=np.eye(D) * 0.1,
            k1=int(z["k1"]), k2=int(z["k2"]), batch_size=int(z["batch_size"]), randstart=True, noise_dtype=torch.float64, noise_seed=seed)
    with contextlib.redirect_stdout(io.StringIO()):
        m.factorize(report=10 ** 9, max_iteration=int(z["iterations"]), detailed_report=False)
    out.append(m.evaluate_test_error()[0])
print(sys.argv[1], "b200 rsme", np.round(out, 4), "median %.4f mean %.4f" % (np.median(out), np.mean(out)))
