#!/usr/bin/env python
"""Print end-to-end test metrics of the B200 build next to the reference's (tests/golden_e2e). Needs a GPU."""
import glob, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ssvi_tf_b200 as pkg
from tests.test_gpu_e2e import _Split
for p in sorted(glob.glob(os.path.join(ROOT, "tests", "golden_e2e", "*.npz"))):
    z = np.load(p); t = _Split(z)
    D = int(z["D"]); cls = {"d": pkg.SSVI_TF_d, "simple": pkg.SSVI_TF_simple, "robust": pkg.SSVI_TF_robust}[str(z["model"])]
    rs, er = [], []
    for seed in (319, 320, 321, 322):
        np.random.seed(seed)
        m = cls(t, D, diag=bool(z["diag"]), mean0=np.ones(D) if t.datatype == "real" else np.zeros(D), cov0=np.eye(D) * 0.1,
                k1=int(z["k1"]), k2=int(z["k2"]), batch_size=int(z["batch_size"]), randstart=True, noise_dtype=torch.float64, noise_seed=seed)
        import io, contextlib
        with contextlib.redirect_stdout(io.StringIO()):
            m.factorize(report=10 ** 9, max_iteration=int(z["iterations"]), detailed_report=False)
        r, e = m.evaluate_test_error(); rs.append(r); er.append(e)
    print("%-20s b200 rsme %s err %s | reference rsme %s err %s" % (os.path.basename(p)[:-4], np.round(rs, 4), np.round(er, 4),
          np.round(z["ref_rsme"], 4), np.round(z["ref_err"], 4)))
