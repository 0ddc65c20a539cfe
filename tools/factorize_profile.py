#!/usr/bin/env python
"""cProfile of the drop-in factorize() loop (reference schedule) on config 1: where the host time of an iteration goes."""
import cProfile, contextlib, io, os, pstats, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ssvi_tf_b200 as pkg
np.random.seed(319)
D, dims = 10, [50, 50, 50]
t = pkg.RV_tensor()
with contextlib.redirect_stdout(io.StringIO()):
    t.synthesize_data(dims, [np.ones(D) * 0.5, np.ones(D), np.ones(D) * 0.2], [np.eye(D) * 0.1] * 3, D, 0.8, 1, 0.1)
    m = pkg.SSVI_TF_d(t, D, mean0=np.ones(D), cov0=np.eye(D) * 0.1, k1=64, k2=128, batch_size=128)
    m.factorize(report=10 ** 9, max_iteration=50, detailed_report=False)
    pr = cProfile.Profile()
    pr.enable()
    m.factorize(report=10 ** 9, max_iteration=1000, detailed_report=False)
    pr.disable()
st = pstats.Stats(pr)
st.sort_stats("cumtime").print_stats(45)
