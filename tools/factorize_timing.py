#!/usr/bin/env python
"""Wall-clock of the drop-in factorize() loop on BASELINE config 1 (synthetic_test.py: real 50x50x50, D = 10, normal
likelihood, SSVI_TF_d, full covariance, k1 = 64, k2 = 128, batch_size = 128), through the reference's own surface:
RV_tensor.synthesize_data -> SSVI_TF_d(...) -> factorize().  Needs a GPU.  Prints one JSON line.

  reference schedule   one row per mode per iteration (SSVI_TF_interface.py:117-133): 3 row updates of <= 128 entries
  rows_per_iteration   "all": every iteration a full pass over every row of every mode (minibatch of 128 per row)
"""
import contextlib, io, json, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ssvi_tf_b200 as pkg


def run(iters, rpi, seed=319):
    np.random.seed(seed)
    D, dims = 10, [50, 50, 50]
    t = pkg.RV_tensor()
    means = [np.ones(D) * 0.5, np.ones(D) * 1.0, np.ones(D) * 0.2]
    covs = [np.eye(D) * 0.1] * 3
    with contextlib.redirect_stdout(io.StringIO()):
        t.synthesize_data(dims, means, covs, D, 0.8, 1, 0.1)
        m = pkg.SSVI_TF_d(t, D, mean0=np.ones(D), cov0=np.eye(D) * 0.1, k1=64, k2=128, batch_size=128, randstart=True,
                          chunk_entries=int(os.environ.get("SSVI_CHUNK", "16")))
        m.device_loop = os.environ.get("SSVI_DEVICE_LOOP", "1") != "0"     # 0: the per-call Python loop
        m.factorize(report=10 ** 9, max_iteration=20, detailed_report=False, rows_per_iteration=rpi)    # warm-up
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        m.factorize(report=10 ** 9, max_iteration=iters, detailed_report=False, rows_per_iteration=rpi)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        rmse, err = m.evaluate_test_error()
    rows = 3 if rpi == 1 else sum(dims)
    return dict(schedule="one row per mode" if rpi == 1 else "all rows", device_loop=bool(m.device_loop and rpi == 1),
                iterations=iters + 1, seconds=round(dt, 3),
                iterations_per_s=round((iters + 1) / dt, 1), row_updates_per_s=round((iters + 1) * rows / dt, 1),
                entry_updates_per_s=round((iters + 1) * rows * 128 / dt, 1), test_rmse=round(float(rmse), 4))


if __name__ == "__main__":
    out = dict(config="c1: real 50x50x50 dense, D=10, normal, SSVI_TF_d full cov, k1=64, batch_size=128",
               runs=[run(2000, 1), run(100, "all")],
               reference_cpu="SURVEY.md 8(d): 650-715 entry-updates/s on one core = ~5.3 row updates/s = ~1.8 iterations/s")
    print(json.dumps(out))
