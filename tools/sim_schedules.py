#!/usr/bin/env python
"""CPU study (oracle only): does the bench's schedule -- full-row Jacobi sweeps of every mode with a pool of P noise slots
shared by a row's entries -- reach the same end-to-end test metric as the reference's schedule (one row per mode per
iteration, own draws per entry)?  Runs oracle/ssvi_oracle.py both ways on a tests/golden_e2e problem and prints the test
RMSE / error beside the reference's own numbers stored in the fixture.  usage: sim_schedules.py <case> [P] [seeds]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ssvi_oracle as so     # noqa: E402

LIK = {"real": "normal", "binary": "bernoulli", "count": "poisson"}


def metric(st, cfg, z, datatype):
    f = so.predict_mean(st, z["test_entries"])
    y = z["test_vals"].astype(float)
    pred = f if datatype == "real" else np.where(f > 0, 1.0, -1.0)
    return float(np.sqrt(np.mean((pred - y) ** 2))), float(np.mean(pred != y)) if datatype == "binary" else float(np.mean(np.abs(pred - y)))


def run(z, seed, schedule, P, sweeps=None):
    datatype, model = str(z["datatype"]), str(z["model"])
    dims = [int(x) for x in z["dims"]]
    D, k1, k2, B = int(z["D"]), int(z["k1"]), int(z["k2"]), int(z["batch_size"])
    N = len(dims)
    rng = np.random.default_rng(seed)
    cfg = so.make_config(model, LIK[datatype], D, diag=bool(z["diag"]), k1=k1, k2=k2, batch_size=B)
    st = so.new_state(cfg, dims, rng=rng, mean0=np.ones(D) if datatype == "real" else np.zeros(D), cov0=np.eye(D) * 0.1)
    coords, vals = z["train_entries"], z["train_vals"].astype(float)
    rowptr, perm = so.build_index(coords, dims)
    iters = int(z["iterations"])
    if schedule == "reference":
        ptr = [0] * N
        for it in range(iters + 1):
            for dim in range(N):
                row = ptr[dim]
                ids = perm[dim][rowptr[dim][row]:rowptr[dim][row + 1]]
                tape = so.make_tape(rng, cfg, N, len(ids), n_obs=len(ids))
                if tape.sub_idx is not None:
                    ids = ids[tape.sub_idx]
                so.row_step(st, cfg, dim, row, coords[ids], vals[ids], tape)
                so.update_hyper_parameter(st, cfg, dim)
            for dim in range(N):
                ptr[dim] += 1
                if ptr[dim] >= dims[dim]:
                    ptr[dim] = 0
                    st["time_step"][dim] += 1
    else:
        n_sweeps = sweeps or max(1, int(round((iters + 1) / np.mean(dims))))
        big = dict(cfg)
        big["batch_size"] = 10 ** 9
        for s in range(n_sweeps):
            for dim in range(N):
                pool = so.make_tape(rng, big, N, P)
                entries, tapes = [], []
                for row in range(dims[dim]):
                    ids = perm[dim][rowptr[dim][row]:rowptr[dim][row + 1]]
                    sl = np.arange(len(ids)) % P
                    tapes.append(so.Tape(pool.z_other[sl], pool.z_own[sl], pool.r_w[sl], pool.z_f[sl]))
                    entries.append((coords[ids], vals[ids]))
                so.rows_step(st, big, dim, np.arange(dims[dim]), entries, tapes)
                so.update_hyper_parameter(st, big, dim)
                st["time_step"][dim] += 1
    return metric(st, cfg, z, datatype)


if __name__ == "__main__":
    name = sys.argv[1]
    P = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    seeds = int(sys.argv[3]) if len(sys.argv) > 3 else 3
    z = np.load(os.path.join(ROOT, "tests", "golden_e2e", name + ".npz"))
    print(name, "reference (fixture): rsme %.4f +- %.4f  err %.4f" % (z["ref_rsme"].mean(), z["ref_rsme"].std(), z["ref_err"].mean()))
    for sched in ("reference", "bench"):
        out = np.array([run(z, 100 + s, sched, P) for s in range(seeds)])
        print("  oracle, %-9s schedule: rsme %.4f +- %.4f   err/mae %.4f   %s" % (sched, out[:, 0].mean(), out[:, 0].std(), out[:, 1].mean(), np.round(out[:, 0], 4)))
