"""Shared test helpers: move oracle-style state dicts in and out of an Engine."""
import glob
import os

import numpy as np
import torch

from oracle import ssvi_oracle as so
from oracle.golden_io import load_case

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden_cases():
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN_DIR, "*.npz")))


def load_golden(name):
    return load_case(os.path.join(GOLDEN_DIR, name + ".npz"))


def engine_load_state(eng, st):
    for d in range(eng.N):
        eng.params[d].copy_(torch.as_tensor(st["params"][d]))
        eng.ada_mean[d].copy_(torch.as_tensor(np.ascontiguousarray(st["ada_mean"][d].T)))
        eng.ada_cov[d].copy_(torch.as_tensor(np.ascontiguousarray(st["ada_cov"][d].T)))
        eng.norm_changes[d].copy_(torch.as_tensor(st["norm_changes"][d]))
        eng.grad_norms[d].copy_(torch.as_tensor(st["grad_norms"][d]))
    eng.pSigma.copy_(torch.as_tensor(np.asarray(st["pSigma"], dtype=np.float64)))
    eng.time_step = [int(t) for t in st["time_step"]]
    eng.w_state.copy_(torch.tensor([st["w_sigma"], st["w_ada_grad"], st["w_tau"]], dtype=torch.float64))
    eng.refresh_sampler()


def engine_dump_state(eng, like):
    out = dict(like)
    out["params"] = [p.cpu().numpy().copy() for p in eng.params]
    out["ada_mean"] = [a.cpu().numpy().T.copy() for a in eng.ada_mean]
    out["ada_cov"] = [a.cpu().numpy().T.copy() for a in eng.ada_cov]
    out["norm_changes"] = [a.cpu().numpy().copy() for a in eng.norm_changes]
    out["grad_norms"] = [a.cpu().numpy().copy() for a in eng.grad_norms]
    out["pSigma"] = eng.pSigma.cpu().numpy().copy()
    w = eng.w_state.cpu().numpy()
    out["w_sigma"], out["w_ada_grad"], out["w_tau"] = float(w[0]), float(w[1]), float(w[2])
    return out


def assert_state_close(got, want, cfg, rtol, what="", atol_frac=1e-3):
    """fp64 parity bar of BASELINE.json: 1e-5 relative on updated means / covariances (elements below atol_frac of the
    array's scale are compared absolutely, at rtol * atol_frac of that scale)."""
    N = len(want["params"])
    keys = ["params", "ada_mean", "norm_changes", "grad_norms"] + (["ada_cov"] if cfg["diag"] else [])
    for d in range(N):
        for k in keys:
            scale = max(1.0, float(np.abs(want[k][d]).max()))
            np.testing.assert_allclose(got[k][d], want[k][d], rtol=rtol, atol=rtol * atol_frac * scale,
                                       err_msg="%s %s[%d]" % (what, k, d))
    if cfg["noise_added"]:
        np.testing.assert_allclose(got["w_sigma"], want["w_sigma"], rtol=rtol, err_msg=what + " w_sigma")
        np.testing.assert_allclose(got["w_ada_grad"], want["w_ada_grad"], rtol=rtol, err_msg=what + " w_ada_grad")


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
