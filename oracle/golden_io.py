"""(De)serialisation of golden cases (tests/golden/*.npz).  TEST INFRASTRUCTURE ONLY.

A case = config + COO tensor + one state snapshot + several independent single-row steps,
each step = (dim, row, noise tape, reference state after the step from that snapshot).
"""
import json

import numpy as np

from oracle import ssvi_oracle as so

_LIST_KEYS = ("params", "ada_mean", "ada_cov", "norm_changes", "grad_norms")
_SCALAR_KEYS = ("w_sigma", "w_ada_grad", "w_tau")


def _pack_state(prefix, st, out):
    for k in _LIST_KEYS:
        for d, a in enumerate(st[k]):
            out["%s_%s_%d" % (prefix, k, d)] = a
    out[prefix + "_pSigma"] = st["pSigma"]
    out[prefix + "_pmu"] = st["pmu"]
    out[prefix + "_time_step"] = np.asarray(st["time_step"], dtype=np.int64)
    out[prefix + "_scalars"] = np.array([st[k] for k in _SCALAR_KEYS], dtype=np.float64)
    out[prefix + "_fixed_covariance"] = np.array(int(bool(st.get("fixed_covariance", False))))


def _unpack_state(prefix, z, N):
    st = {}
    for k in _LIST_KEYS:
        st[k] = [z["%s_%s_%d" % (prefix, k, d)].copy() for d in range(N)]
    st["pSigma"] = z[prefix + "_pSigma"].copy()
    st["pmu"] = z[prefix + "_pmu"].copy()
    st["time_step"] = [int(t) for t in z[prefix + "_time_step"]]
    for k, v in zip(_SCALAR_KEYS, z[prefix + "_scalars"]):
        st[k] = float(v)
    st["fixed_covariance"] = bool(int(z[prefix + "_fixed_covariance"]))
    return st


def save_case(path, cfg, dims, coords, vals, state0, steps):
    out = dict(cfg=np.array(json.dumps(cfg)), dims=np.asarray(dims, dtype=np.int64),
               coords=np.asarray(coords, dtype=np.int64), vals=np.asarray(vals, dtype=np.float64),
               n_steps=np.array(len(steps)))
    _pack_state("s0", state0, out)
    for k, st in enumerate(steps):
        p = "st%d" % k
        out[p + "_dim"] = np.array(st["dim"])
        out[p + "_row"] = np.array(st["row"])
        t = st["tape"]
        out[p + "_z_other"] = t.z_other.astype(np.float32)     # tapes are float32-representable
        out[p + "_z_own"] = t.z_own.astype(np.float32)
        if cfg["noise_added"]:
            out[p + "_r_w"] = t.r_w.astype(np.float32)
            out[p + "_z_f"] = t.z_f.astype(np.float32)
        if t.sub_idx is not None:
            out[p + "_sub_idx"] = np.asarray(t.sub_idx, dtype=np.int64)
        _pack_state(p + "_after", st["after"], out)
    np.savez_compressed(path, **out)


def load_case(path):
    z = np.load(path, allow_pickle=False)
    cfg = json.loads(str(z["cfg"]))
    dims = [int(x) for x in z["dims"]]
    N = len(dims)
    case = dict(cfg=cfg, dims=dims, coords=z["coords"], vals=z["vals"], state0=_unpack_state("s0", z, N),
                steps=[])
    for k in range(int(z["n_steps"])):
        p = "st%d" % k
        f64 = lambda key: z[key].astype(np.float64) if key in z.files else None   # noqa: E731
        tape = so.Tape(f64(p + "_z_other"), f64(p + "_z_own"), f64(p + "_r_w"), f64(p + "_z_f"),
                       z[p + "_sub_idx"] if (p + "_sub_idx") in z.files else None)
        case["steps"].append(dict(dim=int(z[p + "_dim"]), row=int(z[p + "_row"]), tape=tape,
                                  after=_unpack_state(p + "_after", z, N)))
    return case
