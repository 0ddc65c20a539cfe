"""Run the UNMODIFIED reference (/root/reference) under a deterministic noise tape.

TEST INFRASTRUCTURE ONLY.  Nothing in the product package imports this file.  It
is used in the build container (where /root/reference exists) to
  (1) validate oracle/ssvi_oracle.py (the travelling numpy restatement), and
  (2) generate the golden fixtures under tests/golden/ (oracle/gen_golden.py).
It cannot run on the GPU box: /root/reference does not exist there.

Three compatibility shims, none of which edits a reference file:
  shim 1  Tensor.create_random_matrix is called (Tensor/real_tensor.py:19,
          binary_tensor.py:19, count_tensor.py:23) but defined nowhere.
  shim 2  np.take on a ragged list of (coord, val) tuples
          (SSVI/SSVI_TF_interface.py:172) needs an explicit object array on numpy>=1.24.
  shim 3  np.math.factorial (Probability/poisson.py:23): np.math was removed in numpy 2.

The noise tape replaces numpy.random.{choice,multivariate_normal,normal,rayleigh}
for the duration of ONE reference call with functions that pop pre-generated
standard variates in the call order of update_natural_param_batch
(SSVI/SSVI_TF_interface.py:163-264, SSVI_TF_robust.py:28-65, :147-169):
  #0 choice(len, batch_size, replace=False)        -> tape.sub_idx         (only if len > batch_size)
  #1 multivariate_normal(mj, Sj, size=k1)          -> mj + z_other[b, jj] @ chol(Sj).T   for b, for jj
  #2 multivariate_normal(m, S, size=(B, k1))       -> m + z_own @ chol(S).T
  #3 rayleigh(w_sigma**2, size=(B, k1))            -> w_sigma**2 * r_w
  #4 normal(mean_b, scale_b, size=(k2, k1)) per b  -> mean_b + scale_b * z_f[b]
The transform x = m + L z (L = lower Cholesky) is the contract of this build; stock
NumPy factorises by SVD and would give different samples from the same z.
"""
import contextlib
import math
import os
import sys

import numpy as np

REFERENCE_ROOT = os.environ.get("SSVI_REFERENCE_ROOT", "/root/reference")


def available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "SSVI"))


class _MathShim(object):
    """shim 3: stands in for the removed numpy.math; accepts integral floats like py3.5 did."""

    def __getattr__(self, name):
        return getattr(math, name)

    @staticmethod
    def factorial(y):
        return math.factorial(int(y))


_imported = {}


def import_reference():
    """Import the reference packages read-only (no bytecode is written into the tree)."""
    if _imported:
        return _imported
    if not available():
        raise RuntimeError("reference tree not present at %s" % REFERENCE_ROOT)
    sys.dont_write_bytecode = True
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    if not hasattr(np, "math"):
        np.math = _MathShim()
    from Tensor.Tensor import Tensor
    from Tensor.real_tensor import RV_tensor
    from Tensor.binary_tensor import binary_tensor
    from Tensor.count_tensor import count_tensor
    from SSVI.SSVI_TF_d import SSVI_TF_d
    from SSVI.SSVI_TF_simple import SSVI_TF_simple
    from SSVI.SSVI_TF_robust import SSVI_TF_robust

    # shim 1
    if not hasattr(Tensor, "create_random_matrix"):
        def create_random_matrix(self, n, D, mean, cov):
            return np.random.multivariate_normal(mean, cov, size=n)
        Tensor.create_random_matrix = create_random_matrix

    # shim 2
    def find_observed_ui(self, dim, i):
        lst = self.observed_by_id[dim][i]
        out = np.empty((len(lst), 2), dtype=object)
        for r, (coord, val) in enumerate(lst):
            out[r, 0] = coord
            out[r, 1] = val
        return out
    Tensor.find_observed_ui = find_observed_ui

    _imported.update(dict(Tensor=Tensor, RV_tensor=RV_tensor, binary_tensor=binary_tensor,
                          count_tensor=count_tensor, SSVI_TF_d=SSVI_TF_d,
                          SSVI_TF_simple=SSVI_TF_simple, SSVI_TF_robust=SSVI_TF_robust))
    return _imported


class NoiseTape(object):
    """Standard variates for ONE row minibatch, in numpy-natural layout.

    z_other (B, N-1, k1, D)   z_own (B, k1, D)   r_w (B, k1) unit Rayleigh
    z_f (B, k2, k1)           sub_idx (B,) int or None
    """

    def __init__(self, z_other, z_own, r_w=None, z_f=None, sub_idx=None):
        self.z_other, self.z_own, self.r_w, self.z_f, self.sub_idx = z_other, z_own, r_w, z_f, sub_idx
        self.reset()

    def reset(self):
        self._b = 0
        self._jj = 0
        self._fb = 0
        self.calls = []


@contextlib.contextmanager
def patched_rng(tape):
    """Patch numpy.random for one reference step; restores on exit."""
    saved = (np.random.choice, np.random.multivariate_normal, np.random.normal, np.random.rayleigh)
    n_other = tape.z_other.shape[1]

    def choice(n, size=None, replace=True):
        assert tape.sub_idx is not None and replace is False and len(tape.sub_idx) == size
        tape.calls.append("choice")
        return np.asarray(tape.sub_idx)

    def multivariate_normal(mean, cov, size=None):
        L = np.linalg.cholesky(np.asarray(cov, dtype=float))
        if isinstance(size, tuple):        # draw #2: own row, (B, k1)
            z = tape.z_own
            assert z.shape[:2] == tuple(size)
            tape.calls.append("mvn_own")
        else:                              # draw #1: other-mode row for entry b, mode slot jj
            z = tape.z_other[tape._b, tape._jj]
            assert z.shape[0] == size
            tape._jj += 1
            if tape._jj == n_other:
                tape._jj = 0
                tape._b += 1
            tape.calls.append("mvn_other")
        return np.asarray(mean) + z @ L.T

    def normal(loc=0.0, scale=1.0, size=None):
        z = tape.z_f[tape._fb]
        assert z.shape == tuple(size)
        tape._fb += 1
        tape.calls.append("normal")
        return loc + scale * z

    def rayleigh(scale=1.0, size=None):
        assert tape.r_w.shape == tuple(size)
        tape.calls.append("rayleigh")
        return scale * tape.r_w

    np.random.choice, np.random.multivariate_normal = choice, multivariate_normal
    np.random.normal, np.random.rayleigh = normal, rayleigh
    try:
        yield
    finally:
        (np.random.choice, np.random.multivariate_normal, np.random.normal, np.random.rayleigh) = saved


def make_list_tensor(datatype, dims, coords, vals, max_count=0, min_count=0):
    """A reference Tensor object built straight from COO arrays, with observed_by_id built
    exactly as Tensor/Tensor.py:76-83 does (append (coord, val) in entry order, per mode)."""
    ref = import_reference()
    t = ref["Tensor"](datatype=datatype)
    t.dims = list(dims)
    t.train_entries = np.asarray(coords)
    t.train_vals = np.asarray(vals)
    t.test_entries = t.train_entries[:0]
    t.test_vals = t.train_vals[:0]
    t.max_count, t.min_count = max_count, min_count
    t.observed_by_id = [[[] for _ in range(x)] for x in t.dims]
    for i in range(len(t.train_entries)):
        coord = t.train_entries[i]
        val = t.train_vals[i]
        for dim, col in enumerate(coord):
            t.observed_by_id[dim][col].append((coord, val))
    return t


def make_reference_model(model, datatype, dims, coords, vals, D, diag, k1, k2, batch_size,
                         mean_update="S", cov_update="N", eta=1, cov_eta=1, sigma_eta=1, quadrature=False):
    """Build a reference SSVI_TF_{d,simple,robust} object over a COO tensor (randstart=False;
    the caller then overwrites the state with load_state)."""
    ref = import_reference()
    cls = {"d": ref["SSVI_TF_d"], "simple": ref["SSVI_TF_simple"], "robust": ref["SSVI_TF_robust"]}[model]
    vals = np.asarray(vals)
    if datatype == "count":
        vals = vals.astype(np.int64)
    tensor = make_list_tensor(datatype, dims, coords, vals,
                         max_count=int(vals.max()) if datatype == "count" else 0,
                         min_count=int(vals.min()) if datatype == "count" else 0)
    extra = dict(quadrature=True) if quadrature else {}        # SSVI_TF_robust.py:11 (robust only)
    obj = cls(tensor, D, mean_update=mean_update, cov_update=cov_update, diag=diag,
              mean0=np.ones((D,)), cov0=np.eye(D), k1=k1, k2=k2, batch_size=batch_size,
              eta=eta, cov_eta=cov_eta, sigma_eta=sigma_eta, randstart=False, **extra)
    obj.missing = {}
    obj.fixed_covariance = False
    return obj


def load_state(obj, state):
    """Copy an oracle state dict (oracle/ssvi_oracle.py:new_state) into a reference object."""
    for d in range(len(obj.dims)):
        obj.posterior.params[d][...] = state["params"][d]
        obj.ada_acc_grad[d][...] = state["ada_mean"][d]
        if obj.diag:
            obj.ada_acc_grad_cov[d][...] = state["ada_cov"][d]
        obj.norm_changes[d][...] = state["norm_changes"][d]
        obj.grad_norms[d][...] = state["grad_norms"][d]
    obj.pSigma[...] = state["pSigma"]
    obj.time_step = [int(t) for t in state["time_step"]]
    if obj.noise_added:
        obj.w_sigma = float(state["w_sigma"])
        obj.w_ada_grad = float(state["w_ada_grad"])
        obj.w_tau = float(state["w_tau"])
    obj.fixed_covariance = bool(state.get("fixed_covariance", False))


def dump_state(obj, like):
    """Read the reference object's mutable state back into an oracle-style dict."""
    out = dict(like)
    out["params"] = [p.copy() for p in obj.posterior.params]
    out["ada_mean"] = [a.copy() for a in obj.ada_acc_grad]
    out["ada_cov"] = [a.copy() for a in obj.ada_acc_grad_cov] if obj.diag else like["ada_cov"]
    out["norm_changes"] = [a.copy() for a in obj.norm_changes]
    out["grad_norms"] = [a.copy() for a in obj.grad_norms]
    out["pSigma"] = np.array(obj.pSigma, dtype=float)
    out["time_step"] = list(obj.time_step)
    if obj.noise_added:
        out["w_sigma"] = float(obj.w_sigma)
        out["w_ada_grad"] = float(obj.w_ada_grad)
    return out


def reference_row_step(obj, dim, row, tape):
    """ONE call of the reference's update_natural_param_batch(dim, row) under the tape."""
    tape.reset()
    with patched_rng(tape):
        obj.update_natural_param_batch(dim, row)
    return tape.calls
