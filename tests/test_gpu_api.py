"""GPU: the reference-facing Python surface (SSVI_TF_d / _simple / _robust, factorize, predict)."""
import numpy as np
import pytest
import torch

from oracle import ssvi_oracle as so
from tests.helpers import load_golden

pytestmark = pytest.mark.gpu


class _COO(object):
    """Smallest object satisfying the tensor interface the factorizer reads."""

    def __init__(self, datatype, dims, coords, vals):
        self.datatype, self.dims = datatype, list(dims)
        self.train_entries, self.train_vals = np.asarray(coords), np.asarray(vals, dtype=float)
        self.test_entries, self.test_vals = self.train_entries[:0], self.train_vals[:0]
        self.max_count = int(self.train_vals.max()) if datatype == "count" else 0
        self.min_count = int(self.train_vals.min()) if datatype == "count" else 0


@pytest.mark.parametrize("name", ["d_normal_full", "simple_poisson_diag", "robust_bernoulli_full",
                                  "robust_poisson_full_quadrature"])
def test_update_natural_param_batch_is_drop_in(name):
    """model.update_natural_param_batch(dim, i) under a supplied tape == the reference's own call."""
    import ssvi_tf_b200 as pkg
    case = load_golden(name)
    cfg = case["cfg"]
    dt = {"normal": "real", "bernoulli": "binary", "poisson": "count"}[cfg["likelihood"]]
    cls = {"d": pkg.SSVI_TF_d, "simple": pkg.SSVI_TF_simple, "robust": pkg.SSVI_TF_robust}[cfg["model"]]
    extra = dict(quadrature=True) if cfg.get("quadrature") else {}          # SSVI_TF_robust.py:11
    m = cls(_COO(dt, case["dims"], case["coords"], case["vals"]), cfg["D"], diag=cfg["diag"], k1=cfg["k1"],
            k2=cfg["k2"], batch_size=cfg["batch_size"], randstart=False, noise_dtype=torch.float64, **extra)
    if extra:
        assert m.quadrature and m.gauss_degree == 30
    st0 = case["state0"]
    for step in case["steps"]:
        for d in range(m.order):
            for i in range(m.dims[d]):
                if cfg["diag"]:
                    m.posterior.update_vector_distribution(d, i, st0["params"][d][i, 0], st0["params"][d][i, 1])
                else:
                    m.posterior.update_vector_distribution(d, i, st0["params"][d][i, 0], st0["params"][d][i, 1:])
            m.engine.ada_mean[d].copy_(torch.as_tensor(np.ascontiguousarray(st0["ada_mean"][d].T)))
            m.engine.ada_cov[d].copy_(torch.as_tensor(np.ascontiguousarray(st0["ada_cov"][d].T)))
        m.pSigma = st0["pSigma"]
        m.time_step = st0["time_step"]
        m.w_sigma, m.w_ada_grad = st0["w_sigma"], st0["w_ada_grad"]
        dim, row, tape = step["dim"], step["row"], step["tape"]
        m.update_natural_param_batch(dim, row, noise=pkg.NoisePool.from_tape(tape, m.engine.device, torch.float64, diag=cfg["diag"]),
                                     sub_idx=tape.sub_idx)
        m.engine.check_flags()
        mean, cov = m.posterior.get_vector_distribution(dim, row)
        want = step["after"]["params"][dim][row]
        np.testing.assert_allclose(mean, want[0], rtol=1e-5, atol=1e-9)
        np.testing.assert_allclose(cov, want[1] if cfg["diag"] else want[1:], rtol=1e-5, atol=1e-9)
        np.testing.assert_allclose(m.ada_acc_grad[dim].cpu().numpy(), step["after"]["ada_mean"][dim], rtol=1e-5)
        if cfg["noise_added"]:
            np.testing.assert_allclose(m.w_sigma, step["after"]["w_sigma"], rtol=1e-5)


def test_factorize_learns_synthetic_real_tensor():
    import ssvi_tf_b200 as pkg
    np.random.seed(319)
    D = 5
    t = pkg.RV_tensor()
    t.synthesize_data([30, 30, 30], [np.ones(D) * m for m in (0.5, 1.0, 0.2)], [np.eye(D) * 0.1] * 3, D=D,
                      train=0.8, sparsity=0.5, noise=0.05)
    m = pkg.SSVI_TF_d(t, D, mean0=np.ones(D), cov0=np.eye(D) * 0.1, k1=16, k2=16, batch_size=64)
    r0, _ = m.evaluate_test_error()
    m.factorize(report=10 ** 9, max_iteration=40, detailed_report=False, rows_per_iteration="all")
    r1, e1 = m.evaluate_test_error()
    assert np.isfinite(r1) and r1 < 0.5 * r0, (r0, r1)
    assert m.time_step == [41, 41, 41]
    # empty-row bookkeeping (Q11): a row without observations is recorded and skipped
    assert isinstance(m.missing, dict)
    # reference schedule: one row per mode per iteration
    m2 = pkg.SSVI_TF_d(t, D, mean0=np.ones(D), cov0=np.eye(D) * 0.1, k1=16, k2=16, batch_size=64)
    before = m2.posterior.params[0].clone()
    m2.factorize(report=10 ** 9, max_iteration=2, detailed_report=False)
    changed = (m2.posterior.params[0] != before).flatten(1).any(dim=1).cpu().numpy()
    assert list(np.nonzero(changed)[0]) == [0, 1, 2]


def test_reset_and_errors():
    import ssvi_tf_b200 as pkg
    np.random.seed(5)
    t = pkg.binary_tensor()
    t.synthesize_data([10, 9, 8], None, None, D=3, train=0.8, sparsity=0.6)
    with pytest.raises(Exception, match="Unidentified update formula"):
        pkg.SSVI_TF_d(t, 3, mean_update="X")
    m = pkg.SSVI_TF_simple(t, 3, diag=True, k1=8, k2=8, batch_size=16, randstart=False)
    m.factorize(report=10 ** 9, max_iteration=5, detailed_report=False, rows_per_iteration="all")
    assert m.w_ada_grad > 0
    m.reset()
    assert m.w_sigma == 1.0 and m.w_ada_grad == 0.0 and m.time_step == [0, 0, 0]
    assert float(m.ada_acc_grad[0].abs().sum()) == 0.0          # Q9: reset() zeros the accumulator
    np.testing.assert_array_equal(m.posterior.get_vector_distribution(0, 0)[0], np.ones(3))
    rs, err = m.evaluate_train_error()
    assert 0.0 <= err <= 1.0


def test_true_param_reports_and_state_mirrors(capsys):
    """The drivers' post-factorize report (synthetic_test.py:152 -> evaluate_true_params, :1213-1320), the pSigma
    write-through mirror and a stale-free check_stop_cond after a manual update."""
    import ssvi_tf_b200 as pkg
    np.random.seed(7)
    D = 4
    t = pkg.RV_tensor()
    t.synthesize_data([9, 8, 7], [np.ones(D) * m for m in (0.5, 1.0, 0.2)], [np.eye(D) * 0.1] * 3, D=D,
                      train=0.8, sparsity=0.7, noise=0.05)
    m = pkg.SSVI_TF_d(t, D, mean0=np.ones(D), cov0=np.eye(D) * 0.1, k1=16, k2=16, batch_size=32)
    m.factorize(report=10 ** 9, max_iteration=6, detailed_report=False, rows_per_iteration="all")
    out = m.evaluate_true_params()
    assert "Evaluation for true params" in capsys.readouterr().out
    test_rsme, train_rsme = out[0], out[1]
    assert np.isfinite(out).all() and test_rsme < 0.5 and train_rsme < 0.5      # the generating factors fit their own data
    f = t.matrices[0][1] * t.matrices[1][2] * t.matrices[2][3]
    assert m.true_model_predict([1, 2, 3], t.matrices) == pytest.approx(float(f.sum()))
    mean, var = m.compute_posterior_param([1, 2, 3])
    assert mean == pytest.approx(float(m.predict_entry([1, 2, 3]))) and var > 0
    # pSigma: item assignment writes through to the device; the returned array itself is a host copy
    m.pSigma[1] = 2.5
    assert float(m.engine.pSigma[1]) == 2.5 and float(m.pSigma[1]) == 2.5
    # a manual update after factorize must not leave check_stop_cond with the loop's stale reductions
    m.check_stop_cond()
    m.update_natural_param_batch(0, 0)
    m.engine.norm_changes[0][0, 0] = 123.0
    assert m.check_stop_cond()[0] == 123.0


@pytest.mark.parametrize("model,datatype,diag", [("d", "real", False), ("simple", "binary", True), ("robust", "count", False)])
def test_device_loop_equals_per_call_loop_including_the_stop(model, datatype, diag):
    """factorize() in the reference's schedule: the loop enqueued from C (ssvi_factorize_iterations) must leave exactly
    the state of the per-call Python loop -- same minibatch draws, same variates, same pointer / time_step advance
    (:117-133) -- also when check_stop_cond fires in the middle of an enqueued block (:135-142)."""
    import ssvi_tf_b200 as pkg
    np.random.seed(11)
    D = 4
    t = {"real": pkg.RV_tensor, "binary": pkg.binary_tensor, "count": pkg.count_tensor}[datatype]()
    means = None if datatype != "real" else [np.ones(D) * m for m in (0.5, 1.0, 0.2)]
    covs = None if datatype != "real" else [np.eye(D) * 0.1] * 3
    t.synthesize_data([7, 40, 5], means, covs, D=D, train=0.8, sparsity=0.9)
    cls = {"d": pkg.SSVI_TF_d, "simple": pkg.SSVI_TF_simple, "robust": pkg.SSVI_TF_robust}[model]

    def run(device_loop, epsilon, iters, trace=None):
        np.random.seed(3)
        m = cls(t, D, diag=diag, mean0=np.ones(D) if datatype == "real" else np.zeros(D), cov0=np.eye(D) * 0.1,
                k1=16, k2=8, batch_size=32, noise_dtype=torch.float64, noise_seed=5)
        m.device_loop, m.epsilon = device_loop, epsilon
        if trace is not None:
            orig = m.check_stop_cond
            m.check_stop_cond = lambda: trace.append(orig()) or trace[-1]
        m.factorize(report=10 ** 9, max_iteration=iters, detailed_report=False)
        st = [p.cpu().numpy().copy() for p in m.engine.params] + [m.engine.ada_mean[0].cpu().numpy().copy(),
                                                                  m.engine.w_state.cpu().numpy().copy(), np.asarray(m.pSigma)]
        return st, list(m.time_step), (m.d_mean, m.d_cov), dict(m.missing)

    trace = []
    a = run(False, 1e-4, 23, trace)
    b = run(True, 1e-4, 23)
    for x, y in zip(a[0], b[0]):
        np.testing.assert_array_equal(x, y)
    assert a[1:] == b[1:]
    # a stop in the middle of an enqueued block: epsilon just above the smallest change seen in iterations 1..15
    eps = min(max(tr) for tr in trace[1:16]) * (1 + 1e-9)
    stop_at = next(i for i, tr in enumerate(trace) if max(tr) < eps)
    assert stop_at <= 15
    a = run(False, eps, 23)
    b = run(True, eps, 23)
    for x, y in zip(a[0], b[0]):
        np.testing.assert_array_equal(x, y)
    assert a[1:] == b[1:]
    assert a[1] == [(stop_at + 1) // n for n in (7, 40, 5)]
