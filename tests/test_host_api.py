"""CPU: host-side mirrors (likelihood classes, tensor containers, index) against the oracle."""
import numpy as np
import pytest

import ssvi_tf_b200 as pkg
from oracle import ssvi_oracle as so


@pytest.mark.parametrize("lik,cls", [("normal", pkg.NormalDistribution), ("bernoulli", pkg.BernoulliDistribution),
                                     ("poisson", pkg.PoissonDistribution)])
def test_likelihood_classes_match_oracle(lik, cls):
    rng = np.random.default_rng(0)
    obj = cls()
    f = rng.normal(0.3, 2.0, size=(5, 7))
    f[0, 0], f[0, 1] = -40.0, 40.0
    ys = {"normal": [0.7, -1.2], "bernoulli": [1.0, -1.0], "poisson": [0.0, 3.0, 12.0]}[lik]
    s = 1.0
    for y in ys:
        np.testing.assert_allclose(obj.pdf(y, f, s), so.pdf(lik, y, f, s), rtol=1e-13)
        np.testing.assert_allclose(obj.fst_derivative_pdf(y, f, s), so.fst_derivative_pdf(lik, y, f, s), rtol=1e-13)
        np.testing.assert_allclose(obj.snd_derivative_pdf(y, f, s), so.snd_derivative_pdf(lik, y, f, s), rtol=1e-12, atol=1e-300)
        np.testing.assert_allclose(obj.fst_derivative_log_pdf(y, f, s), so.fst_derivative_log_pdf(lik, y, f, s), rtol=1e-13)
        np.testing.assert_allclose(obj.snd_derivative_log_pdf(y, f, s), so.snd_derivative_log_pdf(lik, y, f, s), rtol=1e-13)


def test_tensor_index_matches_reference_order():
    np.random.seed(319)
    t = pkg.RV_tensor()
    D = 4
    t.synthesize_data([6, 5, 4], [np.ones(D) * m for m in (0.5, 1.0, 0.2)], [np.eye(D) * 0.1] * 3, D=D,
                      train=0.8, sparsity=0.5)
    n_obs = int(6 * 5 * 4 * 0.5)
    assert len(t.train_vals) == int(n_obs * 0.8) and len(t.test_vals) == n_obs - len(t.train_vals)
    assert len(np.unique(np.ravel_multi_index(t.observed_entries.T, t.dims))) == n_obs
    rowptr, perm = so.build_index(t.train_entries, t.dims)
    for dim in range(3):
        np.testing.assert_array_equal(t.index()[0][dim], rowptr[dim])
        np.testing.assert_array_equal(t.index()[1][dim], perm[dim])
        for i in range(t.dims[dim]):
            obs = t.find_observed_ui(dim, i)
            assert len(obs) == t.num_observed(dim, i) == len(t.observed_by_id[dim][i])
            # append order of the reference: increasing entry number (Tensor/Tensor.py:160-167)
            want = [e for e in range(len(t.train_vals)) if t.train_entries[e, dim] == i]
            assert [tuple(c) for c, _ in obs] == [tuple(t.train_entries[e]) for e in want]
            assert [v for _, v in obs] == [t.train_vals[e] for e in want]


def test_binary_and_count_generators():
    np.random.seed(1)
    b = pkg.binary_tensor()
    b.synthesize_data([8, 7, 6], None, None, D=3, train=0.8, sparsity=0.4)
    assert set(np.unique(b.train_vals)) <= {-1.0, 1.0}
    c = pkg.count_tensor()
    c.synthesize_data([8, 7, 6], None, None, D=3, train=0.8, sparsity=0.4)
    assert np.all(c.train_vals >= 0) and np.all(c.train_vals == np.rint(c.train_vals))
    assert c.max_count >= c.min_count >= 0


def test_read_from_file(tmp_path):
    p = tmp_path / "t.txt"
    rng = np.random.default_rng(2)
    lines = ["4,3,5"] + ["%d,%d,%d,%d" % (rng.integers(4), rng.integers(3), rng.integers(5), rng.choice([-1, 1]))
                         for _ in range(40)]
    p.write_text("\n".join(lines) + "\n")
    t = pkg.binary_tensor()
    np.random.seed(3)
    t.read_from_file(str(p), 0.8, 0, 0.2)
    assert t.dims == [4, 3, 5] and len(t.train_vals) == 32 and len(t.test_vals) == 8 and len(t.valid_vals) == 0
    assert sum(t.num_observed(0, i) for i in range(4)) == 32
