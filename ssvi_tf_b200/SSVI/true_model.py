"""Host-side report helpers of the reference that sit OFF the hot path: the "true parameters" evaluation the reference's
drivers call right after factorize() (synthetic_test.py:152, test_scripts/test.py:148, learning_curve.py:89) and the
count-prediction helpers (SSVI/SSVI_TF_interface.py:1063-1165, :1213-1320).  Plain NumPy over the host copy of the
posterior; mixed into SSVI_TF.  Vectorised over entries where the reference loops."""
import math

import numpy as np


class TrueModelMixin(object):
    # ---- predict bridge (:1063-1069) -------------------------------------------------------------------------------
    def predict_y_given_m(self, m):
        if self.likelihood_type == "normal":
            return m
        if self.likelihood_type == "bernoulli":
            return 1 if m >= 0 else -1
        raise Exception("Unidentified likelihood type")

    # ---- count prediction by Gauss-Hermite quadrature (:1077-1141) --------------------------------------------------
    def _hermite_rule(self):
        if getattr(self, "hermite_points", None) is None:
            self.hermite_points, self.hermite_weights = np.polynomial.hermite.hermgauss(getattr(self, "gauss_degree", 30))
        return np.asarray(self.hermite_points), np.asarray(self.hermite_weights)

    def compute_fyi(self, k, yi, m, s):
        A = self.link_fun(np.sqrt(2 * s) * yi + m)
        return np.power(A, k) / math.factorial(k) * np.exp(-A) * np.sqrt(2 * s)

    def compute_gauss_hermite(self, k, m, S):
        xs, ws = self._hermite_rule()
        return float(np.sum(ws * self.compute_fyi(k, xs, m, S)) / np.sqrt(2 * np.pi * S))

    def compute_posterior_param(self, entry):
        """Mean and variance of f = sum_D prod_modes x for independent Gaussian rows, as the reference composes them
        (:1093-1124): element-wise products of the means / covariances over the modes plus, per mode, the two
        quadratic forms of the leave-one-out products."""
        rows = [self.posterior.get_vector_distribution(dim, i) for dim, i in enumerate(entry)]
        ms = [np.asarray(m, dtype=np.float64) for m, _ in rows]
        Cs = [np.diag(np.asarray(C, dtype=np.float64)) if self.diag else np.asarray(C, dtype=np.float64) for _, C in rows]
        m_all, S_all = np.prod(ms, axis=0), np.prod(Cs, axis=0)
        mean, var = float(m_all.sum()), float(S_all.sum())
        for d in range(len(entry)):
            m_loo = np.prod([m for j, m in enumerate(ms) if j != d], axis=0) if len(ms) > 1 else np.ones(self.D)
            C_loo = np.prod([C for j, C in enumerate(Cs) if j != d], axis=0) if len(Cs) > 1 else np.ones((self.D, self.D))
            var += float(ms[d] @ (C_loo @ ms[d])) + float(m_loo @ (Cs[d] @ m_loo))
        return mean, var

    def compute_expected_count_quadrature(self, m, S):
        ks = np.arange(self.min_count, self.max_count + 1)
        probs = np.array([self.compute_gauss_hermite(int(k), m, S) for k in ks])
        return float(np.sum(probs * ks) / np.sum(probs))

    def compute_expected_count_sampling(self, entry):
        """(:1144-1165) mean of likelihood draws at f ~ prod of posterior row samples (+ the noise layer)."""
        n = self.predict_num_samples
        ms = np.ones((n, self.D))
        for dim, i in enumerate(entry):
            m, S = self.posterior.get_vector_distribution(dim, i)
            S = np.diag(np.asarray(S)) if self.diag else np.asarray(S)
            ms = ms * np.random.multivariate_normal(np.asarray(m), S, size=n)
        fs = ms.sum(axis=1)
        if self.noise_added:
            fs = np.random.normal(fs, np.random.rayleigh(self.w_sigma ** 2, size=n), size=n)
        return float(np.mean(self.likelihood.sample(fs, n)))

    # ---- evaluation against the generating factors of a synthetic tensor (:1213-1320) -------------------------------
    def _true_f(self, entries, matrices):
        entries = np.asarray(entries, dtype=np.int64).reshape(-1, self.order)
        prod = np.ones((len(entries), self.tensor.D))
        for dim in range(self.order):
            prod = prod * np.asarray(matrices[dim])[entries[:, dim], :]
        return prod.sum(axis=1)

    def true_model_predict_via_sampling(self, f):
        """(:1277-1289) rint(mean link(f + noise)) over 50 draws; noise scale |f| * noise_ratio or tensor.noise."""
        w = np.abs(f) * self.tensor.noise_ratio if self.tensor.noise_ratio else self.tensor.noise
        return np.rint(np.mean(self.link_fun(f + np.random.normal(0, w, size=(50,)))))

    def true_model_predict(self, entry, matrices):
        f = float(self._true_f(np.asarray(entry)[None, :], matrices)[0])
        if self.datatype == "real":
            return f
        p = self.true_model_predict_via_sampling(f)
        return (1 if p > 1 / 2 else -1) if self.datatype == "binary" else p

    def evaluate_true_hidden_vectors(self, entries, vals, matrices):
        """(rsme, error) of the generating model's own predictions (:1240-1258)."""
        vals = np.asarray(vals, dtype=np.float64)
        pred = np.array([self.true_model_predict(e, matrices) for e in np.asarray(entries)], dtype=np.float64)
        diff = pred - vals
        rsme = float(np.sqrt(np.sum(diff ** 2) / len(vals)))
        if self.likelihood_type == "normal":
            error = float(np.sum(np.abs(diff) / np.abs(vals)) / len(vals))
        elif self.likelihood_type == "bernoulli":
            error = float(np.sum(pred != vals) / len(vals))
        else:
            error = float(np.sum(np.abs(diff)) / len(vals))
        return rsme, error

    def evaluate_true_model_nll(self, entries, vals, matrices):
        """(:1292-1320).  As written in the reference the inner loop walks `enumerate(np.arange(order))`, i.e. it reads
        row `dim` of factor matrix `dim` for EVERY entry, so f does not depend on the entry; kept as is."""
        f = float(np.sum(np.prod([np.asarray(self.tensor.matrices[d])[d, :] for d in range(self.order)], axis=0)))
        s = self.likelihood_param
        nll = 0.0
        for y in np.asarray(vals):
            if self.noise_added:
                w = np.abs(f) * self.tensor.noise_ratio if self.tensor.noise_ratio else self.tensor.noise
                fs = np.random.normal(f, w, size=(self.k1,))
            else:
                fs = f
            nll -= np.log(np.mean(self.likelihood.pdf(y, fs, s)))
        return float(nll)

    def evaluate_true_params(self):
        t = self.tensor
        train_rsme, train_error = self.evaluate_true_hidden_vectors(t.train_entries, t.train_vals, t.matrices)
        test_rsme, test_error = self.evaluate_true_hidden_vectors(t.test_entries, t.test_vals, t.matrices)
        test_nll = self.evaluate_true_model_nll(t.test_entries, t.test_vals, t.matrices)
        train_nll = self.evaluate_true_model_nll(t.train_entries, t.train_vals, t.matrices)
        print("Evaluation for true params: ")
        print(" test_rsme | train_rsme | rel-te-err | rel-tr-err |  test_nll  |  train_nll |")
        print("{:^12} {:^12} {:^12} {:^12} {:^12} {:^12}".format(
            np.around(test_rsme, 10), np.around(train_rsme, 10), np.around(test_error, 10),
            np.around(train_error, 10), np.around(test_nll, 2), np.around(train_nll, 2)))
        return test_rsme, train_rsme, test_error, train_error, test_nll, train_nll
