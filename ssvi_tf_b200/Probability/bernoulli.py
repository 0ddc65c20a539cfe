"""BernoulliDistribution (labels y in {-1,+1}): host-side API mirror of Probability/bernoulli.py."""
import numpy as np

from .ProbFun import sigmoid


class BernoulliDistribution(object):
    def __init__(self):
        self.epsilon = 1e-16

    def __str__(self):
        return "BernoulliDistribution"

    def sample(self, m, s=None, k=1):
        return np.random.binomial(1, sigmoid(m), size=k)

    def pdf(self, y, m, s=None):
        """sigma(y m) clipped to [1e-16, 1-1e-16]  (bernoulli.py:16-21)."""
        return sigmoid(np.multiply(y, m), self.epsilon)

    def fst_derivative_pdf(self, y, m, s=None):
        p = self.pdf(y, m)
        return y * p * (1.0 - p)

    def snd_derivative_pdf(self, y, m, s=None):
        p = self.pdf(y, m)
        return p * (1.0 - p) * (1.0 - 2.0 * p)

    def log_pdf(self, y, m, s=None):
        return np.log(self.pdf(y, m, s))

    def fst_derivative_log_pdf(self, y, m, s=None):
        return y * sigmoid(-np.multiply(y, m))

    def snd_derivative_log_pdf(self, y, m, s=None):
        ym = np.multiply(y, m)
        return -sigmoid(ym) * sigmoid(-ym)
