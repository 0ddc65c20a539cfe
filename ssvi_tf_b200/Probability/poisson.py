"""PoissonDistribution with softplus rate: host-side API mirror of Probability/poisson.py."""
import math

import numpy as np
from scipy.stats import poisson as _sp_poisson

from .ProbFun import poisson_link, sigmoid


class PoissonDistribution(object):
    def __init__(self):
        self.epsilon = 1e-8

    def __str__(self):
        return "PoissonDistribution"

    def sample(self, m, s=None, k=1):
        return np.random.poisson(poisson_link(m), size=k)

    def pdf(self, y, m, S=None):
        """A^y / (1 + e^m) / y!, floored at 1e-8  (poisson.py:19-27)."""
        A = poisson_link(m)
        with np.errstate(all="ignore"):
            out = np.power(A, y) / (1.0 + np.exp(m)) / math.factorial(int(y))
        return np.maximum(out, self.epsilon)

    def fst_derivative_pdf(self, y, m, S=None):
        A = poisson_link(m)
        return self.pdf(y, m) * sigmoid(m) * (y / A - 1.0)

    def snd_derivative_pdf(self, y, m, s=None):
        """poisson.py:41-57; the leading pdf is evaluated at poisson_link(m) (quirk Q7)."""
        A = poisson_link(m)
        sg = sigmoid(m)
        t = y / A - 1.0
        inner = -(y / np.square(A)) * np.square(sg) + t * sg * (1.0 - sg)
        return self.fst_derivative_pdf(y, m) * sg * t + self.pdf(y, A) * inner

    def log_pdf(self, y, m, s=None):
        return np.nan_to_num(_sp_poisson.logpmf(y, poisson_link(m)))

    def fst_derivative_log_pdf(self, y, f, s=None):
        with np.errstate(all="ignore"):
            return np.nan_to_num(sigmoid(f) * (y / poisson_link(f) - 1.0))

    def snd_derivative_log_pdf(self, y, f, s=None):
        sg, A = sigmoid(f), poisson_link(f)
        with np.errstate(all="ignore"):
            return np.nan_to_num(sg * (1.0 - sg) * (y / A - 1.0) - y * np.square(sg) / np.square(A))
