"""NormalDistribution: host-side API mirror of Probability/normal.py (device twin: csrc/likelihood.cuh)."""
import math

import numpy as np

_SQRT_2PI = math.sqrt(2.0 * math.pi)


class NormalDistribution(object):
    def __init__(self):
        self.epsilon = 0.00001

    def __str__(self):
        return "NormalDistribution"

    def sample(self, m, s, k):
        return np.random.normal(m, s, size=k)

    def pdf(self, y, m, s):
        """N(y; m, s) floored at epsilon  (normal.py:16-18)."""
        r = (np.asarray(y, float) - m) / s
        return np.maximum(np.exp(-0.5 * r * r) / (_SQRT_2PI * s), self.epsilon)

    def fst_derivative_pdf(self, y, m, s):
        return self.pdf(y, m, s) * (y - m) / np.square(s)

    def snd_derivative_pdf(self, y, m, s):
        """pdf * (((y-m)/s)^2 - 1/s) as the reference returns it (normal.py:27, quirk Q5)."""
        return self.pdf(y, m, s) * (np.square((y - m) / s) - 1.0 / s)

    def log_pdf(self, y, m, s):
        return np.log(self.pdf(y, m, s))

    def fst_derivative_log_pdf(self, y, m, s):
        return (y - m) / np.square(s)

    def snd_derivative_log_pdf(self, y, m, s):
        return -np.ones_like(m, dtype=float) / np.square(s)
