"""Binary (+1/-1) tensor (mirror of Tensor/binary_tensor.py): Bernoulli(sigmoid(f)) observations."""
import numpy as np

from ..Probability import ProbFun as probs
from .Tensor import Tensor


class binary_tensor(Tensor):
    def __init__(self, binary_cutoff=0.):
        super(binary_tensor, self).__init__(datatype="binary")
        self.binary_cutoff = binary_cutoff

    def generate_hidden_matrices(self, mean=None, cov=None):
        """Per-mode factor means spread over linspace(-1, 1, N), covariance 0.1 I (binary_tensor.py:10-21)."""
        centers = np.linspace(-1, 1, len(self.dims))
        return [self.create_random_matrix(n, self.D, np.full((self.D,), centers[i]), 0.1 * np.eye(self.D))
                for i, n in enumerate(self.dims)]

    def data_link_fun(self, m):
        p = probs.sigmoid(m)
        return np.where(np.random.binomial(1, p) == 1, 1, -1)

    def actual_value(self, m):
        return m
