"""Count tensor (mirror of Tensor/count_tensor.py): Poisson(softplus(f)) observations."""
import numpy as np

from ..Probability import ProbFun as probs
from .Tensor import Tensor


class count_tensor(Tensor):
    def __init__(self):
        super(count_tensor, self).__init__(datatype="count")
        self.link_fun = probs.poisson_link
        self.min_count = 100
        self.max_count = 0

    def generate_hidden_matrices(self, mean=None, cov=None):
        """Per-mode factor means spread over linspace(0, 1.5, N), covariance 0.1 I (count_tensor.py:14-25)."""
        centers = np.linspace(0, 1.5, len(self.dims))
        return [self.create_random_matrix(n, self.D, np.full((self.D,), centers[i]), 0.1 * np.eye(self.D))
                for i, n in enumerate(self.dims)]

    def data_link_fun(self, m):
        x = np.random.poisson(self.link_fun(m))
        self.max_count = max(self.max_count, int(np.max(x)))
        self.min_count = min(self.min_count, int(np.min(x)))
        return x

    def actual_value(self, m):
        return self.link_fun(m)
