"""Real-valued tensor (mirror of Tensor/real_tensor.py): identity link, normal likelihood."""
from .Tensor import Tensor


def identity(m):
    return m


class RV_tensor(Tensor):
    def __init__(self):
        super(RV_tensor, self).__init__(datatype="real")
        self.link_fun = identity

    def generate_hidden_matrices(self, means, covariances):
        return [self.create_random_matrix(n, self.D, means[i], covariances[i]) for i, n in enumerate(self.dims)]

    def data_link_fun(self, m):
        return m

    def actual_value(self, m):
        return m
