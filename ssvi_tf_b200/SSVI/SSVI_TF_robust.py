"""SSVI_TF_robust: constructor-compatible mirror of SSVI/SSVI_TF_robust.py; the model-specific expected-derivative
estimator lives in the CUDA kernels (csrc/accumulate_*.cuh, model id "robust")."""
from .SSVI_TF_interface import SSVI_TF


class SSVI_TF_robust(SSVI_TF):
    _model = "robust"

    def __init__(self, tensor, rank, mean_update="S", cov_update="N", noise_update="N",
                 diag=False, mean0=None, cov0=None, sigma0=1,
                 unstable_cov=False, k1=64, k2=64, batch_size=128,
                 eta=1, cov_eta=1, sigma_eta=1, quadrature=False, randstart=True, **device_kwargs):
        # same positional order as SSVI_TF_robust.py:8-11 (quadrature sits before randstart)
        super(SSVI_TF_robust, self).__init__(tensor, rank, mean_update, cov_update, noise_update, diag,
                                             mean0, cov0, sigma0, unstable_cov, k1, k2, batch_size,
                                             eta, cov_eta, sigma_eta, randstart, **device_kwargs)
        self.numerical_epsilon = 1e-16                                      # :21
        self.quadrature = quadrature
        if quadrature:                                                      # :22-25
            self.gauss_degree = 30
            self.engine.set_quadrature(self.gauss_degree)
