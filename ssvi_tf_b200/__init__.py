"""ssvi_tf_b200: B200-native (sm_100a) implementation of SSVI-TF's per-entry stochastic variational
update for Bayesian CP tensor factorization, behind the reference's Python surface.

    from ssvi_tf_b200 import SSVI_TF_d, SSVI_TF_simple, SSVI_TF_robust
    from ssvi_tf_b200 import RV_tensor, binary_tensor, count_tensor

Importing the package needs no GPU; constructing a model does (no CPU fallback).
"""
from .Probability.bernoulli import BernoulliDistribution
from .Probability.normal import NormalDistribution
from .Probability.poisson import PoissonDistribution
from .Tensor.binary_tensor import binary_tensor
from .Tensor.count_tensor import count_tensor
from .Tensor.real_tensor import RV_tensor
from .Tensor.Tensor import Tensor
from .SSVI.SSVI_TF_interface import SSVI_TF
from .SSVI.SSVI_TF_d import SSVI_TF_d
from .SSVI.SSVI_TF_simple import SSVI_TF_simple
from .SSVI.SSVI_TF_robust import SSVI_TF_robust
from .engine import Engine, NoisePool

__all__ = ["SSVI_TF", "SSVI_TF_d", "SSVI_TF_simple", "SSVI_TF_robust", "Tensor", "RV_tensor", "binary_tensor",
           "count_tensor", "NormalDistribution", "BernoulliDistribution", "PoissonDistribution", "Engine",
           "NoisePool"]
