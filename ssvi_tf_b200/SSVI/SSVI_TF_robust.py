"""SSVI_TF_robust: constructor-compatible mirror of SSVI/SSVI_TF_robust.py; the model-specific expected-derivative
estimator lives in the CUDA kernels (csrc/accumulate_*.cuh, model id "robust")."""
from .SSVI_TF_interface import SSVI_TF


class SSVI_TF_robust(SSVI_TF):
    _model = "robust"
