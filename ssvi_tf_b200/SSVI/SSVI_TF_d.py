"""SSVI_TF_d: constructor-compatible mirror of SSVI/SSVI_TF_d.py; the model-specific expected-derivative
estimator lives in the CUDA kernels (csrc/accumulate_*.cuh, model id "d")."""
from .SSVI_TF_interface import SSVI_TF


class SSVI_TF_d(SSVI_TF):
    _model = "d"
