"""SSVI_TF_simple: constructor-compatible mirror of SSVI/SSVI_TF_simple.py; the model-specific expected-derivative
estimator lives in the CUDA kernels (csrc/accumulate_*.cuh, model id "simple")."""
from .SSVI_TF_interface import SSVI_TF


class SSVI_TF_simple(SSVI_TF):
    _model = "simple"
