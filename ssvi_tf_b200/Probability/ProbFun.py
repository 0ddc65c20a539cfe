"""Link functions of the likelihoods (host-side numpy API, mirror of Probability/ProbFun.py).
The device twins live in csrc/likelihood.cuh."""
import numpy as np

epsilon = 1e-16


def sigmoid(x, epsilon=0):
    """Logistic function, NaN mapped to 0, clipped to [epsilon, 1 - epsilon]  (ProbFun.py:22-28)."""
    with np.errstate(all="ignore"):
        out = 1.0 / (1.0 + np.exp(-np.asarray(x, dtype=float)))
    out = np.nan_to_num(out)
    return np.clip(out, epsilon, 1 - epsilon)


def poisson_link(f, epsilon=1e-8):
    """Softplus rate log(1 + e^f), floored at epsilon  (ProbFun.py:52-55; log(1+exp), not log1p)."""
    with np.errstate(all="ignore"):
        out = np.log(1.0 + np.exp(np.asarray(f, dtype=float)))
    return np.maximum(np.nan_to_num(out), epsilon)
