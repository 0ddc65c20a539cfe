"""CPU: the series behind the device's straight-line Poisson terms (csrc/likelihood.cuh: log_1to2, exp_nonpos_c and the
single-exponential form of poisson_pdf_derivs_fast), restated in numpy with the same coefficients and operation order
(no FMA), against libm and against the reference's formulas (Probability/poisson.py:19-57, ProbFun.py:22-28, 52-55)."""
import math

import numpy as np

LOGC = [2.0 / (2 * n + 3) for n in range(10)]
EXPC = [1.0 / math.factorial(n) for n in range(14)]


def log_1to2(x):
    hi = x > math.sqrt(2.0)
    xr = np.where(hi, 0.5 * x, x)
    s = (xr - 1.0) / (xr + 1.0)
    z = s * s
    a01, a23, a45 = z * LOGC[1] + LOGC[0], z * LOGC[3] + LOGC[2], z * LOGC[5] + LOGC[4]
    a67, a89 = z * LOGC[7] + LOGC[6], z * LOGC[9] + LOGC[8]
    z2 = z * z
    b0, b1 = z2 * a23 + a01, z2 * a67 + a45
    z4 = z2 * z2
    pz = (z4 * z4) * a89 + (z4 * b1 + b0)
    l = (s * z) * pz + (s + s)
    return np.where(hi, l + math.log(2.0), l)


def exp_nonpos(x):
    t = x * 1.4426950408889634 + 6755399441055744.0
    nd = t - 6755399441055744.0
    r = nd * -6.93147180369123816490e-01 + x
    r = nd * -1.90821492927058770002e-10 + r
    p = np.zeros_like(r)
    for c in reversed(EXPC):
        p = p * r + c
    return np.ldexp(p, nd.astype(np.int64))


def test_log_series_on_1_to_2():
    x = np.concatenate([1.0 + np.logspace(-17, 0, 100001), np.linspace(1.0, 2.0, 100001)])
    x = x[(x > 1.0) & (x <= 2.0)]
    ref = np.array([math.log(v) for v in x])
    assert np.max(np.abs(log_1to2(x) - ref) / ref) < 1e-15
    assert log_1to2(np.array([1.0]))[0] == 0.0


def test_exp_series_on_nonpositive_arguments():
    x = -np.concatenate([np.logspace(-12, math.log10(708.0), 200001), [0.0]])
    ref = np.exp(x)
    assert np.max(np.abs(exp_nonpos(x) - ref) / ref) < 2e-15


def test_single_exponential_form_equals_reference_poisson_terms():
    """sigmoid(f), log(1 + e^f) and 1 / (1 + e^f) from e = exp(-|f|) against the reference's own expressions."""
    f = np.concatenate([np.linspace(-40.0, 40.0, 80001), [-700.0 + 1e-9, 699.0]])
    e = exp_nonpos(-np.abs(f))
    x = 1.0 + e
    L = log_1to2(x)
    pos = f >= 0
    A = np.maximum(np.where(pos, f + L, L), 1e-8)
    sg = np.where(pos, 1.0 / x, e / x)
    rE1 = np.where(pos, e / x, 1.0 / x)
    with np.errstate(over="ignore"):
        E = np.exp(f)
        A_ref = np.maximum(np.nan_to_num(np.log(1.0 + E)), 1e-8)          # ProbFun.py:52-55
        sg_ref = 1.0 / (1.0 + np.exp(-f))                                   # ProbFun.py:22-28
        rE1_ref = 1.0 / (1.0 + E)
    ok = np.abs(f) < 30.0          # beyond, fl(1 + e^f) itself loses digits the e-form keeps (f >= 0) or reproduces (f < 0)
    np.testing.assert_allclose(A[ok], A_ref[ok], rtol=5e-14)
    np.testing.assert_allclose(A[f < 0], A_ref[f < 0], rtol=5e-14)         # the reference's rounding, digit for digit
    np.testing.assert_allclose(sg, sg_ref, rtol=1e-14)
    np.testing.assert_allclose(rE1[f < 600], rE1_ref[f < 600], rtol=1e-14)


# ---- the fp32 option's operand split (csrc/tc_ops.cuh: tc_store_split8) ---------------------------------

def _split_fp16(x32):
    hi = x32.astype(np.float16)
    lo = (x32 - hi.astype(np.float32)).astype(np.float16)
    return hi, lo


def test_fp16_split_carries_22_bits():
    """hi = rn16(x), lo = rn16(x - hi): hi + lo reproduces a float to 2^-22 relative wherever lo stays a normal half
    (|x| >= 2^-3 here: lo's exponent is 11 below x's and fp16 normals end at 2^-14)."""
    rng = np.random.default_rng(5)
    x = (rng.standard_normal(200000) * np.exp2(rng.integers(-3, 6, 200000))).astype(np.float32)
    x = x[np.abs(x) >= 2.0 ** -3]
    hi, lo = _split_fp16(x)
    back = hi.astype(np.float64) + lo.astype(np.float64)
    assert np.max(np.abs(back - x) / np.abs(x)) <= 2.0 ** -22


def test_split_contraction_meets_the_fp32_bar():
    """V V^T the way the tensor-core kernel forms it (accumulate_tc.cuh) -- the entry's D x k1 Hadamard product in
    fp32, split, three half x half products per K-step accumulated in fp32 (hi hi^T + hi lo^T + lo hi^T; lo lo^T
    dropped), then G += w_b * (V V^T) with the entry's k1-averaged weight (SSVI_TF_d.py:47-56) -- against the fp64
    contraction: far inside the 1e-4 the contract states for the fp32 option, at the ranks of the BASELINE configs."""
    rng = np.random.default_rng(6)
    for D, k1 in ((32, 64), (50, 64), (64, 64), (7, 64)):
        worst = 0.0
        for _ in range(20):
            V = (0.3 + rng.standard_normal((D, k1))).astype(np.float32)             # Hadamard products of sampled rows
            w = np.float32(-abs(rng.standard_normal()) - 0.1)
            hi, lo = _split_fp16(V)
            h32, l32 = hi.astype(np.float32), lo.astype(np.float32)
            P = np.zeros((D, D), dtype=np.float32)
            for s in range(0, k1, 16):                                             # one MMA K-step = 16 samples
                sl = slice(s, s + 16)
                P += h32[:, sl] @ h32[:, sl].T
                P += h32[:, sl] @ l32[:, sl].T
                P += l32[:, sl] @ h32[:, sl].T
            G = w * P
            ref = float(w) * (V.astype(np.float64) @ V.astype(np.float64).T)
            worst = max(worst, np.max(np.abs(G - ref)) / np.max(np.abs(ref)))
        assert worst < 2e-6, (D, k1, worst)


def test_split_overflow_is_what_the_range_flag_reports():
    """A scaled operand beyond fp16's range turns into inf in `hi` (the kernel's accumulators then hold non-finite
    values and the call returns SSVI_FLAG_TC_RANGE); a power-of-two scale (tc_scale_log2) brings it back exactly."""
    x = np.array([7.0e4, -1.0e5, 3.0], dtype=np.float32)
    hi, _ = _split_fp16(x)
    assert np.isinf(hi[:2]).all() and np.isfinite(hi[2])
    a = np.float32(2.0 ** -4)
    hi, lo = _split_fp16(x * a)
    back = (hi.astype(np.float64) + lo.astype(np.float64)) / float(a)
    assert np.all(np.abs(back - x) <= np.abs(x) * 2.0 ** -22)
