// Likelihood terms as device functions: a restatement of the reference's
// Probability/{normal,bernoulli,poisson}.py + ProbFun.py formulas, quirks included
// (Q5 normal snd_derivative_pdf, Q7 poisson double link, clipping floors).
#pragma once
#include "ssvi_common.cuh"

// ProbFun.py:22-28  sigmoid(x, epsilon): 1/(1+exp(-x)), nan_to_num, clip to [eps, 1-eps]
__device__ __forceinline__ double sigmoid_ref(double x, double eps) {
  double r = 1.0 / (1.0 + exp(-x));
  r = nan_to_num(r);
  return fmin(fmax(r, eps), 1.0 - eps);
}

// ProbFun.py:52-55  poisson_link(f): log(1+exp(f)) (not log1p), nan_to_num, floor 1e-8
__device__ __forceinline__ double poisson_link_ref(double f) {
  double r = log(1.0 + exp(f));
  r = nan_to_num(r);
  return fmax(r, 1e-8);
}

// y! for integral y in [0,170] (np.math.factorial, Probability/poisson.py:23); beyond 170 the
// reference's float division overflows to 0 as well (x / inf).
__device__ __forceinline__ double factorial_ref(double y) {
  int n = (int)y;
  if (n > 170) return INFINITY;
  double r = 1.0;
  for (int i = 2; i <= n; ++i) r *= (double)i;   // exact in fp64 up to 22!, correctly rounded products after
  return r;
}

// A^y for the integral counts y of count data (np.power(A, y), poisson.py:23): binary powering, <= 8 squarings for
// y <= 170; a few ulp from the libm pow, and it overflows / underflows where pow does.
__device__ __forceinline__ double ipow_ref(double a, double y) {
  int n = (int)y;
  double r = 1.0;
  while (n > 0) {
    if (n & 1) r *= a;
    a *= a;
    n >>= 1;
  }
  return r;
}

// One exponential serves sigmoid(f) and poisson_link(f) (ProbFun.py:22-28, 52-55): with E = exp(f),
//   poisson_link = max(nan_to_num(log(1 + E)), 1e-8),  sigmoid = 1/(1 + exp(-f)) = E/(1 + E)  (1 when E overflows).
struct PoissonTerms {
  double E1;    // 1 + exp(f)
  double A;     // poisson_link(f)
  double sg;    // sigmoid(f)
};
__device__ __forceinline__ PoissonTerms poisson_terms(double f) {
  PoissonTerms t;
  const double E = exp(f);
  t.E1 = 1.0 + E;
  t.A = fmax(nan_to_num(log(t.E1)), 1e-8);
  t.sg = (E > 1e300) ? 1.0 : E / t.E1;
  if (isnan(f)) t.sg = 0.0;                           // nan_to_num(sigmoid(nan)) = 0
  return t;
}

// d / simple models: first and second derivative of log p(y|f) w.r.t. f
//   normal.py:32-36, bernoulli.py:40-51, poisson.py:70-87
__device__ __forceinline__ void log_pdf_derivs(int lik, double y, double f, double s, double& fst,
                                               double& snd) {
  if (lik == SSVI_LIK_NORMAL) {
    double is2 = 1.0 / (s * s);
    fst = is2 * (y - f);
    snd = -is2;
  } else if (lik == SSVI_LIK_BERNOULLI) {
    // sigmoid(yf) and sigmoid(-yf) (bernoulli.py:40-51) from ONE exponential: with e = exp(-|yf|) in (0, 1],
    // sigmoid(|yf|) = 1/(1+e) and sigmoid(-|yf|) = e/(1+e); equal to the reference's two evaluations to 2 ulp, and
    // saturating to {1, 0} exactly where 1/(1+exp(-x)) does.
    const double yf = y * f;
    const double e = exp(-fabs(yf));
    const double hi = 1.0 / (1.0 + e), lo = e * hi;
    const bool pos = yf >= 0.0;
    double sp = pos ? hi : lo, sm = pos ? lo : hi;
    if (isnan(yf)) sp = sm = 0.0;                      // nan_to_num(sigmoid(nan)) = 0 (ProbFun.py:24)
    fst = y * sm;
    snd = -sp * sm;
  } else {
    const PoissonTerms pt = poisson_terms(f);
    const double sg = pt.sg, A = pt.A;
    double t = y / A - 1.0;
    fst = nan_to_num(sg * t);
    snd = nan_to_num(sg * (1.0 - sg) * t - y * (sg * sg) / (A * A));
  }
}

// robust model: pdf and its first / second derivative w.r.t. f
//   normal.py:16-27, bernoulli.py:16-36, poisson.py:19-57
// (yfact: 1 / y! for the Poisson likelihood, unused otherwise)
__device__ __forceinline__ void pdf_derivs(int lik, double y, double f, double s, double yfact, double& p,
                                           double& p1, double& p2) {
  if (lik == SSVI_LIK_NORMAL) {
    double r = (y - f) / s;
    p = fmax(exp(-0.5 * r * r) / (2.5066282746310002 * s), 1e-5);
    p1 = p * ((y - f) / (s * s));
    p2 = p * (r * r - 1.0 / s);                       // Q5
  } else if (lik == SSVI_LIK_BERNOULLI) {
    double sg = sigmoid_ref(y * f, 1e-16);
    p = sg;
    p1 = (y * sg) * (1.0 - sg);
    p2 = sg * (1.0 - sg) * (1.0 - 2.0 * sg);
  } else {
    // Divisions are the expensive part here (about 25 instructions each in fp64): one reciprocal each of 1+e^f, A and
    // 1+e^A serves every quotient of poisson.py:19-57; yfact arrives as 1/y! (a few ulp from the reference's order).
    const double E = exp(f);
    const double E1 = 1.0 + E, rE1 = 1.0 / E1;
    const double A = fmax(nan_to_num(log(E1)), 1e-8), rA = 1.0 / A;
    double sg = (E > 1e300) ? 1.0 : E * rE1;
    if (isnan(f)) sg = 0.0;
    const double yA = y * rA;
    double t = yA - 1.0;
    p = fmax(ipow_ref(A, y) * rE1 * yfact, 1e-8);     // pdf(y, f)
    p1 = p * (sg * t);
    const double EA1 = 1.0 + exp(A);                  // Q7: pdf(y, poisson_link(m)) -- the link applied twice
    const double A2 = fmax(nan_to_num(log(EA1)), 1e-8);
    const double pl = fmax(ipow_ref(A2, y) / EA1 * yfact, 1e-8);
    double temp = -(yA * rA) * (sg * sg) + t * (sg * (1.0 - sg));
    p2 = p1 * (sg * t) + pl * temp;
  }
}
