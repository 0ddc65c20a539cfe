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

// exp(x) for x <= 0 as straight-line code with a SHORT dependency chain.  The FP64 pipe of a B200 SM is shared with the
// DMMAs of the other warps, so every dependent FP64 instruction of a likelihood evaluation queues behind 16-cycle
// tensor operations: what costs time is the chain depth, not the instruction count.  libm's exp is a 12-deep Horner
// chain plus a range-check branch (which also stops ptxas interleaving two evaluations); here: round-to-int by the
// magic-number add, Cody-Waite reduction, degree-13 Taylor polynomial in Estrin form (4 levels; truncation error
// 0.347^14 / 14! = 4e-18), exponent spliced in with integer arithmetic.  Arguments below -708 return exp(-708) ~ 3e-308
// (the exact value is smaller still; no caller can tell the difference).  About 2 ulp.
__device__ __forceinline__ double exp_nonpos(double x) {
  const double xc = fmax(x, -708.0);
  const double t = fma(xc, 1.4426950408889634, 6755399441055744.0);
  const int n = __double2loint(t);
  const double nd = t - 6755399441055744.0;
  double r = fma(nd, -6.93147180369123816490e-01, xc);
  r = fma(nd, -1.90821492927058770002e-10, r);
  const double r2 = r * r;
  const double p01 = 1.0 + r;
  const double p23 = fma(r, 1.0 / 6.0, 0.5);
  const double p45 = fma(r, 1.0 / 120.0, 1.0 / 24.0);
  const double p67 = fma(r, 1.0 / 5040.0, 1.0 / 720.0);
  const double p89 = fma(r, 1.0 / 362880.0, 1.0 / 40320.0);
  const double pab = fma(r, 1.0 / 39916800.0, 1.0 / 3628800.0);
  const double pcd = fma(r, 1.0 / 6227020800.0, 1.0 / 479001600.0);
  const double r4 = r2 * r2;
  const double q0 = fma(r2, p23, p01);
  const double q1 = fma(r2, p67, p45);
  const double q2 = fma(r2, pab, p89);
  const double r8 = r4 * r4;
  const double q3 = fma(r4, pcd, q2);
  const double s0 = fma(r4, q1, q0);
  const double p = fma(r8, q3, s0);                    // in [0.70, 1.42]: adding n to the exponent field cannot carry
  return __hiloint2double(__double2hiint(p) + (n << 20), __double2loint(p));
}

static __constant__ double SSVI_EXPC[14] = {1.0, 1.0, 0.5, 1.0 / 6.0, 1.0 / 24.0, 1.0 / 120.0, 1.0 / 720.0, 1.0 / 5040.0,
                                     1.0 / 40320.0, 1.0 / 362880.0, 1.0 / 3628800.0, 1.0 / 39916800.0,
                                     1.0 / 479001600.0, 1.0 / 6227020800.0};
// [0..9]: 2/(2n+3), n = 0..9 (atanh series of log);  [10] = ln 2,  [11] = sqrt 2,  [12] = log2(e),  [13] = magic 1.5*2^52
static __constant__ double SSVI_LOGC[16] = {2.0 / 3.0,  2.0 / 5.0,  2.0 / 7.0,  2.0 / 9.0,  2.0 / 11.0, 2.0 / 13.0,
                                     2.0 / 15.0, 2.0 / 17.0, 2.0 / 19.0, 2.0 / 21.0, 6.93147180559945286227e-01,
                                     1.41421356237309514547e+00, 1.4426950408889634, 6755399441055744.0,
                                     -6.93147180369123816490e-01, -1.90821492927058770002e-10};

// exp_nonpos with its coefficients read from CONSTANT MEMORY, for the k2 loops of the simple / robust models: as literals
// ptxas materialises each 64-bit coefficient with two UMOVs before every use, which made those loops issue bound
// (profiles/r01z_ncu_memo_weights.txt).  The sweep kernels' per-entry epilogue keeps the literal version above: there
// the issue slots are idle and the constant loads only add latency (C2: 4.33e8 with literals, 4.27e8 with loads).
// Precondition: -708 <= x <= 0 (the callers bound |f| by 700).
__device__ __forceinline__ double exp_nonpos_c(double x) {
  const double t = fma(x, SSVI_LOGC[12], SSVI_LOGC[13]);
  const int n = __double2loint(t);
  const double nd = t - SSVI_LOGC[13];
  double r = fma(nd, SSVI_LOGC[14], x);
  r = fma(nd, SSVI_LOGC[15], r);
  const double r2 = r * r;
  const double p01 = SSVI_EXPC[0] + r;
  const double p23 = fma(r, SSVI_EXPC[3], SSVI_EXPC[2]);
  const double p45 = fma(r, SSVI_EXPC[5], SSVI_EXPC[4]);
  const double p67 = fma(r, SSVI_EXPC[7], SSVI_EXPC[6]);
  const double p89 = fma(r, SSVI_EXPC[9], SSVI_EXPC[8]);
  const double pab = fma(r, SSVI_EXPC[11], SSVI_EXPC[10]);
  const double pcd = fma(r, SSVI_EXPC[13], SSVI_EXPC[12]);
  const double r4 = r2 * r2;
  const double q0 = fma(r2, p23, p01);
  const double q1 = fma(r2, p67, p45);
  const double q2 = fma(r2, pab, p89);
  const double r8 = r4 * r4;
  const double q3 = fma(r4, pcd, q2);
  const double s0 = fma(r4, q1, q0);
  const double p = fma(r8, q3, s0);
  return __hiloint2double(__double2hiint(p) + (n << 20), __double2loint(p));
}

// 1 / a for a in [1, 2]: hardware seed (MUFU.RCP64H, ~20 bits) + two Newton steps, no special cases to branch on
__device__ __forceinline__ double rcp_1to2(double a) {
  double y;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(a));
  double e = fma(-a, y, 1.0);
  y = fma(y, e, y);
  e = fma(-a, y, 1.0);
  return fma(y, e, y);
}

// 1 / a for a normal, positive a of moderate magnitude: hardware seed + two Newton steps (no special cases)
__device__ __forceinline__ double rcp_pos(double a) {
  double y;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(a));
  double e = fma(-a, y, 1.0);
  y = fma(y, e, y);
  e = fma(-a, y, 1.0);
  return fma(y, e, y);
}

// log(x) for x in [1, 2]: x > sqrt 2 is halved (log x = ln 2 + log(x/2)); with s = (x - 1) / (x + 1), |s| <= 0.1716,
// log x = 2 atanh(s) = 2 s + s^3 (2/3 + 2/5 s^2 + ... + 2/21 s^18), truncation below 1e-18 relative.  x - 1 is exact.
__device__ __forceinline__ double log_1to2(double x) {
  const bool hi = x > SSVI_LOGC[11];
  const double xr = hi ? 0.5 * x : x;
  const double s = (xr - 1.0) * rcp_pos(xr + 1.0);
  const double z = s * s;
  const double a01 = fma(z, SSVI_LOGC[1], SSVI_LOGC[0]);
  const double a23 = fma(z, SSVI_LOGC[3], SSVI_LOGC[2]);
  const double a45 = fma(z, SSVI_LOGC[5], SSVI_LOGC[4]);
  const double a67 = fma(z, SSVI_LOGC[7], SSVI_LOGC[6]);
  const double a89 = fma(z, SSVI_LOGC[9], SSVI_LOGC[8]);
  const double z2 = z * z;
  const double b0 = fma(z2, a23, a01);
  const double b1 = fma(z2, a67, a45);
  const double z4 = z2 * z2;
  const double c0 = fma(z4, b1, b0);
  const double z8 = z4 * z4;
  const double pz = fma(z8, a89, c0);
  const double l = fma(s * z, pz, s + s);
  return hi ? l + SSVI_LOGC[10] : l;
}

// sigmoid(f), A = poisson_link(f) = max(log(1 + e^f), 1e-8) and 1/A for |f| < 700, from ONE exponential of a non-positive
// argument: e = exp(-|f|), x = 1 + e in (1, 2]; f >= 0: sigmoid = 1/x, log(1+e^f) = f + log x; f < 0: e/x, log x
// (ProbFun.py:22-28, 52-55; for f < 0 operation for operation the reference's log(fl(1 + e^f))).
__device__ __forceinline__ void poisson_link_sigmoid_fast(double f, double& sg, double& A, double& rA) {
  const double e = exp_nonpos_c(-fabs(f));
  const double x = 1.0 + e;
  const double r = rcp_1to2(x);
  const double L = log_1to2(x);
  const bool pos = f >= 0.0;
  sg = pos ? r : e * r;
  A = fmax(pos ? f + L : L, 1e-8);
  rA = rcp_pos(A);
}

// bernoulli.py:40-51: sigmoid(yf) and sigmoid(-yf) from ONE exponential: with e = exp(-|yf|) in (0, 1],
// sigmoid(|yf|) = 1/(1+e) and sigmoid(-|yf|) = e/(1+e); equal to the reference's two evaluations to a few ulp, and
// saturating to {1, 0} where 1/(1+exp(-x)) does.
__device__ __forceinline__ void bernoulli_log_derivs(double y, double f, double& fst, double& snd) {
  const double yf = y * f;
  const double e = exp_nonpos(-fabs(yf));
  const double hi = rcp_1to2(1.0 + e), lo = e * hi;
  const bool pos = yf >= 0.0;
  double sp = pos ? hi : lo, sm = pos ? lo : hi;
  if (isnan(yf)) sp = sm = 0.0;                        // nan_to_num(sigmoid(nan)) = 0 (ProbFun.py:24)
  fst = y * sm;
  snd = -sp * sm;
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
    bernoulli_log_derivs(y, f, fst, snd);
  } else {
    if (fabs(f) < 700.0) {                            // the straight-line form of poisson_pdf_derivs_fast (below)
      double sg, A, rA;
      poisson_link_sigmoid_fast(f, sg, A, rA);
      const double yA = y * rA, t = yA - 1.0;
      fst = sg * t;
      snd = sg * (1.0 - sg) * t - (yA * rA) * (sg * sg);
      return;
    }
    const PoissonTerms pt = poisson_terms(f);
    const double sg = pt.sg, A = pt.A;
    double t = y / A - 1.0;
    fst = nan_to_num(sg * t);
    snd = nan_to_num(sg * (1.0 - sg) * t - y * (sg * sg) / (A * A));
  }
}

// Two evaluations at once (the memo / diag sweep kernels hold two samples per lane): the likelihood switch sits OUTSIDE
// the pair, so the two chains are one basic block and ptxas interleaves them.
__device__ __forceinline__ void log_pdf_derivs_pair(int lik, double y, double f0, double f1, double s, double& a0,
                                                    double& b0, double& a1, double& b1) {
  if (lik == SSVI_LIK_BERNOULLI) {
    bernoulli_log_derivs(y, f0, a0, b0);
    bernoulli_log_derivs(y, f1, a1, b1);
  } else if (lik == SSVI_LIK_NORMAL) {
    const double is2 = 1.0 / (s * s);
    a0 = is2 * (y - f0);
    a1 = is2 * (y - f1);
    b0 = b1 = -is2;
  } else {
    log_pdf_derivs(lik, y, f0, s, a0, b0);
    log_pdf_derivs(lik, y, f1, s, a1, b1);
  }
}

// ---- Poisson pdf terms for the robust model's k1 x k2 loop (4096 evaluations per observed entry at config 3) ------------
// That loop is ISSUE bound, not FP64-pipe bound, when written with libm: ptxas materialises every 64-bit literal of the
// inlined exp / log with two UMOVs right before its use (100 of ~490 instructions per evaluation,
// profiles/r01z_ncu_memo_weights.txt).  Here the coefficients sit in constant memory and are read as instruction operands,
// the reciprocals are seed + Newton without the slow-path call of a general division, and everything is straight-line:
//   e = exp(-|f|) in (0, 1],  x = 1 + e in (1, 2],  r = 1/x,  L = log(x)
//   f >= 0:  1/(1+e^f) = e r,  sigmoid(f) = r,    log(1+e^f) = f + L          f < 0:  r,  e r,  L
// (for f < 0 that is the reference's log(fl(1 + e^f)) operation for operation; for f >= 0 it agrees to a few ulp).
// pdf, d pdf / df, d2 pdf / df2 of the Poisson likelihood (poisson.py:19-57, Q7 included) for |f| < 700
__device__ __forceinline__ void poisson_pdf_derivs_fast(double y, double f, double yfact, double& p, double& p1,
                                                        double& p2) {
  const double e = exp_nonpos_c(-fabs(f));
  const double x = 1.0 + e;
  const double r = rcp_1to2(x);
  const double L = log_1to2(x);
  const double er = e * r;
  const bool pos = f >= 0.0;
  const double rE1 = pos ? er : r;                    // 1 / (1 + e^f)
  const double sg = pos ? r : er;                     // sigmoid(f)
  const double A = fmax(pos ? f + L : L, 1e-8);       // poisson_link(f)
  const double rA = rcp_pos(A);
  // Q7: the second derivative evaluates pdf(y, poisson_link(m)) -- the link applied twice (poisson.py:42); A > 0
  const double e2 = exp_nonpos_c(-A);
  const double x2 = 1.0 + e2;
  const double r2 = rcp_1to2(x2);
  const double A2 = fmax(A + log_1to2(x2), 1e-8);
  double pw = 1.0, pw2 = 1.0, b = A, b2 = A2;         // A^y and A2^y for the integral count y, one loop
  for (int n = (int)y; n > 0; n >>= 1) {
    if (n & 1) {
      pw *= b;
      pw2 *= b2;
    }
    b *= b;
    b2 *= b2;
  }
  const double yA = y * rA, t = yA - 1.0;
  p = fmax(pw * rE1 * yfact, 1e-8);
  p1 = p * (sg * t);
  const double pl = fmax(pw2 * (e2 * r2) * yfact, 1e-8);
  const double temp = -(yA * rA) * (sg * sg) + t * (sg * (1.0 - sg));
  p2 = p1 * (sg * t) + pl * temp;
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
    if (fabs(f) < 700.0) {                            // everything a sane run produces; the rest keeps libm's edge cases
      poisson_pdf_derivs_fast(y, f, yfact, p, p1, p2);
      return;
    }
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
