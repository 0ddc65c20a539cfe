// Per-entry Monte-Carlo estimators of the report path.  One sampling pattern serves the
// reference's three estimators:
//   estimate_expected_observation_via_sampling  (SSVI_TF_interface.py:1168-1189)   mean of link(fs)
//   estimate_expectation_term_vlb               (:889-908)                         mean of log_pdf(y, fs)
//   estimate_negative_log_likelihood            (:939-964)                         mean of pdf(y, fs)  (host: -log)
// For every mode d of the entry a sample matrix x_d = m_d + L_d z_d (k draws) is taken, ms_k = sum_n prod_d x_d[n][k];
// SSVI_TF_d uses fs = ms; the simple / robust models add the noise layer fs[i][k] = ms_k + ws_k * z_f[i][k],
// ws_k = w_sigma^2 * r_k (rayleigh(w_sigma^2); normal(ms, ws, size=(k, k)) broadcasts ws over columns).
// Draws come from the bound noise pool: entry e reads slot (slot0 + e) mod P, mode 0 the own-row block z_own, mode
// d >= 1 the block z_other[.][d-1]; k = the pool's k1 draws (the reference hard-codes 20 / k1 / predict_num_samples),
// k2 second-layer draws.  One warp per entry, lanes over the draws; rows of the sampler cache are read through L1/L2
// (every lane reads the same element).  Off the timed path: sized for clarity, not for the FP64 pipe.
#pragma once
#include "likelihood.cuh"

// normal.py:16-18, 29-30; bernoulli.py:16-21, 38-39; poisson.py:19-27, 59-68; link functions :60-75
__device__ __forceinline__ double estimator_term(int what, int lik, double y, double f, double s, double yfact) {
  if (what == SSVI_EST_ROBUST_VLB) {                     // pdf(y, link_fun(fs)): SSVI_TF_robust.py:253
    if (lik == SSVI_LIK_BERNOULLI) f = sigmoid_ref(f, 0.0);
    if (lik == SSVI_LIK_POISSON) f = poisson_link_ref(f);
  }
  if (what == SSVI_EST_EXPECTED_OBS) {
    if (lik == SSVI_LIK_NORMAL) return f;
    if (lik == SSVI_LIK_BERNOULLI) return sigmoid_ref(f, 0.0);
    return poisson_link_ref(f);
  }
  if (lik == SSVI_LIK_NORMAL) {
    const double r = (y - f) / s;
    const double p = fmax(exp(-0.5 * r * r) / (2.5066282746310002 * s), 1e-5);
    return what == SSVI_EST_LOG_PDF ? log(p) : p;
  }
  if (lik == SSVI_LIK_BERNOULLI) {
    const double p = sigmoid_ref(y * f, 1e-16);
    return what == SSVI_EST_LOG_PDF ? log(p) : p;
  }
  if (what != SSVI_EST_LOG_PDF) return fmax(ipow_ref(poisson_link_ref(f), y) / (1.0 + exp(f)) * yfact, 1e-8);
  const double A = poisson_link_ref(f);                  // scipy poisson.logpmf(y, A) = xlogy(y, A) - gammaln(y+1) - A
  const double xl = (y == 0.0) ? 0.0 : y * log(A);
  return nan_to_num(xl - lgamma(y + 1.0) - A);
}

template <typename ZT>
__global__ void __launch_bounds__(128) estimate_entries_kernel(
    DevView v, int what, const int32_t* __restrict__ entries, const double* __restrict__ ys, int64_t n_entries,
    int slot0, const double* __restrict__ w_state, const ZT* __restrict__ z_other, const ZT* __restrict__ z_own,
    const ZT* __restrict__ r_w, const ZT* __restrict__ z_f, double* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int64_t e = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (e >= n_entries) return;
  const int N = v.N, D = v.D, k1 = v.k1, k2 = v.k2, diag = v.diag;
  const int slot = (int)((slot0 + e) % v.pool_slots);
  const double y = ys ? ys[e] : 0.0;
  const double s = v.lik_param;
  const double yfact = (what >= SSVI_EST_PDF && v.lik == SSVI_LIK_POISSON) ? 1.0 / factorial_ref(y) : 1.0;
  const bool layered = v.model != SSVI_MODEL_D;
  const double w_sigma = layered ? w_state[0] : 0.0;
  const double* rows[SSVI_MAX_ORDER];
  const ZT* zb[SSVI_MAX_ORDER];
  for (int d = 0; d < N; ++d) {
    rows[d] = v.sampler[d] + (int64_t)entries[e * N + d] * v.srow_len;
    zb[d] = (d == 0) ? z_own + (int64_t)slot * v.zblk : z_other + ((int64_t)slot * (N - 1) + (d - 1)) * v.zblk;
  }
  double acc = 0.0;
  double msv[4];                                         // this lane's draws k = lane + 32 j (ROBUST_VLB needs them all)
  for (int k0 = 0, j = 0; k0 < k1; k0 += 32, ++j) {
    const int k = k0 + lane;
    double ms = 0.0;
    if (k < k1) {
      for (int n = 0; n < D; ++n) {
        double p = 1.0;
        for (int d = 0; d < N; ++d) {
          const double* r = rows[d];
          double x = r[n];
          if (diag) {
            x += r[D + n] * ldz(zb[d], (int64_t)n * k1 + k);
          } else {
            for (int c = 0; c <= n; ++c) x += r[smp_L_offset(n, c, D)] * ldz(zb[d], smp_z_offset(c, k, D, k1, false));
          }
          p *= x;
        }
        ms += p;
      }
      if (what != SSVI_EST_ROBUST_VLB) {
        if (!layered) {
          acc += estimator_term(what, v.lik, y, ms, s, yfact);
        } else {
          const double ws = (w_sigma * w_sigma) * ldz(r_w, (int64_t)slot * k1 + k);
          const ZT* zf = z_f + (int64_t)slot * k2 * k1 + k;
          double t = 0.0;
          for (int i = 0; i < k2; ++i) t += estimator_term(what, v.lik, y, ms + ws * ldz(zf, (int64_t)i * k1), s, yfact);
          acc += t / (double)k2;
        }
      }
    }
    if (j < 4) msv[j] = ms;
  }
  if (what == SSVI_EST_ROBUST_VLB) {                     // mean_i log mean_k pdf(y, link(ms_k + ws_k z_f[i][k]))
    double tot = 0.0;
    for (int i = 0; i < k2; ++i) {
      double t = 0.0;
      for (int k0 = 0, j = 0; k0 < k1; k0 += 32, ++j) {
        const int k = k0 + lane;
        if (k < k1) {
          const double ws = (w_sigma * w_sigma) * ldz(r_w, (int64_t)slot * k1 + k);
          const double fs = msv[j] + ws * ldz(z_f, ((int64_t)slot * k2 + i) * k1 + k);
          t += estimator_term(what, v.lik, y, fs, s, yfact);
        }
      }
      tot += log(warp_sum(t) / (double)k1);
    }
    if (lane == 0) out[e] = tot / (double)k2;
    return;
  }
  acc = warp_sum(acc);
  if (lane == 0) out[e] = acc / (double)k1;
}
