// Stages 1-3, generic CUDA-core path: any D <= SSVI_MAX_RANK, any k1, full or diagonal covariance.
// One warp owns one row-chunk; lanes run over the k1 Monte-Carlo samples.
//
// Per entry of the chunk (reference: SSVI_TF_interface.py:229-264, 732-750):
//   stage 1  gather the (mean, Cholesky) sampler row of every other mode into shared memory
//   stage 2  x = m + L z  per sample, Hadamard product over modes -> V[n][k]; own-row sample u;
//            predictive mean f_k = u_k . v_k; likelihood weights a_k (for di) and w_k (for Di)
//   stage 3  g[n] += sum_k a_k V[n][k];  G[n1][n2] += sum_k w_k V[n1][k] V[n2][k]
// The chunk record [di | Di | si_sum | count] is written once at the end (no atomics, so the
// result is deterministic run to run).
#pragma once
#include "likelihood.cuh"

// Per-(entry,k) likelihood weights.  Returns through aw/dw; accumulates the sigma statistic.
//   d      (SSVI_TF_d.py:31-57, quirk Q3): derivatives averaged over k1, then broadcast
//   simple (SSVI_TF_simple.py:23-40):      second-layer samples fs = f + ws * z_f
//   robust (SSVI_TF_robust.py:67-169):     pdf ratios, Q6
// MODEL / LIK >= 0: the model and the likelihood as compile-time constants (memo_weights_kernel: the k2 loop is then one
// straight likelihood body whose constants stay in registers); -1: read them from the view.
#ifndef SSVI_W_UNROLL
#define SSVI_W_UNROLL 1                                 // k2 draws of the robust model evaluated side by side (A/B knob)
#endif
template <typename ZT, int MODEL = -1, int LIK = -1>
__device__ __forceinline__ void likelihood_weights(const DevView& v, const ZT* __restrict__ r_w,
                                                   const ZT* __restrict__ z_f, int slot, int k, bool active,
                                                   double y, double yfact, double f, double w_sigma,
                                                   double& aw, double& dw, double& si_part, bool& bad) {
  const int k1 = v.k1, k2 = v.k2;
  const int model = MODEL >= 0 ? MODEL : v.model, lik = LIK >= 0 ? LIK : v.lik;
  const double s = v.lik_param;
  const double inv_k1 = 1.0 / (double)k1;
  if (model == SSVI_MODEL_D) {
    double fst = 0.0, snd = 0.0;
    if (active) log_pdf_derivs(lik, y, f, s, fst, snd);
    // warp-level partial sums are combined by the caller (needs all k); return raw values
    aw = fst;
    dw = snd;
    return;
  }
  aw = 0.0;
  dw = 0.0;
  if (!active) return;
  const double ws = (w_sigma * w_sigma) * ldz(r_w, (int64_t)slot * k1 + k);   // rayleigh(scale=w_sigma^2)
  const ZT* zf = z_f + (int64_t)slot * k2 * k1 + k;
  // second-layer samples fs = f + ws * z_f (Q2: the Rayleigh draw is the standard deviation); the k2 variates are
  // fetched 8 at a time so the loads are in flight together instead of one L2 round trip per sample
  if (model == SSVI_MODEL_SIMPLE) {
    double sf = 0.0, ss = 0.0;
    for (int i0 = 0; i0 < k2; i0 += 8) {
      double zz[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) zz[j] = (i0 + j < k2) ? ldz(zf, (int64_t)(i0 + j) * k1) : 0.0;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (i0 + j < k2) {
          double a, b;
          log_pdf_derivs(lik, y, f + ws * zz[j], s, a, b);
          sf += a;
          ss += b;
        }
      }
    }
    sf /= (double)k2;
    ss /= (double)k2;
    aw = sf * inv_k1;
    dw = 0.5 * ss * inv_k1;
    si_part += ss * ws;
  } else {
    double p = 0.0, p1 = 0.0, p2 = 0.0;
    if (v.quad) {
      // SSVI_TF_robust(quadrature=True), SSVI_TF_robust.py:175-214: the k2 draws are replaced by the Gauss-Hermite
      // rule  E[g] = sum_i w_i g(f + sqrt(2 ws^2) x_i) / sqrt(pi)
      const double sw = sqrt(2.0 * (ws * ws));
      for (int i2 = 0; i2 < v.quad_n; ++i2) {
        const double wq = __ldg(v.quad + v.quad_n + i2);
        double a, b, c;
        pdf_derivs(lik, y, f + sw * __ldg(v.quad + i2), s, yfact, a, b, c);
        p += wq * a;
        p1 += wq * b;
        p2 += wq * c;
      }
      const double sqrt_pi = 1.7724538509055159;
      p /= sqrt_pi;
      p1 /= sqrt_pi;
      p2 /= sqrt_pi;
    } else {
      constexpr int WU = SSVI_W_UNROLL;
#pragma unroll WU
      for (int i2 = 0; i2 < k2; ++i2) {
        double fs = f + ws * ldz(zf, (int64_t)i2 * k1);
        double a, b, c;
        pdf_derivs(lik, y, fs, s, yfact, a, b, c);
        p += a;
        p1 += b;
        p2 += c;
      }
      p /= (double)k2;
      p1 /= (double)k2;
      p2 /= (double)k2;
    }
    if (!(isfinite(p) && isfinite(p1) && isfinite(p2)) || p == 0.0) bad = true;
    aw = (p1 / p) * inv_k1;
    dw = 0.5 / ((double)k1 * p) * (p2 - p2 * p2 / p);                        // Q6
    si_part += ws * p2 / p;
  }
}

// Dynamic shared memory per warp (doubles): own[row_len] + oth[row_len] + V[D*KS] + AW[k1] + DW[k1]
//                                           + G[cov_len] + g[D]
__host__ __device__ inline int generic_ks(int k1) { return k1 | 1; }   // odd stride: conflict-free both ways
__host__ __device__ inline size_t generic_warp_doubles(int D, int k1, int row_len, int cov_len) {
  return (size_t)2 * row_len + (size_t)D * generic_ks(k1) + 2 * (size_t)k1 + cov_len + D;
}

template <typename ZT>
__global__ void __launch_bounds__(128) accumulate_generic_kernel(
    DevView v, int64_t n_chunks, const int32_t* __restrict__ chunk_row, const int64_t* __restrict__ chunk_start,
    const int32_t* __restrict__ chunk_len, const int32_t* __restrict__ chunk_slot0,
    const int64_t* __restrict__ entry_ids, const int64_t* __restrict__ rows, const double* __restrict__ w_state,
    const ZT* __restrict__ z_other, const ZT* __restrict__ z_own, const ZT* __restrict__ r_w,
    const ZT* __restrict__ z_f, double* __restrict__ partial_out) {
  extern __shared__ double smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  const int64_t c = (int64_t)blockIdx.x * wpb + warp;
  if (c >= n_chunks) return;
  const int N = v.N, D = v.D, k1 = v.k1, dim = v.dim, diag = v.diag;
  const int row_len = v.srow_len, cov_len = v.cov_len, KS = generic_ks(k1);
  double* own = smem + (size_t)warp * generic_warp_doubles(D, k1, row_len, cov_len);
  double* oth = own + row_len;
  double* V = oth + row_len;
  double* AW = V + (size_t)D * KS;
  double* DW = AW + k1;
  double* G = DW + k1;
  double* g = G + cov_len;

  const int64_t row = rows[chunk_row[c]];
  const int64_t start = chunk_start[c];
  const int len = chunk_len[c], slot0 = chunk_slot0[c];
  const double w_sigma = (v.model == SSVI_MODEL_D) ? 1.0 : w_state[0];

  {
    const double* src = v.sampler[dim] + row * (int64_t)row_len;
    for (int i = lane; i < row_len; i += 32) own[i] = src[i];
    for (int i = lane; i < cov_len; i += 32) G[i] = 0.0;
    for (int i = lane; i < D; i += 32) g[i] = 0.0;
  }
  double si_part = 0.0;
  bool bad = false;
  __syncwarp();

  for (int e = 0; e < len; ++e) {
    const int64_t eid = entry_ids[start + e];
    const int slot = (slot0 + e) % v.pool_slots;
    const double y = v.vals[eid];
    const double yfact = (v.lik == SSVI_LIK_POISSON && v.model == SSVI_MODEL_ROBUST) ? 1.0 / factorial_ref(y) : 1.0;
    // ---- stage 1+2: other modes ------------------------------------------------------------
    int jj = 0;
    for (int d = 0; d < N; ++d) {
      if (d == dim) continue;
      const int64_t col = v.coords[eid * N + d];
      const double* src = v.sampler[d] + col * (int64_t)row_len;
      for (int i = lane; i < row_len; i += 32) oth[i] = src[i];
      __syncwarp();
      const ZT* z = z_other + ((int64_t)slot * (N - 1) + jj) * v.zblk;
      for (int k = lane; k < k1; k += 32) {
        for (int n = 0; n < D; ++n) {
          double x = oth[n];
          if (diag) {
            x += oth[D + n] * ldz(z, (int64_t)n * k1 + k);
          } else {
            for (int cc = 0; cc <= n; ++cc) x += oth[smp_L_offset(n, cc, D)] * ldz(z, smp_z_offset(cc, k, D, k1, false));
          }
          V[n * KS + k] = (jj == 0) ? x : V[n * KS + k] * x;
        }
      }
      __syncwarp();
      ++jj;
    }
    // ---- own-row sample, predictive mean, likelihood weights ---------------------------------
    const ZT* zo = z_own + (int64_t)slot * v.zblk;
    double sum_a = 0.0, sum_d = 0.0;
    for (int k0 = 0; k0 < k1; k0 += 32) {
      const int k = k0 + lane;
      const bool active = k < k1;
      double f = 0.0;
      if (active) {
        for (int n = 0; n < D; ++n) {
          double u = own[n];
          if (diag) {
            u += own[D + n] * ldz(zo, (int64_t)n * k1 + k);
          } else {
            for (int cc = 0; cc <= n; ++cc) u += own[smp_L_offset(n, cc, D)] * ldz(zo, smp_z_offset(cc, k, D, k1, false));
          }
          f += u * V[n * KS + k];
        }
      }
      double aw, dw;
      likelihood_weights<ZT>(v, r_w, z_f, slot, k, active, y, yfact, f, w_sigma, aw, dw, si_part, bad);
      if (v.model == SSVI_MODEL_D) {
        sum_a += aw;
        sum_d += dw;
      } else if (active) {
        AW[k] = aw;
        DW[k] = dw;
      }
    }
    if (v.model == SSVI_MODEL_D) {       // Q3: average over k1, then broadcast
      sum_a = warp_sum(sum_a) / (double)k1;
      sum_d = warp_sum(sum_d) / (double)k1;
      for (int k = lane; k < k1; k += 32) {
        AW[k] = sum_a / (double)k1;
        DW[k] = 0.5 * sum_d / (double)k1;
      }
    }
    __syncwarp();
    // ---- stage 3: contraction over k ----------------------------------------------------------
    for (int n = lane; n < D; n += 32) {
      double a = 0.0;
      for (int k = 0; k < k1; ++k) a += AW[k] * V[n * KS + k];
      g[n] += a;
    }
    if (diag) {
      for (int n = lane; n < D; n += 32) {
        double a = 0.0;
        for (int k = 0; k < k1; ++k) {
          double x = V[n * KS + k];
          a += DW[k] * x * x;
        }
        G[n] += a;
      }
    } else {
      const int npairs = D * (D + 1) / 2;
      for (int p = lane; p < npairs; p += 32) {
        // unrank p -> (n1 >= n2)
        int n1 = (int)((sqrt(8.0 * p + 1.0) - 1.0) * 0.5);
        while ((n1 + 1) * (n1 + 2) / 2 <= p) ++n1;
        while (n1 * (n1 + 1) / 2 > p) --n1;
        const int n2 = p - n1 * (n1 + 1) / 2;
        double a = 0.0;
        for (int k = 0; k < k1; ++k) a += DW[k] * V[n1 * KS + k] * V[n2 * KS + k];
        G[n1 * D + n2] += a;
      }
    }
    __syncwarp();
  }

  // ---- chunk record -------------------------------------------------------------------------------
  double* out = partial_out + c * (int64_t)v.acc_len;
  for (int n = lane; n < D; n += 32) out[n] = g[n];
  if (diag) {
    for (int n = lane; n < D; n += 32) out[D + n] = G[n];
  } else {
    for (int i = lane; i < D * D; i += 32) {
      int a = i / D, b = i % D;
      out[D + i] = (a >= b) ? G[a * D + b] : G[b * D + a];
    }
  }
  si_part = warp_sum(si_part);
  const unsigned anybad = __ballot_sync(0xffffffffu, bad);
  if (lane == 0) {
    out[D + cov_len] = si_part;
    out[D + cov_len + 1] = (double)len;
    if (anybad) atomicOr(v.flags, SSVI_FLAG_PHI_ZERO);
  }
}
