// Stage 4: the batched per-row update (one CTA per touched row, everything in shared memory).
//
// Reference: update_cov_param (SSVI_TF_interface.py:616-637), update_mean_param (:595-613), their
// diagonal twins (:640-681), the AdaGrad step sizes incl. the aliasing quirk Q1 (:694-713), the
// trackers (:199, :722-729) and the write-back (:219).
//
// Full covariance, cov_update "N":
//     S+ = inv((1-rho) inv(S) + rho (I/sigma - 2 Di)),  then eigenvalues clamped at min_eigenval.
// inv(A) and A share eigenvectors, so ONE symmetric eigendecomposition A = V diag(lam) V^T gives
//     S+ = V diag(max(1/lam, min_eigenval)) V^T
// which equals the reference's inv-then-eigh-clamp for every non-singular A (a negative lam gives a
// negative eigenvalue of inv(A), which the reference clamps to min_eigenval as well).
// inv(S) comes from the cached Cholesky factor: inv(S) = L^-T L^-1.
#pragma once
#include "ssvi_common.cuh"

struct RowUpdateArgs {
  int32_t D, diag, model, lik, k1;
  int32_t mean_update, cov_update, dim, time_step, fixed_cov;
  int32_t row_len, srow_len, cov_len, acc_len;
  double eta, cov_eta, poisson_eta, min_eigenval, cov_floor;
  double* params;
  double* sampler;
  double* ada_mean;
  double* ada_cov;
  double* norm_changes;
  double* grad_norms;
  const int64_t* rows;
  const double* acc;
  const int64_t* n_obs;
  const double* pSigma;
  const double* w_state;
  double* si_out;
  int32_t* flags;
  const int32_t* stop;   // device word of ssvi_factorize_iterations: non-zero = the stop condition has fired, leave the state alone
};

// block-wide sum; `red` holds >= 32 doubles of shared memory. All threads must call.
__device__ __forceinline__ double block_sum(double v, double* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  double t = 0.0;
  for (int i = 0; i < nw; ++i) t += red[i];   // same order in every thread: deterministic
  return t;
}

// Cyclic two-sided Jacobi with the round-robin (tournament) parallel ordering: A <- J^T A J, V <- V J, where J is the
// product of the n/2 disjoint plane rotations of the round.
//   buf   four work matrices of D * LD doubles, LD = D | 1 (an odd stride keeps the column walks of a rotation off a
//         single shared-memory bank: stride D = 32 was a 16-way conflict, 38.7 ms per sweep of 10k rank-32 rows).
//         buf[3] holds the input on entry: D x D row-major with stride D, symmetric; it is copied out first.
//   lam   out  D eigenvalues
//   returns the eigenvector matrix (columns; row-major with stride LD; lam[k] belongs to column k), one of buf[2], buf[3]
//   scratch  3 * (D + 1) + 32 doubles
// One round = every thread forms its elements of J^T A J from four entries of the current A and writes them to the
// OTHER buffer (ping-pong), so a round costs two barriers (rotations known; new matrix written) instead of three passes.
__host__ __device__ inline int jacobi_ld(int D) { return D | 1; }

__device__ const double* jacobi_eigh(double* const buf[4], double* lam, int D, double* scratch, int32_t* flags) {
  const int tid = threadIdx.x, nt = blockDim.x;
  const int LD = jacobi_ld(D);
  const int n = (D + 1) & ~1;                 // players (even); index D is a bye when D is odd
  const int np = n / 2;
  double* rc = scratch;                       // [D] cosine of the rotation that touches index j
  double* rs = rc + (D + 1);                  // [D] signed sine: new_j = rc[j] x_j + rs[j] x_partner
  int* part = (int*)(rs + (D + 1));           // [D] partner index (itself: no rotation)
  double* red = scratch + 3 * (D + 1);
  double* Ac = buf[0];
  double* An = buf[1];
  double* Vc = buf[2];
  double* Vn = buf[3];
  double tot = 0.0;
  for (int idx = tid; idx < D * D; idx += nt) {
    const int i = idx / D, j = idx % D;
    const double a = buf[3][idx];
    Ac[i * LD + j] = a;
    Vc[i * LD + j] = (i == j) ? 1.0 : 0.0;
    tot += a * a;
  }
  tot = block_sum(tot, red);                  // Frobenius norm: invariant under the rotations; buf[3] is free from here
  const int max_sweeps = 40;
  int sweep = 0;
  bool done = (tot == 0.0);
  for (; sweep < max_sweeps && !done; ++sweep) {
    double off = 0.0;
    for (int r = 0; r < n - 1; ++r) {
      // tournament pairing of round r: player n-1 fixed, the others rotate
      if (tid < np) {
        const int a = (tid == 0) ? n - 1 : (r + tid) % (n - 1);
        const int b = (r + n - 1 - tid) % (n - 1);
        const int p = min(a, b), q = max(a, b);
        if (q < D) {
          double c = 1.0, sn = 0.0;
          const double apq = Ac[p * LD + q];
          if (apq != 0.0) {
            const double app = Ac[p * LD + p], aqq = Ac[q * LD + q];
            const double tau = (aqq - app) / (2.0 * apq);
            const double t = (tau >= 0.0 ? 1.0 : -1.0) / (fabs(tau) + sqrt(1.0 + tau * tau));
            c = rsqrt(1.0 + t * t);
            sn = t * c;
          }
          rc[p] = c;
          rc[q] = c;
          rs[p] = -sn;
          rs[q] = sn;
          part[p] = (sn == 0.0) ? p : q;
          part[q] = (sn == 0.0) ? q : p;
        } else {                              // bye
          rc[p] = 1.0;
          rs[p] = 0.0;
          part[p] = p;
        }
      }
      __syncthreads();
      const bool last = (r == n - 2);
      for (int idx = tid; idx < D * D; idx += nt) {
        const int i = idx / D, j = idx % D;
        const int pi = part[i], pj = part[j];
        const double ci = rc[i], si = rs[i], cj = rc[j], sj = rs[j];
        const double b0 = cj * Ac[i * LD + j] + sj * Ac[i * LD + pj];      // (A J)[i][j]
        const double b1 = cj * Ac[pi * LD + j] + sj * Ac[pi * LD + pj];    // (A J)[pi][j]
        const double x = ci * b0 + si * b1;                                // (J^T A J)[i][j]
        An[i * LD + j] = x;
        Vn[i * LD + j] = cj * Vc[i * LD + j] + sj * Vc[i * LD + pj];
        if (last && i != j) off += x * x;
      }
      __syncthreads();
      double* t0 = Ac;
      Ac = An;
      An = t0;
      t0 = Vc;
      Vc = Vn;
      Vn = t0;
    }
    // converged: off-diagonal Frobenius mass at the rounding floor (about D eps^2 of the total after a sweep)
    off = block_sum(off, red);
    done = off <= 1e-27 * tot;
  }
  if (!done && tid == 0) atomicOr(flags, SSVI_FLAG_EIG_NOCONV);
  for (int i = tid; i < D; i += nt) lam[i] = Ac[i * LD + i];
  __syncthreads();
  return Vc;
}

// In-place Cholesky of the symmetric matrix in Sx (shared, row-major) into its lower triangle;
// the strict upper triangle is zeroed.  Returns false (in every thread) on a non-positive pivot.
__device__ bool cholesky_lower(double* Sx, int D) {
  const int tid = threadIdx.x, nt = blockDim.x;
  __shared__ int ok;
  if (tid == 0) ok = 1;
  __syncthreads();
  for (int j = 0; j < D; ++j) {
    if (tid == 0) {
      double d = Sx[j * D + j];
      for (int k = 0; k < j; ++k) d -= Sx[j * D + k] * Sx[j * D + k];
      if (!(d > 0.0)) {
        ok = 0;
        d = 1e-300;
      }
      Sx[j * D + j] = sqrt(d);
    }
    __syncthreads();
    const double piv = Sx[j * D + j];
    for (int i = j + 1 + tid; i < D; i += nt) {
      double a = Sx[i * D + j];
      for (int k = 0; k < j; ++k) a -= Sx[i * D + k] * Sx[j * D + k];
      Sx[i * D + j] = a / piv;
    }
    __syncthreads();
  }
  for (int i = tid; i < D * D; i += nt)
    if (i % D > i / D) Sx[i] = 0.0;
  __syncthreads();
  return ok != 0;
}

// Sampler-cache row (full covariance, layout in ssvi_common.cuh) <-> row-major lower-triangular D x D.
__device__ void sampler_store_full(double* __restrict__ srow, const double* mean, const double* Lsm, int D) {
  const int tid = threadIdx.x, nt = blockDim.x;
  const int NT = smp_nt(D), NS = smp_ns(D), len = smp_row_len(D, false);
  for (int i = tid; i < 8 * NT; i += nt) srow[i] = (i < D) ? mean[i] : 0.0;
  for (int o = 8 * NT + tid; o < len; o += nt) {
    const int e = o - 8 * NT, tile = e >> 5, g = (e >> 2) & 7, q = e & 3;
    int ti = 0, base = 0;
    while (base + smp_tiles_in_row(ti, NS) <= tile) {
      base += smp_tiles_in_row(ti, NS);
      ++ti;
    }
    const int n = 8 * ti + g, c = 4 * (tile - base) + q;
    srow[o] = (n < D && c <= n) ? Lsm[n * D + c] : 0.0;
  }
}

__device__ void sampler_load_L(const double* __restrict__ srow, double* Lsm, int D) {
  for (int idx = threadIdx.x; idx < D * D; idx += blockDim.x) {
    const int n = idx / D, c = idx % D;
    Lsm[idx] = (c <= n) ? srow[smp_L_offset(n, c, D)] : 0.0;
  }
}

// AdaGrad step with the aliasing quirk Q1 (SSVI_TF_interface.py:694-698, 705-713):
// the stored accumulator becomes acc + g^2, the step uses acc + 2 g^2.
__device__ __forceinline__ double adagrad_step(double* acc, double g, double eta) {
  double a = *acc, g2 = g * g;
  *acc = a + g2;
  return eta / sqrt(a + g2 + g2);
}

__device__ __forceinline__ double finalize_si(const RowUpdateArgs& a, double si_sum, double count,
                                              double scale) {
  if (a.model == SSVI_MODEL_D) return 0.0;
  const double w = a.w_state[0];
  double si = si_sum / ((double)a.k1 * 8.0 * w * w);
  if (a.model == SSVI_MODEL_SIMPLE) si /= count;     // SSVI_TF_simple.py:39 mean over (B,k1)
  return si * scale;                                  // update_sigma_param: si_acc *= scale (:685)
}

// ------------------------------------------------------------------------------------------------
// Diagonal covariance: one thread per (row, component); blockDim.x >= D, one row per block.
// ------------------------------------------------------------------------------------------------
__global__ void row_update_diag_kernel(RowUpdateArgs a) {
  __shared__ double red[32];
  const int r = blockIdx.x, n = threadIdx.x, D = a.D;
  if (a.stop && *a.stop) return;
  const double* acc = a.acc + (int64_t)r * a.acc_len;
  const double count = acc[D + a.cov_len + 1];
  if (count == 0.0) return;                                   // :166-168 nothing observed
  const int64_t row = a.rows[r];
  const double scale = (double)a.n_obs[r] / count;            // :181
  const double sigma = a.pSigma[a.dim];
  const bool act = n < D;
  double m = 0, S = 0, di = 0, Di = 0;
  double* prow = a.params + row * a.row_len;
  if (act) {
    m = prow[n];
    S = prow[D + n];
    di = acc[n] * scale;
    Di = acc[D + n] * scale;
  }
  bool bad_acc = act && !(isfinite(di) && isfinite(Di));
  // covariance (:661-681)
  double covGrad = 0, S_next = S;
  if (act) {
    if (a.cov_update == SSVI_UPDATE_S) {                       // Q8: L_next discarded
      double L = sqrt(S);
      covGrad = 1.0 / L - L / sigma + 2.0 * L * Di;
      adagrad_step(a.ada_cov + row * D + n, covGrad, 1.0);
      S_next = L * L;
    } else {
      covGrad = 1.0 / sigma - 2.0 * Di;
      double eta = (a.lik == SSVI_LIK_POISSON) ? a.poisson_eta : a.eta;   // :710-713
      double step = adagrad_step(a.ada_cov + row * D + n, covGrad, eta);
      S_next = 1.0 / ((1.0 - step) * (1.0 / S) + step * covGrad);
    }
    S_next = fmax(S_next, a.cov_floor);                        // :680
  }
  // mean (:640-658), from the OLD S
  double g = 0, m_next = m;
  if (act) {
    if (a.mean_update == SSVI_UPDATE_S) {
      g = (1.0 - m) / sigma + di;                              // pmu = ones (:55)
      double step = adagrad_step(a.ada_mean + row * D + n, g, a.eta);
      m_next = m + step * g;
    } else {
      double C = m / S;
      g = 1.0 / sigma + di - 2.0 * Di * m;
      double step = adagrad_step(a.ada_mean + row * D + n, g, a.eta);
      m_next = S * ((1.0 - step) * C + step * g);
    }
  }
  bool bad_cov = act && !isfinite(S_next);
  if (a.fixed_cov) S_next = S;                                 // :213-214
  double gm = block_sum(g * g, red), gc = block_sum(covGrad * covGrad, red);
  double dm = block_sum((m_next - m) * (m_next - m), red), dc = block_sum((S_next - S) * (S_next - S), red);
  if (act) {
    prow[n] = m_next;
    prow[D + n] = S_next;
    double* srow = a.sampler + row * a.row_len;
    srow[n] = m_next;
    srow[D + n] = sqrt(S_next);
  }
  if (bad_acc) atomicOr(a.flags, SSVI_FLAG_NONFINITE_ACC);
  if (bad_cov) atomicOr(a.flags, SSVI_FLAG_NONFINITE_COV);
  if (n == 0) {
    a.grad_norms[row * 2 + 0] = sqrt(gm);
    a.grad_norms[row * 2 + 1] = sqrt(gc);
    a.norm_changes[row * 2 + 0] = sqrt(dm);
    a.norm_changes[row * 2 + 1] = sqrt(dc);
    if (a.si_out) a.si_out[r] = finalize_si(a, acc[D + a.cov_len], count, scale);
  }
}

// ------------------------------------------------------------------------------------------------
// Full covariance: one CTA per row.  Shared memory (doubles): 5 matrices of D*LD (LD = D | 1: four of them double as the
// eigen-solver's work matrices) + 6*D + scratch.
// ------------------------------------------------------------------------------------------------
__host__ __device__ inline size_t row_update_full_doubles(int D) {
  return (size_t)5 * D * jacobi_ld(D) + 6 * D + 3 * (D + 1) + 32;
}

__global__ void __launch_bounds__(128) row_update_full_kernel(RowUpdateArgs a) {
  extern __shared__ double sm[];
  const int D = a.D, DD = D * D, tid = threadIdx.x, nt = blockDim.x;
  if (a.stop && *a.stop) return;
  const int MS = D * jacobi_ld(D);   // doubles per matrix slot (the code outside the eigen-solver uses stride D inside it)
  double* Sold = sm;             // S before the update
  double* Lm = Sold + MS;        // cached Cholesky factor of S; eigen-solver work; later chol(S_next)
  double* Li = Lm + MS;          // inv(L); eigen-solver work; later S_next
  double* A = Li + MS;           // matrix handed to the eigen-solver
  double* Vm = A + MS;           // eigen-solver work
  double* mv = Vm + MS;          // m
  double* gv = mv + D;           // mean gradient
  double* tv = gv + D;           // temp vector
  double* cv = tv + D;           // C = inv(S) m
  double* wv = cv + D;           // transformed eigenvalues
  double* mn = wv + D;           // m_next
  double* scratch = mn + D;
  double* red = scratch + 3 * (D + 1);
  double* const jbuf[4] = {Lm, Li, Vm, A};

  const int r = blockIdx.x;
  const double* acc = a.acc + (int64_t)r * a.acc_len;
  const double count = acc[D + DD + 1];
  if (count == 0.0) return;                                    // :166-168
  const int64_t row = a.rows[r];
  const double scale = (double)a.n_obs[r] / count;             // :181
  const double sigma = a.pSigma[a.dim];
  const double* di = acc;
  const double* Di = acc + D;
  double* prow = a.params + row * a.row_len;
  double* srow = a.sampler + row * (int64_t)a.srow_len;

  bool bad = false;
  sampler_load_L(srow, Lm, D);
  for (int i = tid; i < DD; i += nt) {
    Sold[i] = prow[D + i];
    if (!isfinite(Di[i])) bad = true;
  }
  for (int i = tid; i < D; i += nt) {
    mv[i] = prow[i];
    if (!isfinite(di[i])) bad = true;
  }
  if (bad) atomicOr(a.flags, SSVI_FLAG_NONFINITE_ACC);          // :185-188
  __syncthreads();

  // inv(L): thread j solves column j by forward substitution
  for (int j = tid; j < D; j += nt) {
    for (int i = 0; i < j; ++i) Li[i * D + j] = 0.0;
    Li[j * D + j] = 1.0 / Lm[j * D + j];
    for (int i = j + 1; i < D; ++i) {
      double s = 0.0;
      for (int k = j; k < i; ++k) s += Lm[i * D + k] * Li[k * D + j];
      Li[i * D + j] = -s / Lm[i * D + i];
    }
  }
  __syncthreads();

  double gc2 = 0.0;   // ||covGrad||_F^2 partial
  if (a.cov_update == SSVI_UPDATE_N) {
    const double rho = a.cov_eta / (double)(a.time_step + 1);   // :701-702
    for (int idx = tid; idx < DD; idx += nt) {
      const int i = idx / D, j = idx % D;
      double p = 0.0;                                           // inv(S)[i][j] = sum_k Li[k][i] Li[k][j]
      for (int k = max(i, j); k < D; ++k) p += Li[k * D + i] * Li[k * D + j];
      const double cg = (i == j ? 1.0 / sigma : 0.0) - 2.0 * (scale * Di[idx]);   // :628
      gc2 += cg * cg;
      A[idx] = (1.0 - rho) * p + rho * cg;                      // :630 (before the outer inv)
    }
  } else {
    // cov_update "S" (quirk Q8): covGrad only feeds grad_norms; S_next = L L^T, then the clamp
    for (int idx = tid; idx < DD; idx += nt) {
      const int i = idx / D, j = idx % D;
      double s = 0.0;
      for (int k = 0; k <= min(i, j); ++k) s += Lm[i * D + k] * Lm[j * D + k];
      A[idx] = s;
      if (j >= i) {                                             // np.triu (:621-622)
        double ld = 0.0;
        for (int k = 0; k <= i; ++k) ld += Lm[i * D + k] * (scale * Di[k * D + j]);
        double cg = (i == j ? 1.0 / Lm[i * D + i] : 0.0) - (i == j ? Lm[i * D + i] / sigma : 0.0) + 2.0 * ld;
        gc2 += cg * cg;
      }
    }
  }
  __syncthreads();
  // symmetrise (the two triangles of A are computed by different threads in different orders)
  for (int idx = tid; idx < DD; idx += nt) {
    const int i = idx / D, j = idx % D;
    if (i < j) {
      double s = 0.5 * (A[i * D + j] + A[j * D + i]);
      A[i * D + j] = s;
      A[j * D + i] = s;
    }
  }
  __syncthreads();

  // mean gradient (needs inv(S) m for the "N" scheme) -- from the OLD m and S (:191-196, Q10)
  if (a.mean_update == SSVI_UPDATE_N) {
    for (int i = tid; i < D; i += nt) {                         // tv = inv(L) m
      double s = 0.0;
      for (int k = 0; k <= i; ++k) s += Li[i * D + k] * mv[k];
      tv[i] = s;
    }
    __syncthreads();
    for (int i = tid; i < D; i += nt) {                         // cv = inv(L)^T tv = inv(S) m   (:605)
      double s = 0.0;
      for (int k = i; k < D; ++k) s += Li[k * D + i] * tv[k];
      cv[i] = s;
    }
    __syncthreads();
  }
  double gm2 = 0.0, dm2 = 0.0;
  for (int i = tid; i < D; i += nt) {
    double g;
    if (a.mean_update == SSVI_UPDATE_S) {
      g = (1.0 - mv[i]) / sigma + scale * di[i];               // :600, pmu = ones (:55)
    } else {
      double s = 0.0;
      for (int k = 0; k < D; ++k) s += (scale * Di[i * D + k]) * mv[k];
      g = 1.0 / sigma + scale * di[i] - 2.0 * s;               // :606
    }
    const double step = adagrad_step(a.ada_mean + row * D + i, g, a.eta);
    gv[i] = g;
    gm2 += g * g;
    if (a.mean_update == SSVI_UPDATE_S) {
      mn[i] = mv[i] + step * g;                                // :602
    } else {
      tv[i] = (1.0 - step) * cv[i] + step * g;                 // :608 C_next
    }
  }
  __syncthreads();
  if (a.mean_update == SSVI_UPDATE_N) {
    for (int i = tid; i < D; i += nt) {                        // m_next = S C_next (:609)
      double s = 0.0;
      for (int k = 0; k < D; ++k) s += Sold[i * D + k] * tv[k];
      mn[i] = s;
    }
    __syncthreads();
  }
  for (int i = tid; i < D; i += nt) dm2 += (mn[i] - mv[i]) * (mn[i] - mv[i]);

  // eigen-decomposition + clamp (:634-636)
  const double* Ev = jacobi_eigh(jbuf, tv, D, scratch, a.flags);   // eigenvectors: stride LD
  const int LD = jacobi_ld(D);
  for (int i = tid; i < D; i += nt) {
    const double lam = tv[i];
    const double w = (a.cov_update == SSVI_UPDATE_N) ? 1.0 / lam : lam;
    wv[i] = fmax(w, a.min_eigenval);                           // NaN -> min_eigenval, like np.maximum? no:
    if (isnan(w)) wv[i] = w;                                   // np.maximum propagates NaN
  }
  __syncthreads();
  double* Sn = Li;                                             // S_next = V diag(w) V^T
  bool badS = false;
  for (int idx = tid; idx < DD; idx += nt) {
    const int i = idx / D, j = idx % D;
    if (j > i) continue;
    double s = 0.0;
    for (int k = 0; k < D; ++k) s += Ev[i * LD + k] * wv[k] * Ev[j * LD + k];
    if (!isfinite(s)) badS = true;
    Sn[i * D + j] = s;
    Sn[j * D + i] = s;
  }
  if (badS) atomicOr(a.flags, SSVI_FLAG_NONFINITE_COV);        // :202-204
  __syncthreads();
  double dc2 = 0.0;
  if (!a.fixed_cov) {
    for (int idx = tid; idx < DD; idx += nt) {
      const double d = Sn[idx] - Sold[idx];
      dc2 += d * d;
      prow[D + idx] = Sn[idx];
    }
  }
  for (int i = tid; i < D; i += nt) prow[i] = mn[i];
  gm2 = block_sum(gm2, red);
  gc2 = block_sum(gc2, red);
  dm2 = block_sum(dm2, red);
  dc2 = block_sum(dc2, red);
  if (tid == 0) {
    a.grad_norms[row * 2 + 0] = sqrt(gm2);                      // :199
    a.grad_norms[row * 2 + 1] = sqrt(gc2);
    a.norm_changes[row * 2 + 0] = sqrt(dm2);                    // :722-729
    a.norm_changes[row * 2 + 1] = sqrt(dc2);
    if (a.si_out) a.si_out[r] = finalize_si(a, acc[D + DD], count, scale);
  }
  __syncthreads();
  if (!a.fixed_cov) {                                           // sampler cache: (m_next, chol(S_next))
    const bool ok = cholesky_lower(Sn, D);
    if (!ok && tid == 0) atomicOr(a.flags, SSVI_FLAG_CHOL_FAIL);
    sampler_store_full(srow, mn, Sn, D);
  } else {
    for (int i = tid; i < D; i += nt) srow[i] = mn[i];
  }
}

// ------------------------------------------------------------------------------------------------
// Sampler refresh: sampler row <- (mean, chol(S)) from params.  One CTA per row (full) / thread per
// component (diag).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) refresh_sampler_full_kernel(const double* __restrict__ params,
                                                                   double* __restrict__ sampler,
                                                                   const int64_t* __restrict__ rows, int D,
                                                                   int32_t* flags) {
  extern __shared__ double sm[];
  const int tid = threadIdx.x, nt = blockDim.x, DD = D * D, row_len = (D + 1) * D;
  const int64_t row = rows ? rows[blockIdx.x] : (int64_t)blockIdx.x;
  const double* p = params + row * row_len;
  double* s = sampler + row * (int64_t)smp_row_len(D, false);
  double* mean = sm + DD;
  for (int i = tid; i < DD; i += nt) sm[i] = p[D + i];
  for (int i = tid; i < D; i += nt) mean[i] = p[i];
  __syncthreads();
  const bool ok = cholesky_lower(sm, D);
  if (!ok && tid == 0) atomicOr(flags, SSVI_FLAG_CHOL_FAIL);
  sampler_store_full(s, mean, sm, D);
}

__global__ void refresh_sampler_diag_kernel(const double* __restrict__ params, double* __restrict__ sampler,
                                            const int64_t* __restrict__ rows, int64_t n_rows, int D) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_rows * D) return;
  const int64_t r = t / D;
  const int n = (int)(t % D);
  const int64_t row = rows ? rows[r] : r;
  sampler[row * 2 * D + n] = params[row * 2 * D + n];
  sampler[row * 2 * D + D + n] = sqrt(params[row * 2 * D + D + n]);
}
