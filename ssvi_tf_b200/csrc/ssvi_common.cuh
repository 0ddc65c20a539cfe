// Shared definitions for the sm_100a SSVI-TF kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <float.h>
#include <math.h>

#include "../../include/ssvi_b200.h"

struct SsviHandle {
  SsviConfig cfg;
  int64_t acc_len;
  // data
  const int32_t* coords;
  const double* vals;
  int64_t nnz;
  // state (device pointers, one per mode)
  double* params[SSVI_MAX_ORDER];
  double* sampler[SSVI_MAX_ORDER];
  double* ada_mean[SSVI_MAX_ORDER];
  double* ada_cov[SSVI_MAX_ORDER];
  double* norm_changes[SSVI_MAX_ORDER];
  double* grad_norms[SSVI_MAX_ORDER];
  bool data_bound, state_bound, noise_bound;
  // noise pool
  const void* z_other;
  const void* z_own;
  const void* r_w;
  const void* z_f;
  int32_t pool_slots;
  // Gauss-Hermite rule of SSVI_TF_robust(quadrature=True): device [nodes (n) | weights (n)], n == 0: Monte Carlo
  const double* quad;
  int32_t quad_n;
  // device flag word
  int32_t* d_flags;
  const int32_t* stop;   // set only while ssvi_factorize_iterations runs: device stop word honoured by the state-writing kernels
  int sm_count;
  int max_smem_optin;
  // sweep path: the sample table covers rows [win_lo, win_lo + win_n) of every mode (ssvi_set_row_windows; default: all)
  int64_t win_lo[SSVI_MAX_ORDER], win_n[SSVI_MAX_ORDER];
  // ssvi_factorize_iterations: the side stream the next row update's minibatch plan is drawn on, created on first use
  cudaStream_t loop_stream;
  cudaEvent_t loop_ev_plan[2], loop_ev_free[2], loop_ev_acc, loop_ev_noise;
  bool loop_ready;
  char err[512];
};

// Kernel-side view of the problem (passed by value).
struct DevView {
  int32_t N, D, k1, k2;
  int32_t diag, model, lik;
  int32_t dim;           // mode being swept
  int32_t row_len;       // doubles per posterior row: (D+1)*D full, 2*D diag
  int32_t srow_len;      // doubles per sampler-cache row (smp_row_len)
  int32_t zblk;          // elements per (slot, mode) noise block (smp_zblk)
  int32_t cov_len;       // D*D or D
  int32_t acc_len;
  int32_t pool_slots;
  int32_t tuning;        // SSVI_TUNE_* bits
  double lik_param;
  const int32_t* coords;
  const double* vals;
  const double* sampler[SSVI_MAX_ORDER];
  int32_t* flags;
  const double* quad;    // SSVI_TF_robust quadrature rule [nodes | weights], NULL: second-layer Monte Carlo
  int32_t quad_n;
};

// ------------------------------------------------------------------------------------------------
// Sampler cache row layout (library-internal; the posterior `params` keep the reference's layout).
//   diagonal covariance : [ mean (D) | sqrt(S) (D) ]
//   full covariance     : [ mean, zero-padded to 8*NT | lower-triangle tiles of L ]
// with NT = ceil(D/8) row tiles and NS = ceil(D/4) column steps.  Tile (i, j) holds the 8x4 block
// L[8i..8i+7][4j..4j+3] as 32 doubles in DMMA A-fragment order (element (g, q) at g*4 + q), so one
// warp reads a fragment as 32 consecutive doubles.  Only tiles touching the lower triangle are stored
// (j <= min(2i+1, NS-1)), in (i, j) order; entries above the diagonal inside a stored tile are zero.
// ------------------------------------------------------------------------------------------------
__host__ __device__ inline int smp_nt(int D) { return (D + 7) >> 3; }
__host__ __device__ inline int smp_ns(int D) { return (D + 3) >> 2; }
__host__ __device__ inline int smp_tiles_in_row(int i, int NS) { return (2 * (i + 1) < NS) ? 2 * (i + 1) : NS; }
__host__ __device__ inline int smp_tile_base(int i, int NS) {
  int s = 0;
  for (int r = 0; r < i; ++r) s += smp_tiles_in_row(r, NS);
  return s;
}
__host__ __device__ inline int smp_row_len(int D, bool diag) {
  if (diag) return 2 * D;
  return 8 * smp_nt(D) + 32 * smp_tile_base(smp_nt(D), smp_ns(D));
}
// offset of L[n][c] (c <= n) inside a full-covariance sampler row
__host__ __device__ inline int smp_L_offset(int n, int c, int D) {
  const int NS = smp_ns(D);
  return 8 * smp_nt(D) + 32 * (smp_tile_base(n >> 3, NS) + (c >> 2)) + (n & 7) * 4 + (c & 3);
}
// Noise block of one (slot, mode): element (c, k).
//   diagonal covariance : [c][k]                       (lanes over k read consecutive elements)
//   full covariance     : DMMA B-fragment order, ((k/8)*NS + c/4)*32 + (k%8)*4 + (c%4)
__host__ __device__ inline int smp_zblk(int D, int k1, bool diag) {
  return diag ? D * k1 : ((k1 + 7) >> 3) * smp_ns(D) * 32;
}
__host__ __device__ inline int smp_z_offset(int c, int k, int D, int k1, bool diag) {
  if (diag) return c * k1 + k;
  return ((k >> 3) * smp_ns(D) + (c >> 2)) * 32 + (k & 7) * 4 + (c & 3);
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// warp sums of TWO values for half the shuffles and additions of two warp_sum calls: after the first exchange even
// lanes carry a's pair sums and odd lanes b's; the butterfly then runs over lane pairs.  Every lane gets both totals.
__device__ __forceinline__ void warp_sum2(double& a, double& b, int lane) {
  const bool odd = lane & 1;
  double keep = odd ? b : a;
  keep += __shfl_xor_sync(0xffffffffu, odd ? a : b, 1);
#pragma unroll
  for (int o = 16; o > 1; o >>= 1) keep += __shfl_xor_sync(0xffffffffu, keep, o);
  const double other = __shfl_xor_sync(0xffffffffu, keep, 1);
  a = odd ? other : keep;
  b = odd ? keep : other;
}

__device__ __forceinline__ double ldz(const double* p, int64_t i) { return __ldg(p + i); }
__device__ __forceinline__ double ldz(const float* p, int64_t i) { return (double)__ldg(p + i); }

// np.nan_to_num for float64: nan -> 0, +-inf -> +-DBL_MAX
__device__ __forceinline__ double nan_to_num(double x) {
  if (isnan(x)) return 0.0;
  if (isinf(x)) return x > 0 ? DBL_MAX : -DBL_MAX;
  return x;
}
