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
  // device flag word
  int32_t* d_flags;
  int sm_count;
  int max_smem_optin;
  char err[512];
};

// Kernel-side view of the problem (passed by value).
struct DevView {
  int32_t N, D, k1, k2;
  int32_t diag, model, lik;
  int32_t dim;           // mode being swept
  int32_t row_len;       // doubles per posterior / sampler row: (D+1)*D full, 2*D diag
  int32_t cov_len;       // D*D or D
  int32_t acc_len;
  int32_t pool_slots;
  double lik_param;
  const int32_t* coords;
  const double* vals;
  const double* sampler[SSVI_MAX_ORDER];
  int32_t* flags;
};

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ double ldz(const double* p, int64_t i) { return __ldg(p + i); }
__device__ __forceinline__ double ldz(const float* p, int64_t i) { return (double)__ldg(p + i); }

// np.nan_to_num for float64: nan -> 0, +-inf -> +-DBL_MAX
__device__ __forceinline__ double nan_to_num(double x) {
  if (isnan(x)) return 0.0;
  if (isinf(x)) return x > 0 ? DBL_MAX : -DBL_MAX;
  return x;
}
