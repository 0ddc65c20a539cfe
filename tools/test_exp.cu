// Accuracy check of the short-chain likelihood helpers (likelihood.cuh) against libm on the device:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I ssvi_tf_b200/csrc tools/test_exp.cu -o tools/test_exp && tools/test_exp
#include <cstdio>
#include <cmath>
#include <cfloat>
#include <cstdint>
#include "likelihood.cuh"
__global__ void k(int n, double lo, double hi, double* err) {
  double me = 0, mr = 0, mb = 0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const double x = lo + (hi - lo) * ((double)i / (double)n);
    const double e = exp_nonpos(x), r = exp(x);
    me = fmax(me, fabs(e - r) / r);
    const double a = 1.0 + e;
    mr = fmax(mr, fabs(rcp_1to2(a) - 1.0 / a) * a);
    // bernoulli derivs against the reference formula (bernoulli.py:40-51): y sigmoid(-y f), -sigmoid(yf) sigmoid(-yf)
    const double y = (i & 1) ? 1.0 : -1.0, f = -x * ((i & 2) ? 1.0 : -1.0) * 0.05;
    double a1, b1;
    bernoulli_log_derivs(y, f, a1, b1);
    const double sp = 1.0 / (1.0 + exp(-y * f)), sm = 1.0 / (1.0 + exp(y * f));
    mb = fmax(mb, fabs(a1 - y * sm) / fmax(fabs(y * sm), 1e-300));
    mb = fmax(mb, fabs(b1 + sp * sm) / fmax(sp * sm, 1e-300));
  }
  atomicMax((unsigned long long*)&err[0], __double_as_longlong(me));
  atomicMax((unsigned long long*)&err[1], __double_as_longlong(mr));
  atomicMax((unsigned long long*)&err[2], __double_as_longlong(mb));
}
int main() {
  double* d; cudaMalloc(&d, 24);
  const double ranges[4][2] = {{-1.0, 0.0}, {-40.0, 0.0}, {-700.0, 0.0}, {-1e-6, 0.0}};
  int bad = 0;
  for (auto& rg : ranges) {
    cudaMemset(d, 0, 24);
    k<<<296, 256>>>(1 << 24, rg[0], rg[1], d);
    double h[3]; cudaMemcpy(h, d, 24, cudaMemcpyDeviceToHost);
    printf("x in [%g, 0]: exp_nonpos max rel err %.3g, rcp_1to2 max rel err %.3g, bernoulli derivs max rel err %.3g\n", rg[0], h[0], h[1], h[2]);
    if (!(h[0] < 1e-15 && h[1] < 5e-16 && h[2] < 1e-14)) bad = 1;
  }
  printf(bad ? "FAIL\n" : "ok\n");
  return bad;
}
