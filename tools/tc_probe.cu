// Probe + micro-benchmark of the tcgen05 contraction that accumulate_tc.cuh is built on ("next round" #6):
// the chunk GEMM  G = SUM_b w_b V_b V_b^T  with V_b = X_j o X_l formed in fp32 from two gathered table records, split into
// fp16 hi / lo pieces, three tcgen05.mma passes (hi hi^T + hi lo^T + lo hi^T) per 16-sample K-step into an fp32
// accumulator in tensor memory, read back with tcgen05.ld and folded into registers with the entry's weight.
//   1. correctness against a double-precision CPU evaluation, per pass / per K-step (which isolates descriptor mistakes);
//   2. throughput of the whole per-entry pipeline and of its parts (no MMA, hi hi^T only, gather only).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -o tools/tc_probe tools/tc_probe.cu
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cuda_runtime.h>
#include "../ssvi_tf_b200/csrc/tc_ops.cuh"

#define CK(x)                                                                              \
  do {                                                                                     \
    cudaError_t e_ = (x);                                                                  \
    if (e_ != cudaSuccess) {                                                               \
      printf("CUDA error %s at %s:%d: %s\n", cudaGetErrorName(e_), __FILE__, __LINE__, cudaGetErrorString(e_)); \
      exit(2);                                                                             \
    }                                                                                      \
  } while (0)

struct ProbeArgs {
  const float4* tab;      // records [R][NT*2*2*32] float4 (tc_rec_index)
  const int32_t* ent;     // [n_cta][len][2] record ids
  const float* y;         // [n_cta][len]
  const int32_t* urec;    // [n_cta] record used as the own-row sample
  float* G;               // [n_cta][32][32]
  float* di;              // [n_cta][32]
  float* raw;             // [128][128] accumulator of CTA 0, quad 0 (or null)
  int32_t len;
  int32_t passes;         // bit 0 hi hi^T, bit 1 hi lo^T, bit 2 lo hi^T
  int32_t ksteps;         // bit s: K-step s issued
  int32_t swap;           // descriptor hypothesis: LBO and SBO exchanged
  int32_t no_tiles;       // timing: skip split + tile stores (and the MMAs)
  float alpha;
};

constexpr int PROBE_SMEM = 4 * TC_TILE_BYTES + 1024;

template <int NT>
__global__ void __launch_bounds__(128, 2) tc_probe_kernel(ProbeArgs a) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int r8 = lane & 7, kq = lane >> 3;
  unsigned char* tiles = smem;                                     // [buf][piece][16 KB]
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + 4 * TC_TILE_BYTES);
  uint64_t* done = full + 2;
  uint32_t* holder = reinterpret_cast<uint32_t*>(done + 2);
  const int len = a.len, nq = (len + 3) >> 2;
  const int64_t cta = blockIdx.x;
  constexpr int REC4 = NT * 2 * 2 * 32;

  for (int i = tid; i < 4 * TC_TILE_BYTES / 16; i += 128) reinterpret_cast<uint4*>(tiles)[i] = make_uint4(0, 0, 0, 0);
  if (tid == 0) {
    tc_mbar_init(&full[0], 4);
    tc_mbar_init(&full[1], 4);
    tc_mbar_init(&done[0], 1);
    tc_mbar_init(&done[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  if (warp == 0) tc_tmem_alloc(holder, 256);
  tc_fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *holder;

  float u[2 * NT][8];
  {
    const float4* p = a.tab + (int64_t)a.urec[cta] * REC4 + lane;
#pragma unroll
    for (int it = 0; it < 2 * NT; ++it) {
      const float4 x0 = __ldg(p + (it * 2) * 32), x1 = __ldg(p + (it * 2 + 1) * 32);
      u[it][0] = x0.x; u[it][1] = x0.y; u[it][2] = x0.z; u[it][3] = x0.w;
      u[it][4] = x1.x; u[it][5] = x1.y; u[it][6] = x1.z; u[it][7] = x1.w;
    }
  }
  float gacc[32], dacc[NT];
#pragma unroll
  for (int c = 0; c < 32; ++c) gacc[c] = 0.f;
#pragma unroll
  for (int i = 0; i < NT; ++i) dacc[i] = 0.f;
  const uint32_t idesc = tc_idesc_f16(128, 128);
  const uint32_t lbo = a.swap ? TC_SBO : TC_LBO, sbo = a.swap ? TC_LBO : TC_SBO;
  float w_prev = 0.f;
  bool prev_valid = false;

  for (int q = 0; q <= nq; ++q) {
    const int buf = q & 1;
    float w_b = 0.f;
    bool valid = false;
    if (q < nq) {
      const int b = 4 * q + warp;
      valid = b < len;
      unsigned char* thi = tiles + (buf * 2) * TC_TILE_BYTES;
      unsigned char* tlo = thi + TC_TILE_BYTES;
      if (valid) {
        const int32_t rj = a.ent[(cta * len + b) * 2], rl = a.ent[(cta * len + b) * 2 + 1];
        const float yb = a.y[cta * len + b];
        const float4* pj = a.tab + (int64_t)rj * REC4 + lane;
        const float4* pl = a.tab + (int64_t)rl * REC4 + lane;
        float fp[2][8], rs[NT];
#pragma unroll
        for (int j = 0; j < 8; ++j) fp[0][j] = fp[1][j] = 0.f;
        float4 c0 = __ldcg(pj), c1 = __ldcg(pj + 32), e0 = __ldcg(pl), e1 = __ldcg(pl + 32);
#pragma unroll
        for (int it = 0; it < 2 * NT; ++it) {
          const int rg = it >> 1, h = it & 1;
          float4 n0 = c0, n1 = c1, m0 = e0, m1 = e1;
          if (it + 1 < 2 * NT) {
            n0 = __ldcg(pj + ((it + 1) * 2) * 32); n1 = __ldcg(pj + ((it + 1) * 2 + 1) * 32);
            m0 = __ldcg(pl + ((it + 1) * 2) * 32); m1 = __ldcg(pl + ((it + 1) * 2 + 1) * 32);
          }
          float v[8] = {c0.x * e0.x, c0.y * e0.y, c0.z * e0.z, c0.w * e0.w, c1.x * e1.x, c1.y * e1.y, c1.z * e1.z, c1.w * e1.w};
          float s = 0.f;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            fp[h][j] = fmaf(u[it][j], v[j], fp[h][j]);
            s += v[j];
          }
          if (h == 0) rs[rg] = s; else rs[rg] += s;
          if (!a.no_tiles) {
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] *= a.alpha;
            tc_store_split8(thi, tlo, 32 * warp + 8 * rg + r8, kq + 4 * h, v);
          }
          c0 = n0; c1 = n1; e0 = m0; e1 = m1;
        }
        // f_k = SUM_d u v: in-lane over the row groups (done), then over the 8 lanes r8 by a transposing reduction
        const bool hb = (lane >> 2) & 1, jb = (lane >> 1) & 1, jc = lane & 1;
        float s8[8], s4[4], f[2];
#pragma unroll
        for (int j = 0; j < 8; ++j) s8[j] = (hb ? fp[1][j] : fp[0][j]) + __shfl_xor_sync(0xffffffffu, hb ? fp[0][j] : fp[1][j], 4);
#pragma unroll
        for (int j = 0; j < 4; ++j) s4[j] = (jb ? s8[4 + j] : s8[j]) + __shfl_xor_sync(0xffffffffu, jb ? s8[j] : s8[4 + j], 2);
#pragma unroll
        for (int j = 0; j < 2; ++j) f[j] = (jc ? s4[2 + j] : s4[j]) + __shfl_xor_sync(0xffffffffu, jc ? s4[j] : s4[2 + j], 1);
        float fs = f[0] + f[1];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) fs += __shfl_xor_sync(0xffffffffu, fs, o);
        const float mf = fs * (1.f / 64.f);
        const float a_b = yb - mf;
        w_b = -1.f / (1.f + mf * mf);
#pragma unroll
        for (int i = 0; i < NT; ++i) {
          float s = rs[i];
          s += __shfl_xor_sync(0xffffffffu, s, 8);
          s += __shfl_xor_sync(0xffffffffu, s, 16);
          dacc[i] = fmaf(a_b, s, dacc[i]);
        }
      }
      tc_fence_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) tc_mbar_arrive(&full[buf]);
      if (warp == 0) {
        tc_mbar_wait(&full[buf], (q >> 1) & 1);
        tc_fence_after();
        if (lane == 0) {
          if (!a.no_tiles) {
            const uint32_t ahi = tc_smem_u32(thi), alo = tc_smem_u32(tlo);
            uint32_t acc = 0;
#pragma unroll
            for (int s = 0; s < 4; ++s) {
              if (!((a.ksteps >> s) & 1)) continue;
              const uint64_t dh = tc_smem_desc(ahi + 256 * s, lbo, sbo), dl = tc_smem_desc(alo + 256 * s, lbo, sbo);
              if (a.passes & 1) { tc_mma_f16(tmem + buf * 128, dh, dh, idesc, acc); acc = 1; }
              if (a.passes & 2) { tc_mma_f16(tmem + buf * 128, dh, dl, idesc, acc); acc = 1; }
              if (a.passes & 4) { tc_mma_f16(tmem + buf * 128, dl, dh, idesc, acc); acc = 1; }
            }
          }
          tc_commit(&done[buf]);
        }
        __syncwarp();
      }
    }
    if (q > 0) {                                        // fold the previous quad's accumulator block into registers
      const int pb = buf ^ 1;
      tc_mbar_wait(&done[pb], ((q - 1) >> 1) & 1);
      tc_fence_after();
      if (!a.no_tiles) {
        float d[32];
        tc_tmem_ld32(tmem + ((uint32_t)(32 * warp) << 16) + pb * 128 + 32 * warp, d);
        if (prev_valid) {
#pragma unroll
          for (int c = 0; c < 32; ++c) gacc[c] = fmaf(w_prev, d[c], gacc[c]);
        }
        if (a.raw != nullptr && cta == 0 && q == 1) {
#pragma unroll
          for (int cb = 0; cb < 4; ++cb) {
            tc_tmem_ld32(tmem + ((uint32_t)(32 * warp) << 16) + pb * 128 + 32 * cb, d);
#pragma unroll
            for (int c = 0; c < 32; ++c) a.raw[(32 * warp + lane) * 128 + 32 * cb + c] = d[c];
          }
        }
      }
      tc_fence_before();
    }
    w_prev = w_b;
    prev_valid = valid;
  }

  __syncthreads();                                      // tiles are dead: reuse them for the CTA reduction
  float* red = reinterpret_cast<float*>(tiles);         // [4][32][33]
#pragma unroll
  for (int c = 0; c < 32; ++c) red[(warp * 32 + lane) * 33 + c] = gacc[c];
  float* dred = red + 4 * 32 * 33;                      // [4][32]
  if (kq == 0) {
#pragma unroll
    for (int i = 0; i < NT; ++i) dred[warp * 32 + 8 * i + r8] = dacc[i];
  }
  __syncthreads();
  const float ia2 = 1.f / (a.alpha * a.alpha);
  for (int i = tid; i < 32 * 32; i += 128) {
    const int d = i >> 5, c = i & 31;
    a.G[cta * 1024 + i] = ia2 * (((red[(0 * 32 + d) * 33 + c] + red[(1 * 32 + d) * 33 + c]) + red[(2 * 32 + d) * 33 + c]) +
                                 red[(3 * 32 + d) * 33 + c]);
  }
  if (tid < 8 * NT) a.di[cta * 32 + tid] = ((dred[tid] + dred[32 + tid]) + dred[64 + tid]) + dred[96 + tid];
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tc_tmem_dealloc(tmem, 256);
}

// the fp64 contraction alone, as accumulate_memo_kernel issues it at NT = 4 (160 DMMAs per entry per warp): comparator
__global__ void __launch_bounds__(128, 3) dmma_contract_kernel(double* out, int entries_per_warp, double seed) {
  double acc[10][2];
#pragma unroll
  for (int p = 0; p < 10; ++p) acc[p][0] = acc[p][1] = 0.0;
  double v[4] = {seed + threadIdx.x * 1e-9, seed * 0.5, seed * 0.25, 1.0 - seed};
  for (int e = 0; e < entries_per_warp; ++e) {
#pragma unroll
    for (int t = 0; t < 16; ++t) {
      int p = 0;
#pragma unroll
      for (int i1 = 0; i1 < 4; ++i1)
#pragma unroll
        for (int i2 = 0; i2 <= i1; ++i2) {
          asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                       : "+d"(acc[p][0]), "+d"(acc[p][1])
                       : "d"(v[i1]), "d"(v[i2]));
          ++p;
        }
    }
  }
  double s = 0.0;
#pragma unroll
  for (int p = 0; p < 10; ++p) s += acc[p][0] + acc[p][1];
  out[blockIdx.x * 128 + threadIdx.x] = s;
}

static float frand(uint64_t& s) {
  s = s * 6364136223846793005ull + 1442695040888963407ull;
  return (float)((s >> 33) & 0xffffff) / 16777216.f;
}

int main() {
  constexpr int NT = 4, D = 32, K = 64, REC = NT * 8 * K;
  const int R = 8192;
  uint64_t seed = 12345;
  std::vector<float> X((size_t)R * D * K), tab((size_t)R * REC);
  for (auto& x : X) x = 3.f * frand(seed) - 1.5f;
  for (int r = 0; r < R; ++r)
    for (int d = 0; d < D; ++d)
      for (int k = 0; k < K; ++k) tab[(size_t)r * REC + tc_rec_index(d, k)] = X[((size_t)r * D + d) * K + k];
  float* d_tab;
  CK(cudaMalloc(&d_tab, tab.size() * 4));
  CK(cudaMemcpy(d_tab, tab.data(), tab.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaFuncSetAttribute(tc_probe_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, PROBE_SMEM));

  auto run = [&](int n_cta, int len, int passes, int ksteps, int swap, int no_tiles, bool want_raw, std::vector<int32_t>& ent,
                 std::vector<float>& y, std::vector<int32_t>& urec, std::vector<float>& G, std::vector<float>& di,
                 std::vector<float>& raw, int reps, float* ms_out) {
    int32_t *d_ent, *d_urec;
    float *d_y, *d_G, *d_di, *d_raw = nullptr;
    CK(cudaMalloc(&d_ent, ent.size() * 4));
    CK(cudaMalloc(&d_urec, urec.size() * 4));
    CK(cudaMalloc(&d_y, y.size() * 4));
    CK(cudaMalloc(&d_G, (size_t)n_cta * 1024 * 4));
    CK(cudaMalloc(&d_di, (size_t)n_cta * 32 * 4));
    if (want_raw) CK(cudaMalloc(&d_raw, 128 * 128 * 4));
    CK(cudaMemcpy(d_ent, ent.data(), ent.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_urec, urec.data(), urec.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_y, y.data(), y.size() * 4, cudaMemcpyHostToDevice));
    ProbeArgs a{reinterpret_cast<const float4*>(d_tab), d_ent, d_y, d_urec, d_G, d_di, d_raw, len, passes, ksteps, swap, no_tiles, 64.f};
    tc_probe_kernel<NT><<<n_cta, 128, PROBE_SMEM>>>(a);
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    if (reps > 0) {
      cudaEvent_t e0, e1;
      CK(cudaEventCreate(&e0));
      CK(cudaEventCreate(&e1));
      CK(cudaEventRecord(e0));
      for (int i = 0; i < reps; ++i) tc_probe_kernel<NT><<<n_cta, 128, PROBE_SMEM>>>(a);
      CK(cudaEventRecord(e1));
      CK(cudaEventSynchronize(e1));
      float ms;
      CK(cudaEventElapsedTime(&ms, e0, e1));
      *ms_out = ms / reps;
    }
    G.resize((size_t)n_cta * 1024);
    di.resize((size_t)n_cta * 32);
    CK(cudaMemcpy(G.data(), d_G, G.size() * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(di.data(), d_di, di.size() * 4, cudaMemcpyDeviceToHost));
    if (want_raw) {
      raw.resize(128 * 128);
      CK(cudaMemcpy(raw.data(), d_raw, raw.size() * 4, cudaMemcpyDeviceToHost));
      cudaFree(d_raw);
    }
    cudaFree(d_ent); cudaFree(d_urec); cudaFree(d_y); cudaFree(d_G); cudaFree(d_di);
  };
  auto make = [&](int n_cta, int len, std::vector<int32_t>& ent, std::vector<float>& y, std::vector<int32_t>& urec) {
    ent.resize((size_t)n_cta * len * 2);
    y.resize((size_t)n_cta * len);
    urec.resize(n_cta);
    for (auto& e : ent) e = (int32_t)(frand(seed) * R) % R;
    for (auto& v : y) v = 2.f * frand(seed) - 1.f;
    for (auto& v : urec) v = (int32_t)(frand(seed) * R) % R;
  };

  // ---- 1. correctness ---------------------------------------------------------------------------------------------------
  {
    const int n_cta = 8, len = 10;
    std::vector<int32_t> ent, urec;
    std::vector<float> y, G, di, raw;
    make(n_cta, len, ent, y, urec);
    auto Vof = [&](int cta, int b, int d, int k) {
      const int rj = ent[((size_t)cta * len + b) * 2], rl = ent[((size_t)cta * len + b) * 2 + 1];
      return X[((size_t)rj * D + d) * K + k] * X[((size_t)rl * D + d) * K + k];          // fp32 product, as the kernel forms it
    };
    struct Case { const char* name; int passes, ksteps, swap; };
    const Case cases[] = {{"all passes, all K-steps", 7, 15, 0}, {"hi hi^T, all K-steps", 1, 15, 0}, {"hi hi^T, K-step 0", 1, 1, 0},
                          {"hi hi^T, K-step 1", 1, 2, 0},        {"hi lo^T, all K-steps", 2, 15, 0}, {"lo hi^T, all K-steps", 4, 15, 0},
                          {"hi hi^T, K-step 0, LBO<->SBO", 1, 1, 1}};
    printf("== raw accumulator of CTA 0, quad 0 (128 x 128, rows/cols 32 w + d <-> entry w) vs CPU ==\n");
    for (const Case& c : cases) {
      float ms;
      run(n_cta, len, c.passes, c.ksteps, c.swap, 0, true, ent, y, urec, G, di, raw, 0, &ms);
      double emax = 0.0, smax = 0.0, eblk = 0.0;
      for (int i = 0; i < 128; ++i)
        for (int j = 0; j < 128; ++j) {
          double ref = 0.0;
          for (int k = 0; k < K; ++k) {
            if (!((c.ksteps >> (k >> 4)) & 1)) continue;
            const float ai = 64.f * Vof(0, i >> 5, i & 31, k), bj = 64.f * Vof(0, j >> 5, j & 31, k);
            const float ah = __half2float(__float2half_rn(ai)), bh = __half2float(__float2half_rn(bj));
            const float al = __half2float(__float2half_rn(ai - ah)), bl = __half2float(__float2half_rn(bj - bh));
            if (c.passes & 1) ref += (double)ah * bh;
            if (c.passes & 2) ref += (double)ah * bl;
            if (c.passes & 4) ref += (double)al * bh;
          }
          const double e = fabs(ref - raw[i * 128 + j]);
          if (e > emax) emax = e;
          if ((i >> 5) == (j >> 5) && e > eblk) eblk = e;
          if (fabs(ref) > smax) smax = fabs(ref);
        }
      printf("  %-32s max|err| %.3e (diagonal blocks %.3e)  max|ref| %.3e  -> rel %.2e %s\n", c.name, emax, eblk, smax, emax / smax,
             emax / smax < 1e-5 ? "OK" : "MISMATCH");
    }
    // final G / di against the exact (double) value of the fp32-formed V
    float ms;
    run(n_cta, len, 7, 15, 0, 0, false, ent, y, urec, G, di, raw, 0, &ms);
    double eg = 0.0, sg = 0.0, ed = 0.0, sd = 0.0;
    for (int cta = 0; cta < n_cta; ++cta) {
      std::vector<double> Gr(1024, 0.0), dr(32, 0.0);
      for (int b = 0; b < len; ++b) {
        double mf = 0.0;
        for (int k = 0; k < K; ++k)
          for (int d = 0; d < D; ++d) mf += (double)X[((size_t)urec[cta] * D + d) * K + k] * Vof(cta, b, d, k);
        mf /= 64.0;
        const double a_b = y[(size_t)cta * len + b] - mf, w_b = -1.0 / (1.0 + mf * mf);
        for (int d = 0; d < D; ++d) {
          double s = 0.0;
          for (int k = 0; k < K; ++k) s += Vof(cta, b, d, k);
          dr[d] += a_b * s;
          for (int e = 0; e < D; ++e) {
            double g = 0.0;
            for (int k = 0; k < K; ++k) g += (double)Vof(cta, b, d, k) * Vof(cta, b, e, k);
            Gr[d * 32 + e] += w_b * g;
          }
        }
      }
      for (int i = 0; i < 1024; ++i) {
        eg = fmax(eg, fabs(Gr[i] - G[(size_t)cta * 1024 + i]));
        sg = fmax(sg, fabs(Gr[i]));
      }
      for (int i = 0; i < 32; ++i) {
        ed = fmax(ed, fabs(dr[i] - di[(size_t)cta * 32 + i]));
        sd = fmax(sd, fabs(dr[i]));
      }
    }
    printf("== chunk result (8 CTAs x 10 entries, ragged last quad) vs double: G rel %.2e  di rel %.2e %s\n", eg / sg, ed / sd,
           (eg / sg < 1e-5 && ed / sd < 1e-5) ? "OK" : "MISMATCH");
  }

  // ---- 2. throughput ------------------------------------------------------------------------------------------------------
  {
    int dev = 0, sms = 0, khz = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    CK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev));
    const int n_cta = sms * 2 * 8, len = 128;
    std::vector<int32_t> ent, urec;
    std::vector<float> y, G, di, raw;
    make(n_cta, len, ent, y, urec);
    struct T { const char* name; int passes, no_tiles; };
    const T ts[] = {{"gather + Hadamard + f only (no tiles, no MMA)", 0, 1}, {"+ split + tile stores (no MMA)", 0, 0},
                    {"+ hi hi^T (1 pass: 4 MMA / quad)", 1, 0}, {"+ hi lo^T (2 passes: 8 MMA / quad)", 3, 0},
                    {"all 3 passes (12 MMA / quad)", 7, 0}};
    printf("== throughput: %d CTAs x %d entries, table %d records (%.0f MB, L2), %d SMs @ %d MHz ==\n", n_cta, len, R,
           R * REC * 4 / 1048576.0, sms, khz / 1000);
    for (const T& t : ts) {
      float ms = 0.f;
      run(n_cta, len, t.passes, 15, 0, t.no_tiles, false, ent, y, urec, G, di, raw, 20, &ms);
      const double eps = (double)n_cta * len / (ms * 1e-3);
      printf("  %-48s %.3f ms  %.3e entries/s  (%.0f SM-cycles per entry at the max clock)\n", t.name, ms, eps,
             (double)sms * khz * 1e3 / eps);
    }
    double* d_out;
    const int epw = 64, nb = sms * 3 * 4;
    CK(cudaMalloc(&d_out, (size_t)nb * 128 * 8));
    dmma_contract_kernel<<<nb, 128>>>(d_out, epw, 0.3);
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    CK(cudaEventRecord(e0));
    for (int i = 0; i < 10; ++i) dmma_contract_kernel<<<nb, 128>>>(d_out, epw, 0.3);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    ms /= 10;
    const double eps = (double)nb * 4 * epw / (ms * 1e-3);
    printf("  %-48s %.3f ms  %.3e entries/s  (%.0f SM-cycles per entry)\n", "fp64 contraction alone: 160 DMMA / entry (NT = 4)", ms, eps,
           (double)sms * khz * 1e3 / eps);
  }
  return 0;
}
