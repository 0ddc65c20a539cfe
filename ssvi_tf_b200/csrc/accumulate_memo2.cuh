// Sample-table sweep for SSVI_TF_simple / SSVI_TF_robust with a full covariance (D <= 32, k1 == 64): three passes per
// slot group over the same table and chunk tables as accumulate_memo.cuh.
//
// For these models the weights of the contraction depend on (entry, sample): a_bk, w_bk come out of k2 second-layer
// evaluations of the likelihood at f_bk (SSVI_TF_simple.py:23-40; SSVI_TF_robust.py:67-169) -- at BASELINE config 3
// that is 4096 Poisson pdf evaluations per entry, two orders of magnitude more work than the contraction itself.  Fused
// into the gather/contract kernel that loop runs at the occupancy the contraction's registers and shared memory allow
// (one 4-warp CTA per SM in the direct kernel); here it is its own element-wise kernel at full occupancy:
//   memo_f_kernel        gather N-1 table records per entry, V = Hadamard product, f_k = u_k . v_k   -> F[entry][k]
//   memo_weights_kernel  one thread per (entry, k): the k2 loop                                     -> (a, w)[entry][k], si
//   memo_c_kernel        gather again (L2 hits), G += V^T diag(w) V by DMMA, g += V^T a             -> chunk records
// V is re-formed in the third pass instead of stored: 2 * D * k1 multiplies against 8 * D * k1 bytes each way.
#pragma once
#include "accumulate_memo.cuh"

struct MemoWork {
  double* F;        // [n][64]      predictive means
  double2* W;       // [n][64]      (a_bk, w_bk)
  double* SI;       // [n][64]      sigma statistic per (entry, sample)
  double* Y;        // [n]          entry value
  int32_t* SL;      // [n]          noise slot
  int64_t e_lo;     // position of the group's first entry in the slot-major entry list
  int64_t n;        // entries in the group
};
__host__ __device__ inline int64_t memo_work_bytes(int64_t n) { return n * (64 * (8 + 16 + 8) + 8 + 4) + 64; }

// ---- shared prologue: stage own row + noise, U(row, slot) -> Us (fragment order) ----------------------------------------
template <int NT>
__device__ __forceinline__ void memo_own_sample(const DevView& v, unsigned char* smem_raw, int64_t row, int slot,
                                                const void* z_own, int z_f32, int tid, int lane, int warp, int g) {
  constexpr int KT = SSVI_MEMO_KT;
  const int NS = smp_ns(v.D);
  double* S = reinterpret_cast<double*>(smem_raw);
  double* zs = S + v.srow_len;
  double2* Us = reinterpret_cast<double2*>(smem_raw);
  {
    const double2* src = reinterpret_cast<const double2*>(v.sampler[v.dim] + row * (int64_t)v.srow_len);
    double2* dst = reinterpret_cast<double2*>(S);
    for (int i = tid; i < (v.srow_len >> 1); i += 128) dst[i] = src[i];
    if (z_f32) {
      const float* zsrc = reinterpret_cast<const float*>(z_own) + (int64_t)slot * v.zblk;
      for (int i = tid; i < v.zblk; i += 128) zs[i] = (double)zsrc[i];
    } else {
      const double* zsrc = reinterpret_cast<const double*>(z_own) + (int64_t)slot * v.zblk;
      for (int i = tid; i < v.zblk; i += 128) zs[i] = zsrc[i];
    }
  }
  __syncthreads();
  double uc[NT][2][2];
#pragma unroll
  for (int i = 0; i < NT; ++i)
#pragma unroll
    for (int h = 0; h < 2; ++h) mma_sample_tile(S, zs, NT, NS, lane, g, i, warp + 4 * h, uc[i][h][0], uc[i][h][1]);
  __syncthreads();
#pragma unroll
  for (int i = 0; i < NT; ++i)
#pragma unroll
    for (int h = 0; h < 2; ++h) Us[(i * KT + warp + 4 * h) * 32 + lane] = make_double2(uc[i][h][0], uc[i][h][1]);
  __syncthreads();
}

__host__ __device__ inline int memo_f_smem(int NT, int D) { return memo_region_bytes(NT, D) + 4 * SSVI_MEMO_KT * 512; }

// ---------------------------------------------------------------------------------------------------------------------
// pass 1: F[entry][k] = u_k . v_k
// ---------------------------------------------------------------------------------------------------------------------
template <int NT, int NOTH>
__global__ void __launch_bounds__(128, 4) memo_f_kernel(DevView v, MemoView mv, MemoWork wk, int64_t n_chunks,
                                                        const int32_t* __restrict__ chunk_row,
                                                        const int64_t* __restrict__ chunk_start,
                                                        const int32_t* __restrict__ chunk_len,
                                                        const int32_t* __restrict__ chunk_slot,
                                                        const int64_t* __restrict__ entry_ids,
                                                        const int64_t* __restrict__ rows,
                                                        const void* __restrict__ z_own, int z_f32) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr int KT = SSVI_MEMO_KT;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, q = lane & 3;
  const int64_t c = blockIdx.x;
  if (c >= n_chunks) return;
  const int N = v.N;
  const int64_t row = rows[chunk_row[c]];
  const int64_t start = chunk_start[c];
  const int len = chunk_len[c], slot = chunk_slot[c];
  const int sl = slot - mv.slot_lo;
  const double2* Us = reinterpret_cast<const double2*>(smem_raw);
  double2* FP = reinterpret_cast<double2*>(smem_raw + memo_region_bytes(NT, v.D)) + warp * (KT * 32);
  const double2* tab2 = reinterpret_cast<const double2*>(mv.tab);
  const int rec2 = mv.rec_len >> 1;
  const int my_cnt = (len - warp + 3) >> 2;

  memo_own_sample<NT>(v, smem_raw, row, slot, z_own, z_f32, tid, lane, warp, g);

  for (int b0 = 0; b0 < my_cnt; b0 += 32) {
    const int nb = min(32, my_cnt - b0);
    double my_y = 0.0;
    int32_t my_rec[NOTH];
#pragma unroll
    for (int jj = 0; jj < NOTH; ++jj) my_rec[jj] = 0;
    if (lane < nb) {
      const int64_t eid = entry_ids[start + (int64_t)(b0 + lane) * 4 + warp];
      my_y = v.vals[eid];
#pragma unroll
      for (int jj = 0; jj < NOTH; ++jj) my_rec[jj] = (int32_t)memo_rec(mv, jj, sl, v.coords[eid * N + mv.od[jj]]);
    }
    double2 cur[NOTH][NT];
#pragma unroll
    for (int jj = 0; jj < NOTH; ++jj) {
      const double2* p = tab2 + (int64_t)__shfl_sync(0xffffffffu, my_rec[jj], 0) * rec2 + lane;
#pragma unroll
      for (int i = 0; i < NT; ++i) cur[jj][i] = ld_tab(p + (i * KT) * 32);
    }
    for (int el = 0; el < nb; ++el) {
      const double y = __shfl_sync(0xffffffffu, my_y, el);
      const int eln = (el + 1 < nb) ? el + 1 : el;
      const double2* pc[NOTH];
      const double2* pn[NOTH];
#pragma unroll
      for (int jj = 0; jj < NOTH; ++jj) {
        pc[jj] = tab2 + (int64_t)__shfl_sync(0xffffffffu, my_rec[jj], el) * rec2 + lane;
        pn[jj] = tab2 + (int64_t)__shfl_sync(0xffffffffu, my_rec[jj], eln) * rec2 + lane;
      }
#pragma unroll
      for (int t = 0; t < KT; ++t) {
        double2 nxt[NOTH][NT];
#pragma unroll
        for (int jj = 0; jj < NOTH; ++jj) {
          const double2* p = (t + 1 < KT) ? pc[jj] + (t + 1) * 32 : pn[jj];
#pragma unroll
          for (int i = 0; i < NT; ++i) nxt[jj][i] = ld_tab(p + (i * KT) * 32);
        }
        double2 fpt = make_double2(0.0, 0.0);
#pragma unroll
        for (int i = 0; i < NT; ++i) {
          double2 x = cur[0][i];
#pragma unroll
          for (int jj = 1; jj < NOTH; ++jj) {
            x.x *= cur[jj][i].x;
            x.y *= cur[jj][i].y;
          }
          const double2 u = Us[(i * KT + t) * 32 + lane];
          fpt.x = fma(u.x, x.x, fpt.x);
          fpt.y = fma(u.y, x.y, fpt.y);
        }
        FP[t * 32 + (lane ^ (t << 2))] = fpt;
#pragma unroll
        for (int jj = 0; jj < NOTH; ++jj)
#pragma unroll
          for (int i = 0; i < NT; ++i) cur[jj][i] = nxt[jj][i];
      }
      __syncwarp();
      double2 a = FP[g * 32 + ((g << 2) | q)];
#pragma unroll
      for (int gg = 1; gg < 8; ++gg) {
        const double2 b = FP[g * 32 + (((gg ^ g) << 2) | q)];
        a.x += b.x;
        a.y += b.y;
      }
      __syncwarp();
      const int64_t loc = start + (int64_t)(b0 + el) * 4 + warp - wk.e_lo;
      *reinterpret_cast<double2*>(wk.F + loc * 64 + 8 * g + 2 * q) = a;     // samples k = 8g + 2q + {0,1}
      if (lane == 0) {
        wk.Y[loc] = y;
        wk.SL[loc] = slot;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// pass 2: one thread per (entry, sample): the second-layer k2 loop
// ---------------------------------------------------------------------------------------------------------------------
#ifndef SSVI_W_MINB
#define SSVI_W_MINB 3                                   // resident 256-thread blocks per SM asked of ptxas (A/B knob)
#endif
template <typename ZT, int MODEL, int LIK>
__global__ void __launch_bounds__(256, SSVI_W_MINB) memo_weights_kernel(DevView v, MemoWork wk, const double* __restrict__ w_state,
                                                              const ZT* __restrict__ r_w, const ZT* __restrict__ z_f) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= wk.n * 64) return;
  const int64_t loc = t >> 6;
  const int k = (int)(t & 63);
  const double y = wk.Y[loc];
  const int slot = wk.SL[loc];
  const double yfact = (LIK == SSVI_LIK_POISSON && MODEL == SSVI_MODEL_ROBUST) ? 1.0 / factorial_ref(y) : 1.0;
  double aw, dw, si = 0.0;
  bool bad = false;
  likelihood_weights<ZT, MODEL, LIK>(v, r_w, z_f, slot, k, true, y, yfact, wk.F[t], w_state[0], aw, dw, si, bad);
  wk.W[t] = make_double2(aw, dw);
  wk.SI[t] = si;
  if (bad) atomicOr(v.flags, SSVI_FLAG_PHI_ZERO);
}

// the instance for the handle's (model, likelihood)
template <typename ZT>
static void launch_memo_weights(const DevView& v, const MemoWork& wk, const double* w_state, const ZT* r_w, const ZT* z_f,
                                cudaStream_t st) {
  const unsigned grid = (unsigned)((wk.n * 64 + 255) / 256);
#define SSVI_W_CASE(M, L)                                                                                  \
  if (v.model == M && v.lik == L) {                                                                        \
    memo_weights_kernel<ZT, M, L><<<grid, 256, 0, st>>>(v, wk, w_state, r_w, z_f);                         \
    return;                                                                                                \
  }
  SSVI_W_CASE(SSVI_MODEL_ROBUST, SSVI_LIK_POISSON)
  SSVI_W_CASE(SSVI_MODEL_ROBUST, SSVI_LIK_BERNOULLI)
  SSVI_W_CASE(SSVI_MODEL_ROBUST, SSVI_LIK_NORMAL)
  SSVI_W_CASE(SSVI_MODEL_SIMPLE, SSVI_LIK_POISSON)
  SSVI_W_CASE(SSVI_MODEL_SIMPLE, SSVI_LIK_BERNOULLI)
  SSVI_W_CASE(SSVI_MODEL_SIMPLE, SSVI_LIK_NORMAL)
#undef SSVI_W_CASE
}

__host__ __device__ inline int memo_c_smem(int NT, int D) { return 4 * (NT * (NT + 1) / 2) * 512 + 4 * (NT * 8 + 1) * 8; }

// ---------------------------------------------------------------------------------------------------------------------
// pass 3: G += V^T diag(w) V, g += V^T a with the stored weights; the DMMA accumulators live across the whole chunk
// ---------------------------------------------------------------------------------------------------------------------
template <int NT, int NOTH>
__global__ void __launch_bounds__(128, 3) memo_c_kernel(DevView v, MemoView mv, MemoWork wk, int64_t n_chunks,
                                                        const int32_t* __restrict__ chunk_row,
                                                        const int64_t* __restrict__ chunk_start,
                                                        const int32_t* __restrict__ chunk_len,
                                                        const int32_t* __restrict__ chunk_slot,
                                                        const int64_t* __restrict__ chunk_out,
                                                        const int64_t* __restrict__ entry_ids,
                                                        double* __restrict__ partial_out) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr int NP = NT * (NT + 1) / 2, KT = SSVI_MEMO_KT;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, q = lane & 3;
  const int64_t c = blockIdx.x;
  if (c >= n_chunks) return;
  const int N = v.N, D = v.D;
  const int64_t start = chunk_start[c];
  const int len = chunk_len[c], slot = chunk_slot[c];
  const int sl = slot - mv.slot_lo;
  const double2* tab2 = reinterpret_cast<const double2*>(mv.tab);
  const int rec2 = mv.rec_len >> 1;
  const int my_cnt = (len - warp + 3) >> 2;

  double acc[NP][2];
#pragma unroll
  for (int p = 0; p < NP; ++p) acc[p][0] = acc[p][1] = 0.0;
  double ga[NT];
#pragma unroll
  for (int i = 0; i < NT; ++i) ga[i] = 0.0;
  double si_part = 0.0;

  for (int b0 = 0; b0 < my_cnt; b0 += 32) {
    const int nb = min(32, my_cnt - b0);
    int32_t my_rec[NOTH];
#pragma unroll
    for (int jj = 0; jj < NOTH; ++jj) my_rec[jj] = 0;
    if (lane < nb) {
      const int64_t eid = entry_ids[start + (int64_t)(b0 + lane) * 4 + warp];
#pragma unroll
      for (int jj = 0; jj < NOTH; ++jj) my_rec[jj] = (int32_t)memo_rec(mv, jj, sl, v.coords[eid * N + mv.od[jj]]);
    }
    double2 cur[NOTH][NT];
#pragma unroll
    for (int jj = 0; jj < NOTH; ++jj) {
      const double2* p = tab2 + (int64_t)__shfl_sync(0xffffffffu, my_rec[jj], 0) * rec2 + lane;
#pragma unroll
      for (int i = 0; i < NT; ++i) cur[jj][i] = ld_tab(p + (i * KT) * 32);
    }
    for (int el = 0; el < nb; ++el) {
      const int eln = (el + 1 < nb) ? el + 1 : el;
      const double2* pc[NOTH];
      const double2* pn[NOTH];
#pragma unroll
      for (int jj = 0; jj < NOTH; ++jj) {
        pc[jj] = tab2 + (int64_t)__shfl_sync(0xffffffffu, my_rec[jj], el) * rec2 + lane;
        pn[jj] = tab2 + (int64_t)__shfl_sync(0xffffffffu, my_rec[jj], eln) * rec2 + lane;
      }
      const int64_t loc = start + (int64_t)(b0 + el) * 4 + warp - wk.e_lo;
      const double2* Wp = wk.W + loc * 64 + 2 * q;       // samples 8t + 2q + {0,1} of k-tile t
      {
        const double2 s2 = *reinterpret_cast<const double2*>(wk.SI + loc * 64 + 2 * lane);
        si_part += s2.x + s2.y;
      }
      double2 w0 = __ldcg(Wp), w1 = __ldcg(Wp + 1);
#pragma unroll
      for (int t = 0; t < KT; ++t) {
        double2 nxt[NOTH][NT];
#pragma unroll
        for (int jj = 0; jj < NOTH; ++jj) {
          const double2* p = (t + 1 < KT) ? pc[jj] + (t + 1) * 32 : pn[jj];
#pragma unroll
          for (int i = 0; i < NT; ++i) nxt[jj][i] = ld_tab(p + (i * KT) * 32);
        }
        double2 n0 = w0, n1 = w1;
        if (t + 1 < KT) {
          n0 = __ldcg(Wp + 8 * (t + 1));
          n1 = __ldcg(Wp + 8 * (t + 1) + 1);
        }
        double2 vv[NT], wv[NT];
#pragma unroll
        for (int i = 0; i < NT; ++i) {
          vv[i] = cur[0][i];
#pragma unroll
          for (int jj = 1; jj < NOTH; ++jj) {
            vv[i].x *= cur[jj][i].x;
            vv[i].y *= cur[jj][i].y;
          }
          wv[i].x = w0.y * vv[i].x;                      // w_k * v
          wv[i].y = w1.y * vv[i].y;
          ga[i] = fma(w0.x, vv[i].x, ga[i]);             // a_k * v
          ga[i] = fma(w1.x, vv[i].y, ga[i]);
        }
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          int p = 0;
#pragma unroll
          for (int i1 = 0; i1 < NT; ++i1)
#pragma unroll
            for (int i2 = 0; i2 <= i1; ++i2) {
              dmma_m8n8k4(acc[p][0], acc[p][1], s ? vv[i1].y : vv[i1].x, s ? wv[i2].y : wv[i2].x);
              ++p;
            }
        }
#pragma unroll
        for (int jj = 0; jj < NOTH; ++jj)
#pragma unroll
          for (int i = 0; i < NT; ++i) cur[jj][i] = nxt[jj][i];
        w0 = n0;
        w1 = n1;
      }
    }
  }

  // ---- CTA reduction in fixed warp order; chunk record [di | Di | si_sum | count] -------------------------------------
  double2* Gs = reinterpret_cast<double2*>(smem_raw) + warp * (NP * 32);
  double* red = reinterpret_cast<double*>(smem_raw + 4 * NP * 512);
#pragma unroll
  for (int p = 0; p < NP; ++p) Gs[p * 32 + lane] = make_double2(acc[p][0], acc[p][1]);
#pragma unroll
  for (int i = 0; i < NT; ++i) {
    double x = ga[i];
    x += __shfl_xor_sync(0xffffffffu, x, 1);
    x += __shfl_xor_sync(0xffffffffu, x, 2);
    if (q == 0) red[warp * (NT * 8 + 1) + 8 * i + g] = x;
  }
  si_part = warp_sum(si_part);
  if (lane == 0) red[warp * (NT * 8 + 1) + NT * 8] = si_part;
  __syncthreads();
  double* out = partial_out + chunk_out[c] * (int64_t)v.acc_len;
  constexpr int RS = NT * 8 + 1;
  for (int n = tid; n < D; n += 128) out[n] = ((red[n] + red[RS + n]) + red[2 * RS + n]) + red[3 * RS + n];
  const double2* Gall = reinterpret_cast<const double2*>(smem_raw);
  for (int idx = tid; idx < NP * 32; idx += 128) {
    const int p = idx >> 5, l = idx & 31;
    const double2 a0 = Gall[idx], a1 = Gall[NP * 32 + idx], a2 = Gall[2 * NP * 32 + idx], a3 = Gall[3 * NP * 32 + idx];
    const double x0 = ((a0.x + a1.x) + a2.x) + a3.x, x1 = ((a0.y + a1.y) + a2.y) + a3.y;
    int i1 = 0;
    while ((i1 + 1) * (i1 + 2) / 2 <= p) ++i1;
    const int i2 = p - i1 * (i1 + 1) / 2;
    const int n1 = 8 * i1 + (l >> 2), n2 = 8 * i2 + 2 * (l & 3);
    if (n1 < D && n2 <= n1) {
      out[D + n1 * D + n2] = x0;
      out[D + n2 * D + n1] = x0;
    }
    if (n1 < D && n2 + 1 <= n1) {
      out[D + n1 * D + n2 + 1] = x1;
      out[D + (n2 + 1) * D + n1] = x1;
    }
  }
  if (tid == 0) {
    out[D + D * D] = ((red[NT * 8] + red[RS + NT * 8]) + red[2 * RS + NT * 8]) + red[3 * RS + NT * 8];
    out[D + D * D + 1] = (double)len;
  }
}
