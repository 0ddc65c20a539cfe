// The fp32 arithmetic option (SSVI_TABLE_F32_TC; BASELINE.json: 1e-4 in fp32) for FULL SWEEPS with a DIAGONAL covariance,
// SSVI_TF_simple + normal likelihood (BASELINE config 4): accumulate_diag_lin_kernel (accumulate_diag.cuh) with every
// per-sample quantity in float.  Same slot-major chunks, same chunk record, same structure -- one CTA per chunk, the
// chunk's N noise blocks staged once in shared memory ([k][DP], now floats: half the bytes, so three CTAs fit an SM
// instead of two), four entries per warp step sharing every noise read, one select-free transposing reduction per
// 4 entries x 4 samples, closed-form weights
//     w_k = (y - f_k - ws_k * mean_i z_f[i][k]) / (s^2 k1)            (SSVI_TF_simple.py:35-38 with normal.py:32-36)
// handed round the warp through shared memory -- on the FP32 pipe, with a lane's two components packed into f32x2
// operands (FFMA2 / FMUL2: Blackwell's two-wide fp32 instructions halve the instruction count of the sample loop), and with
// half the registers.  What stays double: the chunk constants ck / ws (computed once per chunk), y, and the accumulators, which
// are folded into doubles after every group of four entries (256 fp32 terms per component at a time), so the chunk
// record does not carry a long fp32 summation.
#pragma once
#include "accumulate_diag.cuh"

// shared memory of one CTA: N noise blocks [k][DP] floats | ck [k1] | ws [k1] | 4 warps x 16 duplicated weights (float2) | reduction (doubles)
__host__ __device__ inline size_t diag32_cta_smem(int N, int D, int k1) {
  return (size_t)N * k1 * diag_dp(D) * 4 + (size_t)2 * k1 * 4 + 4 * 16 * 8 + 64 + (size_t)4 * (2 * diag_dp(D) + 2) * 8;
}

// transpose_reduce16_perm of accumulate_diag.cuh in float: 4 entries x 4 samples, p[4 e' + kk], the lanes of quarter
// q = lane >> 3 keeping entry e' ^ q in slot e'; on return lane l holds the total of entry (l >> 3), sample (l >> 1) & 3
__device__ __forceinline__ float transpose_reduce16_perm_f(float (&p)[16], int lane) {
  const bool b2 = (lane & 4) != 0, b1 = (lane & 2) != 0;
  float r8[8], r4[4], r2[2];
#pragma unroll
  for (int j = 0; j < 8; ++j) r8[j] = p[j] + __shfl_xor_sync(0xffffffffu, p[j + 8], 16);
#pragma unroll
  for (int j = 0; j < 4; ++j) r4[j] = r8[j] + __shfl_xor_sync(0xffffffffu, r8[j + 4], 8);
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const float give = b2 ? r4[j] : r4[j + 2], keep = b2 ? r4[j + 2] : r4[j];
    r2[j] = keep + __shfl_xor_sync(0xffffffffu, give, 4);
  }
  const float give = b1 ? r2[0] : r2[1], keep = b1 ? r2[1] : r2[0];
  float r = keep + __shfl_xor_sync(0xffffffffu, give, 2);
  r += __shfl_xor_sync(0xffffffffu, r, 1);
  return r;
}

// (mean, sqrt S) of this lane's two components of row `col` of mode d's sampler cache, as floats; zeros when !live / >= D
__device__ __forceinline__ void diag32_load_row(const DevView& v, int d, int32_t col, bool live, int n0, int n1, int D,
                                                float (&m)[2], float (&s)[2]) {
  const double* src = v.sampler[d] + (int64_t)col * v.srow_len;
  m[0] = (live && n0 < D) ? (float)src[n0] : 0.f;
  m[1] = (live && n1 < D) ? (float)src[n1] : 0.f;
  s[0] = (live && n0 < D) ? (float)src[D + n0] : 0.f;
  s[1] = (live && n1 < D) ? (float)src[D + n1] : 0.f;
}

// Stage the chunk's slot in shared memory as FLOATS: the N-1 noise blocks of the other modes, transposed from the pool's
// [c][k] to [k][DP], then the own-row sample U[k][c] = m[c] + sqrt(S)[c] z_own[c][k] (ui samples,
// SSVI_TF_interface.py:251-254), evaluated in double and rounded once.  The caller synchronises.
template <typename ZT>
__device__ __forceinline__ void diag32_stage_chunk(const DevView& v, float* __restrict__ zs, const ZT* __restrict__ z_other,
                                                   const ZT* __restrict__ z_own, int slot, int64_t row, int D, int k1,
                                                   int DP, int tid) {
  const int N = v.N, blk_stride = k1 * DP;
  for (int b = 0; b < N - 1; ++b) {
    const ZT* src = z_other + ((int64_t)slot * (N - 1) + b) * v.zblk;
    float* dst = zs + (size_t)b * blk_stride;
    for (int idx = tid; idx < D * k1; idx += 128) {
      const int cc = idx / k1, k = idx - cc * k1;
      dst[k * DP + cc] = (float)src[idx];
    }
    if (DP != D)
      for (int k = tid; k < k1; k += 128) dst[k * DP + D] = 0.f;
  }
  const ZT* src = z_own + (int64_t)slot * v.zblk;
  const double* own = v.sampler[v.dim] + row * (int64_t)v.srow_len;
  float* dst = zs + (size_t)(N - 1) * blk_stride;
  for (int idx = tid; idx < D * k1; idx += 128) {
    const int cc = idx / k1, k = idx - cc * k1;
    dst[k * DP + cc] = (float)fma(own[D + cc], (double)src[idx], own[cc]);
  }
  if (DP != D)
    for (int k = tid; k < k1; k += 128) dst[k * DP + D] = 0.f;
}

template <int NOTH, typename ZT>
__global__ void __launch_bounds__(128, 3) accumulate_diag_lin32_kernel(
    DevView v, int64_t n_chunks, const int32_t* __restrict__ chunk_row, const int64_t* __restrict__ chunk_start,
    const int32_t* __restrict__ chunk_len, const int32_t* __restrict__ chunk_slot,
    const int64_t* __restrict__ chunk_out, const int64_t* __restrict__ entry_ids, const int64_t* __restrict__ rows,
    const double* __restrict__ w_state, const ZT* __restrict__ z_other, const ZT* __restrict__ z_own,
    const ZT* __restrict__ r_w, const ZT* __restrict__ z_f, double* __restrict__ partial_out) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr int E = 4, T = 4, R = 16;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t c = blockIdx.x;
  if (c >= n_chunks) return;
  const int N = v.N, D = v.D, k1 = v.k1, DP = diag_dp(D), dim = v.dim;
  const int qperm = lane >> 3;
  const int blk_stride = k1 * DP;
  float* zs = reinterpret_cast<float*>(smem_raw);                        // [N][k1][DP]: other modes, then own sample
  float* CK = zs + (size_t)N * blk_stride;
  float* WSK = CK + k1;
  float2* Wt2 = reinterpret_cast<float2*>(smem_raw + (((size_t)N * blk_stride * 4 + 2 * k1 * 4 + 15) & ~(size_t)15)) + warp * R;
  double* red = reinterpret_cast<double*>(smem_raw + (((size_t)N * blk_stride * 4 + 2 * k1 * 4 + 15) & ~(size_t)15) + 4 * R * 8);

  const int64_t row = rows[chunk_row[c]];
  const int64_t start = chunk_start[c];
  const int len = chunk_len[c], slot = chunk_slot[c];
  const double w_sigma = w_state[0];

  int od[NOTH];
  diag_other_modes<NOTH>(N, dim, od);
  const int my_cnt = (len - warp + 3) >> 2;             // this warp's entries: w, w+4, ...
  double my_y;
  int32_t my_col[NOTH];
  diag_fetch_entries<NOTH>(v, entry_ids, start, 0, my_cnt, lane, warp, od, my_y, my_col);

  diag32_stage_chunk<ZT>(v, zs, z_other, z_own, slot, row, D, k1, DP, tid);
  // chunk constants (double arithmetic, once per chunk): ws_k = w_sigma^2 * rayleigh draw (SSVI_TF_interface.py:258-261),
  // ck_k = ws_k * mean_i z_f[i][k]
  for (int k = tid; k < k1; k += 128) {
    const ZT* zf = z_f + (int64_t)slot * v.k2 * k1 + k;
    double t = 0.0;
    for (int i0 = 0; i0 < v.k2; ++i0) t += ldz(zf, (int64_t)i0 * k1);
    const double ws = (w_sigma * w_sigma) * ldz(r_w, (int64_t)slot * k1 + k);
    WSK[k] = (float)ws;
    CK[k] = (float)(ws * (t / (double)v.k2));
  }
  const int n0 = 2 * lane, n1 = 2 * lane + 1;           // lane l owns components 2l, 2l+1
  const int zoff = n0 < DP ? n0 : 0;                    // idle lanes (n0 >= D) read valid noise against m = s = 0
  __syncthreads();

  double g0 = 0.0, g1 = 0.0, G0 = 0.0, G1 = 0.0;        // chunk accumulators of this lane's two components
  const double is2 = 1.0 / (v.lik_param * v.lik_param);
  const float alpha = (float)(is2 / (double)k1);
  const float* zo = zs + (size_t)(N - 1) * blk_stride;
  const int idx_mine = lane >> 1;                       // which of the 16 totals this lane receives
  const int e_mine = idx_mine / T, k_mine = idx_mine % T;
  const bool writer = (lane & 1) == 0;

  for (int b0 = 0; b0 < my_cnt; b0 += 32) {
    const int nb = min(32, my_cnt - b0);
    if (b0 > 0) diag_fetch_entries<NOTH>(v, entry_ids, start, b0, nb, lane, warp, od, my_y, my_col);
    for (int el = 0; el < nb; el += E) {
      const float y_mine = (float)__shfl_sync(0xffffffffu, my_y, min(el + e_mine, nb - 1));
      float2 m[E][NOTH], s[E][NOTH];                    // the lane's two components travel as one f32x2 operand
#pragma unroll
      for (int e = 0; e < E; ++e)
#pragma unroll
        for (int jj = 0; jj < NOTH; ++jj) {
          float mm[2], ss[2];
          diag32_load_row(v, od[jj], __shfl_sync(0xffffffffu, my_col[jj], min(el + (e ^ qperm), 31)),
                          el + (e ^ qperm) < nb, n0, n1, D, mm, ss);
          m[e][jj] = make_float2(mm[0], mm[1]);
          s[e][jj] = make_float2(ss[0], ss[1]);
        }
      float2 fg = make_float2(0.f, 0.f), fG = fg, fH = fg;       // this group's sums (256 terms each)
      for (int tile = 0; tile < k1 / T; ++tile) {
        float p[R];
        float2 vk[E][T];
#pragma unroll
        for (int kk = 0; kk < T; ++kk) {
          const int k = tile * T + kk;
          float2 z[NOTH];
#pragma unroll
          for (int jj = 0; jj < NOTH; ++jj)
            z[jj] = *reinterpret_cast<const float2*>(zs + (size_t)jj * blk_stride + k * DP + zoff);
          const float2 u = *reinterpret_cast<const float2*>(zo + k * DP + zoff);
#pragma unroll
          for (int e = 0; e < E; ++e) {
            float2 vv = __ffma2_rn(s[e][0], z[0], m[e][0]);      // FFMA2: both components per instruction
#pragma unroll
            for (int jj = 1; jj < NOTH; ++jj) vv = __fmul2_rn(vv, __ffma2_rn(s[e][jj], z[jj], m[e][jj]));
            p[e * T + kk] = fmaf(u.y, vv.y, u.x * vv.x);
            // sum_k v^2 (Di): d2 log p / df2 = -1/s^2 for every sample, the weight -0.5 / (s^2 k1) is applied once per chunk
            if (e & 1) fH = __ffma2_rn(vv, vv, fH);
            else fG = __ffma2_rn(vv, vv, fG);
            vk[e][kk] = vv;
          }
        }
        const float f = transpose_reduce16_perm_f(p, lane);
        const float w = alpha * (y_mine - f - CK[tile * T + k_mine]);
        __syncwarp();
        if (writer) Wt2[idx_mine] = make_float2(w, w);  // duplicated: one 16-byte read below yields two ready f32x2 operands
        __syncwarp();
#pragma unroll
        for (int kk = 0; kk < T; kk += 2) {
#pragma unroll
          for (int e = 0; e < E; ++e) {
            const float4 w4 = *reinterpret_cast<const float4*>(Wt2 + (e ^ qperm) * T + kk);
            fg = __ffma2_rn(make_float2(w4.x, w4.y), vk[e][kk], fg);
            fg = __ffma2_rn(make_float2(w4.z, w4.w), vk[e][kk + 1], fg);
          }
        }
      }
      const float fg0 = fg.x, fg1 = fg.y, fG0 = fG.x, fG1 = fG.y, fH0 = fH.x, fH1 = fH.y;
      g0 += (double)fg0;                                // fold the group into the chunk's double accumulators
      g1 += (double)fg1;
      G0 += (double)fG0 + (double)fH0;
      G1 += (double)fG1 + (double)fH1;
    }
  }

  // ---- CTA reduction in fixed warp order; chunk record [di (D) | Di (D) | si_sum | count] -----------------------------
  {
    const double dwc = -0.5 * is2 / (double)k1;
    G0 *= dwc;
    G1 *= dwc;
  }
  // si takes mean_k2 snd * ws = -ws_k / s^2 of every (entry, sample): the same sum for each of this warp's entries
  double wsum = 0.0;
  if (lane == 0)
    for (int k = 0; k < k1; ++k) wsum += (double)WSK[k];
  diag_write_record(v, red, partial_out + chunk_out[c] * (int64_t)v.acc_len, D, DP, len, tid, g0, g1, G0, G1,
                    -is2 * wsum * (double)max(my_cnt, 0));
}

// ---------------------------------------------------------------------------------------------------------------------
// SSVI_TF_d with a diagonal covariance under the fp32 option: accumulate_diag_d_kernel (accumulate_diag.cuh) with float
// samples.  The weights are uniform in k (SSVI_TF_d.py:31-57, quirk Q3: the derivatives are averaged over k1, then
// broadcast), so only sum_k v and sum_k v^2 of each entry are kept (packed f32x2, 64 terms each, folded into the chunk's
// double accumulators with the entry's double weights); the log-likelihood derivative pair of each (entry, sample) is
// evaluated in double on the float f, once per quarter-warp.
// ---------------------------------------------------------------------------------------------------------------------
__host__ __device__ inline size_t diag32_d_cta_smem(int N, int D, int k1) {
  return (size_t)N * k1 * diag_dp(D) * 4 + 64 + (size_t)4 * (2 * diag_dp(D) + 2) * 8;
}

template <int NOTH, typename ZT>
__global__ void __launch_bounds__(128, 3) accumulate_diag_d32_kernel(
    DevView v, int64_t n_chunks, const int32_t* __restrict__ chunk_row, const int64_t* __restrict__ chunk_start,
    const int32_t* __restrict__ chunk_len, const int32_t* __restrict__ chunk_slot,
    const int64_t* __restrict__ chunk_out, const int64_t* __restrict__ entry_ids, const int64_t* __restrict__ rows,
    const ZT* __restrict__ z_other, const ZT* __restrict__ z_own, double* __restrict__ partial_out) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr int E = 4, T = 4;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t c = blockIdx.x;
  if (c >= n_chunks) return;
  const int N = v.N, D = v.D, k1 = v.k1, DP = diag_dp(D), dim = v.dim;
  const int blk_stride = k1 * DP;
  const int qperm = lane >> 3;
  float* zs = reinterpret_cast<float*>(smem_raw);                        // [N][k1][DP]: other modes, then own sample
  double* red = reinterpret_cast<double*>(smem_raw + (((size_t)N * blk_stride * 4 + 15) & ~(size_t)15));

  const int64_t row = rows[chunk_row[c]];
  const int64_t start = chunk_start[c];
  const int len = chunk_len[c], slot = chunk_slot[c];

  int od[NOTH];
  diag_other_modes<NOTH>(N, dim, od);
  const int my_cnt = (len - warp + 3) >> 2;             // this warp's entries: w, w+4, ...
  double my_y;
  int32_t my_col[NOTH];
  diag_fetch_entries<NOTH>(v, entry_ids, start, 0, my_cnt, lane, warp, od, my_y, my_col);
  diag32_stage_chunk<ZT>(v, zs, z_other, z_own, slot, row, D, k1, DP, tid);
  const int n0 = 2 * lane, n1 = 2 * lane + 1;
  const int zoff = n0 < DP ? n0 : 0;
  __syncthreads();

  double g0 = 0.0, g1 = 0.0, G0 = 0.0, G1 = 0.0;
  const double inv_k1 = 1.0 / (double)k1;
  const float* zo = zs + (size_t)(N - 1) * blk_stride;
  const int e_mine = lane >> 3;

  for (int b0 = 0; b0 < my_cnt; b0 += 32) {
    const int nb = min(32, my_cnt - b0);
    if (b0 > 0) diag_fetch_entries<NOTH>(v, entry_ids, start, b0, nb, lane, warp, od, my_y, my_col);
    for (int el = 0; el < nb; el += E) {
      const double y_mine = __shfl_sync(0xffffffffu, my_y, min(el + e_mine, nb - 1));
      float2 m[E][NOTH], s[E][NOTH];
#pragma unroll
      for (int e = 0; e < E; ++e)
#pragma unroll
        for (int jj = 0; jj < NOTH; ++jj) {
          float mm[2], ss[2];
          diag32_load_row(v, od[jj], __shfl_sync(0xffffffffu, my_col[jj], min(el + (e ^ qperm), 31)),
                          el + (e ^ qperm) < nb, n0, n1, D, mm, ss);
          m[e][jj] = make_float2(mm[0], mm[1]);
          s[e][jj] = make_float2(ss[0], ss[1]);
        }
      float2 sv[E], sq[E];
#pragma unroll
      for (int e = 0; e < E; ++e) sv[e] = sq[e] = make_float2(0.f, 0.f);
      double sum_a = 0.0, sum_d = 0.0;
      for (int tile2 = 0; tile2 < k1 / (2 * T); ++tile2) {
        float fsel = 0.f;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          float p[E * T];
#pragma unroll
          for (int kk = 0; kk < T; ++kk) {
            const int k = (2 * tile2 + h) * T + kk;
            float2 z[NOTH];
#pragma unroll
            for (int jj = 0; jj < NOTH; ++jj)
              z[jj] = *reinterpret_cast<const float2*>(zs + (size_t)jj * blk_stride + k * DP + zoff);
            const float2 u = *reinterpret_cast<const float2*>(zo + k * DP + zoff);
#pragma unroll
            for (int e = 0; e < E; ++e) {
              float2 vv = __ffma2_rn(s[e][0], z[0], m[e][0]);
#pragma unroll
              for (int jj = 1; jj < NOTH; ++jj) vv = __fmul2_rn(vv, __ffma2_rn(s[e][jj], z[jj], m[e][jj]));
              p[e * T + kk] = fmaf(u.y, vv.y, u.x * vv.x);
              sv[e] = __fadd2_rn(sv[e], vv);
              sq[e] = __ffma2_rn(vv, vv, sq[e]);
            }
          }
          const float f = transpose_reduce16_perm_f(p, lane);
          if ((lane & 1) == h) fsel = f;
        }
        double fa, fb;
        log_pdf_derivs(v.lik, y_mine, (double)fsel, v.lik_param, fa, fb);
        sum_a += fa;
        sum_d += fb;
      }
      // SSVI_TF_d.py:31-57 (Q3): derivatives averaged over k1 (the 8 lanes of a quarter-warp hold one entry's samples)
#pragma unroll
      for (int o = 1; o < 8; o <<= 1) {
        sum_a += __shfl_xor_sync(0xffffffffu, sum_a, o);
        sum_d += __shfl_xor_sync(0xffffffffu, sum_d, o);
      }
      const double aw_q = (sum_a * inv_k1) * inv_k1, dw_q = 0.5 * (sum_d * inv_k1) * inv_k1;
#pragma unroll
      for (int e = 0; e < E; ++e) {                       // slot e of this lane worked on entry e ^ qperm
        const double aw = __shfl_sync(0xffffffffu, aw_q, (e ^ qperm) << 3);
        const double dw = __shfl_sync(0xffffffffu, dw_q, (e ^ qperm) << 3);
        g0 = fma(aw, (double)sv[e].x, g0);
        g1 = fma(aw, (double)sv[e].y, g1);
        G0 = fma(dw, (double)sq[e].x, G0);
        G1 = fma(dw, (double)sq[e].y, G1);
      }
    }
  }

  diag_write_record(v, red, partial_out + chunk_out[c] * (int64_t)v.acc_len, D, DP, len, tid, g0, g1, G0, G1, 0.0);
}
