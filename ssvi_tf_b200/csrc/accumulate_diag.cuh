// Stages 1-3 for FULL SWEEPS with a DIAGONAL covariance (every model, D <= 64, k1 % 16 == 0, k1 <= 64), slot-major.
//
// With a diagonal covariance a sample is element-wise, x[n][k] = m[n] + sqrt(S)[n] * z[n][k]
// (Posterior_Diag_Covariance, Model/TF_Models.py:72-93; sample_vjs_batch, SSVI_TF_interface.py:732-750), so there is no
// dense contraction and nothing worth memoising: re-evaluating a sample costs D*k1 FMAs against 8*D*k1 bytes to read it
// back.  What IS shared is the noise: in a sweep all entries of a chunk (one row, one noise slot) read the same N noise
// blocks, so one CTA stages them in shared memory ONCE (as doubles: an F2F.F64.F32 costs four DFMAs on this part) and its
// 4 warps stream the chunk's entries against them.  Per entry only the N-1 gathered (mean, sqrt S) rows come from global
// memory -- the algorithmic bytes of SURVEY.md 8(d).
//
// Lanes run over components: lane l owns n = 2l, 2l+1, so the per-row accumulators g[n], G[n] (di, Di) live in
// registers across the whole chunk and need no reduction.  The predictive mean f_k = sum_n u[n][k] v[n][k] is reduced
// over lanes with a transposing shuffle reduction, after which every lane owns a few samples' f.  Three kernels:
//   accumulate_diag_kernel      SSVI_TF_simple / SSVI_TF_robust, any likelihood: the k2 second-layer loop gives per-sample
//                               weights, then a second pass re-evaluates v (one entry per warp step)
//   accumulate_diag_lin_kernel  SSVI_TF_simple + normal likelihood (BASELINE config 4): closed-form weights, four entries
//                               per warp step sharing every shared-memory read
//   accumulate_diag_d_kernel    SSVI_TF_d (weights uniform in k, SSVI_TF_d.py:31-57 quirk Q3), entry-blocked the same way
#pragma once
#include "accumulate_generic.cuh"

__host__ __device__ inline int diag_dp(int D) { return (D + 1) & ~1; }          // components padded to even
// shared memory of one CTA (4 warps): N noise blocks [k][DP] as doubles | per-warp weights [k1] double2 | reduction
__host__ __device__ inline size_t diag_cta_smem(int N, int D, int k1) {
  return (size_t)N * k1 * diag_dp(D) * 8 + (size_t)4 * k1 * 16 + (size_t)4 * (2 * diag_dp(D) + 2) * 8;
}

// ---- pieces shared by the three sweep kernels of this file ---------------------------------------------------------------
// the other modes of the sweep's mode, ascending
template <int NOTH>
__device__ __forceinline__ void diag_other_modes(int N, int dim, int (&od)[NOTH]) {
  int jj = 0;
#pragma unroll
  for (int d = 0; d < SSVI_MAX_ORDER; ++d)
    if (d < N && d != dim && jj < NOTH) od[jj++] = d;
}

// Entries b0 .. b0+31 of warp `warp`'s list (the chunk's entries warp, warp+4, ...): lane l fetches the value and the
// other-mode rows of entry b0 + l; the warp later reads them with shuffles.  Lanes past the end hold y = 0, row 0.
template <int NOTH>
__device__ __forceinline__ void diag_fetch_entries(const DevView& v, const int64_t* __restrict__ entry_ids, int64_t start,
                                                   int b0, int nb, int lane, int warp, const int (&od)[NOTH],
                                                   double& my_y, int32_t (&my_col)[NOTH]) {
  my_y = 0.0;
#pragma unroll
  for (int jj = 0; jj < NOTH; ++jj) my_col[jj] = 0;
  if (lane < nb) {
    const int64_t eid = entry_ids[start + (int64_t)(b0 + lane) * 4 + warp];
    my_y = v.vals[eid];
#pragma unroll
    for (int jj = 0; jj < NOTH; ++jj) my_col[jj] = v.coords[eid * v.N + od[jj]];
  }
}

// Stage the chunk's slot in shared memory as doubles: the N-1 noise blocks of the other modes, transposed from the
// pool's [c][k] to [k][DP], then the own-row sample U[k][c] = m[c] + sqrt(S)[c] z_own[c][k] (ui samples,
// SSVI_TF_interface.py:251-254) -- once per chunk.  The caller synchronises.
template <typename ZT>
__device__ __forceinline__ void diag_stage_chunk(const DevView& v, double* __restrict__ zs, const ZT* __restrict__ z_other,
                                                 const ZT* __restrict__ z_own, int slot, int64_t row, int D, int k1,
                                                 int DP, int tid) {
  const int N = v.N, blk_stride = k1 * DP;
  for (int b = 0; b < N - 1; ++b) {
    const ZT* src = z_other + ((int64_t)slot * (N - 1) + b) * v.zblk;
    double* dst = zs + (size_t)b * blk_stride;
    for (int idx = tid; idx < D * k1; idx += 128) {
      const int cc = idx / k1, k = idx - cc * k1;
      dst[k * DP + cc] = (double)src[idx];
    }
    if (DP != D)
      for (int k = tid; k < k1; k += 128) dst[k * DP + D] = 0.0;
  }
  const ZT* src = z_own + (int64_t)slot * v.zblk;
  const double* own = v.sampler[v.dim] + row * (int64_t)v.srow_len;
  double* dst = zs + (size_t)(N - 1) * blk_stride;
  for (int idx = tid; idx < D * k1; idx += 128) {
    const int cc = idx / k1, k = idx - cc * k1;
    dst[k * DP + cc] = fma(own[D + cc], (double)src[idx], own[cc]);
  }
  if (DP != D)
    for (int k = tid; k < k1; k += 128) dst[k * DP + D] = 0.0;
}

// CTA reduction in fixed warp order and the chunk record [di (D) | Di (D) | si_sum | count]: every lane hands in the
// sums of its two components, lane 0 of each warp the warp's si.
__device__ __forceinline__ void diag_write_record(const DevView& v, double* __restrict__ red, double* __restrict__ out,
                                                  int D, int DP, int len, int tid, double g0, double g1, double G0,
                                                  double G1, double si_warp) {
  const int lane = tid & 31, warp = tid >> 5, n0 = 2 * lane;
  const int rs = 2 * DP + 2;
  double* rw = red + warp * rs;
  if (n0 < DP) {
    rw[n0] = g0;
    rw[n0 + 1] = g1;
    rw[DP + n0] = G0;
    rw[DP + n0 + 1] = G1;
  }
  if (lane == 0) rw[2 * DP] = si_warp;
  __syncthreads();
  for (int i = tid; i < 2 * D + 1; i += 128) {
    const int src = i < D ? i : (i < 2 * D ? DP + (i - D) : 2 * DP);
    out[i] = ((red[src] + red[rs + src]) + red[2 * rs + src]) + red[3 * rs + src];
  }
  if (tid == 0) out[2 * D + 1] = (double)len;
}

// sum over the 32 lanes of p[0..15]; on return lane l holds the total of sample (l >> 1) (both lanes of a pair)
__device__ __forceinline__ double transpose_reduce16(double (&p)[16], int lane) {
  const bool b4 = (lane & 16) != 0, b3 = (lane & 8) != 0, b2 = (lane & 4) != 0, b1 = (lane & 2) != 0;
  double r8[8], r4[4], r2[2];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const double give = b4 ? p[j] : p[j + 8], keep = b4 ? p[j + 8] : p[j];
    r8[j] = keep + __shfl_xor_sync(0xffffffffu, give, 16);
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const double give = b3 ? r8[j] : r8[j + 4], keep = b3 ? r8[j + 4] : r8[j];
    r4[j] = keep + __shfl_xor_sync(0xffffffffu, give, 8);
  }
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const double give = b2 ? r4[j] : r4[j + 2], keep = b2 ? r4[j + 2] : r4[j];
    r2[j] = keep + __shfl_xor_sync(0xffffffffu, give, 4);
  }
  const double give = b1 ? r2[0] : r2[1], keep = b1 ? r2[1] : r2[0];
  double r = keep + __shfl_xor_sync(0xffffffffu, give, 2);
  r += __shfl_xor_sync(0xffffffffu, r, 1);
  return r;                                             // sample index 8*b4 + 4*b3 + 2*b2 + b1 = lane >> 1
}

// v[0..1] = prod_jj (m + s z) for this lane's two components at sample k; z blocks staged as [blk][k][DP]
template <int NOTH>
__device__ __forceinline__ void diag_v(const double* __restrict__ zs, int blk_stride, int k, int DP, int zoff,
                                       const double (&m)[NOTH][2], const double (&s)[NOTH][2], double& v0,
                                       double& v1) {
#pragma unroll
  for (int jj = 0; jj < NOTH; ++jj) {
    const double2 z = *reinterpret_cast<const double2*>(zs + (size_t)jj * blk_stride + k * DP + zoff);
    const double x0 = fma(s[jj][0], z.x, m[jj][0]), x1 = fma(s[jj][1], z.y, m[jj][1]);
    if (jj == 0) {
      v0 = x0;
      v1 = x1;
    } else {
      v0 *= x0;
      v1 *= x1;
    }
  }
}

// One CTA (4 warps) per chunk = run of entries of ONE row that share ONE noise slot.  SSVI_TF_simple / SSVI_TF_robust with
// any likelihood (SSVI_TF_d and the closed-form simple + normal case have the entry-blocked kernels below).
template <int NOTH, typename ZT>
__global__ void __launch_bounds__(128, 2) accumulate_diag_kernel(
    DevView v, int64_t n_chunks, const int32_t* __restrict__ chunk_row, const int64_t* __restrict__ chunk_start,
    const int32_t* __restrict__ chunk_len, const int32_t* __restrict__ chunk_slot,
    const int64_t* __restrict__ chunk_out, const int64_t* __restrict__ entry_ids, const int64_t* __restrict__ rows,
    const double* __restrict__ w_state, const ZT* __restrict__ z_other, const ZT* __restrict__ z_own,
    const ZT* __restrict__ r_w, const ZT* __restrict__ z_f, double* __restrict__ partial_out) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t c = blockIdx.x;
  if (c >= n_chunks) return;
  const int N = v.N, D = v.D, k1 = v.k1, DP = diag_dp(D), dim = v.dim;
  const int blk_stride = k1 * DP;
  double* zs = reinterpret_cast<double*>(smem_raw);                      // [N][k1][DP]: other modes, then own
  double2* AW = reinterpret_cast<double2*>(zs + (size_t)N * blk_stride) + warp * k1;
  double* red = reinterpret_cast<double*>(zs + (size_t)N * blk_stride) + 4 * k1 * 2;

  const int64_t row = rows[chunk_row[c]];
  const int64_t start = chunk_start[c];
  const int len = chunk_len[c], slot = chunk_slot[c];
  const double w_sigma = w_state[0];

  int od[NOTH];
  diag_other_modes<NOTH>(N, dim, od);

  // ---- this warp's entries (w, w+4, ...): ids, values and gathered-row offsets fetched lane-parallel -----------------
  const int my_cnt = (len - warp + 3) >> 2;
  double my_y;
  int32_t my_col[NOTH];
  diag_fetch_entries<NOTH>(v, entry_ids, start, 0, my_cnt, lane, warp, od, my_y, my_col);

  diag_stage_chunk<ZT>(v, zs, z_other, z_own, slot, row, D, k1, DP, tid);
  // lane l owns components 2l, 2l+1 (gathered rows read as zero beyond D: v is then an exact zero there)
  const int n0 = 2 * lane, n1 = 2 * lane + 1;
  const int zoff = n0 < DP ? n0 : 0;                    // idle lanes (n0 >= D) read valid noise against m = s = 0
  __syncthreads();

  double g0 = 0.0, g1 = 0.0, G0 = 0.0, G1 = 0.0, si_part = 0.0;
  bool bad = false;
  const double* zo = zs + (size_t)(N - 1) * blk_stride;

  for (int b0 = 0; b0 < my_cnt; b0 += 32) {
    const int nb = min(32, my_cnt - b0);
    if (b0 > 0) diag_fetch_entries<NOTH>(v, entry_ids, start, b0, nb, lane, warp, od, my_y, my_col);
    // gathered rows of the first entry
    double mn[NOTH][2], sn[NOTH][2];
#pragma unroll
    for (int jj = 0; jj < NOTH; ++jj) {
      const double* src = v.sampler[od[jj]] + (int64_t)__shfl_sync(0xffffffffu, my_col[jj], 0) * v.srow_len;
      mn[jj][0] = n0 < D ? src[n0] : 0.0;
      mn[jj][1] = n1 < D ? src[n1] : 0.0;
      sn[jj][0] = n0 < D ? src[D + n0] : 0.0;
      sn[jj][1] = n1 < D ? src[D + n1] : 0.0;
    }
    for (int el = 0; el < nb; ++el) {
      const double y = __shfl_sync(0xffffffffu, my_y, el);
      double m[NOTH][2], s[NOTH][2];
#pragma unroll
      for (int jj = 0; jj < NOTH; ++jj) {
        m[jj][0] = mn[jj][0];
        m[jj][1] = mn[jj][1];
        s[jj][0] = sn[jj][0];
        s[jj][1] = sn[jj][1];
      }
      {                                                  // prefetch the next entry's rows (last entry: its own again)
        const int eln = (el + 1 < nb) ? el + 1 : el;
#pragma unroll
        for (int jj = 0; jj < NOTH; ++jj) {
          const double* src = v.sampler[od[jj]] + (int64_t)__shfl_sync(0xffffffffu, my_col[jj], eln) * v.srow_len;
          mn[jj][0] = n0 < D ? src[n0] : 0.0;
          mn[jj][1] = n1 < D ? src[n1] : 0.0;
          sn[jj][0] = n0 < D ? src[D + n0] : 0.0;
          sn[jj][1] = n1 < D ? src[D + n1] : 0.0;
        }
      }
      // ---- pass 1: f_k ---------------------------------------------------------------------------------------------
      double fk[2] = {0.0, 0.0};                         // lane owns samples 16*(2a + (lane&1)) + (lane>>1), a = 0, 1
      for (int tile = 0; tile < (k1 >> 4); ++tile) {
        double p[16];
#pragma unroll
        for (int kk = 0; kk < 16; ++kk) {
          const int k = tile * 16 + kk;
          double v0, v1;
          diag_v<NOTH>(zs, blk_stride, k, DP, zoff, m, s, v0, v1);
          const double2 u = *reinterpret_cast<const double2*>(zo + k * DP + zoff);
          p[kk] = fma(u.y, v1, u.x * v0);
        }
        const double f = transpose_reduce16(p, lane);
        if ((tile & 1) == (lane & 1)) {
          if (tile < 2) fk[0] = f; else fk[1] = f;
        }
      }
      {
        // per-sample weights (second-layer k2 loop), then pass 2 re-evaluates v
        const double yfact = (v.lik == SSVI_LIK_POISSON && v.model == SSVI_MODEL_ROBUST) ? 1.0 / factorial_ref(y) : 1.0;
        __syncwarp();
#pragma unroll
        for (int a = 0; a < 2; ++a) {
          const int k = 16 * (2 * a + (lane & 1)) + (lane >> 1);
          const bool active = k < k1;
          double aw, dw;
          likelihood_weights<ZT>(v, r_w, z_f, slot, k, active, y, yfact, fk[a], w_sigma, aw, dw, si_part, bad);
          if (active) AW[k] = make_double2(aw, dw);
        }
        __syncwarp();
#pragma unroll 4
        for (int k = 0; k < k1; ++k) {
          double v0, v1;
          diag_v<NOTH>(zs, blk_stride, k, DP, zoff, m, s, v0, v1);
          const double2 w = AW[k];
          g0 = fma(w.x, v0, g0);
          g1 = fma(w.x, v1, g1);
          G0 = fma(w.y * v0, v0, G0);
          G1 = fma(w.y * v1, v1, G1);
        }
      }
    }
  }

  si_part = warp_sum(si_part);
  const unsigned anybad = __ballot_sync(0xffffffffu, bad);
  if (lane == 0 && anybad) atomicOr(v.flags, SSVI_FLAG_PHI_ZERO);
  diag_write_record(v, red, partial_out + chunk_out[c] * (int64_t)v.acc_len, D, DP, len, tid, g0, g1, G0, G1, si_part);
}

// ---------------------------------------------------------------------------------------------------------------------
// SSVI_TF_simple + normal likelihood (BASELINE config 4), entry-blocked.
//
// The one-entry-at-a-time kernel above reads the N staged noise blocks from shared memory once per ENTRY: N*k1*DP*8 bytes
// (102 KB at config 4) against ~1100 FP64 instructions, so the shared-memory datapath (128 B/clk/SM), not the FP64 pipe,
// sets its pace.  Here a warp walks its entries E at a time: every noise / own-sample double2 read from shared memory
// serves E entries (1/E of the bytes per unit, E independent FMA chains per lane).  The predictive means of T samples
// of the E entries (E*T partial sums per lane) go through ONE transposing shuffle reduction, after which every lane
// holds f of one (entry, sample) of the tile; the closed-form weight  w_k = (y - f_k - ws_k * mean_i z_f[i][k]) / (s^2 k1)
// (SSVI_TF_simple.py:35-38 with normal.py:32-36: the k2 mean is linear in f) is handed round the warp through E*T
// doubles of shared memory and applied to the v kept in registers.  ws_k * zbar_k and sum_k ws_k are constants of the
// chunk.  A group's missing entries (end of the warp's list) run with m = sqrt S = 0, i.e. v = 0: they add exact zeros.
// ---------------------------------------------------------------------------------------------------------------------
#ifndef SSVI_DIAG_LIN_E
#define SSVI_DIAG_LIN_E 4                               // entries per warp step: 2 or 4
#endif
#ifndef SSVI_DIAG_LIN_T
#define SSVI_DIAG_LIN_T 4                               // samples per reduction tile; E*T is 8 or 16, k1 % T == 0
#endif
#ifndef SSVI_DIAG_LIN_PF
#define SSVI_DIAG_LIN_PF 0                              // next group's gathered rows prefetched into registers
#endif

// sum over the 32 lanes of p[0..7]; on return lane l holds the total of index (l >> 2) (all four lanes of a quad)
__device__ __forceinline__ double transpose_reduce8(double (&p)[8], int lane) {
  const bool b4 = (lane & 16) != 0, b3 = (lane & 8) != 0, b2 = (lane & 4) != 0;
  double r4[4], r2[2];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const double give = b4 ? p[j] : p[j + 4], keep = b4 ? p[j + 4] : p[j];
    r4[j] = keep + __shfl_xor_sync(0xffffffffu, give, 16);
  }
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const double give = b3 ? r4[j] : r4[j + 2], keep = b3 ? r4[j + 2] : r4[j];
    r2[j] = keep + __shfl_xor_sync(0xffffffffu, give, 8);
  }
  const double give = b2 ? r2[0] : r2[1], keep = b2 ? r2[1] : r2[0];
  double r = keep + __shfl_xor_sync(0xffffffffu, give, 4);
  r += __shfl_xor_sync(0xffffffffu, r, 2);
  r += __shfl_xor_sync(0xffffffffu, r, 1);
  return r;
}
// transpose_reduce16 for E = 4 entries x T = 4 samples, p[4 e' + kk], when the lanes of quarter q = (lane >> 3) keep entry
// e' ^ q in slot e': the two entry-splitting stages then keep / give the same registers in every lane (no selects);
// on return lane l holds the total of entry (l >> 3), sample (l >> 1) & 3 -- index (l >> 1), as transpose_reduce16.
__device__ __forceinline__ double transpose_reduce16_perm(double (&p)[16], int lane) {
  const bool b2 = (lane & 4) != 0, b1 = (lane & 2) != 0;
  double r8[8], r4[4], r2[2];
#pragma unroll
  for (int j = 0; j < 8; ++j) r8[j] = p[j] + __shfl_xor_sync(0xffffffffu, p[j + 8], 16);
#pragma unroll
  for (int j = 0; j < 4; ++j) r4[j] = r8[j] + __shfl_xor_sync(0xffffffffu, r8[j + 4], 8);
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const double give = b2 ? r4[j] : r4[j + 2], keep = b2 ? r4[j + 2] : r4[j];
    r2[j] = keep + __shfl_xor_sync(0xffffffffu, give, 4);
  }
  const double give = b1 ? r2[0] : r2[1], keep = b1 ? r2[1] : r2[0];
  double r = keep + __shfl_xor_sync(0xffffffffu, give, 2);
  r += __shfl_xor_sync(0xffffffffu, r, 1);
  return r;
}
__device__ __forceinline__ double transpose_reduce_group(double (&p)[16], int lane) { return transpose_reduce16(p, lane); }
__device__ __forceinline__ double transpose_reduce_group(double (&p)[8], int lane) { return transpose_reduce8(p, lane); }

// (mean, sqrt S) of this lane's two components of row `col` of mode-`d` sampler cache; zeros when !live or beyond D
__device__ __forceinline__ void diag_load_row(const DevView& v, int d, int32_t col, bool live, int n0, int n1, int D,
                                              double (&m)[2], double (&s)[2]) {
  const double* src = v.sampler[d] + (int64_t)col * v.srow_len;
  m[0] = (live && n0 < D) ? src[n0] : 0.0;
  m[1] = (live && n1 < D) ? src[n1] : 0.0;
  s[0] = (live && n0 < D) ? src[D + n0] : 0.0;
  s[1] = (live && n1 < D) ? src[D + n1] : 0.0;
}

// CD: the rank as a compile-time constant (with k1 == 64), 0 = read D and k1 from the view: every shared-memory address
// of the sample loop becomes base + immediate.
template <int NOTH, typename ZT, int E, int T, bool PF, int CD>
__global__ void __launch_bounds__(128, 2) accumulate_diag_lin_kernel(
    DevView v, int64_t n_chunks, const int32_t* __restrict__ chunk_row, const int64_t* __restrict__ chunk_start,
    const int32_t* __restrict__ chunk_len, const int32_t* __restrict__ chunk_slot,
    const int64_t* __restrict__ chunk_out, const int64_t* __restrict__ entry_ids, const int64_t* __restrict__ rows,
    const double* __restrict__ w_state, const ZT* __restrict__ z_other, const ZT* __restrict__ z_own,
    const ZT* __restrict__ r_w, const ZT* __restrict__ z_f, double* __restrict__ partial_out) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr int R = E * T;                              // values per lane entering one reduction
  static_assert((E == 2 || E == 4) && (R == 8 || R == 16) && (T % 2) == 0, "E*T partial sums per reduction: 8 or 16");
  constexpr bool PERM = (E == 4 && T == 4);             // slot e' of quarter-warp q works on entry e' ^ q (see above)
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t c = blockIdx.x;
  if (c >= n_chunks) return;
  const int N = v.N, D = CD ? CD : v.D, k1 = CD ? 64 : v.k1, DP = diag_dp(D), dim = v.dim;
  const int qperm = PERM ? (lane >> 3) : 0;
  const int blk_stride = k1 * DP;
  // shared memory (inside the diag_cta_smem budget): noise blocks | ck [k1] | ws [k1] | per-warp weights [R] | reduction
  double* zs = reinterpret_cast<double*>(smem_raw);                      // [N][k1][DP]: other modes, then own sample
  double* CK = zs + (size_t)N * blk_stride;
  double* WSK = CK + k1;
  double* Wt = WSK + k1 + warp * R;
  double* red = zs + (size_t)N * blk_stride + 4 * k1 * 2;

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

  diag_stage_chunk<ZT>(v, zs, z_other, z_own, slot, row, D, k1, DP, tid);
  // chunk constants: ws_k = w_sigma^2 * rayleigh draw (SSVI_TF_interface.py:258-261), ck_k = ws_k * mean_i z_f[i][k]
  for (int k = tid; k < k1; k += 128) {
    const ZT* zf = z_f + (int64_t)slot * v.k2 * k1 + k;
    double t = 0.0;
    for (int i0 = 0; i0 < v.k2; i0 += 8) {
      double zz[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) zz[j] = (i0 + j < v.k2) ? ldz(zf, (int64_t)(i0 + j) * k1) : 0.0;
#pragma unroll
      for (int j = 0; j < 8; ++j) t += zz[j];
    }
    const double ws = (w_sigma * w_sigma) * ldz(r_w, (int64_t)slot * k1 + k);
    WSK[k] = ws;
    CK[k] = ws * (t / (double)v.k2);
  }
  const int n0 = 2 * lane, n1 = 2 * lane + 1;           // lane l owns components 2l, 2l+1
  const int zoff = n0 < DP ? n0 : 0;                    // idle lanes (n0 >= D) read valid noise against m = s = 0
  __syncthreads();

  double g0 = 0.0, g1 = 0.0, G0 = 0.0, G1 = 0.0, H0 = 0.0, H1 = 0.0;
  const double inv_k1 = 1.0 / (double)k1;
  const double is2 = 1.0 / (v.lik_param * v.lik_param);
  const double alpha = is2 * inv_k1;
  const double* zo = zs + (size_t)(N - 1) * blk_stride;
  const int idx_mine = (R == 16) ? (lane >> 1) : (lane >> 2);            // which of the R totals this lane receives
  const int e_mine = idx_mine / T, k_mine = idx_mine % T;
  const bool writer = (R == 16) ? ((lane & 1) == 0) : ((lane & 3) == 0);

  for (int b0 = 0; b0 < my_cnt; b0 += 32) {
    const int nb = min(32, my_cnt - b0);
    if (b0 > 0) diag_fetch_entries<NOTH>(v, entry_ids, start, b0, nb, lane, warp, od, my_y, my_col);
    double mn[PF ? E : 1][NOTH][2], sn[PF ? E : 1][NOTH][2];
    if (PF) {                                           // gathered (mean, sqrt S) rows of the first group
#pragma unroll
      for (int e = 0; e < E; ++e)
#pragma unroll
        for (int jj = 0; jj < NOTH; ++jj)
          diag_load_row(v, od[jj], __shfl_sync(0xffffffffu, my_col[jj], min(e ^ qperm, 31)), (e ^ qperm) < nb, n0, n1,
                        D, mn[e][jj], sn[e][jj]);
    }
    for (int el = 0; el < nb; el += E) {
      const double y_mine = __shfl_sync(0xffffffffu, my_y, min(el + e_mine, nb - 1));
      double m[E][NOTH][2], s[E][NOTH][2];
#pragma unroll
      for (int e = 0; e < E; ++e)
#pragma unroll
        for (int jj = 0; jj < NOTH; ++jj) {
          if (PF) {
            m[e][jj][0] = mn[e][jj][0];
            m[e][jj][1] = mn[e][jj][1];
            s[e][jj][0] = sn[e][jj][0];
            s[e][jj][1] = sn[e][jj][1];
          } else {
            diag_load_row(v, od[jj], __shfl_sync(0xffffffffu, my_col[jj], min(el + (e ^ qperm), 31)),
                          el + (e ^ qperm) < nb, n0, n1, D, m[e][jj], s[e][jj]);
          }
        }
      if (PF) {                                         // the next group's rows travel while this one is worked on
#pragma unroll
        for (int e = 0; e < E; ++e)
#pragma unroll
          for (int jj = 0; jj < NOTH; ++jj)
            diag_load_row(v, od[jj], __shfl_sync(0xffffffffu, my_col[jj], min(el + E + (e ^ qperm), 31)),
                          el + E + (e ^ qperm) < nb, n0, n1, D, mn[e][jj], sn[e][jj]);
      }
      for (int tile = 0; tile < k1 / T; ++tile) {
        double p[R];
        double vk[E][T][2];
#pragma unroll
        for (int kk = 0; kk < T; ++kk) {
          const int k = tile * T + kk;
          double2 z[NOTH];
#pragma unroll
          for (int jj = 0; jj < NOTH; ++jj)
            z[jj] = *reinterpret_cast<const double2*>(zs + (size_t)jj * blk_stride + k * DP + zoff);
          const double2 u = *reinterpret_cast<const double2*>(zo + k * DP + zoff);
#pragma unroll
          for (int e = 0; e < E; ++e) {
            double v0 = fma(s[e][0][0], z[0].x, m[e][0][0]), v1 = fma(s[e][0][1], z[0].y, m[e][0][1]);
#pragma unroll
            for (int jj = 1; jj < NOTH; ++jj) {
              v0 *= fma(s[e][jj][0], z[jj].x, m[e][jj][0]);
              v1 *= fma(s[e][jj][1], z[jj].y, m[e][jj][1]);
            }
            p[e * T + kk] = fma(u.y, v1, u.x * v0);
            // sum_k v^2 (Di): d2 log p / df2 = -1/s^2 for every sample, so the weight -0.5 / (s^2 k1) is applied once,
            // at the end of the chunk; two accumulators per component keep the FMA chains short
            if (e & 1) {
              H0 = fma(v0, v0, H0);
              H1 = fma(v1, v1, H1);
            } else {
              G0 = fma(v0, v0, G0);
              G1 = fma(v1, v1, G1);
            }
            vk[e][kk][0] = v0;
            vk[e][kk][1] = v1;
          }
        }
        double f;
        if constexpr (PERM) f = transpose_reduce16_perm(p, lane); else f = transpose_reduce_group(p, lane);
        const double w = alpha * (y_mine - f - CK[tile * T + k_mine]);
        __syncwarp();
        if (writer) Wt[idx_mine] = w;
        __syncwarp();
#pragma unroll
        for (int kk = 0; kk < T; kk += 2) {
#pragma unroll
          for (int e = 0; e < E; ++e) {
            const double2 w2 = *reinterpret_cast<const double2*>(Wt + (e ^ qperm) * T + kk);
            g0 = fma(w2.x, vk[e][kk][0], g0);
            g1 = fma(w2.x, vk[e][kk][1], g1);
            g0 = fma(w2.y, vk[e][kk + 1][0], g0);
            g1 = fma(w2.y, vk[e][kk + 1][1], g1);
          }
        }
      }
    }
  }

  // ---- CTA reduction in fixed warp order; chunk record [di (D) | Di (D) | si_sum | count] -----------------------------
  {
    const double dwc = -0.5 * alpha;
    G0 = (G0 + H0) * dwc;
    G1 = (G1 + H1) * dwc;
  }
  // si takes mean_k2 snd * ws = -ws_k / s^2 of every (entry, sample): the same sum for each of this warp's entries
  double wsum = 0.0;
  if (lane == 0)
    for (int k = 0; k < k1; ++k) wsum += WSK[k];
  diag_write_record(v, red, partial_out + chunk_out[c] * (int64_t)v.acc_len, D, DP, len, tid, g0, g1, G0, G1,
                    -is2 * wsum * (double)max(my_cnt, 0));
}

// ---------------------------------------------------------------------------------------------------------------------
// SSVI_TF_d with a diagonal covariance, entry-blocked the same way (four entries per warp step, four samples per
// reduction tile, entries permuted per quarter-warp).  The weights are uniform in k (SSVI_TF_d.py:31-57, quirk Q3: the
// derivatives are averaged over k1, then broadcast), so nothing but sum_k v and sum_k v^2 of each entry is kept: after
// the reduction of a tile lane l holds f of entry (l >> 3), sample 4 tile + ((l >> 1) & 3), in both lanes of a pair;
// even lanes keep the even tiles' value, odd lanes the odd tiles', and every second tile each lane evaluates ONE
// likelihood derivative pair -- every (entry, sample) exactly once across the quarter-warp.
// ---------------------------------------------------------------------------------------------------------------------
template <int NOTH, typename ZT>
__global__ void __launch_bounds__(128, 2) accumulate_diag_d_kernel(
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
  double* zs = reinterpret_cast<double*>(smem_raw);                      // [N][k1][DP]: other modes, then own sample
  double* red = zs + (size_t)N * blk_stride + 4 * k1 * 2;

  const int64_t row = rows[chunk_row[c]];
  const int64_t start = chunk_start[c];
  const int len = chunk_len[c], slot = chunk_slot[c];

  int od[NOTH];
  diag_other_modes<NOTH>(N, dim, od);
  const int my_cnt = (len - warp + 3) >> 2;             // this warp's entries: w, w+4, ...
  double my_y;
  int32_t my_col[NOTH];
  diag_fetch_entries<NOTH>(v, entry_ids, start, 0, my_cnt, lane, warp, od, my_y, my_col);
  diag_stage_chunk<ZT>(v, zs, z_other, z_own, slot, row, D, k1, DP, tid);
  const int n0 = 2 * lane, n1 = 2 * lane + 1;
  const int zoff = n0 < DP ? n0 : 0;
  __syncthreads();

  double g0 = 0.0, g1 = 0.0, G0 = 0.0, G1 = 0.0;
  const double inv_k1 = 1.0 / (double)k1;
  const double* zo = zs + (size_t)(N - 1) * blk_stride;
  const int e_mine = lane >> 3;

  for (int b0 = 0; b0 < my_cnt; b0 += 32) {
    const int nb = min(32, my_cnt - b0);
    if (b0 > 0) diag_fetch_entries<NOTH>(v, entry_ids, start, b0, nb, lane, warp, od, my_y, my_col);
    for (int el = 0; el < nb; el += E) {
      const double y_mine = __shfl_sync(0xffffffffu, my_y, min(el + e_mine, nb - 1));
      double m[E][NOTH][2], s[E][NOTH][2];
#pragma unroll
      for (int e = 0; e < E; ++e)
#pragma unroll
        for (int jj = 0; jj < NOTH; ++jj)
          diag_load_row(v, od[jj], __shfl_sync(0xffffffffu, my_col[jj], min(el + (e ^ qperm), 31)),
                        el + (e ^ qperm) < nb, n0, n1, D, m[e][jj], s[e][jj]);
      double sv[E][2], sq[E][2];
#pragma unroll
      for (int e = 0; e < E; ++e) sv[e][0] = sv[e][1] = sq[e][0] = sq[e][1] = 0.0;
      double sum_a = 0.0, sum_d = 0.0;
      for (int tile2 = 0; tile2 < k1 / (2 * T); ++tile2) {
        double fsel = 0.0;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          double p[E * T];
#pragma unroll
          for (int kk = 0; kk < T; ++kk) {
            const int k = (2 * tile2 + h) * T + kk;
            double2 z[NOTH];
#pragma unroll
            for (int jj = 0; jj < NOTH; ++jj)
              z[jj] = *reinterpret_cast<const double2*>(zs + (size_t)jj * blk_stride + k * DP + zoff);
            const double2 u = *reinterpret_cast<const double2*>(zo + k * DP + zoff);
#pragma unroll
            for (int e = 0; e < E; ++e) {
              double v0 = fma(s[e][0][0], z[0].x, m[e][0][0]), v1 = fma(s[e][0][1], z[0].y, m[e][0][1]);
#pragma unroll
              for (int jj = 1; jj < NOTH; ++jj) {
                v0 *= fma(s[e][jj][0], z[jj].x, m[e][jj][0]);
                v1 *= fma(s[e][jj][1], z[jj].y, m[e][jj][1]);
              }
              p[e * T + kk] = fma(u.y, v1, u.x * v0);
              sv[e][0] += v0;
              sv[e][1] += v1;
              sq[e][0] = fma(v0, v0, sq[e][0]);
              sq[e][1] = fma(v1, v1, sq[e][1]);
            }
          }
          const double f = transpose_reduce16_perm(p, lane);
          if ((lane & 1) == h) fsel = f;
        }
        double fa, fb;
        log_pdf_derivs(v.lik, y_mine, fsel, v.lik_param, fa, fb);
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
        g0 = fma(aw, sv[e][0], g0);
        g1 = fma(aw, sv[e][1], g1);
        G0 = fma(dw, sq[e][0], G0);
        G1 = fma(dw, sq[e][1], G1);
      }
    }
  }

  diag_write_record(v, red, partial_out + chunk_out[c] * (int64_t)v.acc_len, D, DP, len, tid, g0, g1, G0, G1, 0.0);
}
