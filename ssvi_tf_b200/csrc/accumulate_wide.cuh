// Stages 1-3 for FULL SWEEPS of SSVI_TF_d with a full covariance and 32 < D <= 64 (NT = 5..8 row tiles), k1 == 64:
// the sample-table path of accumulate_memo.cuh for ranks whose D x D contraction no longer fits one warp's registers
// (the reference's real-data driver defaults to rank 50: real_test.py:27, test_scripts/factorize.py:32).
//
// Same table, same chunk tables, same chunk records as the narrow kernels; what changes is who does what:
//   sample_table_wide_kernel   one warp per table record again, but a row tile's 8 k-tiles are written as soon as they
//                              are sampled (the narrow kernel keeps all NT x 8 tiles of a record in registers)
//   accumulate_wide_kernel     one CTA (4 warps) per chunk, and the WHOLE CTA works on one entry at a time:
//       phase 1  warp w gathers k-tiles {w, w+4} of the entry's N-1 records, forms V = Hadamard product, parks the V
//                tiles in shared memory, takes its share of f_k = u_k . v_k and of the k1-averaged log-likelihood
//                derivatives (SSVI_TF_d.py:31-57, quirk Q3) and of the row sums SUM_k V[n][k];
//       barrier  (one per entry: V and the four partial derivative sums become visible; V is double-buffered)
//       phase 2  the NP = NT(NT+1)/2 tile pairs of G += w V V^T are split over the warps by tile row (warp w owns rows
//                NT-1-w and its mirror, <= 9 pairs): it reads the V tiles it needs back from shared memory and issues
//                its DMMAs with the scalar weight folded into the A operand, so its part of G stays in REGISTERS for
//                the whole chunk -- no shared-memory read-modify-write, no cross-warp sum at the end.
#pragma once
#include "accumulate_memo.cuh"

// tile rows owned by warp w in phase 2: (A, B), B < 0: none.  Even NT: rows (NT-1-w, w) for w < NT/2 -- NT + 1 pairs
// each.  Odd NT: warp 0 takes row NT-1 alone, warp w >= 1 rows (NT-1-w, w-1) while w-1 < NT-1-w.
__host__ __device__ constexpr int wide_row_a(int NT, int w) {
  return (NT % 2 == 0) ? (w < NT / 2 ? NT - 1 - w : -1) : ((w == 0 || (w - 1) < (NT - 1 - w)) ? NT - 1 - w : -1);
}
__host__ __device__ constexpr int wide_row_b(int NT, int w) {
  return (NT % 2 == 0) ? (w < NT / 2 ? w : -1) : ((w >= 1 && (w - 1) < (NT - 1 - w)) ? w - 1 : -1);
}

__host__ __device__ inline int sample_table_wide_smem(int D, int k1) {
  return smp_zblk(D, k1, false) * 8 + 4 * smp_row_len(D, false) * 8;
}

template <typename ZT>
__global__ void __launch_bounds__(128) sample_table_wide_kernel(DevView v, MemoView mv, const ZT* __restrict__ z_other,
                                                                const ZT* __restrict__ z_own) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2;
  const int D = v.D, KT = v.k1 >> 3, NS = smp_ns(D), NT = smp_nt(D);
  int jj = 0;
#pragma unroll
  for (int a = 1; a <= SSVI_MEMO_MAX_OTH; ++a)
    if (a < mv.n_modes && (int64_t)blockIdx.x >= mv.blk_base[a]) jj = a;
  const bool own = jj == mv.n_oth;                                      // the swept mode's own samples (mv.own)
  const int64_t nblk = (mv.n_rows[jj] + 3) >> 2;                        // CTAs per slot of this mode (4 rows each)
  const int64_t local = (int64_t)blockIdx.x - mv.blk_base[jj];
  const int sl = (int)(local / nblk);
  const int64_t jl = (local - (int64_t)sl * nblk) * 4 + warp;           // position inside the mode's row window
  const int64_t j = mv.row_lo[jj] + jl;
  const int slot = mv.slot_lo + sl;
  double* zs = reinterpret_cast<double*>(smem_raw);
  double* S = zs + v.zblk + (size_t)warp * v.srow_len;
  {
    const ZT* zsrc = own ? z_own + (int64_t)slot * v.zblk : z_other + ((int64_t)slot * mv.n_oth + jj) * v.zblk;
    for (int i = tid; i < v.zblk; i += 128) zs[i] = (double)zsrc[i];
  }
  const bool live = jl < mv.n_rows[jj];
  if (live) {
    const double2* src = reinterpret_cast<const double2*>(v.sampler[mv.od[jj]] + j * (int64_t)v.srow_len);
    double2* dst = reinterpret_cast<double2*>(S);
    for (int i = lane; i < (v.srow_len >> 1); i += 32) dst[i] = src[i];
  }
  __syncthreads();
  if (!live) return;
  double2* out = reinterpret_cast<double2*>(mv.tab + memo_rec(mv, jj, sl, j) * (int64_t)mv.rec_len);
  for (int i = 0; i < NT; ++i) {                        // same DMMA order per (row tile, k-tile) as mma_sample_mode
    const double mean = S[8 * i + g];
    const double* At = S + 8 * NT + 32 * (i * (i + 1)) + lane;
    double c[8][2];
#pragma unroll
    for (int t = 0; t < 8; ++t) c[t][0] = c[t][1] = mean;
    const int nj = min(2 * (i + 1), NS);
    for (int jx = 0; jx < nj; ++jx) {
      const double a = At[32 * jx];
#pragma unroll
      for (int t = 0; t < 8; ++t)
        if (t < KT) dmma_m8n8k4(c[t][0], c[t][1], a, zs[(t * NS + jx) * 32 + lane]);
    }
    if (mv.tab32 == 2) {                                // SSVI_TABLE_F32_TC: floats in the tensor-core path's order
      float* outf = reinterpret_cast<float*>(mv.tab) + memo_rec(mv, jj, sl, j) * (int64_t)mv.rec_len;
      const double a = own ? 1.0 / (double)mv.tc_alpha : (jj == 0 ? (double)mv.tc_alpha : 1.0);
#pragma unroll
      for (int t = 0; t < 8; ++t)
        if (t < KT)
          *reinterpret_cast<float2*>(outf + tc_rec_index(8 * i + g, 8 * t + 2 * (lane & 3))) =
              make_float2((float)(a * c[t][0]), (float)(a * c[t][1]));
      continue;
    }
#pragma unroll
    for (int t = 0; t < 8; ++t)
      if (t < KT) out[(i * KT + t) * 32 + lane] = make_double2(c[t][0], c[t][1]);
  }
}

// shared memory of one CTA, bytes: [ V double buffer | U | 8 partial sums ] with the prologue's staging (row + noise block
// as doubles) aliased on the V buffers
__host__ __device__ inline int wide_vbytes(int NT, int D) {
  const int vbuf = 2 * NT * SSVI_MEMO_KT * 512;
  const int stage = smp_row_len(D, false) * 8 + smp_zblk(D, 8 * SSVI_MEMO_KT, false) * 8;
  return ((vbuf > stage ? vbuf : stage) + 15) & ~15;
}
__host__ __device__ inline int wide_cta_smem(int NT, int D) { return wide_vbytes(NT, D) + NT * SSVI_MEMO_KT * 512 + 256; }

// Phase 2 of one entry for warp W: its tile rows A (and B) of G += dw V V^T, V read back from shared memory.
// acc[0 .. A] = pairs (A, i2), acc[A+1 .. A+1+B] = pairs (B, i2).
template <int NT, int W>
__device__ __forceinline__ void wide_contract(const double2* __restrict__ Vb, int lane, double dw, double (&acc)[9][2]) {
  constexpr int A = wide_row_a(NT, W), B = wide_row_b(NT, W), KT = SSVI_MEMO_KT;
  if constexpr (A >= 0) {
#pragma unroll
    for (int t = 0; t < KT; ++t) {
      double2 vt[A + 1];
#pragma unroll
      for (int i = 0; i <= A; ++i) vt[i] = Vb[(t * NT + i) * 32 + lane];
#pragma unroll
      for (int s = 0; s < 2; ++s) {
        const double aa = dw * (s ? vt[A].y : vt[A].x);
#pragma unroll
        for (int i2 = 0; i2 <= A; ++i2) dmma_m8n8k4(acc[i2][0], acc[i2][1], aa, s ? vt[i2].y : vt[i2].x);
        if constexpr (B >= 0) {
          const double ab = dw * (s ? vt[B].y : vt[B].x);
#pragma unroll
          for (int i2 = 0; i2 <= B; ++i2)
            dmma_m8n8k4(acc[A + 1 + i2][0], acc[A + 1 + i2][1], ab, s ? vt[i2].y : vt[i2].x);
        }
      }
    }
  }
}

// warp W's pairs -> the chunk record's Di block (lower triangle incl. diagonal, mirrored: exactly symmetric)
template <int NT, int W>
__device__ __forceinline__ void wide_store(double* __restrict__ out, int D, int lane, const double (&acc)[9][2]) {
  constexpr int A = wide_row_a(NT, W), B = wide_row_b(NT, W);
  const int g = lane >> 2, q = lane & 3;
  if constexpr (A >= 0) {
#pragma unroll
    for (int p = 0; p < A + 1 + (B >= 0 ? B + 1 : 0); ++p) {
      const int i1 = p <= A ? A : B, i2 = p <= A ? p : p - A - 1;
      const int n1 = 8 * i1 + g, n2 = 8 * i2 + 2 * q;
      if (n1 < D && n2 <= n1) {
        out[D + n1 * D + n2] = acc[p][0];
        out[D + n2 * D + n1] = acc[p][0];
      }
      if (n1 < D && n2 + 1 <= n1) {
        out[D + n1 * D + n2 + 1] = acc[p][1];
        out[D + (n2 + 1) * D + n1] = acc[p][1];
      }
    }
  }
}

template <int NT, int NOTH>
__global__ void __launch_bounds__(128, 2) accumulate_wide_kernel(
    DevView v, MemoView mv, int64_t n_chunks, const int32_t* __restrict__ chunk_row,
    const int64_t* __restrict__ chunk_start, const int32_t* __restrict__ chunk_len,
    const int32_t* __restrict__ chunk_slot, const int64_t* __restrict__ chunk_out,
    const int64_t* __restrict__ entry_ids, const int64_t* __restrict__ rows, const void* __restrict__ z_own,
    int z_f32, double* __restrict__ partial_out) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr int KT = SSVI_MEMO_KT;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, q = lane & 3;
  const int64_t c = blockIdx.x;
  if (c >= n_chunks) return;
  const int N = v.N, D = v.D, NS = smp_ns(D);
  const int64_t row = rows[chunk_row[c]];
  const int64_t start = chunk_start[c];
  const int len = chunk_len[c], slot = chunk_slot[c];
  const int sl = slot - mv.slot_lo;

  const int vbytes = wide_vbytes(NT, D);
  double* S = reinterpret_cast<double*>(smem_raw);                       // prologue staging, aliased on the V buffers
  double* zs = S + v.srow_len;
  double2* Vs = reinterpret_cast<double2*>(smem_raw);                    // [2][KT][NT][32]
  double2* Us = reinterpret_cast<double2*>(smem_raw + vbytes);           // [NT][KT][32]
  double2* part = reinterpret_cast<double2*>(smem_raw + vbytes + NT * KT * 512);   // [2][4] (sum_a, sum_d)
  double* red = reinterpret_cast<double*>(smem_raw);                     // [4][NT * 8] at the very end

  // ---- prologue: U(row, slot) = m + L z_own, once per CTA (ui samples, SSVI_TF_interface.py:251-254) --------------------
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
  for (int i = 0; i < NT; ++i)
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      double c0, c1;
      mma_sample_tile(S, zs, NT, NS, lane, g, i, warp + 4 * h, c0, c1);
      Us[(i * KT + warp + 4 * h) * 32 + lane] = make_double2(c0, c1);
    }
  __syncthreads();                                      // U complete; the staging region becomes the V buffers

  const double2* tab2 = reinterpret_cast<const double2*>(mv.tab);
  const int rec2 = mv.rec_len >> 1;
  const double inv_k1 = 1.0 / (double)(8 * KT);
  double acc[9][2];
#pragma unroll
  for (int p = 0; p < 9; ++p) acc[p][0] = acc[p][1] = 0.0;
  double dacc[NT];
#pragma unroll
  for (int i = 0; i < NT; ++i) dacc[i] = 0.0;

  for (int b0 = 0; b0 < len; b0 += 32) {
    const int nb = min(32, len - b0);
    double my_y = 0.0;
    int32_t my_rec[NOTH];
#pragma unroll
    for (int jj = 0; jj < NOTH; ++jj) my_rec[jj] = 0;
    if (lane < nb) {
      const int64_t eid = entry_ids[start + b0 + lane];
      my_y = v.vals[eid];
#pragma unroll
      for (int jj = 0; jj < NOTH; ++jj) my_rec[jj] = (int32_t)memo_rec(mv, jj, sl, v.coords[eid * N + mv.od[jj]]);
    }
    for (int el = 0; el < nb; ++el) {
      const int buf = (b0 + el) & 1;
      const double y = __shfl_sync(0xffffffffu, my_y, el);
      const double2* pc[NOTH];
#pragma unroll
      for (int jj = 0; jj < NOTH; ++jj)
        pc[jj] = tab2 + (int64_t)__shfl_sync(0xffffffffu, my_rec[jj], el) * rec2 + lane;
      // ---- phase 1: this warp's two k-tiles of V, its share of f and of the derivative sums -----------------------------
      double rs[NT];
#pragma unroll
      for (int i = 0; i < NT; ++i) rs[i] = 0.0;
      double sa = 0.0, sd = 0.0;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int t = warp + 4 * h;
        double2 x[NOTH][NT];
#pragma unroll
        for (int jj = 0; jj < NOTH; ++jj)
#pragma unroll
          for (int i = 0; i < NT; ++i) x[jj][i] = ld_tab(pc[jj] + (i * KT + t) * 32);
        double2 fp = make_double2(0.0, 0.0);
#pragma unroll
        for (int i = 0; i < NT; ++i) {
          double2 vv = x[0][i];
#pragma unroll
          for (int jj = 1; jj < NOTH; ++jj) {
            vv.x *= x[jj][i].x;
            vv.y *= x[jj][i].y;
          }
          Vs[((buf * KT + t) * NT + i) * 32 + lane] = vv;
          const double2 u = Us[(i * KT + t) * 32 + lane];
          fp.x = fma(u.x, vv.x, fp.x);
          fp.y = fma(u.y, vv.y, fp.y);
          rs[i] += vv.x + vv.y;
        }
#pragma unroll
        for (int o = 4; o < 32; o <<= 1) {              // sum over the 8 row groups: f_k for k = 8t + 2q + {0, 1}
          fp.x += __shfl_xor_sync(0xffffffffu, fp.x, o);
          fp.y += __shfl_xor_sync(0xffffffffu, fp.y, o);
        }
        double a0, b0d, a1, b1d;
        log_pdf_derivs_pair(v.lik, y, fp.x, fp.y, v.lik_param, a0, b0d, a1, b1d);
        sa += a0 + a1;
        sd += b0d + b1d;
      }
      if (g != 0) sa = sd = 0.0;                        // every row group holds the same f: count each sample once
      warp_sum2(sa, sd, lane);
      if (lane == 0) part[buf * 4 + warp] = make_double2(sa, sd);
      __syncthreads();                                  // V[buf] and the four partial sums are visible
      // ---- SSVI_TF_d.py:31-57 (Q3): derivatives averaged over k1, then broadcast ------------------------------------------
      const double2 p0 = part[buf * 4], p1 = part[buf * 4 + 1], p2 = part[buf * 4 + 2], p3 = part[buf * 4 + 3];
      const double sum_a = (((p0.x + p1.x) + p2.x) + p3.x) * inv_k1, sum_d = (((p0.y + p1.y) + p2.y) + p3.y) * inv_k1;
      const double aw = sum_a * inv_k1, dw = 0.5 * sum_d * inv_k1;
#pragma unroll
      for (int i = 0; i < NT; ++i) dacc[i] = fma(aw, rs[i], dacc[i]);
      // ---- phase 2: this warp's tile rows of G += dw V V^T ----------------------------------------------------------------
      const double2* Vb = Vs + buf * KT * NT * 32;
      switch (warp) {
        case 0: wide_contract<NT, 0>(Vb, lane, dw, acc); break;
        case 1: wide_contract<NT, 1>(Vb, lane, dw, acc); break;
        case 2: wide_contract<NT, 2>(Vb, lane, dw, acc); break;
        default: wide_contract<NT, 3>(Vb, lane, dw, acc); break;
      }
    }
  }

  // ---- chunk record [di | Di | si_sum | count] ------------------------------------------------------------------------
  double* out = partial_out + chunk_out[c] * (int64_t)v.acc_len;
  switch (warp) {
    case 0: wide_store<NT, 0>(out, D, lane, acc); break;
    case 1: wide_store<NT, 1>(out, D, lane, acc); break;
    case 2: wide_store<NT, 2>(out, D, lane, acc); break;
    default: wide_store<NT, 3>(out, D, lane, acc); break;
  }
  __syncthreads();                                      // every warp is done reading V: the region becomes `red`
#pragma unroll
  for (int i = 0; i < NT; ++i) {
    double xs = dacc[i];
    xs += __shfl_xor_sync(0xffffffffu, xs, 1);
    xs += __shfl_xor_sync(0xffffffffu, xs, 2);
    if (q == 0) red[warp * (NT * 8) + 8 * i + g] = xs;
  }
  __syncthreads();
  for (int n = tid; n < D; n += 128)
    out[n] = ((red[n] + red[NT * 8 + n]) + red[2 * NT * 8 + n]) + red[3 * NT * 8 + n];
  if (tid == 0) {
    out[D + D * D] = 0.0;
    out[D + D * D + 1] = (double)len;
  }
}
