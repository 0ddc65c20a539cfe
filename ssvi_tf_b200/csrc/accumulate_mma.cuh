// Stages 1-3, full covariance, D <= 32, k1 % 8 == 0, k1 <= 64: the FP64 tensor-pipe path.
//
// Measured on B200 (profiles/r01_microbench_fp64_pipes.txt): DFMA 35.8 TFLOP/s, DMMA (mma.sync m8n8k4
// f64) 37.1 TFLOP/s, interleaved 33 -- ONE pipe.  The two dense contractions of this path
//     sampling     X^T (D x k1)  = L (D x D, lower) . Z^T (D x k1) + m         per (entry, mode)
//     contraction  G   (D x D)  += V^T (D x k1) . diag(w) . V (k1 x D)          per entry
// are therefore issued as DMMA: one instruction does 256 FMAs with operands that never leave the
// register file, where a DFMA formulation spends most issue slots on shared-memory operand traffic.
// tcgen05 has no FP64 kind, so the 5th-gen tensor path does not apply to an fp64 contract.
//
// One warp owns one row-chunk.  With g = lane>>2, q = lane&3 the m8n8k4 fragments are
//     A (8x4,row)  a  = A[g][q]        B (4x8,col)  b = B[q][g]        C (8x8)  c0,c1 = C[g][2q], C[g][2q+1]
// Sampling tile (n-tile i, k-tile t):  A = L[8i+g][4j+q]  (straight from the gathered sampler row in
// global/L2: every element is needed by exactly one lane), B = z[4j+q][8t+g] (noise pool, L1/L2),
// C = x[n = 8i+g][k = 8t+2q+s].  Only the k-steps j <= 2i+1 touch the lower triangle.
// The Hadamard product over modes is an element-wise multiply of C fragments: V lives in registers
// as vv[i][t][s].  For the contraction the SAME registers are both operands: with the k-slot
// permutation (slot q of step (t,s)) <-> k = 8t+2q+s,  A = V^T[8*i1+g][slot q] = vv[i1][t][s] and
// B = (wV)[slot q][8*i2+g] = w_k * vv[i2][t][s] -- no shuffles, no shared memory.
#pragma once
#include "accumulate_generic.cuh"

__device__ __forceinline__ void dmma_m8n8k4(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

// x = m + L z for every (n, k) of the warp's fragment layout.
//   OWN = false: vv <- (first ? x : vv * x)            (sample_vjs_batch, SSVI_TF_interface.py:732-750)
//   OWN = true : fp[t][s] += x * vv                     (ui samples and mean_batch, :251-257)
template <int NT, typename ZT, bool OWN>
__device__ __forceinline__ void mma_sample_mode(const double* __restrict__ S, const ZT* __restrict__ z, int D,
                                                int k1, int KT, int g, int q, bool first,
                                                double (&vv)[NT][8][2], double (&fp)[8][2]) {
  const int NS = (D + 3) >> 2;
#pragma unroll
  for (int i = 0; i < NT; ++i) {
    const int n = 8 * i + g;
    const bool nvalid = n < D;
    const double mean = nvalid ? S[n] : 0.0;
    const double* Ln = S + D + n * D;
    double a[2 * NT];
#pragma unroll
    for (int j = 0; j < 2 * (i + 1); ++j) {
      const int c = 4 * j + q;
      a[j] = (nvalid && c < D) ? Ln[c] : 0.0;
    }
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      if (t < KT) {
        double c0 = mean, c1 = mean;
#pragma unroll
        for (int j = 0; j < 2 * (i + 1); ++j) {
          if (j < NS) {
            const int c = 4 * j + q;
            const double b = (c < D) ? ldz(z, (int64_t)c * k1 + 8 * t + g) : 0.0;
            dmma_m8n8k4(c0, c1, a[j], b);
          }
        }
        if (OWN) {
          fp[t][0] += c0 * vv[i][t][0];
          fp[t][1] += c1 * vv[i][t][1];
        } else if (first) {
          vv[i][t][0] = c0;
          vv[i][t][1] = c1;
        } else {
          vv[i][t][0] *= c0;
          vv[i][t][1] *= c1;
        }
      }
    }
  }
}

// shared memory per warp (doubles): G accumulators [NP][64] + F[64] + AW[64] + DW[64]
// + a parking area for V (NT*8*2*32): V is parked while the likelihood code runs, which frees ~100 registers
template <int NT, bool SECOND>
__host__ __device__ constexpr int mma_warp_doubles() {
  return (NT * (NT + 1) / 2) * 64 + 192 + NT * 16 * 32;
}

// SECOND = false: SSVI_TF_d (no second sampling layer);  true: SSVI_TF_simple / SSVI_TF_robust
template <int NT, typename ZT, bool SECOND>
__global__ void __launch_bounds__(128) accumulate_mma_kernel(
    DevView v, int64_t n_chunks, const int32_t* __restrict__ chunk_row, const int64_t* __restrict__ chunk_start,
    const int32_t* __restrict__ chunk_len, const int32_t* __restrict__ chunk_slot0,
    const int64_t* __restrict__ entry_ids, const int64_t* __restrict__ rows, const double* __restrict__ w_state,
    const ZT* __restrict__ z_other, const ZT* __restrict__ z_own, const ZT* __restrict__ r_w,
    const ZT* __restrict__ z_f, double* __restrict__ partial_out) {
  extern __shared__ double smem[];
  constexpr int NP = NT * (NT + 1) / 2;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  const int g = lane >> 2, q = lane & 3;
  const int64_t c = (int64_t)blockIdx.x * wpb + warp;
  if (c >= n_chunks) return;
  const int N = v.N, D = v.D, k1 = v.k1, KT = v.k1 >> 3, dim = v.dim, row_len = v.row_len;
  double* Gs = smem + (size_t)warp * mma_warp_doubles<NT, SECOND>();
  double* F = Gs + NP * 64;
  double* AW = F + 64;
  double* DW = AW + 64;
  double* park = DW + 64;

  const int64_t row = rows[chunk_row[c]];
  const int64_t start = chunk_start[c];
  const int len = chunk_len[c], slot0 = chunk_slot0[c];
  const double w_sigma = (v.model == SSVI_MODEL_D) ? 1.0 : w_state[0];
  const double* Sown = v.sampler[dim] + row * row_len;

#pragma unroll
  for (int p = 0; p < NP; ++p) {
    Gs[p * 64 + lane * 2] = 0.0;
    Gs[p * 64 + lane * 2 + 1] = 0.0;
  }
  double gp[NT];
#pragma unroll
  for (int i = 0; i < NT; ++i) gp[i] = 0.0;
  double si_part = 0.0;
  bool bad = false;

  for (int e = 0; e < len; ++e) {
    const int64_t eid = entry_ids[start + e];
    const int slot = (slot0 + e) % v.pool_slots;
    const double y = v.vals[eid];
    const double yfact = (v.lik == SSVI_LIK_POISSON && v.model == SSVI_MODEL_ROBUST) ? factorial_ref(y) : 1.0;
    double vv[NT][8][2];
    double fp[8][2];
#pragma unroll
    for (int t = 0; t < 8; ++t) fp[t][0] = fp[t][1] = 0.0;
    // ---- stages 1+2: gather + sample every other mode, Hadamard product in registers ---------------
    int jj = 0;
    for (int d = 0; d < N; ++d) {
      if (d == dim) continue;
      const int64_t col = v.coords[eid * N + d];
      const double* S = v.sampler[d] + col * row_len;
      const ZT* z = z_other + ((int64_t)slot * (N - 1) + jj) * D * k1;
      mma_sample_mode<NT, ZT, false>(S, z, D, k1, KT, g, q, jj == 0, vv, fp);
      ++jj;
    }
    // ---- own-row sample and predictive mean f_k = u_k . v_k -----------------------------------------
    mma_sample_mode<NT, ZT, true>(Sown, z_own + (int64_t)slot * D * k1, D, k1, KT, g, q, false, vv, fp);
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      if (t < KT) {
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          double x = fp[t][s];
          x += __shfl_xor_sync(0xffffffffu, x, 4);
          x += __shfl_xor_sync(0xffffffffu, x, 8);
          x += __shfl_xor_sync(0xffffffffu, x, 16);
          fp[t][s] = x;
        }
        if (g == 0) *reinterpret_cast<double2*>(&F[8 * t + 2 * q]) = make_double2(fp[t][0], fp[t][1]);
      }
    }
    {                                     // park V: the likelihood code below wants the register file
#pragma unroll
      for (int i = 0; i < NT; ++i)
#pragma unroll
        for (int t = 0; t < 8; ++t)
          *reinterpret_cast<double2*>(&park[((i * 8 + t) * 32 + lane) * 2]) = make_double2(vv[i][t][0], vv[i][t][1]);
    }
    __syncwarp();
    // ---- likelihood weights, lanes over k (same code as the generic path) ---------------------------
    double sum_a = 0.0, sum_d = 0.0;
    for (int k0 = 0; k0 < k1; k0 += 32) {
      const int k = k0 + lane;
      const bool active = k < k1;
      const double f = active ? F[k] : 0.0;
      double aw, dw;
      likelihood_weights<ZT>(v, r_w, z_f, slot, k, active, y, yfact, f, w_sigma, aw, dw, si_part, bad);
      if (!SECOND) {
        sum_a += aw;
        sum_d += dw;
      } else if (active) {
        AW[k] = aw;
        DW[k] = dw;
      }
    }
    if (!SECOND) {                        // Q3: average over k1, then broadcast
      sum_a = warp_sum(sum_a) / (double)k1;
      sum_d = warp_sum(sum_d) / (double)k1;
      for (int k = lane; k < k1; k += 32) {
        AW[k] = sum_a / (double)k1;
        DW[k] = 0.5 * sum_d / (double)k1;
      }
    }
    __syncwarp();
    {
#pragma unroll
      for (int i = 0; i < NT; ++i)
#pragma unroll
        for (int t = 0; t < 8; ++t) {
          const double2 x = *reinterpret_cast<const double2*>(&park[((i * 8 + t) * 32 + lane) * 2]);
          vv[i][t][0] = x.x;
          vv[i][t][1] = x.y;
        }
    }
    // ---- stage 3: G += V^T diag(w) V, g += V^T a, operands straight from the V registers -------------
    double acc[NP][2];
#pragma unroll
    for (int p = 0; p < NP; ++p) acc[p][0] = acc[p][1] = 0.0;
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      if (t < KT) {
        const double2 aw2 = *reinterpret_cast<const double2*>(&AW[8 * t + 2 * q]);
        const double2 dw2 = *reinterpret_cast<const double2*>(&DW[8 * t + 2 * q]);
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          const double aw = s ? aw2.y : aw2.x, dw = s ? dw2.y : dw2.x;
          double wv[NT];
#pragma unroll
          for (int i = 0; i < NT; ++i) {
            wv[i] = dw * vv[i][t][s];
            gp[i] += aw * vv[i][t][s];
          }
          int p = 0;
#pragma unroll
          for (int i1 = 0; i1 < NT; ++i1)
#pragma unroll
            for (int i2 = 0; i2 <= i1; ++i2) {
              dmma_m8n8k4(acc[p][0], acc[p][1], vv[i1][t][s], wv[i2]);
              ++p;
            }
        }
      }
    }
#pragma unroll
    for (int p = 0; p < NP; ++p) {
      double2* gs = reinterpret_cast<double2*>(&Gs[p * 64 + lane * 2]);
      double2 o = *gs;
      o.x += acc[p][0];
      o.y += acc[p][1];
      *gs = o;
    }
    __syncwarp();
  }

  // ---- chunk record [di | Di | si_sum | count] ---------------------------------------------------------
  double* out = partial_out + c * (int64_t)v.acc_len;
#pragma unroll
  for (int i = 0; i < NT; ++i) {
    double x = gp[i];
    x += __shfl_xor_sync(0xffffffffu, x, 1);
    x += __shfl_xor_sync(0xffffffffu, x, 2);
    const int n = 8 * i + g;
    if (q == 0 && n < D) out[n] = x;
  }
  {
    int p = 0;
#pragma unroll
    for (int i1 = 0; i1 < NT; ++i1)
#pragma unroll
      for (int i2 = 0; i2 <= i1; ++i2) {
        const int n1 = 8 * i1 + g;
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          const int n2 = 8 * i2 + 2 * q + s;
          if (n1 < D && n2 <= n1) {          // lower triangle (incl. diagonal), mirrored: exactly symmetric
            const double x = Gs[p * 64 + lane * 2 + s];
            out[D + n1 * D + n2] = x;
            out[D + n2 * D + n1] = x;
          }
        }
        ++p;
      }
  }
  si_part = warp_sum(si_part);
  const unsigned anybad = __ballot_sync(0xffffffffu, bad);
  if (lane == 0) {
    out[D + D * D] = si_part;
    out[D + D * D + 1] = (double)len;
    if (anybad) atomicOr(v.flags, SSVI_FLAG_PHI_ZERO);
  }
}
