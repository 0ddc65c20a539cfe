// Stages 1-3, full covariance, D <= 32, k1 % 8 == 0, k1 <= 64: the FP64 tensor-pipe path.
//
// Measured on B200 (profiles/r01_microbench_fp64_pipes.txt): DFMA 35.8 TFLOP/s, DMMA (mma.sync m8n8k4
// f64) 37.1 TFLOP/s, interleaved 33 -- ONE pipe.  The two dense contractions of this path
//     sampling     X^T (D x k1)  = L (D x D, lower) . Z^T (D x k1) + m         per (entry, mode)
//     contraction  G   (D x D)  += V^T (D x k1) . diag(w) . V (k1 x D)          per entry
// are issued as DMMA: one instruction does 256 FMAs, and the operand layouts below make every
// fragment one conflict-free 32-element shared-memory read (sampling) or a value already sitting in
// the lane's registers (contraction).  tcgen05 has no FP64 kind, so the 5th-gen tensor path does not
// apply to an fp64 contract; TMA does: operands arrive by cp.async.bulk + mbarrier.
//
// One warp owns one row-chunk and runs its own 2-stage pipeline over "mode steps":
//     step (e, jj): gathered sampler row of other-mode jj of entry e + its noise block   (bulk copies)
//     step (e, own): the own-row noise block                  (the own row is loaded once per chunk)
// Lane 0 issues the bulk copies of step u+1 before the warp computes step u; completion is tracked
// by one mbarrier per stage.  Entry ids / coordinates are prefetched one entry ahead.
//
// With g = lane>>2, q = lane&3 the m8n8k4 fragments are
//     A (8x4,row)  a  = A[g][q]        B (4x8,col)  b = B[q][g]        C (8x8)  c0,c1 = C[g][2q], C[g][2q+1]
// Sampling tile (n-tile i, k-tile t):  A = L[8i+g][4j+q], B = z[4j+q][8t+g], C = x[n=8i+g][k=8t+2q+s];
// the sampler cache stores L tile by tile in A-fragment order and the noise pool stores z in
// B-fragment order (ssvi_common.cuh), so both are read as base + lane.  Only k-steps j <= 2i+1 touch
// the lower triangle.  The Hadamard product over modes is an element-wise product of C fragments: V
// lives in registers as vv[i][t][s].  For the contraction the SAME registers are both operands: with
// the k-slot permutation (slot q of step (t,s)) <-> k = 8t+2q+s,  A = V^T[8*i1+g][slot q] = vv[i1][t][s]
// and B = (wV)[slot q][8*i2+g] = w_k * vv[i2][t][s].
#pragma once
#include "accumulate_generic.cuh"

__device__ __forceinline__ void dmma_m8n8k4(double& c0, double& c1, double a, double b) {
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
      : "+d"(c0), "+d"(c1)
      : "d"(a), "d"(b));
}

// ---- mbarrier / bulk-copy (TMA 1-D) helpers ------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "LAB_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra DONE;\n"
      "bra LAB_WAIT;\n"
      "DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}

__device__ __forceinline__ double zfrag(const double* p, int i) { return p[i]; }
__device__ __forceinline__ double zfrag(const float* p, int i) { return (double)p[i]; }

// x = m + L z for every (n, k) of the warp's fragment layout, operands from shared memory.
//   OWN = false: vv <- (first ? x : vv * x)            (sample_vjs_batch, SSVI_TF_interface.py:732-750)
//   OWN = true : fp[t][s] += x * vv                     (ui samples and mean_batch, :251-257)
template <int NT, typename ZT, bool OWN>
__device__ __forceinline__ void mma_sample_mode(const double* __restrict__ S, const ZT* __restrict__ z, int NS,
                                                int KT, int lane, int g, bool first, double (&vv)[NT][8][2],
                                                double (&fp)[8][2]) {
#pragma unroll
  for (int i = 0; i < NT; ++i) {
    const double mean = S[8 * i + g];
    const double* At = S + 8 * NT + 32 * (i * (i + 1)) + lane;     // tiles of row-tile i start at i(i+1)
    double c[8][2];
#pragma unroll
    for (int t = 0; t < 8; ++t) c[t][0] = c[t][1] = mean;
#pragma unroll
    for (int j = 0; j < 2 * (i + 1); ++j) {
      if (j < NS) {
        const double a = At[32 * j];
#pragma unroll
        for (int t = 0; t < 8; ++t)
          if (t < KT) dmma_m8n8k4(c[t][0], c[t][1], a, zfrag(z, (t * NS + j) * 32 + lane));
      }
    }
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      if (OWN) {
        fp[t][0] += c[t][0] * vv[i][t][0];
        fp[t][1] += c[t][1] * vv[i][t][1];
      } else if (first) {
        vv[i][t][0] = c[t][0];
        vv[i][t][1] = c[t][1];
      } else {
        vv[i][t][0] *= c[t][0];
        vv[i][t][1] *= c[t][1];
      }
    }
  }
}

// Shared memory per warp, bytes (every section a multiple of 16):
//   2 mbarriers | own row | 2 x gathered row | 2 x noise block | G [NP][64] | F, AW, DW [64] | V parking
struct MmaSmem {
  int row_bytes, z_bytes, g_bytes, park_bytes, total;
};
__host__ __device__ inline MmaSmem mma_smem_layout(int NT, int D, int k1, int zsize, bool park) {
  MmaSmem m;
  m.row_bytes = smp_row_len(D, false) * 8;
  m.z_bytes = ((smp_zblk(D, k1, false) * zsize + 15) / 16) * 16;
  m.g_bytes = (NT * (NT + 1) / 2) * 64 * 8;
  m.park_bytes = park ? NT * 16 * 32 * 8 : 0;
  m.total = 16 + 3 * m.row_bytes + 2 * m.z_bytes + m.g_bytes + 3 * 64 * 8 + m.park_bytes;
  return m;
}

// SECOND = false: SSVI_TF_d (no second sampling layer);  true: SSVI_TF_simple / SSVI_TF_robust
template <int NT, typename ZT, bool SECOND>
__global__ void __launch_bounds__(256) accumulate_mma_kernel(
    DevView v, int64_t n_chunks, const int32_t* __restrict__ chunk_row, const int64_t* __restrict__ chunk_start,
    const int32_t* __restrict__ chunk_len, const int32_t* __restrict__ chunk_slot0,
    const int64_t* __restrict__ entry_ids, const int64_t* __restrict__ rows, const double* __restrict__ w_state,
    const ZT* __restrict__ z_other, const ZT* __restrict__ z_own, const ZT* __restrict__ r_w,
    const ZT* __restrict__ z_f, double* __restrict__ partial_out) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr int NP = NT * (NT + 1) / 2;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  const int g = lane >> 2, q = lane & 3;
  const int64_t c = (int64_t)blockIdx.x * wpb + warp;
  if (c >= n_chunks) return;
  const int N = v.N, D = v.D, k1 = v.k1, KT = v.k1 >> 3, dim = v.dim, NS = smp_ns(D);
  const MmaSmem L = mma_smem_layout(NT, D, k1, (int)sizeof(ZT), SECOND);
  unsigned char* base = smem_raw + (size_t)warp * L.total;
  uint64_t* bars = reinterpret_cast<uint64_t*>(base);
  double* own = reinterpret_cast<double*>(base + 16);
  unsigned char* rowbuf = base + 16 + L.row_bytes;                 // 2 stages
  unsigned char* zbuf = rowbuf + 2 * L.row_bytes;                  // 2 stages
  double* Gs = reinterpret_cast<double*>(zbuf + 2 * L.z_bytes);
  double* F = Gs + NP * 64;
  double* AW = F + 64;
  double* DW = AW + 64;
  double* park = DW + 64;

  const int64_t row = rows[chunk_row[c]];
  const int64_t start = chunk_start[c];
  const int len = chunk_len[c], slot0 = chunk_slot0[c];
  const double w_sigma = SECOND ? w_state[0] : 1.0;
  const uint32_t row_bytes = (uint32_t)L.row_bytes, z_bytes = (uint32_t)(v.zblk * (int)sizeof(ZT));

  if (lane == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  {
    const double* src = v.sampler[dim] + row * (int64_t)v.srow_len;
    for (int i = lane; i < v.srow_len; i += 32) own[i] = src[i];
  }
#pragma unroll
  for (int p = 0; p < NP; ++p) {
    Gs[p * 64 + lane * 2] = 0.0;
    Gs[p * 64 + lane * 2 + 1] = 0.0;
  }
  __syncwarp();

  // other modes in ascending order
  int od[SSVI_MAX_ORDER - 1];
  {
    int jj = 0;
#pragma unroll
    for (int d = 0; d < SSVI_MAX_ORDER; ++d)
      if (d < N && d != dim) od[jj++] = d;
  }

  // entry prefetch registers (one entry ahead): id, value, coordinates of the other modes
  int64_t eid_n = entry_ids[start];
  double y_n = v.vals[eid_n];
  int32_t col_n[SSVI_MAX_ORDER - 1];
#pragma unroll
  for (int jj = 0; jj < SSVI_MAX_ORDER - 1; ++jj)
    if (jj < N - 1) col_n[jj] = v.coords[eid_n * N + od[jj]];

  uint32_t u = 0;                                    // global step counter: stage = u & 1, parity = (u >> 1) & 1
  if (lane == 0) {                                   // issue step 0: (entry 0, other-mode 0)
    const int slot = slot0 % v.pool_slots;
    mbar_expect_tx(&bars[0], row_bytes + z_bytes);
    bulk_g2s(rowbuf, v.sampler[od[0]] + (int64_t)col_n[0] * v.srow_len, row_bytes, &bars[0]);
    bulk_g2s(zbuf, z_other + ((int64_t)slot * (N - 1)) * v.zblk, z_bytes, &bars[0]);
  }

  double gp[NT];
#pragma unroll
  for (int i = 0; i < NT; ++i) gp[i] = 0.0;
  double si_part = 0.0;
  bool bad = false;

  for (int e = 0; e < len; ++e) {
    const int slot = (slot0 + e) % v.pool_slots;
    const double y = y_n;
    int32_t col[SSVI_MAX_ORDER - 1];
#pragma unroll
    for (int jj = 0; jj < SSVI_MAX_ORDER - 1; ++jj) col[jj] = col_n[jj];
    const bool more = e + 1 < len;
    if (more) {                                      // prefetch the next entry's indices
      eid_n = entry_ids[start + e + 1];
      y_n = v.vals[eid_n];
#pragma unroll
      for (int jj = 0; jj < SSVI_MAX_ORDER - 1; ++jj)
        if (jj < N - 1) col_n[jj] = v.coords[eid_n * N + od[jj]];
    }
    const double yfact =
        (SECOND && v.lik == SSVI_LIK_POISSON && v.model == SSVI_MODEL_ROBUST) ? 1.0 / factorial_ref(y) : 1.0;
    double vv[NT][8][2];
    double fp[8][2];
#pragma unroll
    for (int t = 0; t < 8; ++t) fp[t][0] = fp[t][1] = 0.0;

    // ---- stages 1+2: N-1 other modes, then the own row ------------------------------------------------
    for (int jj = 0; jj < N; ++jj) {
      const uint32_t st = u & 1u, nst = st ^ 1u;
      __syncwarp();                                  // every lane is done reading stage `nst` (step u-1)
      if (lane == 0) {                               // issue step u+1 into the other stage
        if (jj + 1 < N - 1) {
          mbar_expect_tx(&bars[nst], row_bytes + z_bytes);
          bulk_g2s(rowbuf + nst * L.row_bytes, v.sampler[od[jj + 1]] + (int64_t)col[jj + 1] * v.srow_len, row_bytes,
                   &bars[nst]);
          bulk_g2s(zbuf + nst * L.z_bytes, z_other + ((int64_t)slot * (N - 1) + jj + 1) * v.zblk, z_bytes,
                   &bars[nst]);
        } else if (jj + 1 == N - 1) {
          mbar_expect_tx(&bars[nst], z_bytes);
          bulk_g2s(zbuf + nst * L.z_bytes, z_own + (int64_t)slot * v.zblk, z_bytes, &bars[nst]);
        } else if (more) {
          const int nslot = (slot0 + e + 1) % v.pool_slots;
          mbar_expect_tx(&bars[nst], row_bytes + z_bytes);
          bulk_g2s(rowbuf + nst * L.row_bytes, v.sampler[od[0]] + (int64_t)col_n[0] * v.srow_len, row_bytes,
                   &bars[nst]);
          bulk_g2s(zbuf + nst * L.z_bytes, z_other + ((int64_t)nslot * (N - 1)) * v.zblk, z_bytes, &bars[nst]);
        }
      }
      mbar_wait(&bars[st], (u >> 1) & 1u);
      const ZT* zs = reinterpret_cast<const ZT*>(zbuf + st * L.z_bytes);
      if (jj < N - 1) {
        const double* S = reinterpret_cast<const double*>(rowbuf + st * L.row_bytes);
        mma_sample_mode<NT, ZT, false>(S, zs, NS, KT, lane, g, jj == 0, vv, fp);
      } else {
        mma_sample_mode<NT, ZT, true>(own, zs, NS, KT, lane, g, false, vv, fp);
      }
      ++u;
    }
    // ---- predictive mean f_k = u_k . v_k: reduce over the 8 row groups -----------------------------------
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
    if (SECOND) {                                    // park V: the k2 loop below wants the register file
#pragma unroll
      for (int i = 0; i < NT; ++i)
#pragma unroll
        for (int t = 0; t < 8; ++t)
          *reinterpret_cast<double2*>(&park[((i * 8 + t) * 32 + lane) * 2]) = make_double2(vv[i][t][0], vv[i][t][1]);
    }
    __syncwarp();
    // ---- likelihood weights, lanes over k -------------------------------------------------------------------
    if (SECOND) {
      for (int k0 = 0; k0 < k1; k0 += 32) {
        const int k = k0 + lane;
        const bool active = k < k1;
        const double f = active ? F[k] : 0.0;
        double aw, dw;
        likelihood_weights<ZT>(v, r_w, z_f, slot, k, active, y, yfact, f, w_sigma, aw, dw, si_part, bad);
        if (active) {
          AW[k] = aw;
          DW[k] = dw;
        }
      }
    } else {                                         // SSVI_TF_d.py:31-57, Q3: average over k1, then broadcast
      double sum_a = 0.0, sum_d = 0.0;
      for (int k = lane; k < k1; k += 32) {
        double a, b;
        log_pdf_derivs(v.lik, y, F[k], v.lik_param, a, b);
        sum_a += a;
        sum_d += b;
      }
      sum_a = warp_sum(sum_a) / (double)k1;
      sum_d = warp_sum(sum_d) / (double)k1;
      for (int k = lane; k < k1; k += 32) {
        AW[k] = sum_a / (double)k1;
        DW[k] = 0.5 * sum_d / (double)k1;
      }
    }
    __syncwarp();
    if (SECOND) {
#pragma unroll
      for (int i = 0; i < NT; ++i)
#pragma unroll
        for (int t = 0; t < 8; ++t) {
          const double2 x = *reinterpret_cast<const double2*>(&park[((i * 8 + t) * 32 + lane) * 2]);
          vv[i][t][0] = x.x;
          vv[i][t][1] = x.y;
        }
    }
    // ---- stage 3: G += V^T diag(w) V, g += V^T a, operands straight from the V registers -------------------
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
  }

  // ---- chunk record [di | Di | si_sum | count] ---------------------------------------------------------------
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
