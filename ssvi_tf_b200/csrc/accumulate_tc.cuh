// Stages 1-3 for FULL SWEEPS of SSVI_TF_d, full covariance, D <= 32, k1 == 64 -- the fp32 option of the contract
// (SSVI_TABLE_F32_TC; BASELINE.json: 1e-4 in fp32) with the contraction on the tcgen05 tensor cores.
//
// Same slot-major sample-table scheme, chunk tables and chunk records as accumulate_memo.cuh; what differs is the
// arithmetic.  The table holds float samples X(j, s) in the lane order of tc_rec_index (tc_ops.cuh).  One CTA (4 warps)
// per chunk; warp w takes entries w, w + 4, ...  ("quad" q = the four entries 4q .. 4q + 3).  Per entry a warp
//   * gathers the N - 1 records (two coalesced 16-byte loads per lane and item, one item ahead), forms V = Hadamard
//     product in fp32 registers, f_k = u_k . v_k against the own-row sample U(row, slot) held in registers (evaluated once
//     per CTA with the DMMA sequence of every other kernel), and the row sums SUM_k V[n][k]      (SSVI_TF_d.py:31-46);
//   * writes rows 32 w .. 32 w + 31 of the quad's operand tile: a V split into fp16 hi + lo (22 significand bits), K-major,
//     no swizzle, straight from registers;
//   * evaluates the k1-averaged log-likelihood derivatives in fp64 (two samples per lane; quirk Q3) -> weights aw, dw.
// When the four warps have arrived, one thread issues 12 tcgen05.mma (kind::f16, M = N = 128, K = 16): per K-step
// hi hi^T + hi lo^T + lo hi^T into a 128 x 128 fp32 accumulator in tensor memory whose diagonal 32 x 32 blocks are the
// entries' V V^T (the off-diagonal blocks are the price of the 128-row minimum of the instruction; the tensor pipe has the
// time: profiles/r3a_tc_probe.txt).  While the next quad is being formed, each warp reads ITS block of the previous
// accumulator (tcgen05.ld: lane n gets row n) and folds it into registers with the entry's weight:
//     G += dw_b V_b V_b^T                                                           (SSVI_TF_d.py:47-56)
// Tiles and accumulators are double-buffered; two mbarriers per buffer (tile written by 4 warps / MMAs complete).
// fp16 range: a |V| > 65504 overflows to inf, which lands in the chunk record and raises SSVI_FLAG_NONFINITE_ACC |
// SSVI_FLAG_TC_RANGE; a (MemoView::tc_alpha, a power of two) moves the window.
#pragma once
#include "accumulate_memo.cuh"
#include "tc_ops.cuh"

#define SSVI_TC_SMEM (4 * TC_TILE_BYTES + 64)

template <int NT, int NOTH>
__global__ void __launch_bounds__(128, 2) accumulate_tc_kernel(
    DevView v, MemoView mv, int64_t n_chunks, const int32_t* __restrict__ chunk_row,
    const int64_t* __restrict__ chunk_start, const int32_t* __restrict__ chunk_len,
    const int32_t* __restrict__ chunk_slot, const int64_t* __restrict__ chunk_out,
    const int64_t* __restrict__ entry_ids, const int64_t* __restrict__ rows, const void* __restrict__ z_own,
    int z_f32, double* __restrict__ partial_out) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  unsigned char* smem = smem_raw;
  constexpr int KT = SSVI_MEMO_KT, NI = 2 * NT, REC4 = NT * 2 * 2 * 32;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2;
  const int r8 = lane & 7, kq = lane >> 3;
  const int64_t c = blockIdx.x;
  if (c >= n_chunks) return;
  const int N = v.N, D = v.D, NS = smp_ns(D);
  const int64_t row = rows[chunk_row[c]];
  const int64_t start = chunk_start[c];
  const int len = chunk_len[c], slot = chunk_slot[c];
  const int sl = slot - mv.slot_lo;
  const int nq = (len + 3) >> 2;

  unsigned char* tiles = smem;                                     // [buffer][hi | lo][16 KB]
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + 4 * TC_TILE_BYTES);
  uint64_t* done = full + 2;
  uint32_t* holder = reinterpret_cast<uint32_t*>(done + 2);
  // the prologue's staging (sampler row + noise block as doubles) and U overlay the tiles, which are idle until the loop
  double* S = reinterpret_cast<double*>(smem);
  double* zs = S + v.srow_len;
  double2* Us = reinterpret_cast<double2*>(smem + 2 * TC_TILE_BYTES);

  if (tid == 0) {
    tc_mbar_init(&full[0], 4);
    tc_mbar_init(&full[1], 4);
    tc_mbar_init(&done[0], 1);
    tc_mbar_init(&done[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  if (warp == 0) tc_tmem_alloc(holder, 256);

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
#pragma unroll
  for (int i = 0; i < NT; ++i)
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      double c0, c1;
      mma_sample_tile(S, zs, NT, NS, lane, g, i, warp + 4 * h, c0, c1);
      Us[(i * KT + warp + 4 * h) * 32 + lane] = make_double2(c0, c1);
    }
  __syncthreads();
  float u[NI][8];                                       // this lane's share of U: rows 8 rg + r8, samples 8 (kq + 4 h) + j
#pragma unroll
  for (int it = 0; it < NI; ++it) {
    const int rg = it >> 1, h = it & 1;
#pragma unroll
    for (int jp = 0; jp < 4; ++jp) {
      const double2 x = Us[(rg * KT + kq + 4 * h) * 32 + r8 * 4 + jp];
      u[it][2 * jp] = (float)x.x;
      u[it][2 * jp + 1] = (float)x.y;
    }
  }
  __syncthreads();                                      // staging and U are dead: the region becomes the operand tiles
  for (int i = tid; i < 4 * TC_TILE_BYTES / 16; i += 128) reinterpret_cast<uint4*>(tiles)[i] = make_uint4(0, 0, 0, 0);
  tc_fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *holder;

  // ---- main loop -----------------------------------------------------------------------------------------------------
  const float4* tab4 = reinterpret_cast<const float4*>(mv.tab);
  const float alpha = mv.tc_alpha;
  const double inv_k1 = 1.0 / (double)(8 * KT);
  const uint32_t idesc = tc_idesc_f16(128, 128);
  float gacc[32], dacc[NT];
#pragma unroll
  for (int cc = 0; cc < 32; ++cc) gacc[cc] = 0.f;
#pragma unroll
  for (int i = 0; i < NT; ++i) dacc[i] = 0.f;
  double my_y = 0.0;                                    // lane l: entry (q0 + l) of this warp, q0 = the quad batch's first
  int32_t my_rec[NOTH];
#pragma unroll
  for (int jj = 0; jj < NOTH; ++jj) my_rec[jj] = 0;
  float4 cur[NOTH][2];                                  // item 0 of the entry about to be processed
  bool have_cur = false;
  float w_prev = 0.f;
  bool prev_valid = false;

  for (int q = 0; q <= nq; ++q) {
    const int buf = q & 1;
    float w_b = 0.f;
    bool valid = false;
    if (q < nq) {
      valid = 4 * q + warp < len;                       // warp-uniform
      unsigned char* thi = tiles + (buf * 2) * TC_TILE_BYTES;
      unsigned char* tlo = thi + TC_TILE_BYTES;
      if (valid) {
        const int el = q & 31;
        if (el == 0) {                                  // next 32 entries of this warp: ids, values, table records
          my_y = 0.0;
#pragma unroll
          for (int jj = 0; jj < NOTH; ++jj) my_rec[jj] = 0;
          const int b = 4 * (q + lane) + warp;
          if (b < len) {
            const int64_t eid = entry_ids[start + b];
            my_y = v.vals[eid];
#pragma unroll
            for (int jj = 0; jj < NOTH; ++jj) my_rec[jj] = (int32_t)memo_rec(mv, jj, sl, v.coords[eid * N + mv.od[jj]]);
          }
          have_cur = false;
        }
        const double y = __shfl_sync(0xffffffffu, my_y, el);
        const bool next_valid = (el + 1 < 32) && (4 * (q + 1) + warp < len);
        const float4* pc[NOTH];
        const float4* pn[NOTH];
#pragma unroll
        for (int jj = 0; jj < NOTH; ++jj) {
          pc[jj] = tab4 + (int64_t)__shfl_sync(0xffffffffu, my_rec[jj], el) * REC4 + lane;
          pn[jj] = tab4 + (int64_t)__shfl_sync(0xffffffffu, my_rec[jj], next_valid ? el + 1 : el) * REC4 + lane;
        }
        if (!have_cur) {
#pragma unroll
          for (int jj = 0; jj < NOTH; ++jj) {
            cur[jj][0] = __ldcg(pc[jj]);
            cur[jj][1] = __ldcg(pc[jj] + 32);
          }
        }
        float fp[2][8], rs[NT];
#pragma unroll
        for (int j = 0; j < 8; ++j) fp[0][j] = fp[1][j] = 0.f;
#pragma unroll
        for (int it = 0; it < NI; ++it) {
          const int rg = it >> 1, h = it & 1;
          float4 nxt[NOTH][2];                          // one item ahead (the next entry's first item at the end)
#pragma unroll
          for (int jj = 0; jj < NOTH; ++jj) {
            const float4* p = (it + 1 < NI) ? pc[jj] + ((it + 1) * 2) * 32 : pn[jj];
            nxt[jj][0] = __ldcg(p);
            nxt[jj][1] = __ldcg(p + 32);
          }
          float x[8] = {cur[0][0].x, cur[0][0].y, cur[0][0].z, cur[0][0].w, cur[0][1].x, cur[0][1].y, cur[0][1].z, cur[0][1].w};
#pragma unroll
          for (int jj = 1; jj < NOTH; ++jj) {
            x[0] *= cur[jj][0].x; x[1] *= cur[jj][0].y; x[2] *= cur[jj][0].z; x[3] *= cur[jj][0].w;
            x[4] *= cur[jj][1].x; x[5] *= cur[jj][1].y; x[6] *= cur[jj][1].z; x[7] *= cur[jj][1].w;
          }
          float s = 0.f;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            fp[h][j] = fmaf(u[it][j], x[j], fp[h][j]);
            s += x[j];
          }
          if (h == 0) rs[rg] = s; else rs[rg] += s;
          tc_store_split8(thi, tlo, 32 * warp + 8 * rg + r8, kq + 4 * h, x, alpha);
#pragma unroll
          for (int jj = 0; jj < NOTH; ++jj) {
            cur[jj][0] = nxt[jj][0];
            cur[jj][1] = nxt[jj][1];
          }
        }
        have_cur = next_valid;
        // f_k = SUM_n u v: summed in-lane over the row groups, then over the 8 lanes r8 by a transposing reduction that
        // leaves lane (hb, jb, jc | kq) with the two samples k = 8 (kq + 4 hb) + 4 jb + 2 jc + {0, 1}
        const bool hb = (lane >> 2) & 1, jb = (lane >> 1) & 1, jc = lane & 1;
        float s8[8], s4[4], f[2];
#pragma unroll
        for (int j = 0; j < 8; ++j) s8[j] = (hb ? fp[1][j] : fp[0][j]) + __shfl_xor_sync(0xffffffffu, hb ? fp[0][j] : fp[1][j], 4);
#pragma unroll
        for (int j = 0; j < 4; ++j) s4[j] = (jb ? s8[4 + j] : s8[j]) + __shfl_xor_sync(0xffffffffu, jb ? s8[j] : s8[4 + j], 2);
#pragma unroll
        for (int j = 0; j < 2; ++j) f[j] = (jc ? s4[2 + j] : s4[j]) + __shfl_xor_sync(0xffffffffu, jc ? s4[j] : s4[2 + j], 1);
        // ---- SSVI_TF_d.py:31-57 (Q3): derivatives averaged over k1, then broadcast -----------------------------------
        double a0, b0, a1, b1;
        log_pdf_derivs_pair(v.lik, y, (double)f[0], (double)f[1], v.lik_param, a0, b0, a1, b1);
        double sum_a = a0 + a1, sum_d = b0 + b1;
        warp_sum2(sum_a, sum_d, lane);
        const float aw = (float)(sum_a * inv_k1 * inv_k1);
        w_b = (float)(0.5 * sum_d * inv_k1 * inv_k1);
#pragma unroll
        for (int i = 0; i < NT; ++i) {                  // row sums: over this lane's 16 samples (done), then over kq
          float s = rs[i];
          s += __shfl_xor_sync(0xffffffffu, s, 8);
          s += __shfl_xor_sync(0xffffffffu, s, 16);
          dacc[i] = fmaf(aw, s, dacc[i]);
        }
      }
      tc_fence_async_smem();                            // tile stores -> visible to the tensor core
      tc_fence_before();                                // and this warp's reads of the accumulator it is about to release
      __syncwarp();
      if (lane == 0) tc_mbar_arrive(&full[buf]);
      if (warp == 0) {
        tc_mbar_wait(&full[buf], (q >> 1) & 1);
        tc_fence_after();
        if (lane == 0) {
          const uint32_t ahi = tc_smem_u32(thi), alo = tc_smem_u32(tlo), dst = tmem + buf * 128;
#pragma unroll
          for (int s = 0; s < 4; ++s) {
            const uint64_t dh = tc_smem_desc(ahi + 256 * s, TC_LBO, TC_SBO), dl = tc_smem_desc(alo + 256 * s, TC_LBO, TC_SBO);
            tc_mma_f16(dst, dh, dh, idesc, s > 0);
            tc_mma_f16(dst, dh, dl, idesc, 1);
            tc_mma_f16(dst, dl, dh, idesc, 1);
          }
          tc_commit(&done[buf]);
        }
        __syncwarp();
      }
    }
    if (q > 0) {                                        // the previous quad's V V^T: this warp's block, row n in lane n
      const int pb = buf ^ 1;
      tc_mbar_wait(&done[pb], ((q - 1) >> 1) & 1);
      tc_fence_after();
      float d[32];
      tc_tmem_ld32(tmem + ((uint32_t)(32 * warp) << 16) + pb * 128 + 32 * warp, d);
      if (prev_valid) {
#pragma unroll
        for (int cc = 0; cc < 32; ++cc) gacc[cc] = fmaf(w_prev, d[cc], gacc[cc]);
      }
      tc_fence_before();
    }
    w_prev = w_b;
    prev_valid = valid;
  }

  // ---- CTA reduction (fixed warp order => deterministic) and the chunk record [di | Di | si_sum | count] ------------
  __syncthreads();                                      // every MMA has completed and been read: the tiles are dead
  float* red = reinterpret_cast<float*>(tiles);         // [4][32][33]
  float* dred = red + 4 * 32 * 33;                      // [4][32]
#pragma unroll
  for (int cc = 0; cc < 32; ++cc) red[(warp * 32 + lane) * 33 + cc] = gacc[cc];
  if (kq == 0) {
#pragma unroll
    for (int i = 0; i < NT; ++i) dred[warp * 32 + 8 * i + r8] = dacc[i];
  }
  __syncthreads();
  double* out = partial_out + chunk_out[c] * (int64_t)v.acc_len;
  const double ia2 = 1.0 / ((double)alpha * (double)alpha);
  bool bad = false;
  for (int n = tid; n < D; n += 128) {
    const float x = ((dred[n] + dred[32 + n]) + dred[64 + n]) + dred[96 + n];
    out[n] = (double)x;
    bad |= !isfinite(x);
  }
  for (int i = tid; i < 32 * 32; i += 128) {
    const int n1 = i >> 5, n2 = i & 31;
    if (n1 < D && n2 <= n1) {                           // lower triangle incl. diagonal, mirrored: exactly symmetric
      const float x = ((red[n1 * 33 + n2] + red[(32 + n1) * 33 + n2]) + red[(64 + n1) * 33 + n2]) + red[(96 + n1) * 33 + n2];
      const double xd = ia2 * (double)x;
      out[D + n1 * D + n2] = xd;
      out[D + n2 * D + n1] = xd;
      bad |= !isfinite(x);
    }
  }
  if (bad) atomicOr(v.flags, SSVI_FLAG_NONFINITE_ACC | SSVI_FLAG_TC_RANGE);
  if (tid == 0) {
    out[D + D * D] = 0.0;
    out[D + D * D + 1] = (double)len;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tc_tmem_dealloc(tmem, 256);
}
