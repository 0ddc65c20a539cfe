// Stages 1-3 for FULL SWEEPS of SSVI_TF_d, full covariance, D <= 32, k1 == 64 -- the fp32 option of the contract
// (SSVI_TABLE_F32_TC; BASELINE.json: 1e-4 in fp32) with the contraction on the tcgen05 tensor cores.
//
// Same slot-major sample-table scheme, chunk tables and chunk records as accumulate_memo.cuh; what differs is the
// arithmetic.  The table holds float samples X(j, s) in the lane order of tc_rec_index (tc_layout.cuh).  One CTA (4 warps)
// per chunk; warp w takes entries w, w + 4, ...  ("quad" q = the four entries 4q .. 4q + 3).  Per entry a warp
//   * gathers the N - 1 records (two coalesced 16-byte loads per lane and item, two items in flight), forms V = Hadamard
//     product in fp32 registers, f_k = u_k . v_k against the own-row sample U(row, slot) (a record of the table like the
//     others: sample_table_kernel samples the swept mode too, so a chunk's prologue is one 8 KB copy into shared memory),
//     and the row sums SUM_k V[n][k]                                                             (SSVI_TF_d.py:31-46);
//   * writes rows 32 w .. 32 w + 31 of the quad's operand tile: a V split into fp16 hi + lo (22 significand bits), K-major,
//     no swizzle, straight from registers;
//   * evaluates the k1-averaged log-likelihood derivatives in fp64 (two samples per lane; quirk Q3) -> weights aw, dw.
// The last of the four warps to arrive (a shared-memory counter) issues, from one thread, 12 tcgen05.mma (kind::f16,
// M = N = 128, K = 16): per K-step hi hi^T + hi lo^T + lo hi^T into a 128 x 128 fp32 accumulator in tensor memory whose diagonal 32 x 32 blocks are the
// entries' V V^T (the off-diagonal blocks are the price of the 128-row minimum of the instruction; the tensor pipe has the
// time: profiles/r3a_tc_probe.txt).  While the next quad is being formed, each warp reads ITS block of the previous
// accumulator (tcgen05.ld: lane n gets row n) and folds it into registers with the entry's weight:
//     G += dw_b V_b V_b^T                                                           (SSVI_TF_d.py:47-56)
// A warp folds quad q - 1 after it has formed quad q's tile and before it arrives for quad q, so ONE accumulator (128
// tensor-memory columns per CTA) suffices; the operand tiles are double-buffered (MMA q reads one while the warps write
// q + 1).  One mbarrier, done (tcgen05.commit, one phase per quad), tells the warps a quad is ready.  U sits in shared memory
// as floats pre-divided by a; the table records of the first other mode are pre-multiplied by a (sample_table_kernel), so
// x = a V needs no multiply.  72 KB of shared memory and <= 168 registers: three CTAs per SM.
// fp16 range: a |V| > 65504 overflows to inf, which lands in the chunk record and raises SSVI_FLAG_NONFINITE_ACC |
// SSVI_FLAG_TC_RANGE; a (MemoView::tc_alpha, a power of two) moves the window.
#pragma once
#include "accumulate_memo.cuh"
#include "tc_ops.cuh"

// RING (SSVI_TUNE_TC_RING; a measured alternative, not the default): the records reach the warp through a per-warp ring
// of stages in shared memory, one stage = one item (1 KB of each of the N - 1 records), filled by cp.async.bulk (one lane
// issues, an mbarrier per stage counts the bytes): four items in flight per warp at no register cost, 104 KB of shared
// memory, two CTAs per SM.  It LOSES: config 5 at full size 2.11e8 against 3.37e8 entry-updates/s on the same plans
// (profiles/r3g_bench_c5tc_ring.json): 1 KB bulk copies are served far more slowly than the 32 LDG.128 of the default path
// move the same bytes.  An L2 prefetch of the next entry's records (prefetch.global.L2, an entry ahead) was measured too:
// 3.19e8 against 3.32e8 (profiles/r3h).  Default: two register buffers refilled in place (PF below), 72 KB, <= 168
// registers, three CTAs per SM.
#define SSVI_TC_U_BYTES(NT) ((NT) * 2048)
__host__ __device__ constexpr int tc_ring_stages(int NOTH, bool RING) { return RING ? (NOTH == 3 ? 2 : 4) : 0; }
__host__ __device__ constexpr int tc_ring_bytes(int NOTH, bool RING) { return 4 * tc_ring_stages(NOTH, RING) * NOTH * 1024; }
__host__ __device__ constexpr int tc_cta_smem(int NOTH, bool RING) {
  return 4 * TC_TILE_BYTES + SSVI_TC_U_BYTES(4) + tc_ring_bytes(NOTH, RING) + 256;
}

// PF: register buffers for table items (1 KB of each record) = items a warp has in flight: item it of ANY entry lives in
// buffer it % PF (2 NT is a multiple of PF), and as soon as the Hadamard product has consumed an item its buffer is
// refilled with item it + PF (of the next entry at the end).  PF == 2: 168 registers, three CTAs per SM.  PF > 2
// (SSVI_TUNE_TC_PF): ~200 registers, two CTAs per SM.
__host__ __device__ constexpr int tc_deep_pf(int NT) { return NT == 3 ? 3 : (NT == 1 ? 2 : 4); }

template <int NT, int NOTH, bool RING, int PF = 2>
__global__ void __launch_bounds__(128, (RING || PF > 2) ? 2 : 3) accumulate_tc_kernel(
    DevView v, MemoView mv, int64_t n_chunks, const int32_t* __restrict__ chunk_row,
    const int64_t* __restrict__ chunk_start, const int32_t* __restrict__ chunk_len,
    const int32_t* __restrict__ chunk_slot, const int64_t* __restrict__ chunk_out,
    const int64_t* __restrict__ entry_ids, const int64_t* __restrict__ rows, const void* __restrict__ z_own,
    int z_f32, double* __restrict__ partial_out) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  unsigned char* smem = smem_raw;
  constexpr int KT = SSVI_MEMO_KT, NI = 2 * NT, REC4 = NT * 2 * 2 * 32;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int r8 = lane & 7, kq = lane >> 3;
  const int64_t c = blockIdx.x;
  if (c >= n_chunks) return;
  const int N = v.N, D = v.D;
  const int64_t row = rows[chunk_row[c]];
  const int64_t start = chunk_start[c];
  const int len = chunk_len[c], slot = chunk_slot[c];
  const int sl = slot - mv.slot_lo;
  const int nq = (len + 3) >> 2;

  unsigned char* tiles = smem;                                     // [buffer][hi | lo][16 KB]
  float4* Uf = reinterpret_cast<float4*>(smem + 4 * TC_TILE_BYTES);        // [item][p][lane]: U / a in the lanes' order
  constexpr int RS = tc_ring_stages(NOTH, RING), STAGE_BYTES = NOTH * 1024;
  unsigned char* wring = smem + 4 * TC_TILE_BYTES + SSVI_TC_U_BYTES(4) + warp * (RS * STAGE_BYTES);   // this warp's ring
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + 4 * TC_TILE_BYTES + SSVI_TC_U_BYTES(4) + tc_ring_bytes(NOTH, RING));
  uint64_t* done = full + 2;
  uint64_t* rbar = done + 1 + warp * 4;                 // this warp's stage barriers (RS <= 4)
  uint32_t* holder = reinterpret_cast<uint32_t*>(done + 1 + 16);
  uint32_t* cnt = holder + 1;                           // [2] arrival counters of the operand-tile buffers

  if (tid == 0) {
    cnt[0] = cnt[1] = 0;
    tc_mbar_init(done, 1);
    for (int i = 0; i < 16; ++i) tc_mbar_init(done + 1 + i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  if (warp == 0) tc_tmem_alloc(holder, 128);

  // ---- this warp's entries (w, w + 4, ...): lane l holds value and table records of entry q0 + l of the current batch of
  //      32; fetched before the prologue so that the index -> coordinate chain overlaps the own-row sampling --------------
  const int my_cnt = (len - warp + 3) >> 2;             // may be 0 for high warps of a short chunk
  const unsigned char* tabb = reinterpret_cast<const unsigned char*>(mv.tab);
  constexpr int REC_BYTES = NT * 2048;
  double my_y = 0.0;
  int32_t my_rec[NOTH];
  auto load_batch = [&](int q0) {
    my_y = 0.0;
#pragma unroll
    for (int jj = 0; jj < NOTH; ++jj) my_rec[jj] = 0;
    const int b = 4 * (q0 + lane) + warp;
    if (b < len) {
      const int64_t eid = __ldcs(entry_ids + start + b);         // touched once: keep the L2 for the table
      my_y = __ldcs(v.vals + eid);
#pragma unroll
      for (int jj = 0; jj < NOTH; ++jj) my_rec[jj] = (int32_t)memo_rec(mv, jj, sl, __ldcs(v.coords + eid * N + mv.od[jj]));
    }
  };
  load_batch(0);
  // ring producer: items [issued, min(limit, consumed + S)) -> stages (item % S); every lane calls, lane 0 issues
  int issued = 0;
  auto top_up = [&](int consumed, int limit) {
    if constexpr (RING) {
      while (issued < limit && issued < consumed + RS) {
        const int e2 = issued / NI, it2 = issued - e2 * NI;
        int32_t r[NOTH];
#pragma unroll
        for (int jj = 0; jj < NOTH; ++jj) r[jj] = __shfl_sync(0xffffffffu, my_rec[jj], e2 & 31);
        if (lane == 0) {
          const int st = issued % RS;
          mbar_expect_tx(&rbar[st], STAGE_BYTES);
#pragma unroll
          for (int jj = 0; jj < NOTH; ++jj)
            bulk_g2s(wring + st * STAGE_BYTES + jj * 1024, tabb + (int64_t)r[jj] * REC_BYTES + it2 * 1024, 1024, &rbar[st]);
        }
        ++issued;
      }
    }
  };

  // ---- prologue: U(row, slot) / a (ui samples, SSVI_TF_interface.py:251-254) is a record of the table too (the swept mode
  //      is sampled by sample_table_kernel as one more "mode", MemoView::own): one 2 NT KB copy per chunk instead of staging
  //      the sampler row and the noise block and re-sampling U in every chunk of the (row, slot) pair ----------------------------
  const float alpha = mv.tc_alpha;
  const double inv_alpha = 1.0 / (double)alpha;
  {
    const float4* usrc = reinterpret_cast<const float4*>(mv.tab) + memo_rec(mv, NOTH, sl, row) * (int64_t)REC4;
    for (int i = tid; i < REC4; i += 128) Uf[i] = __ldcg(usrc + i);
  }
  constexpr int R = RING ? 1 : PF;                      // register buffers: item it of ANY entry lives in rb[it % R]
  static_assert(RING || NI % R == 0, "2 NT must be a multiple of PF");
  float4 rb[R][NOTH][2];
  bool have_cur = false;                                // items 0 .. PF-1 of the entry about to be processed are in flight
  if (!RING && my_cnt > 0) {                            // the first entry's items travel across the prologue's barrier
#pragma unroll
    for (int jj = 0; jj < NOTH; ++jj) {
      const float4* p0 = reinterpret_cast<const float4*>(mv.tab) + (int64_t)__shfl_sync(0xffffffffu, my_rec[jj], 0) * REC4 + lane;
#pragma unroll
      for (int d = 0; d < R; ++d) {
        rb[d][jj][0] = __ldcg(p0 + (d * 2) * 32);
        rb[d][jj][1] = __ldcg(p0 + (d * 2 + 1) * 32);
      }
    }
    have_cur = true;
  }
  __syncthreads();                                      // (also: the mbarriers are initialised)
  top_up(0, min(my_cnt, 32) * NI);
  // The operand tiles are NOT cleared: row n of an accumulator block
  // depends on row n of the operands only, and rows / blocks that no entry wrote (n >= 8 NT, warps without an entry in
  // the last quad) are never folded into an output.
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *holder;

  // ---- main loop -----------------------------------------------------------------------------------------------------
  const float4* tab4 = reinterpret_cast<const float4*>(mv.tab);
  const double inv_k1 = 1.0 / (double)(8 * KT);
  const uint32_t idesc = tc_idesc_f16(128, 128);
  float gacc[32], dacc[NT];
#pragma unroll
  for (int cc = 0; cc < 32; ++cc) gacc[cc] = 0.f;
#pragma unroll
  for (int i = 0; i < NT; ++i) dacc[i] = 0.f;
  float w_prev = 0.f;
  bool prev_valid = false;

  // this warp's block of the accumulator (quad p's V V^T, row n in lane n) folded into registers with the entry's weight
  auto fold = [&](int p) {
    tc_mbar_wait(done, p & 1);
    tc_fence_after();
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      float d[16];
      tc_tmem_ld16(tmem + ((uint32_t)(32 * warp) << 16) + 32 * warp + 16 * half, d);
      if (prev_valid) {
#pragma unroll
        for (int cc = 0; cc < 16; ++cc) gacc[16 * half + cc] = fmaf(w_prev, d[cc], gacc[16 * half + cc]);
      }
    }
    tc_fence_before();
  };

  for (int q = 0; q < nq; ++q) {
    const int buf = q & 1;
    float w_b = 0.f;
    bool valid = false;
    {
      valid = 4 * q + warp < len;                       // warp-uniform
      unsigned char* thi = tiles + (buf * 2) * TC_TILE_BYTES;
      unsigned char* tlo = thi + TC_TILE_BYTES;
      if (valid) {
        const int el = q & 31;
        if (el == 0 && q > 0) {                         // next 32 entries of this warp (the ring is drained here)
          load_batch(q);
          have_cur = false;
        }
        const int lim = min(my_cnt, q - el + 32) * NI;  // the producer stops at the batch's end
        const double y = __shfl_sync(0xffffffffu, my_y, el);
        const bool next_valid = (el + 1 < 32) && (4 * (q + 1) + warp < len);
        const float4* pc[NOTH];
        const float4* pn[NOTH];
#pragma unroll
        for (int jj = 0; jj < NOTH; ++jj) {
          pc[jj] = tab4 + (int64_t)__shfl_sync(0xffffffffu, my_rec[jj], el) * REC4 + lane;
          pn[jj] = tab4 + (int64_t)__shfl_sync(0xffffffffu, my_rec[jj], next_valid ? el + 1 : el) * REC4 + lane;
        }
        if (!RING && !have_cur) {
#pragma unroll
          for (int d = 0; d < R; ++d)
#pragma unroll
            for (int jj = 0; jj < NOTH; ++jj) {
              rb[d][jj][0] = __ldcg(pc[jj] + (d * 2) * 32);
              rb[d][jj][1] = __ldcg(pc[jj] + (d * 2 + 1) * 32);
            }
        }
        float fp[2][8], rs[NT];
#pragma unroll
        for (int j = 0; j < 8; ++j) fp[0][j] = fp[1][j] = 0.f;
#pragma unroll
        for (int it = 0; it < NI; ++it) {
          const int rg = it >> 1, h = it & 1;
          float4 (&cur)[NOTH][2] = rb[RING ? 0 : it % R];
          if constexpr (RING) {
            const int gi = q * NI + it, st = gi % RS;
            top_up(gi, lim);
            tc_mbar_wait(&rbar[st], (gi / RS) & 1);
            const float4* rp = reinterpret_cast<const float4*>(wring + st * STAGE_BYTES) + lane;
#pragma unroll
            for (int jj = 0; jj < NOTH; ++jj) {
              cur[jj][0] = rp[jj * 64];
              cur[jj][1] = rp[jj * 64 + 32];
            }
          }
          float x[8] = {cur[0][0].x, cur[0][0].y, cur[0][0].z, cur[0][0].w, cur[0][1].x, cur[0][1].y, cur[0][1].z, cur[0][1].w};
#pragma unroll
          for (int jj = 1; jj < NOTH; ++jj) {
            x[0] *= cur[jj][0].x; x[1] *= cur[jj][0].y; x[2] *= cur[jj][0].z; x[3] *= cur[jj][0].w;
            x[4] *= cur[jj][1].x; x[5] *= cur[jj][1].y; x[6] *= cur[jj][1].z; x[7] *= cur[jj][1].w;
          }
          if constexpr (RING) {                         // every lane's reads of the stage have returned: refill it
            __syncwarp();
            top_up(q * NI + it + 1, lim);
          } else {                                      // the buffer is free: item it + PF (of the next entry at the end)
            const int nit = it + PF;
#pragma unroll
            for (int jj = 0; jj < NOTH; ++jj) {
              const float4* p = (nit < NI) ? pc[jj] + (nit * 2) * 32 : pn[jj] + ((nit - NI) * 2) * 32;
              cur[jj][0] = __ldcg(p);
              cur[jj][1] = __ldcg(p + 32);
            }
          }
          const float4 u0 = Uf[(it * 2) * 32 + lane], u1 = Uf[(it * 2 + 1) * 32 + lane];
          const float uu[8] = {u0.x, u0.y, u0.z, u0.w, u1.x, u1.y, u1.z, u1.w};
          float s = 0.f;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            fp[h][j] = fmaf(uu[j], x[j], fp[h][j]);
            s += x[j];
          }
          if (h == 0) rs[rg] = s; else rs[rg] += s;
          tc_store_split8(thi, tlo, 32 * warp + 8 * rg + r8, kq + 4 * h, x);
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
        const float aw = (float)(sum_a * inv_k1 * inv_k1 * inv_alpha);   // the row sums below are of x = a V
        w_b = (float)(0.5 * sum_d * inv_k1 * inv_k1);
#pragma unroll
        for (int i = 0; i < NT; ++i) {                  // row sums: over this lane's 16 samples (done), then over kq
          float s = rs[i];
          s += __shfl_xor_sync(0xffffffffu, s, 8);
          s += __shfl_xor_sync(0xffffffffu, s, 16);
          dacc[i] = fmaf(aw, s, dacc[i]);
        }
      }
      if (q > 0) fold(q - 1);                           // before this warp licenses the MMAs that overwrite the accumulator
      tc_fence_async_smem();                            // tile stores -> visible to the tensor core
      tc_fence_before();
      __syncwarp();
      // the LAST warp to arrive issues the quad's MMAs (no warp waits for the others here): arrivals for quad q go to
      // counter q & 1, which has received all four of quad q - 2 before any of quad q (a warp folds q - 1 first, and
      // MMA q - 1 needed everybody's arrival for q - 1, which follows the arrival for q - 2 in program order)
      uint32_t last = 0;
      if (lane == 0) last = (tc_arrive_count(&cnt[buf]) & 3u) == 3u;
      last = __shfl_sync(0xffffffffu, last, 0);
      if (last) {
        tc_fence_after();
        if (lane == 0) {
          const uint32_t ahi = tc_smem_u32(thi), alo = tc_smem_u32(tlo), dst = tmem;
#pragma unroll
          for (int s = 0; s < 4; ++s) {
            const uint64_t dh = tc_smem_desc(ahi + 256 * s, TC_LBO, TC_SBO), dl = tc_smem_desc(alo + 256 * s, TC_LBO, TC_SBO);
            tc_mma_f16(dst, dh, dh, idesc, s > 0);
            tc_mma_f16(dst, dh, dl, idesc, 1);
            tc_mma_f16(dst, dl, dh, idesc, 1);
          }
          tc_commit(done);
        }
        __syncwarp();
      }
    }
    w_prev = w_b;
    prev_valid = valid;
  }
  if (nq > 0) fold(nq - 1);

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
    __stcs(out + n, (double)x);
    bad |= !isfinite(x);
  }
  for (int i = tid; i < 32 * 32; i += 128) {
    const int n1 = i >> 5, n2 = i & 31;
    if (n1 < D && n2 <= n1) {                           // lower triangle incl. diagonal, mirrored: exactly symmetric
      const float x = ((red[n1 * 33 + n2] + red[(32 + n1) * 33 + n2]) + red[(64 + n1) * 33 + n2]) + red[(96 + n1) * 33 + n2];
      const double xd = ia2 * (double)x;
      __stcs(out + D + n1 * D + n2, xd);
      __stcs(out + D + n2 * D + n1, xd);
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
  if (warp == 0) tc_tmem_dealloc(tmem, 128);
}

// =====================================================================================================================
// Ranks 33..64 (NT = 5..8 row tiles; the reference's real-data default is rank 50, real_test.py:27): the same pipeline
// with TWO warps per entry.  An entry's V is 64 rows x 64 samples, so a 128-row operand tile holds two entries ("pair"
// p = entries 2p, 2p + 1) and the M = N = 128 accumulator carries their V V^T in its two diagonal 64 x 64 blocks.  Warp
// w = 2e + h works on entry slot e and on the entry's rows 32h .. 32h + 31 (row groups 4h .. min(4h + 4, NT) - 1): it
// gathers only those rows of the records, writes tile rows 32w .. 32w + 31 exactly as the narrow kernel does, and sums
// f over ITS rows; the two partial f vectors meet in shared memory across a 64-thread named barrier, after which both
// warps evaluate the (identical) weights.  Each warp folds its 32 x 64 block of the accumulator (64 registers).
// 82 KB of shared memory, <= 255 registers: two CTAs per SM.
// =====================================================================================================================
__host__ __device__ constexpr int tc_wide_smem() { return 4 * TC_TILE_BYTES + 16384 + 2 * 4 * 64 * 4 + 256; }

template <int NT, int NOTH>
__global__ void __launch_bounds__(128, 2) accumulate_tc_wide_kernel(
    DevView v, MemoView mv, int64_t n_chunks, const int32_t* __restrict__ chunk_row,
    const int64_t* __restrict__ chunk_start, const int32_t* __restrict__ chunk_len,
    const int32_t* __restrict__ chunk_slot, const int64_t* __restrict__ chunk_out,
    const int64_t* __restrict__ entry_ids, const int64_t* __restrict__ rows, const void* __restrict__ z_own,
    int z_f32, double* __restrict__ partial_out) {
  static_assert(NT >= 5 && NT <= 8, "two warps per entry: 33 <= D <= 64");
  extern __shared__ __align__(16) unsigned char smem_raw[];
  unsigned char* smem = smem_raw;
  constexpr int KT = SSVI_MEMO_KT, REC4 = NT * 2 * 2 * 32;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int r8 = lane & 7, kq = lane >> 3;
  const int e = warp >> 1, h = warp & 1;                // entry slot of the pair, half of the entry's rows
  const int ngw = h ? NT - 4 : 4;                       // row groups of this warp, 2 items (k halves) each
  const int nil = 2 * ngw;
  const int64_t c = blockIdx.x;
  if (c >= n_chunks) return;
  const int N = v.N, D = v.D;
  const int64_t row = rows[chunk_row[c]];
  const int64_t start = chunk_start[c];
  const int len = chunk_len[c], slot = chunk_slot[c];
  const int sl = slot - mv.slot_lo;
  const int np = (len + 1) >> 1;

  unsigned char* tiles = smem;                                     // [buffer][hi | lo][16 KB]
  float* Uf = reinterpret_cast<float*>(smem + 4 * TC_TILE_BYTES);  // U / a, NT row groups in the table's record order
  float* fx = reinterpret_cast<float*>(smem + 4 * TC_TILE_BYTES + 16384);   // [2][4 warps][64] partial f of the pair's halves
  uint64_t* done = reinterpret_cast<uint64_t*>(smem + 4 * TC_TILE_BYTES + 16384 + 2048);
  uint32_t* holder = reinterpret_cast<uint32_t*>(done + 1);
  uint32_t* cnt = holder + 1;

  if (tid == 0) {
    cnt[0] = cnt[1] = 0;
    tc_mbar_init(done, 1);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  if (warp == 0) tc_tmem_alloc(holder, 128);

  double my_y = 0.0;
  int32_t my_rec[NOTH];
  auto load_batch = [&](int p0) {
    my_y = 0.0;
#pragma unroll
    for (int jj = 0; jj < NOTH; ++jj) my_rec[jj] = 0;
    const int b = 2 * (p0 + lane) + e;
    if (b < len) {
      const int64_t eid = __ldcs(entry_ids + start + b);
      my_y = __ldcs(v.vals + eid);
#pragma unroll
      for (int jj = 0; jj < NOTH; ++jj) my_rec[jj] = (int32_t)memo_rec(mv, jj, sl, __ldcs(v.coords + eid * N + mv.od[jj]));
    }
  };
  load_batch(0);

  // ---- prologue: U(row, slot) / a is a record of the table (MemoView::own): one copy per chunk --------------------------------
  const float alpha = mv.tc_alpha;
  const double inv_alpha = 1.0 / (double)alpha;
  {
    const float4* usrc = reinterpret_cast<const float4*>(mv.tab) + memo_rec(mv, NOTH, sl, row) * (int64_t)REC4;
    float4* Uw = reinterpret_cast<float4*>(Uf);
    for (int i = tid; i < REC4; i += 128) Uw[i] = __ldcg(usrc + i);
  }
  tc_fence_before();
  __syncthreads();                                      // U complete (and the mbarrier initialised)
  tc_fence_after();
  const uint32_t tmem = *holder;

  // ---- main loop -----------------------------------------------------------------------------------------------------
  const float4* tab4 = reinterpret_cast<const float4*>(mv.tab);
  const float4* Uf4 = reinterpret_cast<const float4*>(Uf);
  const double inv_k1 = 1.0 / (double)(8 * KT);
  const uint32_t idesc = tc_idesc_f16(128, 128);
  float gacc[64], dacc[4];
#pragma unroll
  for (int cc = 0; cc < 64; ++cc) gacc[cc] = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) dacc[i] = 0.f;
  float4 rb[2][NOTH][2];                                // item it of ANY entry lives in rb[it & 1] (nil is even)
  bool have_cur = false;
  float w_prev = 0.f;
  bool prev_valid = false;

  auto fold = [&](int p) {
    tc_mbar_wait(done, p & 1);
    tc_fence_after();
#pragma unroll
    for (int part = 0; part < 4; ++part) {
      float d[16];
      tc_tmem_ld16(tmem + ((uint32_t)(32 * warp) << 16) + 64 * e + 16 * part, d);
      if (prev_valid) {
#pragma unroll
        for (int cc = 0; cc < 16; ++cc) gacc[16 * part + cc] = fmaf(w_prev, d[cc], gacc[16 * part + cc]);
      }
    }
    tc_fence_before();
  };

  for (int p = 0; p < np; ++p) {
    const int buf = p & 1;
    float w_b = 0.f;
    const bool valid = 2 * p + e < len;                 // uniform over the pair of warps
    unsigned char* thi = tiles + (buf * 2) * TC_TILE_BYTES;
    unsigned char* tlo = thi + TC_TILE_BYTES;
    if (valid) {
      const int el = p & 31;
      if (el == 0 && p > 0) {
        load_batch(p);
        have_cur = false;
      }
      const double y = __shfl_sync(0xffffffffu, my_y, el);
      const bool next_valid = (el + 1 < 32) && (2 * (p + 1) + e < len);
      const float4* pc[NOTH];
      const float4* pn[NOTH];
#pragma unroll
      for (int jj = 0; jj < NOTH; ++jj) {               // this warp's rows start at row group 4h: item 8h of the record
        pc[jj] = tab4 + (int64_t)__shfl_sync(0xffffffffu, my_rec[jj], el) * REC4 + (8 * h) * 64 + lane;
        pn[jj] = tab4 + (int64_t)__shfl_sync(0xffffffffu, my_rec[jj], next_valid ? el + 1 : el) * REC4 + (8 * h) * 64 + lane;
      }
      if (!have_cur) {
#pragma unroll
        for (int d = 0; d < 2; ++d)
#pragma unroll
          for (int jj = 0; jj < NOTH; ++jj) {
            rb[d][jj][0] = __ldcg(pc[jj] + (d * 2) * 32);
            rb[d][jj][1] = __ldcg(pc[jj] + (d * 2 + 1) * 32);
          }
      }
      float fp[2][8], rs[4];
#pragma unroll
      for (int j = 0; j < 8; ++j) fp[0][j] = fp[1][j] = 0.f;
#pragma unroll
      for (int i = 0; i < 4; ++i) rs[i] = 0.f;
#pragma unroll
      for (int it = 0; it < 8; ++it) {
        if (it < nil) {                                 // warp-uniform
          const int rgl = it >> 1, hh = it & 1;
          float4 (&cur)[NOTH][2] = rb[it & 1];
          float x[8] = {cur[0][0].x, cur[0][0].y, cur[0][0].z, cur[0][0].w, cur[0][1].x, cur[0][1].y, cur[0][1].z, cur[0][1].w};
#pragma unroll
          for (int jj = 1; jj < NOTH; ++jj) {
            x[0] *= cur[jj][0].x; x[1] *= cur[jj][0].y; x[2] *= cur[jj][0].z; x[3] *= cur[jj][0].w;
            x[4] *= cur[jj][1].x; x[5] *= cur[jj][1].y; x[6] *= cur[jj][1].z; x[7] *= cur[jj][1].w;
          }
          {                                             // the buffer is free: item it + 2 (of the next entry at the end)
            const int nit = it + 2;
#pragma unroll
            for (int jj = 0; jj < NOTH; ++jj) {
              const float4* q4 = (nit < nil) ? pc[jj] + (nit * 2) * 32 : pn[jj] + ((nit - nil) * 2) * 32;
              cur[jj][0] = __ldcg(q4);
              cur[jj][1] = __ldcg(q4 + 32);
            }
          }
          const float4 u0 = Uf4[((8 * h + it) * 2) * 32 + lane], u1 = Uf4[((8 * h + it) * 2 + 1) * 32 + lane];
          const float uu[8] = {u0.x, u0.y, u0.z, u0.w, u1.x, u1.y, u1.z, u1.w};
          float s = 0.f;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            fp[hh][j] = fmaf(uu[j], x[j], fp[hh][j]);
            s += x[j];
          }
          rs[rgl] += s;
          tc_store_split8(thi, tlo, 32 * warp + 8 * rgl + r8, kq + 4 * hh, x);
        }
      }
      have_cur = next_valid;
      // partial f over this warp's rows -> lane (hb, jb, jc | kq) holds samples k = 8 (kq + 4 hb) + 4 jb + 2 jc + {0, 1}
      const bool hb = (lane >> 2) & 1, jb = (lane >> 1) & 1, jc = lane & 1;
      float s8[8], s4[4], f[2];
#pragma unroll
      for (int j = 0; j < 8; ++j) s8[j] = (hb ? fp[1][j] : fp[0][j]) + __shfl_xor_sync(0xffffffffu, hb ? fp[0][j] : fp[1][j], 4);
#pragma unroll
      for (int j = 0; j < 4; ++j) s4[j] = (jb ? s8[4 + j] : s8[j]) + __shfl_xor_sync(0xffffffffu, jb ? s8[j] : s8[4 + j], 2);
#pragma unroll
      for (int j = 0; j < 2; ++j) f[j] = (jc ? s4[2 + j] : s4[j]) + __shfl_xor_sync(0xffffffffu, jc ? s4[j] : s4[2 + j], 1);
      // the other half of the entry's rows: the partner warp's partial, through shared memory (double-buffered by pair)
      float2* mine = reinterpret_cast<float2*>(fx + ((p & 1) * 4 + warp) * 64) + lane;
      const float2* theirs = reinterpret_cast<const float2*>(fx + ((p & 1) * 4 + (warp ^ 1)) * 64) + lane;
      *mine = make_float2(f[0], f[1]);
      if (e == 0) asm volatile("bar.sync 1, 64;\n" ::: "memory");      // the pair's own named barrier (ids fixed at compile
      else asm volatile("bar.sync 2, 64;\n" ::: "memory");             // time: a register id would reserve all sixteen)
      const float2 o = *theirs;
      const float f0 = h ? o.x + f[0] : f[0] + o.x, f1 = h ? o.y + f[1] : f[1] + o.y;   // same order in both warps
      // ---- SSVI_TF_d.py:31-57 (Q3): derivatives averaged over k1 (both warps of the pair evaluate the same numbers) ---
      double a0, b0, a1, b1;
      log_pdf_derivs_pair(v.lik, y, (double)f0, (double)f1, v.lik_param, a0, b0, a1, b1);
      double sum_a = a0 + a1, sum_d = b0 + b1;
      warp_sum2(sum_a, sum_d, lane);
      const float aw = (float)(sum_a * inv_k1 * inv_k1 * inv_alpha);
      w_b = (float)(0.5 * sum_d * inv_k1 * inv_k1);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float s = rs[i];
        s += __shfl_xor_sync(0xffffffffu, s, 8);
        s += __shfl_xor_sync(0xffffffffu, s, 16);
        dacc[i] = fmaf(aw, s, dacc[i]);
      }
    }
    if (p > 0) fold(p - 1);
    tc_fence_async_smem();
    tc_fence_before();
    __syncwarp();
    uint32_t last = 0;
    if (lane == 0) last = (tc_arrive_count(&cnt[buf]) & 3u) == 3u;
    last = __shfl_sync(0xffffffffu, last, 0);
    if (last) {
      tc_fence_after();
      if (lane == 0) {
        const uint32_t ahi = tc_smem_u32(thi), alo = tc_smem_u32(tlo);
#pragma unroll
        for (int s = 0; s < 4; ++s) {
          const uint64_t dh = tc_smem_desc(ahi + 256 * s, TC_LBO, TC_SBO), dl = tc_smem_desc(alo + 256 * s, TC_LBO, TC_SBO);
          tc_mma_f16(tmem, dh, dh, idesc, s > 0);
          tc_mma_f16(tmem, dh, dl, idesc, 1);
          tc_mma_f16(tmem, dl, dh, idesc, 1);
        }
        tc_commit(done);
      }
      __syncwarp();
    }
    w_prev = w_b;
    prev_valid = valid;
  }
  if (np > 0) fold(np - 1);

  // ---- CTA reduction and the chunk record [di | Di | si_sum | count] -------------------------------------------------
  __syncthreads();
  float* red = reinterpret_cast<float*>(tiles);         // [4 warps][32 lanes][65]
  float* dred = red + 4 * 32 * 65;                      // [4 warps][32]
#pragma unroll
  for (int cc = 0; cc < 64; ++cc) red[(warp * 32 + lane) * 65 + cc] = gacc[cc];
  if (kq == 0) {
#pragma unroll
    for (int i = 0; i < 4; ++i) dred[warp * 32 + 8 * i + r8] = dacc[i];
  }
  __syncthreads();
  double* out = partial_out + chunk_out[c] * (int64_t)v.acc_len;
  const double ia2 = 1.0 / ((double)alpha * (double)alpha);
  bool bad = false;
  for (int n = tid; n < D; n += 128) {                  // row n: half n >> 5, lane n & 31 of the warps 0 + half and 2 + half
    const int hh = n >> 5, l = n & 31;
    const float x = dred[hh * 32 + l] + dred[(2 + hh) * 32 + l];
    __stcs(out + n, (double)x);
    bad |= !isfinite(x);
  }
  for (int i = tid; i < 64 * 64; i += 128) {
    const int n1 = i >> 6, n2 = i & 63;
    if (n1 < D && n2 <= n1) {                           // lower triangle incl. diagonal, mirrored: exactly symmetric
      const int hh = n1 >> 5, l = n1 & 31;
      const float x = red[(hh * 32 + l) * 65 + n2] + red[((2 + hh) * 32 + l) * 65 + n2];
      const double xd = ia2 * (double)x;
      __stcs(out + D + n1 * D + n2, xd);
      __stcs(out + D + n2 * D + n1, xd);
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
  if (warp == 0) tc_tmem_dealloc(tmem, 128);
}
