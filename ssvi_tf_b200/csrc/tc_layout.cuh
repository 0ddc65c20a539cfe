// Record layout of the sample table of the tensor-core sweep path (SSVI_TABLE_F32_TC): shared by the kernel that writes
// the table (sample_table_kernel, accumulate_memo.cuh) and the one that gathers it (accumulate_tc.cuh).
#pragma once

// Table record of the fp32 tensor-core path: NT row groups x 64 samples as float4 vectors
//     [rg][h][p][lane] float4,   lane = (d & 7) + 8 * ((k >> 3) & 3),  h = k >> 5,  p = (k >> 2) & 1,  element k & 3
// so lane (r8 = lane & 7, kq = lane >> 3) reads, for row group rg and half h, the 8 samples k = 8 (kq + 4 h) .. + 7 of row
// 8 rg + r8 with two fully coalesced 16-byte loads (p = 0, 1).
__host__ __device__ inline int tc_rec_index(int d, int k) {
  const int rg = d >> 3, r8 = d & 7, k8 = k >> 3, h = k8 >> 2, kq = k8 & 3, p = (k >> 2) & 1, e = k & 3;
  return ((((rg * 2 + h) * 2 + p) * 32 + (r8 + 8 * kq)) << 2) + e;
}
