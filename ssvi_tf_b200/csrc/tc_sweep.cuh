// Launcher of the tensor-core sweep kernel (accumulate_tc.cuh); defined in tc_sweep.cu, its own translation unit.
#pragma once
#include "accumulate_memo.cuh"

// Returns a cudaError_t as int.  nt = ceil(D / 8) in 1..8 (5..8: two warps per entry), mv.n_oth in 1..3, mv.tab32 == 2.
int ssvi_launch_accumulate_tc(const DevView& v, const MemoView& mv, int nt, int64_t n_chunks, const int32_t* chunk_row,
                              const int64_t* chunk_start, const int32_t* chunk_len, const int32_t* chunk_slot,
                              const int64_t* chunk_out, const int64_t* entry_ids, const int64_t* rows, const void* z_own,
                              int z_f32, double* partial_out, cudaStream_t st);

// The fp32 kernels of the diagonal-covariance sweep (accumulate_diag32.cuh, under SSVI_TABLE_F32_TC).  Returns a cudaError_t as int; cudaErrorInvalidValue when the shared memory does not fit.
// lin != 0: SSVI_TF_simple + normal likelihood (closed-form weights); lin == 0: SSVI_TF_d (w_state, r_w, z_f unused).
int ssvi_launch_accumulate_diag32(const DevView& v, int lin, int n_oth, int noise_f32, int max_smem, int64_t n_chunks,
                                  const int32_t* chunk_row, const int64_t* chunk_start, const int32_t* chunk_len,
                                  const int32_t* chunk_slot, const int64_t* chunk_out, const int64_t* entry_ids,
                                  const int64_t* rows, const double* w_state, const void* z_other, const void* z_own,
                                  const void* r_w, const void* z_f, double* partial_out, cudaStream_t st);
