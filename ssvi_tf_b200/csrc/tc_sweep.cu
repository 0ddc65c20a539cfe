// Translation unit of the tcgen05 sweep kernel (SSVI_TABLE_F32_TC): kept apart from ssvi_abi.cu so that the two compile
// in parallel.  No extern "C" entry point lives here: the kernel is reached through ssvi_sweep_accumulate.
#include "tc_sweep.cuh"
#include "accumulate_tc.cuh"
#include "accumulate_diag32.cuh"

template <int NT, int NOTH, bool RING, int PF>
static int launch_tc1(const DevView& v, const MemoView& mv, int64_t n_chunks, const int32_t* chunk_row,
                     const int64_t* chunk_start, const int32_t* chunk_len, const int32_t* chunk_slot,
                     const int64_t* chunk_out, const int64_t* entry_ids, const int64_t* rows, const void* z_own, int z_f32,
                     double* partial_out, cudaStream_t st) {
  auto kern = accumulate_tc_kernel<NT, NOTH, RING, PF>;
  constexpr int smem = tc_cta_smem(NOTH, RING);
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return (int)e;
  kern<<<(unsigned)n_chunks, 128, smem, st>>>(v, mv, n_chunks, chunk_row, chunk_start, chunk_len, chunk_slot, chunk_out,
                                                      entry_ids, rows, z_own, z_f32, partial_out);
  return (int)cudaGetLastError();
}

template <int NT, int NOTH>
static int launch_tc(const DevView& v, const MemoView& mv, int64_t n_chunks, const int32_t* chunk_row,
                     const int64_t* chunk_start, const int32_t* chunk_len, const int32_t* chunk_slot,
                     const int64_t* chunk_out, const int64_t* entry_ids, const int64_t* rows, const void* z_own, int z_f32,
                     double* partial_out, cudaStream_t st) {
#define TC_L(RING_, PF_)                                                                                                 \
  launch_tc1<NT, NOTH, RING_, PF_>(v, mv, n_chunks, chunk_row, chunk_start, chunk_len, chunk_slot, chunk_out, entry_ids, rows, \
                                   z_own, z_f32, partial_out, st)
  if (v.tuning & SSVI_TUNE_TC_RING) return TC_L(true, 2);
  if (v.tuning & SSVI_TUNE_TC_PF) return TC_L(false, tc_deep_pf(NT));
  return TC_L(false, 2);
#undef TC_L
}

template <int NT, int NOTH>
static int launch_tc_wide(const DevView& v, const MemoView& mv, int64_t n_chunks, const int32_t* chunk_row,
                          const int64_t* chunk_start, const int32_t* chunk_len, const int32_t* chunk_slot,
                          const int64_t* chunk_out, const int64_t* entry_ids, const int64_t* rows, const void* z_own,
                          int z_f32, double* partial_out, cudaStream_t st) {
  auto kern = accumulate_tc_wide_kernel<NT, NOTH>;
  constexpr int smem = tc_wide_smem();
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return (int)e;
  kern<<<(unsigned)n_chunks, 128, smem, st>>>(v, mv, n_chunks, chunk_row, chunk_start, chunk_len, chunk_slot, chunk_out,
                                              entry_ids, rows, z_own, z_f32, partial_out);
  return (int)cudaGetLastError();
}

int ssvi_launch_accumulate_tc(const DevView& v, const MemoView& mv, int nt, int64_t n_chunks, const int32_t* chunk_row,
                              const int64_t* chunk_start, const int32_t* chunk_len, const int32_t* chunk_slot,
                              const int64_t* chunk_out, const int64_t* entry_ids, const int64_t* rows, const void* z_own,
                              int z_f32, double* partial_out, cudaStream_t st) {
#define TC_ARGS v, mv, n_chunks, chunk_row, chunk_start, chunk_len, chunk_slot, chunk_out, entry_ids, rows, z_own, z_f32, partial_out, st
#define TC_NT(NT_)                                   \
  switch (mv.n_oth) {                                \
    case 1: return launch_tc<NT_, 1>(TC_ARGS);       \
    case 2: return launch_tc<NT_, 2>(TC_ARGS);       \
    default: return launch_tc<NT_, 3>(TC_ARGS);      \
  }
#define TC_WIDE(NT_)                                                                                            \
  switch (mv.n_oth) {                                                                                           \
    case 1: return launch_tc_wide<NT_, 1>(TC_ARGS);                                                             \
    case 2: return launch_tc_wide<NT_, 2>(TC_ARGS);                                                             \
    default: return launch_tc_wide<NT_, 3>(TC_ARGS);                                                            \
  }
  switch (nt) {
    case 1: TC_NT(1)
    case 2: TC_NT(2)
    case 3: TC_NT(3)
    case 4: TC_NT(4)
    case 5: TC_WIDE(5)
    case 6: TC_WIDE(6)
    case 7: TC_WIDE(7)
    default: TC_WIDE(8)
  }
#undef TC_WIDE
#undef TC_NT
#undef TC_ARGS
}

template <int NOTH, typename ZT>
static int launch_diag32(const DevView& v, int max_smem, int64_t n_chunks, const int32_t* chunk_row,
                         const int64_t* chunk_start, const int32_t* chunk_len, const int32_t* chunk_slot,
                         const int64_t* chunk_out, const int64_t* entry_ids, const int64_t* rows, const double* w_state,
                         const void* z_other, const void* z_own, const void* r_w, const void* z_f, double* partial_out,
                         cudaStream_t st) {
  const size_t smem = diag32_cta_smem(v.N, v.D, v.k1);
  if (smem > (size_t)max_smem) return (int)cudaErrorInvalidValue;
  auto kern = accumulate_diag_lin32_kernel<NOTH, ZT>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  kern<<<(unsigned)n_chunks, 128, smem, st>>>(v, n_chunks, chunk_row, chunk_start, chunk_len, chunk_slot, chunk_out, entry_ids,
                                              rows, w_state, (const ZT*)z_other, (const ZT*)z_own, (const ZT*)r_w,
                                              (const ZT*)z_f, partial_out);
  return (int)cudaGetLastError();
}

template <int NOTH, typename ZT>
static int launch_diag_d32(const DevView& v, int max_smem, int64_t n_chunks, const int32_t* chunk_row,
                           const int64_t* chunk_start, const int32_t* chunk_len, const int32_t* chunk_slot,
                           const int64_t* chunk_out, const int64_t* entry_ids, const int64_t* rows, const double*,
                           const void* z_other, const void* z_own, const void*, const void*, double* partial_out,
                           cudaStream_t st) {
  const size_t smem = diag32_d_cta_smem(v.N, v.D, v.k1);
  if (smem > (size_t)max_smem) return (int)cudaErrorInvalidValue;
  auto kern = accumulate_diag_d32_kernel<NOTH, ZT>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  kern<<<(unsigned)n_chunks, 128, smem, st>>>(v, n_chunks, chunk_row, chunk_start, chunk_len, chunk_slot, chunk_out, entry_ids,
                                              rows, (const ZT*)z_other, (const ZT*)z_own, partial_out);
  return (int)cudaGetLastError();
}

int ssvi_launch_accumulate_diag32(const DevView& v, int lin, int n_oth, int noise_f32, int max_smem, int64_t n_chunks,
                                  const int32_t* chunk_row, const int64_t* chunk_start, const int32_t* chunk_len,
                                  const int32_t* chunk_slot, const int64_t* chunk_out, const int64_t* entry_ids,
                                  const int64_t* rows, const double* w_state, const void* z_other, const void* z_own,
                                  const void* r_w, const void* z_f, double* partial_out, cudaStream_t st) {
#define D32_ARGS v, max_smem, n_chunks, chunk_row, chunk_start, chunk_len, chunk_slot, chunk_out, entry_ids, rows, w_state, z_other, z_own, r_w, z_f, partial_out, st
  if (!lin) {
    if (noise_f32) {
      switch (n_oth) {
        case 1: return launch_diag_d32<1, float>(D32_ARGS);
        case 2: return launch_diag_d32<2, float>(D32_ARGS);
        default: return launch_diag_d32<3, float>(D32_ARGS);
      }
    }
    switch (n_oth) {
      case 1: return launch_diag_d32<1, double>(D32_ARGS);
      case 2: return launch_diag_d32<2, double>(D32_ARGS);
      default: return launch_diag_d32<3, double>(D32_ARGS);
    }
  }
  if (noise_f32) {
    switch (n_oth) {
      case 1: return launch_diag32<1, float>(D32_ARGS);
      case 2: return launch_diag32<2, float>(D32_ARGS);
      default: return launch_diag32<3, float>(D32_ARGS);
    }
  }
  switch (n_oth) {
    case 1: return launch_diag32<1, double>(D32_ARGS);
    case 2: return launch_diag32<2, double>(D32_ARGS);
    default: return launch_diag32<3, double>(D32_ARGS);
  }
#undef D32_ARGS
}
