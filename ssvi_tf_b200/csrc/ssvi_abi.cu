// libssvi_b200.so -- C ABI (include/ssvi_b200.h) over the sm_100a kernels.
#include <stdio.h>
#include <string.h>
#include <new>

#include "accumulate_memo2.cuh"
#include "accumulate_wide.cuh"
#include "accumulate_diag.cuh"
#include "estimate.cuh"
#include "row_update.cuh"
#include "tc_sweep.cuh"
#include <vector>

#define SSVI_VERSION_STR "ssvi_b200 0.1 sm_100a"

static char g_create_err[512] = "";

static int fail(SsviHandle* h, int code, const char* fmt, const char* detail = "") {
  char* dst = h ? h->err : g_create_err;
  snprintf(dst, 512, fmt, detail);
  return code;
}

#define CUDA_TRY(h, expr)                                                                 \
  do {                                                                                    \
    cudaError_t e_ = (expr);                                                              \
    if (e_ != cudaSuccess) return fail((h), SSVI_E_CUDA, #expr ": %s", cudaGetErrorString(e_)); \
  } while (0)

static inline int cov_len_of(const SsviConfig& c) { return c.cov_kind == SSVI_COV_DIAG ? c.rank : c.rank * c.rank; }
static inline int row_len_of(const SsviConfig& c) { return c.rank + cov_len_of(c); }

extern "C" const char* ssvi_version(void) { return SSVI_VERSION_STR; }

extern "C" int64_t ssvi_acc_len(const SsviConfig* cfg) {
  if (!cfg) return 0;
  return (int64_t)cfg->rank + cov_len_of(*cfg) + 2;
}

extern "C" int64_t ssvi_sampler_row_len(const SsviConfig* cfg) {
  if (!cfg) return 0;
  return smp_row_len(cfg->rank, cfg->cov_kind == SSVI_COV_DIAG);
}

extern "C" int64_t ssvi_noise_block_len(const SsviConfig* cfg) {
  if (!cfg) return 0;
  return smp_zblk(cfg->rank, cfg->k1, cfg->cov_kind == SSVI_COV_DIAG);
}

extern "C" int64_t ssvi_noise_offset(const SsviConfig* cfg, int32_t c, int32_t k) {
  if (!cfg) return -1;
  return smp_z_offset(c, k, cfg->rank, cfg->k1, cfg->cov_kind == SSVI_COV_DIAG);
}

extern "C" const char* ssvi_last_error(const SsviHandle* h) { return h ? h->err : g_create_err; }

extern "C" int ssvi_create(const SsviConfig* cfg, SsviHandle** out) {
  if (!cfg || !out) return fail(nullptr, SSVI_E_INVALID, "ssvi_create: null argument%s");
  if (cfg->order < 2 || cfg->order > SSVI_MAX_ORDER) return fail(nullptr, SSVI_E_INVALID, "order must be in [2,8]%s");
  if (cfg->rank < 1 || cfg->rank > SSVI_MAX_RANK) return fail(nullptr, SSVI_E_INVALID, "rank must be in [1,64]%s");
  if (cfg->k1 < 1 || cfg->k2 < 1) return fail(nullptr, SSVI_E_INVALID, "k1, k2 must be >= 1%s");
  if (cfg->cov_kind != SSVI_COV_FULL && cfg->cov_kind != SSVI_COV_DIAG) return fail(nullptr, SSVI_E_INVALID, "bad cov_kind%s");
  if (cfg->model < SSVI_MODEL_D || cfg->model > SSVI_MODEL_ROBUST) return fail(nullptr, SSVI_E_INVALID, "bad model%s");
  if (cfg->likelihood < SSVI_LIK_NORMAL || cfg->likelihood > SSVI_LIK_POISSON) return fail(nullptr, SSVI_E_INVALID, "bad likelihood%s");
  // the reference raises Exception("Unidentified update formula ...") (SSVI_TF_interface.py:612, 632)
  if (cfg->mean_update != SSVI_UPDATE_S && cfg->mean_update != SSVI_UPDATE_N) return fail(nullptr, SSVI_E_INVALID, "Unidentified update formula for mean param%s");
  if (cfg->cov_update != SSVI_UPDATE_S && cfg->cov_update != SSVI_UPDATE_N) return fail(nullptr, SSVI_E_INVALID, "Unidentified update formula for covariance param%s");
  if (cfg->noise_dtype != SSVI_NOISE_F64 && cfg->noise_dtype != SSVI_NOISE_F32) return fail(nullptr, SSVI_E_INVALID, "bad noise_dtype%s");
  if (cfg->table_dtype != SSVI_TABLE_F64 && cfg->table_dtype != SSVI_TABLE_F32 && cfg->table_dtype != SSVI_TABLE_F32_TC)
    return fail(nullptr, SSVI_E_INVALID, "bad table_dtype%s");
  if (cfg->tc_scale_log2 < -14 || cfg->tc_scale_log2 > 14) return fail(nullptr, SSVI_E_INVALID, "tc_scale_log2 outside [-14, 14]%s");
  for (int d = 0; d < cfg->order; ++d)
    if (cfg->dims[d] < 1) return fail(nullptr, SSVI_E_INVALID, "dims must be >= 1%s");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
    return fail(nullptr, SSVI_E_CUDA, "no CUDA device: libssvi_b200 has no CPU fallback%s");
  SsviHandle* h = new (std::nothrow) SsviHandle();
  if (!h) return fail(nullptr, SSVI_E_INVALID, "out of host memory%s");
  memset(h, 0, sizeof(*h));
  h->cfg = *cfg;
  h->acc_len = ssvi_acc_len(cfg);
  for (int d = 0; d < cfg->order; ++d) h->win_n[d] = cfg->dims[d];      // sample-table row windows: every row
  int dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, dev);
  cudaDeviceGetAttribute(&h->max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  cudaError_t e = cudaMalloc(&h->d_flags, sizeof(int32_t));
  if (e == cudaSuccess) e = cudaMemset(h->d_flags, 0, sizeof(int32_t));
  if (e != cudaSuccess) {
    fail(nullptr, SSVI_E_CUDA, "cudaMalloc(flags): %s", cudaGetErrorString(e));
    delete h;
    return SSVI_E_CUDA;
  }
  *out = h;
  return SSVI_OK;
}

extern "C" void ssvi_destroy(SsviHandle* h) {
  if (!h) return;
  if (h->d_flags) cudaFree(h->d_flags);
  if (h->loop_ready) {
    cudaStreamDestroy(h->loop_stream);
    for (int i = 0; i < 2; ++i) {
      cudaEventDestroy(h->loop_ev_plan[i]);
      cudaEventDestroy(h->loop_ev_free[i]);
    }
    cudaEventDestroy(h->loop_ev_acc);
    cudaEventDestroy(h->loop_ev_noise);
  }
  delete h;
}

extern "C" int ssvi_set_row_windows(SsviHandle* h, const int64_t* lo, const int64_t* n) {
  if (!h || !lo || !n) return SSVI_E_INVALID;
  for (int d = 0; d < h->cfg.order; ++d)
    if (lo[d] < 0 || n[d] < 1 || lo[d] + n[d] > h->cfg.dims[d])
      return fail(h, SSVI_E_INVALID, "set_row_windows: window outside the mode's index range%s");
  for (int d = 0; d < h->cfg.order; ++d) {
    h->win_lo[d] = lo[d];
    h->win_n[d] = n[d];
  }
  return SSVI_OK;
}

extern "C" int ssvi_set_tuning(SsviHandle* h, int32_t tuning) {
  if (!h) return SSVI_E_INVALID;
  h->cfg.tuning = tuning;
  return SSVI_OK;
}

extern "C" int ssvi_bind_data(SsviHandle* h, const int32_t* coords, const double* vals, int64_t nnz) {
  if (!h) return SSVI_E_INVALID;
  if (nnz < 0 || (nnz > 0 && (!coords || !vals))) return fail(h, SSVI_E_INVALID, "bind_data: null pointer%s");
  h->coords = coords;
  h->vals = vals;
  h->nnz = nnz;
  h->data_bound = true;
  return SSVI_OK;
}

extern "C" int ssvi_bind_state(SsviHandle* h, double* const* params, double* const* sampler,
                               double* const* ada_mean, double* const* ada_cov, double* const* norm_changes,
                               double* const* grad_norms) {
  if (!h) return SSVI_E_INVALID;
  if (!params || !sampler || !ada_mean || !norm_changes || !grad_norms)
    return fail(h, SSVI_E_INVALID, "bind_state: null pointer array%s");
  if (h->cfg.cov_kind == SSVI_COV_DIAG && !ada_cov) return fail(h, SSVI_E_INVALID, "bind_state: ada_cov required for diagonal covariance%s");
  for (int d = 0; d < h->cfg.order; ++d) {
    if (!params[d] || !sampler[d] || !ada_mean[d] || !norm_changes[d] || !grad_norms[d])
      return fail(h, SSVI_E_INVALID, "bind_state: null device pointer%s");
    if (h->cfg.cov_kind == SSVI_COV_DIAG && !ada_cov[d]) return fail(h, SSVI_E_INVALID, "bind_state: null ada_cov pointer%s");
    h->params[d] = params[d];
    h->sampler[d] = sampler[d];
    h->ada_mean[d] = ada_mean[d];
    h->ada_cov[d] = ada_cov ? ada_cov[d] : nullptr;
    h->norm_changes[d] = norm_changes[d];
    h->grad_norms[d] = grad_norms[d];
  }
  h->state_bound = true;
  return SSVI_OK;
}

extern "C" int ssvi_bind_noise(SsviHandle* h, const void* z_other, const void* z_own, const void* r_w,
                               const void* z_f, int32_t pool_slots) {
  if (!h) return SSVI_E_INVALID;
  if (!z_other || !z_own || pool_slots < 1) return fail(h, SSVI_E_INVALID, "bind_noise: null pointer or empty pool%s");
  if (h->cfg.model != SSVI_MODEL_D && (!r_w || !z_f)) return fail(h, SSVI_E_INVALID, "bind_noise: r_w and z_f required for simple/robust%s");
  h->z_other = z_other;
  h->z_own = z_own;
  h->r_w = r_w;
  h->z_f = z_f;
  h->pool_slots = pool_slots;
  h->noise_bound = true;
  return SSVI_OK;
}

extern "C" int ssvi_bind_quadrature(SsviHandle* h, const double* rule, int32_t n) {
  if (!h) return SSVI_E_INVALID;
  if (n < 0 || (n > 0 && !rule)) return fail(h, SSVI_E_INVALID, "bind_quadrature: null rule%s");
  if (n > 0 && h->cfg.model != SSVI_MODEL_ROBUST) return fail(h, SSVI_E_INVALID, "bind_quadrature: only SSVI_TF_robust has a quadrature option%s");
  h->quad = n > 0 ? rule : nullptr;
  h->quad_n = n;
  return SSVI_OK;
}

static DevView make_view(const SsviHandle* h, int dim) {
  DevView v;
  memset(&v, 0, sizeof(v));
  const SsviConfig& c = h->cfg;
  v.N = c.order;
  v.D = c.rank;
  v.k1 = c.k1;
  v.k2 = c.k2;
  v.diag = c.cov_kind == SSVI_COV_DIAG;
  v.model = c.model;
  v.lik = c.likelihood;
  v.dim = dim;
  v.row_len = row_len_of(c);
  v.srow_len = smp_row_len(c.rank, c.cov_kind == SSVI_COV_DIAG);
  v.zblk = smp_zblk(c.rank, c.k1, c.cov_kind == SSVI_COV_DIAG);
  v.cov_len = cov_len_of(c);
  v.acc_len = (int32_t)h->acc_len;
  v.pool_slots = h->pool_slots;
  v.tuning = h->cfg.tuning;
  v.lik_param = c.likelihood_param;
  v.coords = h->coords;
  v.vals = h->vals;
  for (int d = 0; d < c.order; ++d) v.sampler[d] = h->sampler[d];
  v.flags = h->d_flags;
  v.quad = h->quad;
  v.quad_n = h->quad_n;
  return v;
}

extern "C" int ssvi_refresh_sampler(SsviHandle* h, int32_t dim, const int64_t* rows, int64_t n_rows, void* stream) {
  if (!h) return SSVI_E_INVALID;
  if (!h->state_bound) return fail(h, SSVI_E_STATE, "refresh_sampler: state not bound%s");
  if (dim < 0 || dim >= h->cfg.order) return fail(h, SSVI_E_INVALID, "refresh_sampler: bad dim%s");
  if (!rows) n_rows = h->cfg.dims[dim];
  if (n_rows <= 0) return SSVI_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const int D = h->cfg.rank;
  if (h->cfg.cov_kind == SSVI_COV_DIAG) {
    const int64_t total = n_rows * D;
    refresh_sampler_diag_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(h->params[dim], h->sampler[dim], rows, n_rows, D);
  } else {
    const size_t smem = (size_t)(D * D + D) * sizeof(double);
    refresh_sampler_full_kernel<<<(unsigned)n_rows, 128, smem, st>>>(h->params[dim], h->sampler[dim], rows, D, h->d_flags);
  }
  CUDA_TRY(h, cudaGetLastError());
  return SSVI_OK;
}

template <typename ZT>
static int launch_accumulate_generic(SsviHandle* h, const DevView& v, int64_t n_chunks, const int32_t* chunk_row,
                                     const int64_t* chunk_start, const int32_t* chunk_len, const int32_t* chunk_slot0,
                                     const int64_t* entry_ids, const int64_t* rows, const double* w_state,
                                     double* partial_out, cudaStream_t st) {
  const size_t per_warp = generic_warp_doubles(v.D, v.k1, v.srow_len, v.cov_len) * sizeof(double);
  int wpb = 4;
  while (wpb > 1 && per_warp * wpb > (size_t)h->max_smem_optin) wpb >>= 1;
  const size_t smem = per_warp * wpb;
  if (smem > (size_t)h->max_smem_optin) return fail(h, SSVI_E_INVALID, "accumulate: D*k1 too large for shared memory%s");
  auto kern = accumulate_generic_kernel<ZT>;
  CUDA_TRY(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const unsigned grid = (unsigned)((n_chunks + wpb - 1) / wpb);
  kern<<<grid, wpb * 32, smem, st>>>(v, n_chunks, chunk_row, chunk_start, chunk_len, chunk_slot0, entry_ids, rows,
                                     w_state, (const ZT*)h->z_other, (const ZT*)h->z_own, (const ZT*)h->r_w,
                                     (const ZT*)h->z_f, partial_out);
  CUDA_TRY(h, cudaGetLastError());
  return SSVI_OK;
}

template <int NT, typename ZT, bool SECOND>
static int launch_accumulate_mma(SsviHandle* h, const DevView& v, int64_t n_chunks, const int32_t* chunk_row,
                                 const int64_t* chunk_start, const int32_t* chunk_len, const int32_t* chunk_slot0,
                                 const int64_t* entry_ids, const int64_t* rows, const double* w_state,
                                 double* partial_out, cudaStream_t st) {
  const MmaSmem L = mma_smem_layout(NT, v.D, v.k1, (int)sizeof(ZT), SECOND);
  int wpb = 8;                                        // one CTA per SM; as many warps as shared memory holds
  while (wpb > 1 && (size_t)L.total * wpb > (size_t)h->max_smem_optin) --wpb;
  const size_t smem = (size_t)L.total * wpb;
  if (smem > (size_t)h->max_smem_optin) return fail(h, SSVI_E_INVALID, "accumulate: shared memory%s");
  auto kern = accumulate_mma_kernel<NT, ZT, SECOND>;
  CUDA_TRY(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const unsigned grid = (unsigned)((n_chunks + wpb - 1) / wpb);
  kern<<<grid, wpb * 32, smem, st>>>(v, n_chunks, chunk_row, chunk_start, chunk_len, chunk_slot0, entry_ids, rows,
                                     w_state, (const ZT*)h->z_other, (const ZT*)h->z_own, (const ZT*)h->r_w,
                                     (const ZT*)h->z_f, partial_out);
  CUDA_TRY(h, cudaGetLastError());
  return SSVI_OK;
}

template <typename ZT>
static int dispatch_accumulate(SsviHandle* h, const DevView& v, int64_t n_chunks, const int32_t* chunk_row,
                               const int64_t* chunk_start, const int32_t* chunk_len, const int32_t* chunk_slot0,
                               const int64_t* entry_ids, const int64_t* rows, const double* w_state,
                               double* partial_out, cudaStream_t st) {
#define SSVI_ARGS h, v, n_chunks, chunk_row, chunk_start, chunk_len, chunk_slot0, entry_ids, rows, w_state, partial_out, st
  const bool mma_ok = !v.diag && v.D <= 32 && (v.k1 % 8) == 0 && v.k1 <= 64 && h->cfg.kernel_path != 1;
  if (mma_ok) {
    const int nt = (v.D + 7) / 8;
    if (v.model == SSVI_MODEL_D) {
      switch (nt) {
        case 1: return launch_accumulate_mma<1, ZT, false>(SSVI_ARGS);
        case 2: return launch_accumulate_mma<2, ZT, false>(SSVI_ARGS);
        case 3: return launch_accumulate_mma<3, ZT, false>(SSVI_ARGS);
        default: return launch_accumulate_mma<4, ZT, false>(SSVI_ARGS);
      }
    }
    switch (nt) {
      case 1: return launch_accumulate_mma<1, ZT, true>(SSVI_ARGS);
      case 2: return launch_accumulate_mma<2, ZT, true>(SSVI_ARGS);
      case 3: return launch_accumulate_mma<3, ZT, true>(SSVI_ARGS);
      default: return launch_accumulate_mma<4, ZT, true>(SSVI_ARGS);
    }
  }
  return launch_accumulate_generic<ZT>(SSVI_ARGS);
#undef SSVI_ARGS
}

extern "C" int ssvi_accumulate(SsviHandle* h, int32_t dim, int64_t n_chunks, const int32_t* chunk_row,
                               const int64_t* chunk_start, const int32_t* chunk_len, const int32_t* chunk_slot0,
                               const int64_t* entry_ids, const int64_t* rows, const double* w_state,
                               double* partial_out, void* stream) {
  if (!h) return SSVI_E_INVALID;
  if (!h->data_bound || !h->state_bound || !h->noise_bound) return fail(h, SSVI_E_STATE, "accumulate: bind data, state and noise first%s");
  if (dim < 0 || dim >= h->cfg.order) return fail(h, SSVI_E_INVALID, "accumulate: bad dim%s");
  if (n_chunks == 0) return SSVI_OK;
  if (n_chunks < 0 || !chunk_row || !chunk_start || !chunk_len || !chunk_slot0 || !entry_ids || !rows || !partial_out)
    return fail(h, SSVI_E_INVALID, "accumulate: null pointer%s");
  if (h->cfg.model != SSVI_MODEL_D && !w_state) return fail(h, SSVI_E_INVALID, "accumulate: w_state required for simple/robust%s");
  const DevView v = make_view(h, dim);
  cudaStream_t st = (cudaStream_t)stream;
  if (h->cfg.noise_dtype == SSVI_NOISE_F32)
    return dispatch_accumulate<float>(h, v, n_chunks, chunk_row, chunk_start, chunk_len, chunk_slot0, entry_ids, rows, w_state, partial_out, st);
  return dispatch_accumulate<double>(h, v, n_chunks, chunk_row, chunk_start, chunk_len, chunk_slot0, entry_ids, rows, w_state, partial_out, st);
}

// ---- sample-table ("memoised") sweep path: accumulate_memo.cuh ------------------------------------------------------
static bool sweep_diag_ok(const SsviConfig& c, int max_smem) {
  return c.cov_kind == SSVI_COV_DIAG && c.rank <= 64 && (c.k1 % 16) == 0 && c.k1 <= 64 && c.order >= 2 &&
         c.order - 1 <= SSVI_MEMO_MAX_OTH && c.kernel_path != 1 &&
         diag_cta_smem(c.order, c.rank, c.k1) <= (size_t)max_smem;
}
static bool sweep_full_ok(const SsviConfig& c) {
  // rank <= 32: one warp holds an entry's whole contraction (accumulate_memo*.cuh, every model); 32 < rank <= 64: the
  // contraction split over the CTA's warps (accumulate_wide.cuh, SSVI_TF_d)
  return c.cov_kind == SSVI_COV_FULL && (c.rank <= 32 || (c.rank <= 64 && c.model == SSVI_MODEL_D)) &&
         c.k1 == 8 * SSVI_MEMO_KT && c.order >= 2 && c.order - 1 <= SSVI_MEMO_MAX_OTH && c.kernel_path != 1;
}
// the float table is built for the kernels of SSVI_TF_d (the three-pass path of simple / robust keeps doubles)
// ... or in the tensor-core path's order, for accumulate_tc[_wide]_kernel (fp32 arithmetic, tcgen05 contraction; rank <= 64)
static bool table_tc(const SsviConfig& c) {
  return c.table_dtype == SSVI_TABLE_F32_TC && c.model == SSVI_MODEL_D && c.cov_kind == SSVI_COV_FULL && c.rank <= 64;
}
static bool table_f32(const SsviConfig& c) {
  return table_tc(c) || ((c.table_dtype == SSVI_TABLE_F32 || c.table_dtype == SSVI_TABLE_F32_TC) && c.model == SSVI_MODEL_D &&
                         c.cov_kind == SSVI_COV_FULL && c.rank <= 32);
}
static bool sweep_ok(const SsviHandle* h) { return sweep_full_ok(h->cfg) || sweep_diag_ok(h->cfg, h->max_smem_optin); }

extern "C" int ssvi_sweep_supported(const SsviHandle* h) { return (h && sweep_ok(h)) ? 1 : 0; }

extern "C" int64_t ssvi_sweep_table_bytes(const SsviHandle* h, int32_t dim, int32_t n_slots) {
  if (!h || !sweep_full_ok(h->cfg) || dim < 0 || dim >= h->cfg.order || n_slots < 1) return 0;   // diagonal: no table
  int64_t recs = 0;
  for (int d = 0; d < h->cfg.order; ++d)
    if (d != dim || table_tc(h->cfg)) recs += h->win_n[d] * (int64_t)n_slots;   // tensor-core path: + the swept mode's own samples
  return recs * memo_rec_len(h->cfg.rank, h->cfg.k1) * (int64_t)(table_f32(h->cfg) ? sizeof(float) : sizeof(double));
}

extern "C" int64_t ssvi_sweep_work_bytes(const SsviHandle* h, int64_t n_entries) {
  if (!h || n_entries < 0 || !sweep_full_ok(h->cfg) || h->cfg.model == SSVI_MODEL_D) return 0;
  return memo_work_bytes(n_entries);
}

template <int NT, typename ZT>
static int launch_sweep(SsviHandle* h, const DevView& v, const MemoView& mv, int64_t n_chunks, const int32_t* chunk_row,
                        const int64_t* chunk_start, const int32_t* chunk_len, const int32_t* chunk_slot,
                        const int64_t* chunk_out, const int64_t* entry_ids, const int64_t* rows, double* partial_out,
                        const double* w_state, const MemoWork& wk, cudaStream_t st) {
  {
    const size_t smem = (size_t)sample_table_smem(v.D, v.k1);
    auto kern = sample_table_kernel<NT, ZT>;
    CUDA_TRY(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(unsigned)mv.blk_base[mv.n_modes], 256, smem, st>>>(v, mv, (const ZT*)h->z_other, (const ZT*)h->z_own);
    CUDA_TRY(h, cudaGetLastError());
  }
  if (n_chunks == 0) return SSVI_OK;
  const int z_f32 = sizeof(ZT) == 4;
  if (h->cfg.model != SSVI_MODEL_D) {                   // simple / robust: F pass, weights, contraction pass
    const size_t smf = (size_t)memo_f_smem(NT, v.D), smc = (size_t)memo_c_smem(NT, v.D);
#define SSVI_MEMO2_LAUNCH(NOTH)                                                                                    \
    do {                                                                                                           \
      auto kf = memo_f_kernel<NT, NOTH>;                                                                           \
      auto kc = memo_c_kernel<NT, NOTH>;                                                                           \
      CUDA_TRY(h, cudaFuncSetAttribute(kf, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smf));                \
      CUDA_TRY(h, cudaFuncSetAttribute(kc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smc));                \
      kf<<<(unsigned)n_chunks, 128, smf, st>>>(v, mv, wk, n_chunks, chunk_row, chunk_start, chunk_len, chunk_slot, \
                                               entry_ids, rows, h->z_own, z_f32);                                  \
      launch_memo_weights<ZT>(v, wk, w_state, (const ZT*)h->r_w, (const ZT*)h->z_f, st);                           \
      kc<<<(unsigned)n_chunks, 128, smc, st>>>(v, mv, wk, n_chunks, chunk_row, chunk_start, chunk_len, chunk_slot, \
                                               chunk_out, entry_ids, partial_out);                                 \
    } while (0)
    switch (mv.n_oth) {
      case 1: SSVI_MEMO2_LAUNCH(1); break;
      case 2: SSVI_MEMO2_LAUNCH(2); break;
      default: SSVI_MEMO2_LAUNCH(3); break;
    }
#undef SSVI_MEMO2_LAUNCH
    CUDA_TRY(h, cudaGetLastError());
    return SSVI_OK;
  }
  const size_t smem = (size_t)memo_cta_smem(NT, v.D);
#define SSVI_MEMO_LAUNCH_K(KERN)                                                                                   \
  do {                                                                                                             \
    auto kern = KERN;                                                                                              \
    CUDA_TRY(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));               \
    kern<<<(unsigned)n_chunks, 128, smem, st>>>(v, mv, n_chunks, chunk_row, chunk_start, chunk_len, chunk_slot,    \
                                                chunk_out, entry_ids, rows, h->z_own, z_f32, partial_out);         \
  } while (0)
#define SSVI_MEMO_LAUNCH(NOTH, ONES, LIN) SSVI_MEMO_LAUNCH_K((accumulate_memo_kernel<NT, NOTH, ONES, LIN>))
  // config 5's instances (four row tiles, two other modes) with a deeper register prefetch at two CTAs per SM
  if constexpr (NT == 4) {
    const int pfv = (h->cfg.tuning >> 2) & 3;
    if (mv.n_oth == 2 && pfv && (v.D & 7) == 0 && !mv.tab32) {
      const bool lin4 = h->cfg.likelihood == SSVI_LIK_NORMAL && (h->cfg.tuning & SSVI_TUNE_LIN);
      if (lin4 && pfv == 1) SSVI_MEMO_LAUNCH_K((accumulate_memo_kernel<4, 2, false, true, 2, 2>));
      else if (lin4) SSVI_MEMO_LAUNCH_K((accumulate_memo_kernel<4, 2, false, true, 3, 2>));
      else if (pfv == 1) SSVI_MEMO_LAUNCH_K((accumulate_memo_kernel<4, 2, false, false, 2, 2>));
      else SSVI_MEMO_LAUNCH_K((accumulate_memo_kernel<4, 2, false, false, 3, 2>));
      CUDA_TRY(h, cudaGetLastError());
      return SSVI_OK;
    }
  }
  if (mv.tab32 == 2) {                                  // SSVI_TABLE_F32_TC: tc_sweep.cu
    const cudaError_t e = (cudaError_t)ssvi_launch_accumulate_tc(v, mv, NT, n_chunks, chunk_row, chunk_start, chunk_len,
                                                                 chunk_slot, chunk_out, entry_ids, rows, h->z_own, z_f32,
                                                                 partial_out, st);
    CUDA_TRY(h, e);
    return SSVI_OK;
  }
  if (mv.tab32) {                                       // SSVI_TABLE_F32: the base kernel over a float table
    constexpr int MB = (NT <= 3 ? SSVI_MEMO_MINB : 3);
    const bool ones32 = (v.D & 7) != 0;
    switch (mv.n_oth * 2 + (ones32 ? 1 : 0)) {
      case 2: SSVI_MEMO_LAUNCH_K((accumulate_memo_kernel<NT, 1, false, false, 1, MB, true>)); break;
      case 3: SSVI_MEMO_LAUNCH_K((accumulate_memo_kernel<NT, 1, true, false, 1, MB, true>)); break;
      case 4: SSVI_MEMO_LAUNCH_K((accumulate_memo_kernel<NT, 2, false, false, 1, MB, true>)); break;
      case 5: SSVI_MEMO_LAUNCH_K((accumulate_memo_kernel<NT, 2, true, false, 1, MB, true>)); break;
      case 6: SSVI_MEMO_LAUNCH_K((accumulate_memo_kernel<NT, 3, false, false, 1, MB, true>)); break;
      default: SSVI_MEMO_LAUNCH_K((accumulate_memo_kernel<NT, 3, true, false, 1, MB, true>)); break;
    }
    CUDA_TRY(h, cudaGetLastError());
    return SSVI_OK;
  }
  const bool ones = (v.D & 7) != 0;
  // normal likelihood: weights linear in f, contraction accumulated across the chunk (accumulate_memo.cuh, LIN)
  const bool lin = h->cfg.likelihood == SSVI_LIK_NORMAL && (h->cfg.tuning & SSVI_TUNE_LIN);
  switch (lin ? 10 + mv.n_oth : mv.n_oth * 2 + (ones ? 1 : 0)) {
    case 2: SSVI_MEMO_LAUNCH(1, false, false); break;
    case 3: SSVI_MEMO_LAUNCH(1, true, false); break;
    case 4: SSVI_MEMO_LAUNCH(2, false, false); break;
    case 5: SSVI_MEMO_LAUNCH(2, true, false); break;
    case 6: SSVI_MEMO_LAUNCH(3, false, false); break;
    case 7: SSVI_MEMO_LAUNCH(3, true, false); break;
    case 11: SSVI_MEMO_LAUNCH(1, false, true); break;
    case 12: SSVI_MEMO_LAUNCH(2, false, true); break;
    default: SSVI_MEMO_LAUNCH(3, false, true); break;
  }
#undef SSVI_MEMO_LAUNCH
#undef SSVI_MEMO_LAUNCH_K
  CUDA_TRY(h, cudaGetLastError());
  return SSVI_OK;
}

template <typename ZT>
static int launch_sweep_wide(SsviHandle* h, const DevView& v, const MemoView& mv, int nt, int64_t n_chunks,
                             const int32_t* chunk_row, const int64_t* chunk_start, const int32_t* chunk_len,
                             const int32_t* chunk_slot, const int64_t* chunk_out, const int64_t* entry_ids,
                             const int64_t* rows, double* partial_out, cudaStream_t st) {
  {
    const size_t smem = (size_t)sample_table_wide_smem(v.D, v.k1);
    auto kern = sample_table_wide_kernel<ZT>;
    CUDA_TRY(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(unsigned)mv.blk_base[mv.n_modes], 128, smem, st>>>(v, mv, (const ZT*)h->z_other, (const ZT*)h->z_own);
    CUDA_TRY(h, cudaGetLastError());
  }
  if (n_chunks == 0) return SSVI_OK;
  const int z_f32 = sizeof(ZT) == 4;
  if (mv.tab32 == 2) {                                  // SSVI_TABLE_F32_TC: tc_sweep.cu (two warps per entry)
    CUDA_TRY(h, (cudaError_t)ssvi_launch_accumulate_tc(v, mv, nt, n_chunks, chunk_row, chunk_start, chunk_len, chunk_slot,
                                                       chunk_out, entry_ids, rows, h->z_own, z_f32, partial_out, st));
    return SSVI_OK;
  }
  const size_t smem = (size_t)wide_cta_smem(nt, v.D);
  if (smem > (size_t)h->max_smem_optin) return fail(h, SSVI_E_INVALID, "sweep_accumulate: shared memory%s");
#define SSVI_WIDE_LAUNCH(NT_, NOTH_)                                                                               \
  do {                                                                                                             \
    auto kern = accumulate_wide_kernel<NT_, NOTH_>;                                                                \
    CUDA_TRY(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));               \
    kern<<<(unsigned)n_chunks, 128, smem, st>>>(v, mv, n_chunks, chunk_row, chunk_start, chunk_len, chunk_slot,    \
                                                chunk_out, entry_ids, rows, h->z_own, z_f32, partial_out);         \
  } while (0)
#define SSVI_WIDE_NT(NT_)                                                                                          \
  switch (mv.n_oth) {                                                                                              \
    case 1: SSVI_WIDE_LAUNCH(NT_, 1); break;                                                                       \
    case 2: SSVI_WIDE_LAUNCH(NT_, 2); break;                                                                       \
    default: SSVI_WIDE_LAUNCH(NT_, 3); break;                                                                      \
  }
  switch (nt) {
    case 5: SSVI_WIDE_NT(5) break;
    case 6: SSVI_WIDE_NT(6) break;
    case 7: SSVI_WIDE_NT(7) break;
    default: SSVI_WIDE_NT(8) break;
  }
#undef SSVI_WIDE_NT
#undef SSVI_WIDE_LAUNCH
  CUDA_TRY(h, cudaGetLastError());
  return SSVI_OK;
}

template <int NOTH, typename ZT, bool SECOND, bool LIN>
static int launch_sweep_diag(SsviHandle* h, const DevView& v, int64_t n_chunks, const int32_t* chunk_row,
                             const int64_t* chunk_start, const int32_t* chunk_len, const int32_t* chunk_slot,
                             const int64_t* chunk_out, const int64_t* entry_ids, const int64_t* rows,
                             const double* w_state, double* partial_out, cudaStream_t st) {
  const size_t smem = diag_cta_smem(v.N, v.D, v.k1);
  if constexpr (LIN) {                                  // closed-form weights, SSVI_DIAG_LIN_E entries per warp step
    // BASELINE config 4's shape (4-way, D = 50, k1 = 64) has an instance with the rank compiled in
    const bool c4 = NOTH == 3 && v.D == 50 && v.k1 == 64;
    auto kern = c4 ? accumulate_diag_lin_kernel<NOTH, ZT, SSVI_DIAG_LIN_E, SSVI_DIAG_LIN_T, (SSVI_DIAG_LIN_PF != 0), (NOTH == 3 ? 50 : 0)>
                   : accumulate_diag_lin_kernel<NOTH, ZT, SSVI_DIAG_LIN_E, SSVI_DIAG_LIN_T, (SSVI_DIAG_LIN_PF != 0), 0>;
    CUDA_TRY(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(unsigned)n_chunks, 128, smem, st>>>(v, n_chunks, chunk_row, chunk_start, chunk_len, chunk_slot, chunk_out,
                                                entry_ids, rows, w_state, (const ZT*)h->z_other, (const ZT*)h->z_own,
                                                (const ZT*)h->r_w, (const ZT*)h->z_f, partial_out);
  } else if constexpr (!SECOND) {                       // SSVI_TF_d: four entries per warp step
    auto kern = accumulate_diag_d_kernel<NOTH, ZT>;
    CUDA_TRY(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(unsigned)n_chunks, 128, smem, st>>>(v, n_chunks, chunk_row, chunk_start, chunk_len, chunk_slot, chunk_out,
                                                entry_ids, rows, (const ZT*)h->z_other, (const ZT*)h->z_own, partial_out);
  } else {
    auto kern = accumulate_diag_kernel<NOTH, ZT>;
    CUDA_TRY(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(unsigned)n_chunks, 128, smem, st>>>(v, n_chunks, chunk_row, chunk_start, chunk_len, chunk_slot, chunk_out,
                                                entry_ids, rows, w_state, (const ZT*)h->z_other, (const ZT*)h->z_own,
                                                (const ZT*)h->r_w, (const ZT*)h->z_f, partial_out);
  }
  CUDA_TRY(h, cudaGetLastError());
  return SSVI_OK;
}

static int sweep_diag(SsviHandle* h, int32_t dim, int64_t n_chunks, const int32_t* chunk_row, const int64_t* chunk_start,
                      const int32_t* chunk_len, const int32_t* chunk_slot, const int64_t* chunk_out,
                      const int64_t* entry_ids, const int64_t* rows, const double* w_state, double* partial_out,
                      cudaStream_t st) {
  if (n_chunks == 0) return SSVI_OK;
  if (!chunk_row || !chunk_start || !chunk_len || !chunk_slot || !chunk_out || !entry_ids || !rows || !partial_out)
    return fail(h, SSVI_E_INVALID, "sweep_accumulate: null pointer%s");
  const bool second = h->cfg.model != SSVI_MODEL_D;
  if (second && !w_state) return fail(h, SSVI_E_INVALID, "sweep_accumulate: w_state required for simple/robust%s");
  const DevView v = make_view(h, dim);
#define SSVI_DIAG_ARGS h, v, n_chunks, chunk_row, chunk_start, chunk_len, chunk_slot, chunk_out, entry_ids, rows, w_state, partial_out, st
  // closed-form weights: SSVI_TF_simple + normal likelihood (BASELINE config 4), see accumulate_diag.cuh
  const bool lin = h->cfg.model == SSVI_MODEL_SIMPLE && h->cfg.likelihood == SSVI_LIK_NORMAL;
  // the fp32 arithmetic option (tc_sweep.cu): the closed-form kernel and SSVI_TF_d's; the general simple / robust kernel
  // (k1 x k2 likelihood evaluations per entry) has no fp32 instance
  if ((lin || !second) && h->cfg.table_dtype == SSVI_TABLE_F32_TC && (h->cfg.k1 % 8) == 0) {
    CUDA_TRY(h, (cudaError_t)ssvi_launch_accumulate_diag32(
                    v, lin ? 1 : 0, h->cfg.order - 1, h->cfg.noise_dtype == SSVI_NOISE_F32, h->max_smem_optin, n_chunks, chunk_row,
                    chunk_start, chunk_len, chunk_slot, chunk_out, entry_ids, rows, w_state, h->z_other, h->z_own, h->r_w,
                    h->z_f, partial_out, st));
    return SSVI_OK;
  }
#define SSVI_DIAG_CASE(NOTH)                                                                              \
  if (h->cfg.noise_dtype == SSVI_NOISE_F32)                                                               \
    return lin ? launch_sweep_diag<NOTH, float, true, true>(SSVI_DIAG_ARGS)                               \
               : second ? launch_sweep_diag<NOTH, float, true, false>(SSVI_DIAG_ARGS)                     \
                        : launch_sweep_diag<NOTH, float, false, false>(SSVI_DIAG_ARGS);                   \
  return lin ? launch_sweep_diag<NOTH, double, true, true>(SSVI_DIAG_ARGS)                                \
             : second ? launch_sweep_diag<NOTH, double, true, false>(SSVI_DIAG_ARGS)                      \
                      : launch_sweep_diag<NOTH, double, false, false>(SSVI_DIAG_ARGS);
  switch (h->cfg.order - 1) {
    case 1: SSVI_DIAG_CASE(1)
    case 2: SSVI_DIAG_CASE(2)
    default: SSVI_DIAG_CASE(3)
  }
#undef SSVI_DIAG_CASE
#undef SSVI_DIAG_ARGS
}

extern "C" int ssvi_sweep_accumulate(SsviHandle* h, int32_t dim, int32_t slot_lo, int32_t n_slots, int64_t n_chunks,
                                     const int32_t* chunk_row, const int64_t* chunk_start, const int32_t* chunk_len,
                                     const int32_t* chunk_slot, const int64_t* chunk_out, const int64_t* entry_ids,
                                     const int64_t* rows, const double* w_state, void* table_ws,
                                     int64_t table_ws_bytes, void* work_ws, int64_t work_ws_bytes, int64_t entry_lo,
                                     int64_t n_entries, double* partial_out, void* stream) {
  if (!h) return SSVI_E_INVALID;
  if (!h->data_bound || !h->state_bound || !h->noise_bound) return fail(h, SSVI_E_STATE, "sweep_accumulate: bind data, state and noise first%s");
  if (!sweep_ok(h)) return fail(h, SSVI_E_INVALID, "sweep_accumulate: configuration not supported by the sample-table path (see ssvi_sweep_supported)%s");
  if (dim < 0 || dim >= h->cfg.order) return fail(h, SSVI_E_INVALID, "sweep_accumulate: bad dim%s");
  if (n_slots < 1 || slot_lo < 0 || slot_lo + n_slots > h->pool_slots) return fail(h, SSVI_E_INVALID, "sweep_accumulate: slot range outside the noise pool%s");
  if (n_chunks < 0) return fail(h, SSVI_E_INVALID, "sweep_accumulate: negative chunk count%s");
  if (h->cfg.cov_kind == SSVI_COV_DIAG)
    return sweep_diag(h, dim, n_chunks, chunk_row, chunk_start, chunk_len, chunk_slot, chunk_out, entry_ids, rows,
                      w_state, partial_out, (cudaStream_t)stream);
  if (!table_ws) return fail(h, SSVI_E_INVALID, "sweep_accumulate: null workspace%s");
  if (n_chunks > 0 && (!chunk_row || !chunk_start || !chunk_len || !chunk_slot || !chunk_out || !entry_ids || !rows || !partial_out))
    return fail(h, SSVI_E_INVALID, "sweep_accumulate: null pointer%s");
  if (table_ws_bytes < ssvi_sweep_table_bytes(h, dim, n_slots)) return fail(h, SSVI_E_INVALID, "sweep_accumulate: table workspace too small%s");
  const DevView v = make_view(h, dim);
  MemoView mv;
  memset(&mv, 0, sizeof(mv));
  mv.n_oth = h->cfg.order - 1;
  mv.rec_len = memo_rec_len(h->cfg.rank, h->cfg.k1);
  mv.slot_lo = slot_lo;
  mv.n_slots = n_slots;
  mv.tab = (double*)table_ws;
  mv.tab32 = table_tc(h->cfg) ? 2 : (table_f32(h->cfg) ? 1 : 0);
  mv.tc_alpha = ldexpf(1.f, h->cfg.tc_scale_log2 ? h->cfg.tc_scale_log2 : 6);
  int jj = 0;
  int64_t base = 0, blocks = 0;
  for (int d = 0; d < h->cfg.order; ++d) {
    if (d == dim) continue;
    mv.od[jj] = d;
    mv.n_rows[jj] = h->win_n[d];
    mv.row_lo[jj] = h->win_lo[d];
    mv.rec_base[jj] = base;
    mv.blk_base[jj] = blocks;
    base += h->win_n[d] * (int64_t)n_slots;
    blocks += ((h->win_n[d] + 7) / 8) * (int64_t)n_slots;
    ++jj;
  }
  mv.own = mv.tab32 == 2 ? 1 : 0;
  mv.n_modes = mv.n_oth + mv.own;
  if (mv.own) {                                         // the swept mode's own samples as one more "mode" of the table
    mv.od[jj] = dim;
    mv.n_rows[jj] = h->win_n[dim];
    mv.row_lo[jj] = h->win_lo[dim];
    mv.rec_base[jj] = base;
    mv.blk_base[jj] = blocks;
    base += h->win_n[dim] * (int64_t)n_slots;
    blocks += ((h->win_n[dim] + 7) / 8) * (int64_t)n_slots;
    ++jj;
  }
  mv.rec_base[jj] = base;
  mv.blk_base[jj] = blocks;
  if (base >= ((int64_t)1 << 31)) return fail(h, SSVI_E_INVALID, "sweep_accumulate: too many table records; use fewer slots per call%s");
  MemoWork wk;
  memset(&wk, 0, sizeof(wk));
  if (h->cfg.model != SSVI_MODEL_D && n_chunks > 0) {
    if (!w_state) return fail(h, SSVI_E_INVALID, "sweep_accumulate: w_state required for simple/robust%s");
    if (n_entries < 0 || entry_lo < 0 || !work_ws || work_ws_bytes < memo_work_bytes(n_entries))
      return fail(h, SSVI_E_INVALID, "sweep_accumulate: work workspace missing or too small (ssvi_sweep_work_bytes)%s");
    unsigned char* b = (unsigned char*)work_ws;
    wk.F = (double*)b;
    wk.W = (double2*)(b + n_entries * 512);
    wk.SI = (double*)(b + n_entries * 1536);
    wk.Y = (double*)(b + n_entries * 2048);
    wk.SL = (int32_t*)(b + n_entries * 2056);
    wk.e_lo = entry_lo;
    wk.n = n_entries;
  }
  cudaStream_t st = (cudaStream_t)stream;
#define SSVI_SWEEP_ARGS h, v, mv, n_chunks, chunk_row, chunk_start, chunk_len, chunk_slot, chunk_out, entry_ids, rows, partial_out, w_state, wk, st
  const int nt = (v.D + 7) / 8;
  if (nt > 4) {                                         // 32 < rank <= 64: accumulate_wide.cuh (4 table rows per CTA)
    int64_t blocks = 0;
    for (int a = 0; a < mv.n_modes; ++a) {
      mv.blk_base[a] = blocks;
      blocks += ((mv.n_rows[a] + 3) / 4) * (int64_t)n_slots;
    }
    mv.blk_base[mv.n_modes] = blocks;
    if (h->cfg.noise_dtype == SSVI_NOISE_F32)
      return launch_sweep_wide<float>(h, v, mv, nt, n_chunks, chunk_row, chunk_start, chunk_len, chunk_slot, chunk_out,
                                      entry_ids, rows, partial_out, st);
    return launch_sweep_wide<double>(h, v, mv, nt, n_chunks, chunk_row, chunk_start, chunk_len, chunk_slot, chunk_out,
                                     entry_ids, rows, partial_out, st);
  }
  if (h->cfg.noise_dtype == SSVI_NOISE_F32) {
    switch (nt) {
      case 1: return launch_sweep<1, float>(SSVI_SWEEP_ARGS);
      case 2: return launch_sweep<2, float>(SSVI_SWEEP_ARGS);
      case 3: return launch_sweep<3, float>(SSVI_SWEEP_ARGS);
      default: return launch_sweep<4, float>(SSVI_SWEEP_ARGS);
    }
  }
  switch (nt) {
    case 1: return launch_sweep<1, double>(SSVI_SWEEP_ARGS);
    case 2: return launch_sweep<2, double>(SSVI_SWEEP_ARGS);
    case 3: return launch_sweep<3, double>(SSVI_SWEEP_ARGS);
    default: return launch_sweep<4, double>(SSVI_SWEEP_ARGS);
  }
#undef SSVI_SWEEP_ARGS
}

// acc_out[r][i] = sum over the row's chunk records, in chunk order (deterministic)
__global__ void reduce_chunks_kernel(int64_t n_rows, int acc_len, const int64_t* __restrict__ row_chunk_ptr,
                                     const double* __restrict__ partial, double* __restrict__ acc_out, int accumulate) {
  const int64_t r = blockIdx.x;
  const int64_t c0 = row_chunk_ptr[r], c1 = row_chunk_ptr[r + 1];
  if (accumulate && c0 == c1) return;
  for (int i = threadIdx.x; i < acc_len; i += blockDim.x) {
    double s = 0.0;
    int64_t c = c0;
    for (; c + 8 <= c1; c += 8) {                        // eight loads in flight, added in chunk order (same sum, bit for bit)
      double x[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) x[j] = partial[(c + j) * acc_len + i];
#pragma unroll
      for (int j = 0; j < 8; ++j) s += x[j];
    }
    for (; c < c1; ++c) s += partial[c * acc_len + i];
    acc_out[r * acc_len + i] = accumulate ? acc_out[r * acc_len + i] + s : s;
  }
}

extern "C" int ssvi_reduce_chunks(SsviHandle* h, int64_t n_rows, const int64_t* row_chunk_ptr, const double* partial,
                                  double* acc_out, void* stream) {
  if (!h) return SSVI_E_INVALID;
  if (n_rows == 0) return SSVI_OK;
  if (n_rows < 0 || !row_chunk_ptr || !partial || !acc_out) return fail(h, SSVI_E_INVALID, "reduce_chunks: null pointer%s");
  reduce_chunks_kernel<<<(unsigned)n_rows, 128, 0, (cudaStream_t)stream>>>(n_rows, (int)h->acc_len, row_chunk_ptr, partial, acc_out, 0);
  CUDA_TRY(h, cudaGetLastError());
  return SSVI_OK;
}

extern "C" int ssvi_reduce_chunks_into(SsviHandle* h, int64_t n_rows, const int64_t* row_chunk_ptr, const double* partial,
                                       double* acc_inout, int32_t accumulate, void* stream) {
  if (!h) return SSVI_E_INVALID;
  if (n_rows == 0) return SSVI_OK;
  if (n_rows < 0 || !row_chunk_ptr || !partial || !acc_inout)
    return fail(h, SSVI_E_INVALID, "reduce_chunks_into: null pointer%s");
  reduce_chunks_kernel<<<(unsigned)n_rows, 128, 0, (cudaStream_t)stream>>>(n_rows, (int)h->acc_len, row_chunk_ptr, partial,
                                                                           acc_inout, accumulate ? 1 : 0);
  CUDA_TRY(h, cudaGetLastError());
  return SSVI_OK;
}

extern "C" int ssvi_row_update(SsviHandle* h, int32_t dim, int64_t n_rows, const int64_t* rows, const double* acc,
                               const int64_t* n_obs, const double* pSigma, const double* w_state, int32_t time_step,
                               int32_t fixed_covariance, double* si_out, void* stream) {
  if (!h) return SSVI_E_INVALID;
  if (!h->state_bound) return fail(h, SSVI_E_STATE, "row_update: state not bound%s");
  if (dim < 0 || dim >= h->cfg.order) return fail(h, SSVI_E_INVALID, "row_update: bad dim%s");
  if (n_rows == 0) return SSVI_OK;
  if (n_rows < 0 || !rows || !acc || !n_obs || !pSigma) return fail(h, SSVI_E_INVALID, "row_update: null pointer%s");
  if (h->cfg.model != SSVI_MODEL_D && !w_state) return fail(h, SSVI_E_INVALID, "row_update: w_state required for simple/robust%s");
  const SsviConfig& c = h->cfg;
  RowUpdateArgs a;
  memset(&a, 0, sizeof(a));
  a.D = c.rank;
  a.diag = c.cov_kind == SSVI_COV_DIAG;
  a.model = c.model;
  a.lik = c.likelihood;
  a.k1 = c.k1;
  a.mean_update = c.mean_update;
  a.cov_update = c.cov_update;
  a.dim = dim;
  a.time_step = time_step;
  a.fixed_cov = fixed_covariance;
  a.row_len = row_len_of(c);
  a.srow_len = smp_row_len(c.rank, c.cov_kind == SSVI_COV_DIAG);
  a.cov_len = cov_len_of(c);
  a.acc_len = (int32_t)h->acc_len;
  a.eta = c.eta;
  a.cov_eta = c.cov_eta;
  a.poisson_eta = c.poisson_eta;
  a.min_eigenval = c.min_eigenval;
  a.cov_floor = c.cov_floor;
  a.params = h->params[dim];
  a.sampler = h->sampler[dim];
  a.ada_mean = h->ada_mean[dim];
  a.ada_cov = h->ada_cov[dim];
  a.norm_changes = h->norm_changes[dim];
  a.grad_norms = h->grad_norms[dim];
  a.rows = rows;
  a.acc = acc;
  a.n_obs = n_obs;
  a.pSigma = pSigma;
  a.w_state = w_state;
  a.si_out = si_out;
  a.flags = h->d_flags;
  a.stop = h->stop;
  cudaStream_t st = (cudaStream_t)stream;
  if (a.diag) {
    const int threads = ((c.rank + 31) / 32) * 32;
    row_update_diag_kernel<<<(unsigned)n_rows, threads, 0, st>>>(a);
  } else {
    const size_t smem = row_update_full_doubles(c.rank) * sizeof(double);
    if (smem > (size_t)h->max_smem_optin) return fail(h, SSVI_E_INVALID, "row_update: rank too large for shared memory%s");
    CUDA_TRY(h, cudaFuncSetAttribute(row_update_full_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    row_update_full_kernel<<<(unsigned)n_rows, 128, smem, st>>>(a);
  }
  CUDA_TRY(h, cudaGetLastError());
  return SSVI_OK;
}

// update_sigma_param (SSVI_TF_interface.py:684-691; robust: SSVI_TF_robust.py:216-232) replayed in row order.
__global__ void sigma_update_kernel(int64_t n_rows, int acc_len, int model, double sigma_eta,
                                    const double* __restrict__ si, const double* __restrict__ acc, double* w_state,
                                    const int32_t* stop) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  if (stop && *stop) return;
  double w_sigma = w_state[0], w_ada = w_state[1];
  const double w_tau = w_state[2];
  for (int64_t r = 0; r < n_rows; ++r) {
    if (acc[r * acc_len + acc_len - 1] == 0.0) continue;
    const double w_grad = -1.0 / (2.0 * w_tau * w_tau) + si[r];
    w_ada += w_grad * w_grad;                                  // :716-719
    const double w_step = sigma_eta / w_ada;
    const double update = (1.0 - w_step) * (-0.5 / (w_sigma * w_sigma)) + w_step * w_grad;
    const double nxt = sqrt(-0.5 / update);
    if (model == SSVI_MODEL_ROBUST && isnan(nxt)) continue;    // SSVI_TF_robust.py:229-230
    w_sigma = nxt;
  }
  w_state[0] = w_sigma;
  w_state[1] = w_ada;
}

extern "C" int ssvi_sigma_update(SsviHandle* h, int64_t n_rows, const double* si, const double* acc, double* w_state,
                                 void* stream) {
  if (!h) return SSVI_E_INVALID;
  if (h->cfg.model == SSVI_MODEL_D || n_rows == 0) return SSVI_OK;
  if (n_rows < 0 || !si || !acc || !w_state) return fail(h, SSVI_E_INVALID, "sigma_update: null pointer%s");
  sigma_update_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(n_rows, (int)h->acc_len, h->cfg.model, h->cfg.sigma_eta, si, acc, w_state,
                                                          h->stop);
  CUDA_TRY(h, cudaGetLastError());
  return SSVI_OK;
}

// update_hyper_parameter (:146-160) + per-mode check_stop_cond (:967-984).  Single CTA: deterministic.
__global__ void __launch_bounds__(1024) hyper_reduce_kernel(const double* __restrict__ params,
                                                            const double* __restrict__ norm_changes, int64_t n_rows,
                                                            int D, int diag, double* out, double* psigma_out,
                                                            const int32_t* stop) {
  __shared__ double s_sum[32], s_m[32], s_c[32];
  if (stop && *stop) return;
  const int row_len = diag ? 2 * D : (D + 1) * D;
  double sum = 0.0, mm = 0.0, mc = 0.0;
  for (int64_t t = threadIdx.x; t < n_rows * D; t += blockDim.x) {
    const int64_t r = t / D;
    const int n = (int)(t % D);
    const double* p = params + r * row_len;
    const double m = p[n];
    const double s = diag ? p[D + n] : p[D + n * D + n];
    sum += s + m * m;
  }
  for (int64_t r = threadIdx.x; r < n_rows; r += blockDim.x) {
    mm = fmax(mm, norm_changes[r * 2]);
    mc = fmax(mc, norm_changes[r * 2 + 1]);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  sum = warp_sum(sum);
  for (int o = 16; o > 0; o >>= 1) {
    mm = fmax(mm, __shfl_xor_sync(0xffffffffu, mm, o));
    mc = fmax(mc, __shfl_xor_sync(0xffffffffu, mc, o));
  }
  if (lane == 0) {
    s_sum[warp] = sum;
    s_m[warp] = mm;
    s_c[warp] = mc;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0, a = 0.0, b = 0.0;
    for (int i = 0; i < (int)(blockDim.x >> 5); ++i) {
      t += s_sum[i];
      a = fmax(a, s_m[i]);
      b = fmax(b, s_c[i]);
    }
    out[0] = t / ((double)n_rows * (double)D);
    out[1] = a;
    out[2] = b;
    if (psigma_out) *psigma_out = out[0];
  }
}

static int hyper_reduce_impl(SsviHandle* h, int32_t dim, double* out, double* psigma_out, void* stream) {
  if (!h) return SSVI_E_INVALID;
  if (!h->state_bound) return fail(h, SSVI_E_STATE, "hyper_reduce: state not bound%s");
  if (dim < 0 || dim >= h->cfg.order || !out) return fail(h, SSVI_E_INVALID, "hyper_reduce: bad argument%s");
  hyper_reduce_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(h->params[dim], h->norm_changes[dim], h->cfg.dims[dim], h->cfg.rank,
                                                            h->cfg.cov_kind == SSVI_COV_DIAG, out, psigma_out, h->stop);
  CUDA_TRY(h, cudaGetLastError());
  return SSVI_OK;
}

extern "C" int ssvi_hyper_reduce(SsviHandle* h, int32_t dim, double* out, void* stream) {
  return hyper_reduce_impl(h, dim, out, nullptr, stream);
}

// predict_entry, posterior-mean predictor (:1039-1052): sum_D prod_modes m
__global__ void predict_mean_kernel(DevView v, const double* const* params_unused, const int32_t* __restrict__ entries,
                                    int64_t n, double* __restrict__ out, const double* p0, const double* p1,
                                    const double* p2, const double* p3, const double* p4, const double* p5,
                                    const double* p6, const double* p7) {
  const double* P[SSVI_MAX_ORDER] = {p0, p1, p2, p3, p4, p5, p6, p7};
  const int lane = threadIdx.x & 31;
  const int64_t e = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (e >= n) return;
  double s = 0.0;
  for (int c = lane; c < v.D; c += 32) {
    double u = 1.0;
    for (int d = 0; d < v.N; ++d) u *= P[d][(int64_t)entries[e * v.N + d] * v.row_len + c];
    s += u;
  }
  s = warp_sum(s);
  if (lane == 0) out[e] = s;
}

extern "C" int ssvi_predict_mean(SsviHandle* h, const int32_t* entries, int64_t n, double* out, void* stream) {
  if (!h) return SSVI_E_INVALID;
  if (!h->state_bound) return fail(h, SSVI_E_STATE, "predict_mean: state not bound%s");
  if (n == 0) return SSVI_OK;
  if (n < 0 || !entries || !out) return fail(h, SSVI_E_INVALID, "predict_mean: null pointer%s");
  DevView v = make_view(h, 0);
  const int threads = 256;
  const int64_t blocks = (n * 32 + threads - 1) / threads;
  double** P = h->params;
  predict_mean_kernel<<<(unsigned)blocks, threads, 0, (cudaStream_t)stream>>>(v, nullptr, entries, n, out, P[0], P[1], P[2], P[3], P[4], P[5], P[6], P[7]);
  CUDA_TRY(h, cudaGetLastError());
  return SSVI_OK;
}

// ---- minibatch selection on the device (SSVI_TF_interface.py:164-178) -------------------------------------------------------
// One warp per row.  A row holding more than B entries keeps a uniform random B-subset of its list, drawn without
// replacement by Floyd's algorithm (for j = len-B .. len-1: t uniform in [0, j]; take t, or j if t was taken already):
// B sequential steps, the membership test spread over the lanes.  Counter-based RNG (splitmix64 of seed, row, step).
__host__ __device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

// one warp draws one row's minibatch into dst; picks: >= 2 B int32 of shared memory owned by the warp (picks | draws)
__device__ __forceinline__ void draw_row_minibatch(int lane, int32_t* picks, int64_t row, const int64_t* __restrict__ rowptr,
                                                   const int64_t* __restrict__ perm, int B, uint64_t seed,
                                                   int64_t* __restrict__ dst) {
  const int64_t lo = rowptr[row], len = rowptr[row + 1] - lo;
  if (len <= B) {                                       // kept whole, in append order (:170-174)
    for (int64_t i = lane; i < len; i += 32) dst[i] = perm[lo + i];
    return;
  }
  const uint64_t key = splitmix64(seed ^ splitmix64((uint64_t)row));
  int32_t* draws = picks + B;                           // the B uniform draws, evaluated lane-parallel up front: the
  for (int c = lane; c < B; c += 32)                    // sequential part below is then only the membership test
    draws[c] = (int32_t)__umul64hi(splitmix64(key + (uint64_t)c), (uint64_t)(len - B + c + 1));   // uniform in [0, j]
  __syncwarp();
  int cnt = 0;
  for (int64_t j = len - B; j < len; ++j, ++cnt) {
    const int32_t t = draws[cnt];
    bool hit = false;
    for (int i = lane; i < cnt; i += 32) hit |= (picks[i] == t);
    const bool taken = __any_sync(0xffffffffu, hit);
    if (lane == 0) picks[cnt] = taken ? (int32_t)j : t;
    __syncwarp();
  }
  for (int i = lane; i < B; i += 32) dst[i] = perm[lo + picks[i]];
}

__global__ void __launch_bounds__(128) draw_minibatches_kernel(int64_t n_rows, const int64_t* __restrict__ rows,
                                                               const int64_t* __restrict__ rowptr,
                                                               const int64_t* __restrict__ perm, int B, uint64_t seed,
                                                               const int64_t* __restrict__ seg_start,
                                                               int64_t* __restrict__ out) {
  extern __shared__ int32_t picks_all[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t r = (int64_t)blockIdx.x * 4 + warp;
  if (r >= n_rows) return;
  draw_row_minibatch(lane, picks_all + warp * 2 * B, rows[r], rowptr, perm, B, seed, out + seg_start[r]);
}

extern "C" int ssvi_draw_minibatches(SsviHandle* h, int64_t n_rows, const int64_t* rows, const int64_t* rowptr,
                                     const int64_t* perm, int32_t batch_size, uint64_t seed, const int64_t* seg_start,
                                     int64_t* out, void* stream) {
  if (!h) return SSVI_E_INVALID;
  if (n_rows == 0) return SSVI_OK;
  if (n_rows < 0 || !rows || !rowptr || !perm || !seg_start || !out) return fail(h, SSVI_E_INVALID, "draw_minibatches: null pointer%s");
  if (batch_size < 1 || batch_size > 2048) return fail(h, SSVI_E_INVALID, "draw_minibatches: batch_size must be in [1, 2048]%s");
  const size_t smem = (size_t)4 * 2 * batch_size * sizeof(int32_t);
  if (smem > 48 * 1024)
    CUDA_TRY(h, cudaFuncSetAttribute(draw_minibatches_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  draw_minibatches_kernel<<<(unsigned)((n_rows + 3) / 4), 128, smem, (cudaStream_t)stream>>>(
      n_rows, rows, rowptr, perm, batch_size, seed, seg_start, out);
  CUDA_TRY(h, cudaGetLastError());
  return SSVI_OK;
}

// ---- variate generation on the device ---------------------------------------------------------------------------------
// Stand-in for the reference's draws from numpy's global RNG (np.random.multivariate_normal SSVI_TF_interface.py:252-254,
// :743-745; np.random.rayleigh :260 / SSVI_TF_robust.py:53; np.random.normal SSVI_TF_simple.py:35 / SSVI_TF_robust.py:163):
// fills a noise pool with STANDARD variates.  Counter-based (splitmix64 of key + element index, Box-Muller), so a
// (seed, draw) pair always gives the same pool and nothing but the two counters crosses the bus.
template <typename ZT>
__global__ void fill_noise_kernel(ZT* __restrict__ z_other, int64_t n_other, ZT* __restrict__ z_own, int64_t n_own,
                                  ZT* __restrict__ r_w, int64_t n_rw, ZT* __restrict__ z_f, int64_t n_zf, uint64_t key) {
  const int64_t total = n_other + n_own + n_rw + n_zf;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const uint64_t a = splitmix64(key + 2ull * (uint64_t)i), b = splitmix64(key + 2ull * (uint64_t)i + 1ull);
    const double u1 = ((double)(a >> 11) + 0.5) * (1.0 / 9007199254740992.0);      // (0, 1)
    const double u2 = ((double)(b >> 11) + 0.5) * (1.0 / 9007199254740992.0);
    const double r = sqrt(-2.0 * log(u1));
    int64_t j = i;
    if (j < n_other) { z_other[j] = (ZT)(r * cospi(2.0 * u2)); continue; }
    j -= n_other;
    if (j < n_own) { z_own[j] = (ZT)(r * cospi(2.0 * u2)); continue; }
    j -= n_own;
    if (j < n_rw) { r_w[j] = (ZT)r; continue; }                                    // unit Rayleigh
    j -= n_rw;
    z_f[j] = (ZT)(r * cospi(2.0 * u2));
  }
}

extern "C" int ssvi_fill_noise(SsviHandle* h, void* z_other, void* z_own, void* r_w, void* z_f, int32_t pool_slots,
                               uint64_t seed, uint64_t draw, void* stream) {
  if (!h) return SSVI_E_INVALID;
  if (!z_other || !z_own || pool_slots < 1) return fail(h, SSVI_E_INVALID, "fill_noise: null pointer or empty pool%s");
  const SsviConfig& c = h->cfg;
  const bool second = c.model != SSVI_MODEL_D;
  if (second && (!r_w || !z_f)) return fail(h, SSVI_E_INVALID, "fill_noise: r_w and z_f required for simple/robust%s");
  const int64_t blk = smp_zblk(c.rank, c.k1, c.cov_kind == SSVI_COV_DIAG);
  const int64_t n_own = (int64_t)pool_slots * blk, n_other = n_own * (c.order - 1);
  const int64_t n_rw = second ? (int64_t)pool_slots * c.k1 : 0, n_zf = second ? n_rw * c.k2 : 0;
  const int64_t total = n_other + n_own + n_rw + n_zf;
  const uint64_t key = splitmix64(seed ^ splitmix64(draw + 0x5851F42D4C957F2Dull));
  const unsigned grid = (unsigned)((total + 255) / 256 < 4096 ? (total + 255) / 256 : 4096);
  cudaStream_t st = (cudaStream_t)stream;
  if (c.noise_dtype == SSVI_NOISE_F32)
    fill_noise_kernel<float><<<grid, 256, 0, st>>>((float*)z_other, n_other, (float*)z_own, n_own, (float*)r_w, n_rw,
                                                   (float*)z_f, n_zf, key);
  else
    fill_noise_kernel<double><<<grid, 256, 0, st>>>((double*)z_other, n_other, (double*)z_own, n_own, (double*)r_w, n_rw,
                                                    (double*)z_f, n_zf, key);
  CUDA_TRY(h, cudaGetLastError());
  return SSVI_OK;
}

extern "C" int ssvi_estimate_entries(SsviHandle* h, int32_t what, const int32_t* entries, const double* ys, int64_t n,
                                     int32_t slot0, const double* w_state, double* out, void* stream) {
  if (!h) return SSVI_E_INVALID;
  if (!h->state_bound) return fail(h, SSVI_E_STATE, "estimate_entries: state not bound%s");
  if (!h->noise_bound) return fail(h, SSVI_E_STATE, "estimate_entries: noise not bound%s");
  if (what < SSVI_EST_EXPECTED_OBS || what > SSVI_EST_ROBUST_VLB) return fail(h, SSVI_E_INVALID, "estimate_entries: bad estimator%s");
  if (what == SSVI_EST_ROBUST_VLB && (h->cfg.model == SSVI_MODEL_D || h->cfg.k1 > 128))
    return fail(h, SSVI_E_INVALID, "estimate_entries: SSVI_EST_ROBUST_VLB needs a noise-layer model and k1 <= 128%s");
  if (n == 0) return SSVI_OK;
  if (n < 0 || !entries || !out || slot0 < 0) return fail(h, SSVI_E_INVALID, "estimate_entries: bad argument%s");
  if (what != SSVI_EST_EXPECTED_OBS && !ys) return fail(h, SSVI_E_INVALID, "estimate_entries: ys required%s");
  if (h->cfg.model != SSVI_MODEL_D && (!w_state || !h->r_w || !h->z_f))
    return fail(h, SSVI_E_INVALID, "estimate_entries: w_state and the noise-layer pool are required for simple/robust%s");
  const DevView v = make_view(h, 0);
  const int64_t blocks = (n + 3) / 4;
  cudaStream_t st = (cudaStream_t)stream;
  if (h->cfg.noise_dtype == SSVI_NOISE_F32)
    estimate_entries_kernel<float><<<(unsigned)blocks, 128, 0, st>>>(
        v, what, entries, ys, n, slot0, w_state, (const float*)h->z_other, (const float*)h->z_own, (const float*)h->r_w,
        (const float*)h->z_f, out);
  else
    estimate_entries_kernel<double><<<(unsigned)blocks, 128, 0, st>>>(
        v, what, entries, ys, n, slot0, w_state, (const double*)h->z_other, (const double*)h->z_own,
        (const double*)h->r_w, (const double*)h->z_f, out);
  CUDA_TRY(h, cudaGetLastError());
  return SSVI_OK;
}

extern "C" int ssvi_check_flags(SsviHandle* h, int32_t* flags_out, void* stream) {
  if (!h || !flags_out) return SSVI_E_INVALID;
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(h, cudaMemcpyAsync(flags_out, h->d_flags, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(h, cudaMemsetAsync(h->d_flags, 0, sizeof(int32_t), st));
  CUDA_TRY(h, cudaStreamSynchronize(st));
  return SSVI_OK;
}

// ---- factorize()'s inner loop on the host side of the library (SSVI_TF_interface.py:117-143) --------------------------------
// The reference's schedule updates ONE row per mode per iteration; driven from Python that is ~10 library calls, a chunk
// table upload and a blocking read of the stop condition per row.  Here a whole block of iterations is enqueued from C:
// per (iteration, mode) one warp plans the row's minibatch on the device (draw + chunk tables), then the same kernels as
// the per-call path run (variates, stages 1-3, chunk reduction, stage 4, w_sigma, pSigma / max-change reduction), and a
// one-thread kernel evaluates check_stop_cond (:967-984) after every iteration.  When it fires it latches a device word
// that the state-writing kernels honour, so the iterations enqueued behind it leave the state exactly as the reference's
// `break` (:142) does; the caller reads the word whenever it likes.  Nothing crosses the bus inside the loop.
__global__ void plan_row_kernel(int64_t row, const int64_t* __restrict__ rowptr, const int64_t* __restrict__ perm, int B,
                                int W, uint64_t seed, int64_t* __restrict__ entry_ids, int32_t* __restrict__ chunk_row,
                                int64_t* __restrict__ chunk_start, int32_t* __restrict__ chunk_len,
                                int32_t* __restrict__ chunk_slot0, int64_t* __restrict__ rcp, int64_t* __restrict__ rows,
                                int64_t* __restrict__ n_obs) {
  extern __shared__ int32_t picks[];
  const int lane = threadIdx.x;
  const int64_t len = rowptr[row + 1] - rowptr[row];
  const int n = (int)(len < B ? len : B);
  draw_row_minibatch(lane, picks, row, rowptr, perm, B, seed, entry_ids);
  const int nch = (n + W - 1) / W;
  for (int c = lane; c < nch; c += 32) {
    chunk_row[c] = 0;
    chunk_start[c] = (int64_t)c * W;
    chunk_len[c] = min(W, n - c * W);
    chunk_slot0[c] = c * W;
  }
  if (lane == 0) {
    rcp[0] = 0;
    rcp[1] = nch;
    rows[0] = row;
    n_obs[0] = len;
  }
}

__global__ void stop_check_kernel(const double* __restrict__ hyper_out, int N, double eps, int64_t iteration,
                                  int32_t* stop) {
  if (threadIdx.x != 0 || blockIdx.x != 0 || stop[0]) return;
  double dm = 0.0, dc = 0.0;
  for (int d = 0; d < N; ++d) {
    dm = fmax(dm, hyper_out[d * 3 + 1]);
    dc = fmax(dc, hyper_out[d * 3 + 2]);
  }
  if (fmax(dm, dc) < eps) {                            // :142
    stop[0] = 1;
    stop[1] = (int32_t)iteration;
  }
}

static inline int64_t align16(int64_t x) { return (x + 15) & ~(int64_t)15; }

struct LoopPlan {                                        // one row update's minibatch plan (plan_row_kernel writes it)
  int64_t* entry_ids;
  int32_t *chunk_row, *chunk_len, *chunk_slot0;
  int64_t *chunk_start, *rcp, *rows, *n_obs;
};
struct LoopWs {
  LoopPlan plan[2];                                      // double-buffered: update u + 1 is planned while update u runs
  double *partial, *acc, *si;
  int64_t bytes;
};

static LoopWs loop_ws_layout(const SsviHandle* h, int32_t B, unsigned char* base) {
  const int64_t nch = (B + 1) / 2;                      // chunks are >= 2 entries long
  LoopWs w;
  int64_t o = 0;
  auto take = [&](int64_t bytes) { unsigned char* p = base + o; o += align16(bytes); return p; };
  for (int b = 0; b < 2; ++b) {
    LoopPlan& p = w.plan[b];
    p.entry_ids = (int64_t*)take(8 * (int64_t)B);
    p.chunk_row = (int32_t*)take(4 * nch);
    p.chunk_len = (int32_t*)take(4 * nch);
    p.chunk_slot0 = (int32_t*)take(4 * nch);
    p.chunk_start = (int64_t*)take(8 * nch);
    p.rcp = (int64_t*)take(16);
    p.rows = (int64_t*)take(8);
    p.n_obs = (int64_t*)take(8);
  }
  w.partial = (double*)take(8 * nch * h->acc_len);
  w.acc = (double*)take(8 * h->acc_len);
  w.si = (double*)take(8);
  w.bytes = o;
  return w;
}

extern "C" int64_t ssvi_loop_workspace_bytes(const SsviHandle* h, int32_t batch_size) {
  if (!h || batch_size < 1) return 0;
  return loop_ws_layout(h, batch_size, nullptr).bytes;
}

extern "C" int ssvi_factorize_iterations(SsviHandle* h, SsviLoop* lp, int64_t first_iteration, int64_t n_iterations,
                                         void* stream) {
  if (!h || !lp) return SSVI_E_INVALID;
  if (!h->data_bound || !h->state_bound) return fail(h, SSVI_E_STATE, "factorize_iterations: bind data and state first%s");
  const SsviConfig& c = h->cfg;
  const int N = c.order;
  if (lp->batch_size < 1 || lp->batch_size > 2048) return fail(h, SSVI_E_INVALID, "factorize_iterations: batch_size must be in [1, 2048]%s");
  if (!lp->pSigma || !lp->hyper_out || !lp->stop || !lp->workspace || !lp->z_other || !lp->z_own)
    return fail(h, SSVI_E_INVALID, "factorize_iterations: null pointer%s");
  if (c.model != SSVI_MODEL_D && !lp->w_state) return fail(h, SSVI_E_INVALID, "factorize_iterations: w_state required for simple/robust%s");
  if (lp->workspace_bytes < ssvi_loop_workspace_bytes(h, lp->batch_size)) return fail(h, SSVI_E_INVALID, "factorize_iterations: workspace too small (ssvi_loop_workspace_bytes)%s");
  for (int d = 0; d < N; ++d)
    if (!lp->rowptr[d] || !lp->perm[d] || !lp->rowptr_host[d]) return fail(h, SSVI_E_INVALID, "factorize_iterations: index of every mode required%s");
  int rc = ssvi_bind_noise(h, lp->z_other, lp->z_own, lp->r_w, lp->z_f, lp->pool_slots);
  if (rc != SSVI_OK) return rc;
  const LoopWs w = loop_ws_layout(h, lp->batch_size, (unsigned char*)lp->workspace);
  cudaStream_t st = (cudaStream_t)stream;
  const int B = lp->batch_size;
  if (!h->loop_ready) {                                 // side stream + events of the plan pipeline, once per handle
    CUDA_TRY(h, cudaStreamCreateWithFlags(&h->loop_stream, cudaStreamNonBlocking));
    for (int i = 0; i < 2; ++i) {
      CUDA_TRY(h, cudaEventCreateWithFlags(&h->loop_ev_plan[i], cudaEventDisableTiming));
      CUDA_TRY(h, cudaEventCreateWithFlags(&h->loop_ev_free[i], cudaEventDisableTiming));
    }
    CUDA_TRY(h, cudaEventCreateWithFlags(&h->loop_ev_acc, cudaEventDisableTiming));
    CUDA_TRY(h, cudaEventCreateWithFlags(&h->loop_ev_noise, cudaEventDisableTiming));
    h->loop_ready = true;
  }
  // The minibatch plan of a row update (the reference's np.random.choice, :170-174, + the chunk tables) depends on the
  // index and the draw counter only, not on the state: it is drawn on a side stream one update AHEAD, into the other of two
  // plan buffers, while the current update's kernels run (the sequential without-replacement draw was 18 % of a row
  // update on the critical path).  ev_plan[b]: plan in buffer b written; ev_free[b]: the update that read buffer b done.
  struct Upd { int d; int64_t row, len; int n, W, nch; uint64_t seed, noise_draw; };
  auto next_upd = [&](int64_t it, int d, const int64_t* col, uint64_t draws) {
    Upd u;
    u.d = d;
    u.row = col[d];
    u.len = lp->rowptr_host[d][u.row + 1] - lp->rowptr_host[d][u.row];
    u.n = (int)(u.len < B ? u.len : B);
    // same chunking rule as the per-call path (engine.py:_chunk_tables): short plans are cut fine enough to spread
    int per = (u.n + 8 * h->sm_count - 1) / (8 * h->sm_count);
    if (per < 2) per = 2;
    u.W = lp->chunk_entries < per ? lp->chunk_entries : per;
    u.nch = u.n > 0 ? (u.n + u.W - 1) / u.W : 0;
    u.seed = lp->minibatch_seed + draws;
    (void)it;
    return u;
  };
  // the schedule of this call, in order: (iteration, mode) -> row (update_column_pointer, :115, :128-133)
  std::vector<Upd> sched;
  sched.reserve((size_t)n_iterations * N);
  {
    int64_t col[SSVI_MAX_ORDER];
    for (int d = 0; d < N; ++d) col[d] = lp->col[d];
    uint64_t draws = lp->minibatch_draws;
    for (int64_t it = first_iteration; it < first_iteration + n_iterations; ++it) {
      for (int d = 0; d < N; ++d) {
        sched.push_back(next_upd(it, d, col, ++draws));
        sched.back().noise_draw = lp->noise_draws + (uint64_t)sched.size();   // every (iteration, mode) advances the counter
      }
      for (int d = 0; d < N; ++d)
        if (++col[d] >= c.dims[d]) col[d] = 0;
    }
  }
  int64_t planned = 0, n_planned_real = 0;              // schedule position planned so far; plans issued (buffer parity)
  std::vector<int> buf_of(sched.size(), -1);
  auto plan_ahead = [&](int64_t upto) -> int {          // issue the plans of schedule entries [planned, upto]
    for (; planned <= upto && planned < (int64_t)sched.size(); ++planned) {
      const Upd& u = sched[planned];
      if (u.n <= 0) continue;
      const int b = (int)(n_planned_real & 1);
      if (n_planned_real >= 2) CUDA_TRY(h, cudaStreamWaitEvent(h->loop_stream, h->loop_ev_free[b], 0));
      const LoopPlan& p = w.plan[b];
      plan_row_kernel<<<1, 32, (size_t)2 * B * sizeof(int32_t), h->loop_stream>>>(
          u.row, lp->rowptr[u.d], lp->perm[u.d], B, u.W, u.seed, p.entry_ids, p.chunk_row, p.chunk_start, p.chunk_len,
          p.chunk_slot0, p.rcp, p.rows, p.n_obs);
      CUDA_TRY(h, cudaEventRecord(h->loop_ev_plan[b], h->loop_stream));
      buf_of[planned] = b;
      ++n_planned_real;
    }
    return SSVI_OK;
  };
  // the plan buffers may still be read by the tail of the previous call on `st`: the side stream starts behind it
  CUDA_TRY(h, cudaEventRecord(h->loop_ev_free[0], st));
  CUDA_TRY(h, cudaStreamWaitEvent(h->loop_stream, h->loop_ev_free[0], 0));
  h->stop = lp->stop;
  struct StopGuard {                                     // the stop word is honoured only while this call enqueues work
    SsviHandle* h;
    ~StopGuard() { h->stop = nullptr; }
  } stop_guard{h};
  rc = SSVI_OK;
  int64_t pos = 0;
  bool noise_ahead = false;                              // the next real update's variates are being drawn on the side stream
  for (int64_t it = first_iteration; it < first_iteration + n_iterations && rc == SSVI_OK; ++it) {
    for (int d = 0; d < N && rc == SSVI_OK; ++d, ++pos) {
      const Upd& u = sched[pos];
      ++lp->minibatch_draws;
      ++lp->noise_draws;                                 // every call of the reference draws, observed row or not
      if (u.n > 0) {
        // this update's plan and the next REAL update's (the side stream runs one plan ahead of the kernels below)
        int64_t ahead = pos + 1;
        while (ahead < (int64_t)sched.size() && sched[ahead].n <= 0) ++ahead;
        rc = plan_ahead(ahead);
        if (rc != SSVI_OK) break;
        const int b = buf_of[pos];
        const LoopPlan& p = w.plan[b];
        CUDA_TRY(h, cudaStreamWaitEvent(st, h->loop_ev_plan[b], 0));
        // this update's variates: drawn behind the previous update's accumulate on the side stream (below), except for the
        // first real update of the call
        if (!noise_ahead) rc = ssvi_fill_noise(h, lp->z_other, lp->z_own, lp->r_w, lp->z_f, lp->pool_slots, lp->noise_seed, u.noise_draw, stream);
        else CUDA_TRY(h, cudaStreamWaitEvent(st, h->loop_ev_noise, 0));
        if (rc == SSVI_OK)
          rc = ssvi_accumulate(h, d, u.nch, p.chunk_row, p.chunk_start, p.chunk_len, p.chunk_slot0, p.entry_ids, p.rows,
                               lp->w_state, w.partial, stream);
        noise_ahead = false;
        if (rc == SSVI_OK && ahead < (int64_t)sched.size()) {  // the pool is free once accumulate has read it: next draws now
          CUDA_TRY(h, cudaEventRecord(h->loop_ev_acc, st));
          CUDA_TRY(h, cudaStreamWaitEvent(h->loop_stream, h->loop_ev_acc, 0));
          rc = ssvi_fill_noise(h, lp->z_other, lp->z_own, lp->r_w, lp->z_f, lp->pool_slots, lp->noise_seed,
                               sched[ahead].noise_draw, (void*)h->loop_stream);
          CUDA_TRY(h, cudaEventRecord(h->loop_ev_noise, h->loop_stream));
          noise_ahead = true;
        }
        if (rc == SSVI_OK) rc = ssvi_reduce_chunks(h, 1, p.rcp, w.partial, w.acc, stream);
        if (rc == SSVI_OK)
          rc = ssvi_row_update(h, d, 1, p.rows, w.acc, p.n_obs, lp->pSigma, lp->w_state, lp->time_step[d],
                               lp->fixed_covariance, w.si, stream);
        if (rc == SSVI_OK) CUDA_TRY(h, cudaEventRecord(h->loop_ev_free[b], st));
        if (rc == SSVI_OK && c.model != SSVI_MODEL_D) rc = ssvi_sigma_update(h, 1, w.si, w.acc, lp->w_state, stream);
        lp->launches += 5 + (c.model != SSVI_MODEL_D ? 1 : 0);
      }
      if (rc == SSVI_OK) rc = hyper_reduce_impl(h, d, lp->hyper_out + 3 * d, lp->pSigma + d, stream);   // :125
      lp->launches += 1;
    }
    for (int d = 0; d < N; ++d) {                        // :128-133
      if (++lp->col[d] >= c.dims[d]) {
        lp->col[d] = 0;
        ++lp->time_step[d];
      }
    }
    stop_check_kernel<<<1, 32, 0, st>>>(lp->hyper_out, N, lp->epsilon, it, lp->stop);                    // :135-142
    lp->launches += 1;
  }
  h->stop = nullptr;
  if (rc == SSVI_OK) CUDA_TRY(h, cudaGetLastError());
  return rc;
}

// ---- stages 1-3 + chunk reduction with NO caller workspace and NO allocation -------------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).  ssvi_accumulate needs a
// scratch buffer of n_chunks chunk records (`partial_out`) that ssvi_reduce_chunks then folds into the row accumulators;
// here that scratch is a file-scope device array of SSVI_STATIC_SCRATCH_BYTES (every GPU that loads the library holds
// it until the process exits).  In HOST code the array's name is only a stand-in for the symbol, not a device address:
// the address is taken with cudaGetSymbolAddress once per device, with that device current, and passed to the kernels.
// Every call shares the array and it keeps the last call's contents; nothing in it has to start at zero, because
// every accumulate kernel writes each chunk record in full before reduce_chunks_kernel reads it.  Two calls that
// overlap on one GPU (different streams, or different handles / host threads) race on it.
static __device__ __align__(256) double g_static_scratch[SSVI_STATIC_SCRATCH_BYTES / sizeof(double)];
#define SSVI_MAX_DEVICES 64
static double* g_static_scratch_ptr[SSVI_MAX_DEVICES];   // per device, filled on the first call made with it current

extern "C" int64_t ssvi_static_scratch_bytes(void) { return (int64_t)SSVI_STATIC_SCRATCH_BYTES; }

extern "C" int ssvi_accumulate_rows_static(SsviHandle* h, int32_t dim, int64_t n_chunks, const int32_t* chunk_row,
                                           const int64_t* chunk_start, const int32_t* chunk_len,
                                           const int32_t* chunk_slot0, const int64_t* entry_ids, const int64_t* rows,
                                           const double* w_state, int64_t n_rows, const int64_t* row_chunk_ptr,
                                           double* acc_out, void* stream) {
  if (!h) return SSVI_E_INVALID;
  if (n_rows < 0 || n_chunks < 0) return fail(h, SSVI_E_INVALID, "accumulate_rows_static: negative count%s");
  if (n_rows == 0) return SSVI_OK;
  if (!row_chunk_ptr || !acc_out) return fail(h, SSVI_E_INVALID, "accumulate_rows_static: null pointer%s");
  const int64_t need = n_chunks * h->acc_len * (int64_t)sizeof(double);
  if (need > (int64_t)SSVI_STATIC_SCRATCH_BYTES) {
    char msg[160];
    snprintf(msg, sizeof(msg), "%lld chunk records of %lld doubles need %lld bytes, the static scratch holds %lld",
             (long long)n_chunks, (long long)h->acc_len, (long long)need, (long long)SSVI_STATIC_SCRATCH_BYTES);
    return fail(h, SSVI_E_INVALID, "accumulate_rows_static: %s (use ssvi_accumulate with a caller workspace)", msg);
  }
  int dev = -1;
  CUDA_TRY(h, cudaGetDevice(&dev));
  if (dev < 0 || dev >= SSVI_MAX_DEVICES) return fail(h, SSVI_E_INVALID, "accumulate_rows_static: device ordinal out of range%s");
  if (!g_static_scratch_ptr[dev]) {                     // set-up, once per device, with that device current
    double* p = nullptr;
    CUDA_TRY(h, cudaGetSymbolAddress((void**)&p, g_static_scratch));
    if (!p) return fail(h, SSVI_E_CUDA, "accumulate_rows_static: cudaGetSymbolAddress returned a null address%s");
    g_static_scratch_ptr[dev] = p;
  }
  double* scratch = g_static_scratch_ptr[dev];
  if (n_chunks > 0) {
    const int rc = ssvi_accumulate(h, dim, n_chunks, chunk_row, chunk_start, chunk_len, chunk_slot0, entry_ids, rows,
                                   w_state, scratch, stream);
    if (rc != SSVI_OK) return rc;
  }
  return ssvi_reduce_chunks(h, n_rows, row_chunk_ptr, scratch, acc_out, stream);
}
