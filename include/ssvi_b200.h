/*
 * ssvi_b200.h -- C ABI of the B200-native SSVI-TF hot path (libssvi_b200.so).
 *
 * The reference (dnguyen1196/SSVI-TF) is pure Python: it has NO FFI layer.  The boundary this
 * library replaces is the Python method SSVI_TF.update_natural_param_batch(dim, col)
 * (SSVI/SSVI_TF_interface.py:163-219) with its model hooks estimate_di_Di_si_batch
 * (SSVI_TF_interface.py:222-264, SSVI_TF_robust.py:28-169, SSVI_TF_d.py:31-57,
 * SSVI_TF_simple.py:23-40), the row updates (:595-719), update_hyper_parameter (:146-160),
 * check_stop_cond (:967-984) and predict_entry (:1037-1059).  Every entry point below cites the
 * reference lines it stands in for.  INTEGRATION.md shows the ctypes binding a maintainer of the
 * reference would add.
 *
 * Conventions
 *   - extern "C", plain pointers and sizes only.  Unless marked HOST, every pointer is a DEVICE
 *     pointer owned by the caller; the library borrows it for the duration of the call (or, for
 *     the bind_* calls, until rebinding / destroy) and never frees it.
 *   - All work is enqueued on the caller's CUDA stream (`stream` = cudaStream_t passed as void*;
 *     NULL = legacy default stream).  No call synchronises the device except ssvi_check_flags.
 *   - Return value: 0 on success, negative SSVI_E_* otherwise; ssvi_last_error() gives the text.
 *     No C++ exception crosses the ABI.  Device-side guards (the reference's asserts on
 *     NaN/Inf/phi==0, SSVI_TF_interface.py:185-188, 202-205; SSVI_TF_robust.py:87-92, 106) raise
 *     bits in a device flag word read back by ssvi_check_flags().
 *   - One handle per host thread / GPU; a handle is not thread-safe.
 *   - Arithmetic is IEEE fp64 (the reference computes in numpy float64).
 *
 * Noise contract ("identical noise draws").  The reference draws from numpy's global RNG; here the
 * caller supplies STANDARD variates in a pool of `pool_slots` slots.  Entry number b of a row's
 * minibatch uses slot (b mod pool_slots).  Layouts (k fastest, so 32 lanes read 32 samples):
 *     z_other [slot][jj][block]  jj = other modes in ascending order; block = ssvi_noise_block_len
 *     z_own   [slot][block]      elements, element (c < D, k < k1) at ssvi_noise_offset(c, k)
 *     r_w     [slot][k]          unit Rayleigh  (ws = w_sigma^2 * r_w;  simple/robust only)
 *     z_f     [slot][k2][k]      (fs = mean + ws * z_f;                 simple/robust only)
 * Transforms: x = m + L z with L the lower Cholesky factor of the row covariance
 * (diagonal covariance: x = m + sqrt(S) * z).  The oracle applies the same transform inside its
 * patched multivariate_normal (stock NumPy uses an SVD factor).
 */
#ifndef SSVI_B200_H
#define SSVI_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SSVI_MAX_ORDER 8
#define SSVI_MAX_RANK 64

enum { SSVI_COV_FULL = 0, SSVI_COV_DIAG = 1 };
enum { SSVI_MODEL_D = 0, SSVI_MODEL_SIMPLE = 1, SSVI_MODEL_ROBUST = 2 };
enum { SSVI_LIK_NORMAL = 0, SSVI_LIK_BERNOULLI = 1, SSVI_LIK_POISSON = 2 };
enum { SSVI_UPDATE_S = 0, SSVI_UPDATE_N = 1 };
enum { SSVI_NOISE_F64 = 0, SSVI_NOISE_F32 = 1 };
/* Sample table of ssvi_sweep_accumulate.  F64 (default): the tabulated samples X(j, s) = m_j + L_j Z(s) are the doubles the
 * direct kernels compute.  F32: they are rounded to float when they are written (relative 6e-8) and widened again when
 * read; every sum, product and likelihood term is still fp64.  Halves the bytes a sweep gathers.  The fp32 option of
 * the contract (tolerance 1e-4); measured deviation of the updated rows from the F64 table: see DESIGN.md 3.5. */
enum { SSVI_TABLE_F64 = 0, SSVI_TABLE_F32 = 1, SSVI_TABLE_F32_TC = 2 };
/* F32_TC: the fp32 ARITHMETIC option (tolerance 1e-4).  Float table in the tensor-core path's lane order; V, f and the row
 * sums are formed in fp32, the contraction G = SUM_b w_b V_b V_b^T (SSVI_TF_d.py:47-56) runs on the tcgen05 tensor cores
 * (V split into two fp16 pieces scaled by 2^tc_scale_log2, hi hi^T + hi lo^T + lo hi^T accumulated in fp32 in tensor
 * memory: csrc/accumulate_tc.cuh); likelihood derivatives, chunk records and everything from stage 3 on stay fp64.
 * SSVI_TF_d, full covariance, rank <= 64, k1 == 64.  With a DIAGONAL covariance the flag selects the fp32 instances of the
 * sweep kernels of SSVI_TF_d and of SSVI_TF_simple + normal likelihood (csrc/accumulate_diag32.cuh).  Other configurations treat it
 * as SSVI_TABLE_F32 / F64. */

enum {
  SSVI_OK = 0,
  SSVI_E_INVALID = -1,   /* bad argument / unsupported configuration */
  SSVI_E_STATE = -2,     /* call order: data / state / noise not bound */
  SSVI_E_CUDA = -3       /* CUDA runtime error (text in ssvi_last_error) */
};

/* bits of the device flag word (ssvi_check_flags) */
enum {
  SSVI_FLAG_NONFINITE_ACC = 1,   /* di / Di not finite          (SSVI_TF_interface.py:185-188) */
  SSVI_FLAG_NONFINITE_COV = 2,   /* S_next not finite           (SSVI_TF_interface.py:202-204) */
  SSVI_FLAG_PHI_ZERO = 4,        /* phi == 0 or not finite      (SSVI_TF_robust.py:87-92, 106) */
  SSVI_FLAG_CHOL_FAIL = 8,       /* covariance not positive definite in the sampler refresh     */
  SSVI_FLAG_EIG_NOCONV = 16,     /* Jacobi eigen-iteration hit its sweep limit                  */
  SSVI_FLAG_TC_RANGE = 32        /* SSVI_TABLE_F32_TC: an operand left the fp16 range (|2^tc_scale_log2 V| > 65504); lower
                                    cfg.tc_scale_log2 or use the fp64 path */
};

/* Constructor kwargs + hidden constants of SSVI_TF.__init__ (SSVI_TF_interface.py:21-98). */
typedef struct SsviConfig {
  int32_t order;            /* N = len(tensor.dims)                              (:29)  */
  int32_t rank;             /* D                                                 (:31)  */
  int32_t cov_kind;         /* SSVI_COV_*      `diag`                            (:36)  */
  int32_t model;            /* SSVI_MODEL_*    SSVI_TF_d / _simple / _robust            */
  int32_t likelihood;       /* SSVI_LIK_*      from tensor.datatype              (:59-77) */
  int32_t k1;               /* first-layer samples                               (:50)  */
  int32_t k2;               /* second-layer samples                              (:51)  */
  int32_t mean_update;      /* SSVI_UPDATE_*   "S" | "N"                         (:32)  */
  int32_t cov_update;       /* SSVI_UPDATE_*   "S" | "N"                         (:33)  */
  int32_t noise_dtype;      /* SSVI_NOISE_*    element type of the noise pool           */
  int32_t kernel_path;      /* 0 = auto; 1 = force the generic CUDA-core accumulate path (testing)  */
  int32_t tuning;           /* bit set of SSVI_TUNE_* kernel options (0 = defaults); results do not depend on it */
  int32_t table_dtype;      /* SSVI_TABLE_*: element type of the sweep path's sample table (SSVI_TF_d, full covariance) */
  int32_t tc_scale_log2;    /* SSVI_TABLE_F32_TC: log2 of the operand scale of the fp16 split; 0 = default (6) */
  double eta;               /* (:85)  */
  double cov_eta;           /* (:86)  */
  double sigma_eta;         /* (:88)  */
  double poisson_eta;       /* 0.1   (:87)  */
  double min_eigenval;      /* 0.01  (:48)  */
  double cov_floor;         /* 1e-4  self.epsilon, floor of the diagonal covariance (:80, :680) */
  double likelihood_param;  /* s = 1 for normal (:62) */
  int64_t dims[SSVI_MAX_ORDER];
} SsviConfig;

typedef struct SsviHandle SsviHandle;

/* Length (in doubles) of one per-row / per-chunk accumulator record:
 *   [ di (D) | Di (D*D full, D diag) | si_sum | count ]                                        */
int64_t ssvi_acc_len(const SsviConfig* cfg);

/* Doubles per row of the sampler cache (library-internal layout: mean + Cholesky factor in DMMA
 * fragment order for full covariance; mean + sqrt(S) for diagonal).  sampler[d] holds n_d such rows. */
int64_t ssvi_sampler_row_len(const SsviConfig* cfg);

/* Noise block of one (slot, mode): its length in elements and the offset of element (c, k)
 * (c < D component, k < k1 sample).  Diagonal covariance: c*k1 + k.  Full covariance: DMMA B-fragment
 * order ((k/8)*ceil(D/4) + c/4)*32 + (k%8)*4 + c%4, zero-padded to multiples of 4 (c) and 8 (k). */
int64_t ssvi_noise_block_len(const SsviConfig* cfg);
int64_t ssvi_noise_offset(const SsviConfig* cfg, int32_t c, int32_t k);

/* Kernel options (cfg.tuning / ssvi_set_tuning): launch-order and prefetch choices that never change a result. */
enum {
  SSVI_TUNE_L2_PREFETCH = 1,     /* sweep kernels: ask the next entry's table records into L2 an entry ahead */
  SSVI_TUNE_LIN = 2,             /* SSVI_TF_d + normal likelihood sweeps: the linear-weight kernel (contraction accumulated
                                    across the chunk) instead of the general one */
  SSVI_TUNE_PF2 = 4,             /* rank 25..32, order 3 sweeps: two k-tiles of table records in flight, 2 CTAs per SM */
  SSVI_TUNE_PF3 = 8,             /* ... three k-tiles in flight */
  SSVI_TUNE_TC_RING = 16,        /* SSVI_TABLE_F32_TC sweeps: table records through a per-warp cp.async.bulk ring in shared
                                    memory (2 CTAs per SM) instead of the register prefetch (3 CTAs per SM); measured 1.6x
                                    SLOWER on config 5: 1 KB bulk copies are served too slowly */
  SSVI_TUNE_TC_PF = 32           /* SSVI_TABLE_F32_TC sweeps: 2-3 KB of each table record in flight per warp (a ring of
                                    register buffers, ~200 registers, 2 CTAs per SM) instead of 1 KB (3 CTAs per SM) */
};

/* The library itself allocates only the device flag word of a handle. */
int ssvi_create(const SsviConfig* cfg /*HOST*/, SsviHandle** out /*HOST*/);
void ssvi_destroy(SsviHandle* h);
int ssvi_set_tuning(SsviHandle* h, int32_t tuning);   /* replaces cfg.tuning for the calls that follow */
/* No reference counterpart (the reference holds one tensor in one process).  The sweep path's sample table covers rows
 * [lo[d], lo[d] + n[d]) of mode d only: the index range this handle's entries touch (a shard cut along index-space blocks
 * touches a sub-range of every mode, and sampling the rest is wasted work: 5.5 % of a rank's sweep at 8 ranks on config 5).
 * lo, n HOST [order].  EVERY coordinate of the bound entries, and every row swept, must lie inside its mode's window.
 * Default: all rows.  Changes what ssvi_sweep_table_bytes returns. */
int ssvi_set_row_windows(SsviHandle* h, const int64_t* lo /*HOST*/, const int64_t* n /*HOST*/);
const char* ssvi_last_error(const SsviHandle* h); /* HOST string; h may be NULL for create errors */

/* Observed TRAIN entries in COO form: Tensor.train_entries / train_vals (Tensor/Tensor.py:66-67,
 * 117-118).  coords int32 [nnz][N] row-major, vals double [nnz]. */
int ssvi_bind_data(SsviHandle* h, const int32_t* coords, const double* vals, int64_t nnz);

/* Posterior and optimiser state, one device pointer per mode (HOST arrays of N device pointers).
 *   params[d]        (n_d, D+1, D) full | (n_d, 2, D) diag   == Posterior.params[d] (Model/TF_Models.py:53-93)
 *   sampler[d]       (n_d, ssvi_sampler_row_len): mean + lower Cholesky factor (diag: sqrt(S)); a cache
 *                    the library maintains (ssvi_refresh_sampler / ssvi_row_update)
 *   ada_mean[d]      (n_d, D)   transposed ada_acc_grad[d]      (:83, :694-698)
 *   ada_cov[d]       (n_d, D)   transposed ada_acc_grad_cov[d]  (:41, :705-713); may be NULL for full cov
 *   norm_changes[d]  (n_d, 2)   (:91, :722-729)
 *   grad_norms[d]    (n_d, 2)   (:95, :199) */
int ssvi_bind_state(SsviHandle* h, double* const* params /*HOST*/, double* const* sampler /*HOST*/,
                    double* const* ada_mean /*HOST*/, double* const* ada_cov /*HOST*/,
                    double* const* norm_changes /*HOST*/, double* const* grad_norms /*HOST*/);

/* Noise pool (see "Noise contract").  Element type per cfg.noise_dtype.  r_w / z_f may be NULL for
 * SSVI_MODEL_D. */
int ssvi_bind_noise(SsviHandle* h, const void* z_other, const void* z_own, const void* r_w,
                    const void* z_f, int32_t pool_slots);

/* SSVI_TF_robust(quadrature=True) (SSVI_TF_robust.py:11, :22-25, :175-214): replace the k2 second-layer draws of
 * estimate_expected_derivatives_pdf_batch by the Gauss-Hermite rule
 *   E[g(f)] = sum_i w_i g(mean + sqrt(2 ws^2) x_i) / sqrt(pi)   (estimate_expected_pdf_quadrature, :198-214).
 * rule: DEVICE double[2 n] = [nodes x_i | weights w_i] (np.polynomial.hermite.hermgauss(n); the reference uses
 * n = 30), borrowed until rebound.  n == 0 returns to Monte Carlo.  z_f is not read while a rule is bound. */
int ssvi_bind_quadrature(SsviHandle* h, const double* rule, int32_t n);

/* Recompute sampler[dim] rows from params[dim] (Cholesky / sqrt).  rows == NULL: all n_dim rows.
 * Replaces the per-draw factorisation inside np.random.multivariate_normal at
 * SSVI_TF_interface.py:743-745, 252-254. */
int ssvi_refresh_sampler(SsviHandle* h, int32_t dim, const int64_t* rows, int64_t n_rows, void* stream);

/* Stages 1-3 for a list of row-chunks of mode `dim`
 * (estimate_di_Di_si_batch: SSVI_TF_interface.py:229-264 + SSVI_TF_d.py:31-57 / :562-592 +
 *  SSVI_TF_simple.py:23-40 / SSVI_TF_robust.py:28-169).
 * A chunk is a run of consecutive minibatch entries of ONE row:
 *   chunk_row[c]    int32  index into rows[]
 *   chunk_start[c]  int64  offset into entry_ids[] of the chunk's first entry
 *   chunk_len[c]    int32  entries in the chunk (>= 1)
 *   chunk_slot0[c]  int32  position of the chunk's first entry inside its row's minibatch
 *                          (entry t of the chunk uses noise slot (chunk_slot0 + t) mod pool_slots)
 *   entry_ids[]     int64  indices into the bound COO arrays (the per-mode CSR permutation of
 *                          Tensor.find_observed_ui, Tensor/Tensor.py:251-258, or a subsample of it,
 *                          SSVI_TF_interface.py:170-172)
 *   rows[]          int64  row ids of mode `dim`
 *   w_state         double[3] = { w_sigma, w_ada_grad, w_tau } (simple/robust; may be NULL for d)
 *   partial_out     double [n_chunks][ssvi_acc_len]   UNSCALED sums over the chunk's entries */
int ssvi_accumulate(SsviHandle* h, int32_t dim, int64_t n_chunks, const int32_t* chunk_row,
                    const int64_t* chunk_start, const int32_t* chunk_len, const int32_t* chunk_slot0,
                    const int64_t* entry_ids, const int64_t* rows, const double* w_state,
                    double* partial_out, void* stream);

/* ---- The sample-table path for full sweeps (slot-major; full covariance, rank <= 32, k1 == 64 -- the
 * reference's default, SSVI_TF_interface.py:23 -- and order <= 4).  Same mathematics and the same outputs as ssvi_accumulate, organised so that the sample
 * X(j, s) = m_j + L_j Z(s, jj) of sample_vjs_batch (SSVI_TF_interface.py:732-750) is evaluated once per (other-mode
 * row j, noise slot s) instead of once per (entry, mode): with P pool slots a sweep asks for each pair nnz / (n_j P)
 * times.  The caller runs a sweep as a sequence of calls over slot groups [slot_lo, slot_lo + n_slots); each call
 * first fills the table workspace (n_slots * sum_{d != dim} dims[d] records) and then processes the chunks of those
 * slots.  A chunk here is a run of entries of ONE row that all read ONE noise slot:
 *   chunk_slot[c]   int32  that slot (must lie inside the call's slot group)
 *   chunk_out[c]    int64  index of the chunk's record inside partial_out (any order; lets the caller keep the
 *                          records row-major for ssvi_reduce_chunks while the launch order is slot-major)
 * chunk_row / chunk_start / chunk_len / entry_ids / rows / partial_out as for ssvi_accumulate.
 * table_ws: device workspace of at least ssvi_sweep_table_bytes(h, dim, n_slots) bytes, 16-byte aligned.
 *
 * Diagonal covariance (every model, rank <= 64, k1 % 16 == 0, k1 <= 64, order <= 4) takes the same slot-major chunk
 * tables: samples are element-wise there, so nothing is tabulated (ssvi_sweep_table_bytes == 0, table_ws may be NULL,
 * any slot range in one call); each CTA stages its chunk's noise blocks in shared memory once instead.
 * w_state as for ssvi_accumulate (simple / robust; may be NULL for SSVI_TF_d).
 *
 * SSVI_TF_simple / SSVI_TF_robust with a full covariance run a slot group in three passes (predictive means F ->
 * per-(entry, sample) likelihood weights at full occupancy -> contraction) and need a second workspace of
 * ssvi_sweep_work_bytes(h, n_entries) bytes, where the call's chunks cover the n_entries consecutive positions
 * [entry_lo, entry_lo + n_entries) of entry_ids (true by construction for slot-major chunk tables).  For SSVI_TF_d and
 * for diagonal covariances work_ws may be NULL and entry_lo / n_entries are ignored. */
int ssvi_sweep_supported(const SsviHandle* h);                                   /* 1 if this path applies to h's config */
int64_t ssvi_sweep_table_bytes(const SsviHandle* h, int32_t dim, int32_t n_slots);
int64_t ssvi_sweep_work_bytes(const SsviHandle* h, int64_t n_entries);           /* 0 unless simple/robust + full cov */
int ssvi_sweep_accumulate(SsviHandle* h, int32_t dim, int32_t slot_lo, int32_t n_slots, int64_t n_chunks,
                          const int32_t* chunk_row, const int64_t* chunk_start, const int32_t* chunk_len,
                          const int32_t* chunk_slot, const int64_t* chunk_out, const int64_t* entry_ids,
                          const int64_t* rows, const double* w_state, void* table_ws, int64_t table_ws_bytes,
                          void* work_ws, int64_t work_ws_bytes, int64_t entry_lo, int64_t n_entries,
                          double* partial_out, void* stream);

/* Deterministic in-order sum of the chunk records of each row:
 *   acc_out[r] = sum_{c = row_chunk_ptr[r] .. row_chunk_ptr[r+1]-1} partial[c]
 * acc_out double [n_rows][ssvi_acc_len] is the payload of the multi-GPU all-reduce. */
int ssvi_reduce_chunks(SsviHandle* h, int64_t n_rows, const int64_t* row_chunk_ptr,
                       const double* partial, double* acc_out, void* stream);
/* The same sum folded into a running accumulator (accumulate != 0: acc[r] += sum, else acc[r] = sum).  Lets a sweep
 * be reduced slot group by slot group -- the same accumulation loops as above, split by group -- so the chunk-record
 * buffer holds ONE group's records (config 5 at full size would need > 100 GB of records otherwise) and the records
 * are still in L2 when they are summed.  Deterministic for a fixed sequence of calls. */
int ssvi_reduce_chunks_into(SsviHandle* h, int64_t n_rows, const int64_t* row_chunk_ptr,
                            const double* partial, double* acc_inout, int32_t accumulate, void* stream);

/* Stage 4 for the listed rows of mode `dim`: scale = n_obs[r] / count[r] (:181-183), covariance
 * then mean update (:191-196, 595-681, 694-713), trackers (:199, :217, 722-729), write-back (:219)
 * and the sampler-cache rows.  Rows whose count is 0 are left untouched (:166-168).
 *   n_obs        int64 [n_rows]   len(find_observed_ui(dim,row)) summed over all shards
 *   pSigma       double [N]       (:56, :146-160)
 *   time_step    cov step = cov_eta / (time_step + 1)  (:701-702)
 *   si_out       double [n_rows]  si * scale per row, finalised per model (SSVI_TF_simple.py:39 mean,
 *                                 SSVI_TF_robust.py:137 sum); NaN-free zero for SSVI_MODEL_D; may be NULL */
int ssvi_row_update(SsviHandle* h, int32_t dim, int64_t n_rows, const int64_t* rows, const double* acc,
                    const int64_t* n_obs, const double* pSigma, const double* w_state, int32_t time_step,
                    int32_t fixed_covariance, double* si_out, void* stream);

/* The scalar w_sigma recurrence replayed over the rows in list order
 * (update_sigma_param :684-691, robust override SSVI_TF_robust.py:216-232, step :716-719).
 * Rows with count 0 (acc record) are skipped.  w_state double[3] is updated in place. */
int ssvi_sigma_update(SsviHandle* h, int64_t n_rows, const double* si, const double* acc,
                      double* w_state, void* stream);

/* update_hyper_parameter(dim) (:146-160) and the per-mode part of check_stop_cond (:967-984):
 *   out[0] = sum_rows (tr S + m.m) / (n_dim * D)      (the new pSigma[dim])
 *   out[1] = max_rows norm_changes[dim][:,0]   out[2] = max_rows norm_changes[dim][:,1] */
int ssvi_hyper_reduce(SsviHandle* h, int32_t dim, double* out /*double[3]*/, void* stream);

/* predict_entry for the posterior-mean predictor (:1039-1052): out[e] = sum_D prod_modes m.
 * entries int32 [n][N]. */
int ssvi_predict_mean(SsviHandle* h, const int32_t* entries, int64_t n, double* out, void* stream);

/* Minibatch selection of update_natural_param_batch (:164-178) for a list of rows of ONE mode, drawn on the device:
 * a row holding more than batch_size entries keeps batch_size of them -- a uniform random subset without replacement,
 * np.random.choice(len, batch_size, replace=False) (:170-172) -- and a shorter row keeps all, in append order.
 *   rowptr int64 [n_dim + 1], perm int64 [nnz]: the mode's inverted index (Tensor.find_observed_ui, Tensor.py:251-258)
 *   rows int64 [n_rows];  seg_start int64 [n_rows]: where each row's picks begin in out;  out int64: entry ids
 * Counter-based generator keyed by (seed, row): a call is reproducible, a new seed gives a fresh draw.
 * batch_size <= 2048. */
int ssvi_draw_minibatches(SsviHandle* h, int64_t n_rows, const int64_t* rows, const int64_t* rowptr,
                          const int64_t* perm, int32_t batch_size, uint64_t seed, const int64_t* seg_start,
                          int64_t* out, void* stream);

/* Variate generation on the device: the stand-in for the reference's draws from numpy's global RNG
 * (np.random.multivariate_normal SSVI_TF_interface.py:252-254, :743-745; np.random.rayleigh :260 and
 * SSVI_TF_robust.py:53; np.random.normal SSVI_TF_simple.py:35, SSVI_TF_robust.py:163).  Fills the four pool arrays
 * (layouts and element type as in ssvi_bind_noise; r_w / z_f may be NULL for SSVI_MODEL_D) with standard normal
 * variates (r_w: unit Rayleigh) from a counter-based generator keyed by (seed, draw): the same pair always gives the
 * same pool, a new `draw` a fresh one.  The pool still has to be bound with ssvi_bind_noise. */
int ssvi_fill_noise(SsviHandle* h, void* z_other, void* z_own, void* r_w, void* z_f, int32_t pool_slots,
                    uint64_t seed, uint64_t draw, void* stream);

/* Per-entry Monte-Carlo estimators of the report path, from the bound noise pool (entry e reads slot
 * (slot0 + e) mod P; mode 0 draws from z_own, mode d >= 1 from z_other[.][d-1]; k1 draws, k2 noise-layer draws):
 *   what = SSVI_EST_EXPECTED_OBS (0)  out[e] = mean link(fs)        estimate_expected_observation_via_sampling (:1168-1189)
 *   what = SSVI_EST_LOG_PDF      (1)  out[e] = mean log_pdf(y, fs)  estimate_expectation_term_vlb (:889-908)
 *   what = SSVI_EST_PDF          (2)  out[e] = mean pdf(y, fs)      estimate_negative_log_likelihood (:939-964) before -log
 *   what = SSVI_EST_ROBUST_VLB   (3)  out[e] = mean_i log mean_k pdf(y, link(fs[i][k]))   SSVI_TF_robust's override of the VLB
 *                                     term (SSVI_TF_robust.py:234-262: link before the pdf, log of the first-layer mean);
 *                                     simple / robust models only, k1 <= 128
 * with ms_k = sum_D prod_modes x_mode[k], fs = ms for SSVI_TF_d and ms_k + w_sigma^2 r_k z_f[i][k] otherwise.
 * entries int32 [n][N]; ys double [n] (may be NULL for what = 0); w_state as in ssvi_accumulate. */
#define SSVI_EST_EXPECTED_OBS 0
#define SSVI_EST_LOG_PDF 1
#define SSVI_EST_PDF 2
#define SSVI_EST_ROBUST_VLB 3
int ssvi_estimate_entries(SsviHandle* h, int32_t what, const int32_t* entries, const double* ys, int64_t n,
                          int32_t slot0, const double* w_state, double* out, void* stream);

/* A block of iterations of factorize()'s loop (SSVI_TF_interface.py:117-143) in the reference's schedule: per
 * iteration, for every mode, update_natural_param_batch(dim, col[dim]) (:124) + update_hyper_parameter(dim) (:125), then
 * the pointer / time_step advance (:128-133) and check_stop_cond (:135-142).  Everything is enqueued on `stream` from C:
 * the row's minibatch is planned on the device (the draw of ssvi_draw_minibatches keyed by minibatch_seed +
 * minibatch_draws, chunk tables), fresh variates are drawn into the pool (ssvi_fill_noise keyed by noise_seed,
 * noise_draws) and the same kernels as ssvi_accumulate / ssvi_reduce_chunks / ssvi_row_update / ssvi_sigma_update /
 * ssvi_hyper_reduce run.  When the stop condition fires at iteration i the device word stop[0] is latched to 1
 * (stop[1] = i) and the state-writing kernels of the iterations enqueued behind it do nothing: the final state is the
 * one the reference's `break` leaves.  The call never synchronises; the caller reads `stop` when it wants to.
 * HOST struct; the counters, col and time_step are advanced in place (for every enqueued iteration: after a stop the
 * caller recomputes them from stop[1]).  Single GPU (no process group). */
typedef struct SsviLoop {
  const int64_t* rowptr[SSVI_MAX_ORDER];        /* DEVICE int64 [n_d + 1] per mode: the inverted index (Tensor.py:251-258) */
  const int64_t* perm[SSVI_MAX_ORDER];          /* DEVICE int64 [nnz] per mode */
  const int64_t* rowptr_host[SSVI_MAX_ORDER];   /* HOST copies of rowptr */
  void* z_other; void* z_own; void* r_w; void* z_f;   /* DEVICE noise pool (layouts of ssvi_bind_noise); bound by the call */
  double* pSigma;                               /* DEVICE [N]    (:56, :146-160) */
  double* w_state;                              /* DEVICE [3]    simple / robust */
  double* hyper_out;                            /* DEVICE [N][3] ssvi_hyper_reduce outputs of every mode */
  int32_t* stop;                                /* DEVICE int32[2], zeroed by the caller before the first block */
  void* workspace;                              /* DEVICE, >= ssvi_loop_workspace_bytes(h, batch_size) */
  int64_t workspace_bytes;
  double epsilon;                               /* stop when max(d_mean, d_cov) < epsilon (:142; the reference uses 1e-4) */
  uint64_t noise_seed, noise_draws;             /* in/out */
  uint64_t minibatch_seed, minibatch_draws;     /* in/out */
  int64_t col[SSVI_MAX_ORDER];                  /* in/out: update_column_pointer (:115) */
  int64_t launches;                             /* out: kernels enqueued (added to) */
  int32_t time_step[SSVI_MAX_ORDER];            /* in/out (:79, :130) */
  int32_t pool_slots, batch_size, chunk_entries, fixed_covariance;
} SsviLoop;
int64_t ssvi_loop_workspace_bytes(const SsviHandle* h, int32_t batch_size);
int ssvi_factorize_iterations(SsviHandle* h, SsviLoop* loop /*HOST*/, int64_t first_iteration, int64_t n_iterations,
                              void* stream);

/* Stages 1-3 + the in-order chunk reduction in ONE call that needs no workspace and allocates nothing.
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).  Equivalent to
 *   ssvi_accumulate(h, dim, n_chunks, ..., partial, stream);  ssvi_reduce_chunks(h, n_rows, row_chunk_ptr, partial, acc_out, stream);
 * with `partial` -- n_chunks records of ssvi_acc_len doubles -- held in a file-scope static device array of
 * SSVI_STATIC_SCRATCH_BYTES (8 MiB; every GPU that loads the library holds it until the process exits) instead of a
 * caller buffer.  Inputs and the output acc_out [n_rows][ssvi_acc_len] are the caller's, as for the two calls above.
 * Fails with SSVI_E_INVALID and a message naming both sizes when n_chunks * ssvi_acc_len * 8 exceeds the maximum.
 * Calls no cudaMalloc / cudaFree / allocating library routine.  All calls on a GPU share the one array, which keeps
 * the previous call's contents (nothing in it needs clearing: every chunk record is written in full before it is
 * read); calls that overlap on the GPU -- other streams, handles or host threads -- RACE on it: serialise them. */
#define SSVI_STATIC_SCRATCH_BYTES (8u << 20)
int64_t ssvi_static_scratch_bytes(void);
int ssvi_accumulate_rows_static(SsviHandle* h, int32_t dim, int64_t n_chunks, const int32_t* chunk_row,
                                const int64_t* chunk_start, const int32_t* chunk_len, const int32_t* chunk_slot0,
                                const int64_t* entry_ids, const int64_t* rows, const double* w_state, int64_t n_rows,
                                const int64_t* row_chunk_ptr, double* acc_out, void* stream);

/* Reads and clears the device flag word (synchronises `stream`).  flags_out HOST. */
int ssvi_check_flags(SsviHandle* h, int32_t* flags_out /*HOST*/, void* stream);

/* Library / kernel build info, e.g. "ssvi_b200 0.1 sm_100a". HOST string. */
const char* ssvi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* SSVI_B200_H */
