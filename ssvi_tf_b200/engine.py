"""Host-side driver of libssvi_b200: device data model + sweep plans + the C-ABI calls.

PyTorch is used for device memory, streams and torch.distributed only; every computation of the hot
path happens in the hand-written kernels behind include/ssvi_b200.h.
"""
import ctypes

import numpy as np
import torch

from . import _lib

MODEL_IDS = {"d": _lib.MODEL_D, "simple": _lib.MODEL_SIMPLE, "robust": _lib.MODEL_ROBUST}
LIK_IDS = {"normal": _lib.LIK_NORMAL, "bernoulli": _lib.LIK_BERNOULLI, "poisson": _lib.LIK_POISSON}
UPDATE_IDS = {"S": _lib.UPDATE_S, "N": _lib.UPDATE_N}
FLAG_TEXT = {
    _lib.FLAG_NONFINITE_ACC: "di/Di not finite (reference asserts SSVI_TF_interface.py:185-188)",
    _lib.FLAG_NONFINITE_COV: "S_next not finite (reference asserts SSVI_TF_interface.py:202-204)",
    _lib.FLAG_PHI_ZERO: "phi == 0 or not finite (reference asserts SSVI_TF_robust.py:87-92, 106)",
    _lib.FLAG_CHOL_FAIL: "covariance not positive definite",
    _lib.FLAG_EIG_NOCONV: "Jacobi eigen-iteration did not converge",
    _lib.FLAG_TC_RANGE: "tensor-core sweep (table_dtype='f32_tc'): an operand left the fp16 range; lower tc_scale_log2",
}


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def build_csr(coords, dims):
    """Per-mode inverted index in COO insertion order (Tensor/Tensor.py:76-83, 152-169, 251-258):
    perm[d][rowptr[d][i]:rowptr[d][i+1]] = entries touching row i of mode d, in append order.
    Host build (stable sort); returns lists of int64 numpy arrays."""
    coords = np.asarray(coords)
    rowptr, perm = [], []
    for d, n in enumerate(dims):
        col = coords[:, d]
        if len(col) and (col.min() < 0 or col.max() >= n):
            raise ValueError("coordinate out of range in mode %d" % d)
        perm.append(np.argsort(col, kind="stable").astype(np.int64))
        rp = np.zeros(n + 1, dtype=np.int64)
        np.cumsum(np.bincount(col, minlength=n), out=rp[1:])
        rowptr.append(rp)
    return rowptr, perm


def noise_block_len(D, k1, diag):
    """Elements of one (slot, mode) noise block == ssvi_noise_block_len (include/ssvi_b200.h)."""
    return D * k1 if diag else ((k1 + 7) // 8) * ((D + 3) // 4) * 32


def noise_offsets(D, k1, diag):
    """off[c, k] = position of element (component c, sample k) inside a block == ssvi_noise_offset."""
    c = np.arange(D)[:, None]
    k = np.arange(k1)[None, :]
    if diag:
        return c * k1 + k
    return ((k // 8) * ((D + 3) // 4) + c // 4) * 32 + (k % 8) * 4 + (c % 4)


class NoisePool(object):
    """The supplied standard variates (see include/ssvi_b200.h "Noise contract").

    Device layouts: z_other [P][N-1][block], z_own [P][block], r_w [P][k1], z_f [P][k2][k1]; element
    (c, k) of a block sits at noise_offsets(D, k1, diag)[c, k] (plain [c][k] for a diagonal covariance,
    DMMA B-fragment order for a full one).  Entry number b of a row's minibatch reads slot b mod P."""

    def __init__(self, z_other, z_own, r_w, z_f):
        self.z_other, self.z_own, self.r_w, self.z_f = z_other, z_own, r_w, z_f
        self.slots = int(z_own.shape[0])
        self.dtype = z_own.dtype

    @classmethod
    def random(cls, slots, N, D, k1, k2, need_second_layer, device, dtype=torch.float32, seed=319, diag=False):
        g = torch.Generator(device=device)
        g.manual_seed(seed)
        blk = noise_block_len(D, k1, diag)
        z_other = torch.randn((slots, N - 1, blk), generator=g, device=device, dtype=dtype)
        z_own = torch.randn((slots, blk), generator=g, device=device, dtype=dtype)
        r_w = z_f = None
        if need_second_layer:
            u = torch.rand((slots, k1), generator=g, device=device, dtype=torch.float64).clamp_(min=1e-300)
            r_w = torch.sqrt(-2.0 * torch.log(u)).to(dtype)          # unit Rayleigh
            z_f = torch.randn((slots, k2, k1), generator=g, device=device, dtype=dtype)
        return cls(z_other, z_own, r_w, z_f)

    @classmethod
    def empty(cls, slots, N, D, k1, k2, need_second_layer, device, dtype=torch.float32, diag=False):
        """Uninitialised pool of the right shapes (Engine.fill_noise draws into it on the device)."""
        blk = noise_block_len(D, k1, diag)
        z_other = torch.empty((slots, N - 1, blk), device=device, dtype=dtype)
        z_own = torch.empty((slots, blk), device=device, dtype=dtype)
        r_w = torch.empty((slots, k1), device=device, dtype=dtype) if need_second_layer else None
        z_f = torch.empty((slots, k2, k1), device=device, dtype=dtype) if need_second_layer else None
        return cls(z_other, z_own, r_w, z_f)

    @classmethod
    def from_tape(cls, tape, device, dtype=torch.float64, diag=False):
        """tape: numpy-natural layout (z_other (B,N-1,k1,D), z_own (B,k1,D), r_w (B,k1), z_f (B,k2,k1))."""
        k1, D = tape.z_own.shape[1], tape.z_own.shape[2]
        off = noise_offsets(D, k1, diag).reshape(-1)                  # [c, k] flattened
        blk = noise_block_len(D, k1, diag)

        def block(a):                                                 # (..., k1, D) -> (..., blk)
            src = np.swapaxes(a, -1, -2).reshape(a.shape[:-2] + (D * k1,))
            out = np.zeros(a.shape[:-2] + (blk,), dtype=np.float64)
            out[..., off] = src
            return torch.as_tensor(out).to(device=device, dtype=dtype).contiguous()

        def dev(a):
            return None if a is None else torch.as_tensor(np.ascontiguousarray(a)).to(device=device, dtype=dtype)
        return cls(block(tape.z_other), block(tape.z_own), dev(tape.r_w), dev(tape.z_f))


class SweepPlan(object):
    """Chunk tables for one call of ssvi_accumulate over a list of rows of one mode."""

    def __init__(self, dim, rows, n_obs, entry_ids, chunk_row, chunk_start, chunk_len, chunk_slot0,
                 row_chunk_ptr, batch_sizes):
        self.dim = dim
        self.rows = rows                    # int64 [n_rows] device
        self.n_obs = n_obs                  # int64 [n_rows] device: len(find_observed_ui(dim,row))
        self.entry_ids = entry_ids          # int64 device
        self.chunk_row, self.chunk_start = chunk_row, chunk_start
        self.chunk_len, self.chunk_slot0 = chunk_len, chunk_slot0
        self.row_chunk_ptr = row_chunk_ptr  # int64 [n_rows+1] device
        self.n_rows = int(rows.numel())
        self.n_chunks = int(chunk_row.numel())
        self.batch_sizes = batch_sizes      # host int64 [n_rows]: entries in each row's minibatch
        self.n_entries = int(batch_sizes.sum())


class SlotSweepPlan(object):
    """Slot-major chunk tables for ssvi_sweep_accumulate (the sample-table path of a full sweep).

    Chunks are sorted by (slot, [tile,] row, segment) -- the launch order -- and carry the index of their record
    in their slot group's row-major record buffer (chunk_out), so ssvi_reduce_chunks_into sums each row's records
    of the group in one contiguous, fixed-order run.  groups: [(slot_lo, n_slots, chunk_lo, chunk_hi)]."""

    def __init__(self, dim, rows, n_obs, entry_ids, chunk_row, chunk_start, chunk_len, chunk_slot, chunk_out,
                 group_rcp, batch_sizes, groups, table_bytes):
        self.dim = dim
        self.rows, self.n_obs, self.entry_ids = rows, n_obs, entry_ids
        self.chunk_row, self.chunk_start, self.chunk_len = chunk_row, chunk_start, chunk_len
        self.chunk_slot, self.chunk_out = chunk_slot, chunk_out     # chunk_out: record index inside the batch's buffer
        self.group_rcp = group_rcp                                  # int64 [n_batches][n_rows + 1] device
        self.batch_of = self.batch_last = None                      # host [n_groups]: reduction batch of every group
        self.max_batch_records = 0
        self.n_rows = int(rows.numel())
        self.n_chunks = int(chunk_row.numel())
        self.batch_sizes = batch_sizes
        self.n_entries = int(batch_sizes.sum())
        self.groups = groups
        self.table_bytes = int(table_bytes)
        self.tile_blocks = 1                # L2 tile blocks per other mode (_plan_sweep_tiled)
        self.chunk_start_h = None           # host int64 [n_chunks + 1]: first entry position of every chunk (+ total)


class Engine(object):
    """One GPU's view of the problem: COO shard + CSR index + posterior state + noise pool."""

    def __init__(self, cfg, dims, coords, vals, device="cuda", noise_dtype=torch.float32, chunk_entries=16,
                 kernel_path=0, table_budget_bytes=48 << 20, sweep_chunk_entries=64, sweep_streams=2,
                 sweep_tile_blocks=0, sweep_tile_min_entries=24.0, record_budget_bytes=256 << 20, tuning=0,
                 table_dtype=torch.float64, tc_scale_log2=0):
        """cfg: dict with model, likelihood, D, diag, k1, k2, batch_size, mean_update, cov_update, eta,
        cov_eta, sigma_eta (+ hidden constants poisson_eta, min_eigenval, epsilon, likelihood_param)."""
        if not torch.cuda.is_available():
            raise RuntimeError("ssvi_tf_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
        self.lib = _lib.load()
        self.cfg = dict(cfg)
        self.dims = [int(x) for x in dims]
        self.N = len(self.dims)
        self.D = int(cfg["D"])
        self.diag = bool(cfg["diag"])
        self.device = torch.device(device)
        self._dev_index = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self.sm_count = torch.cuda.get_device_properties(self.device).multi_processor_count   # 148 on a B200
        self.chunk_entries = int(chunk_entries)
        self.table_budget_bytes = int(table_budget_bytes)      # L2-resident sample table of the sweep path
        self.sweep_chunk_entries = int(sweep_chunk_entries)
        self.sweep_streams = int(sweep_streams)
        self.sweep_tile_blocks = int(sweep_tile_blocks)        # 0: chosen from the table budget (_sweep_tile_blocks)
        self.sweep_tile_min_entries = float(sweep_tile_min_entries)
        self.record_budget_bytes = int(record_budget_bytes)    # chunk records kept per stream before they are folded
        self._side_streams = []
        self._pin_ring = None
        self.noise_dtype = noise_dtype
        if cfg["mean_update"] not in UPDATE_IDS or cfg["cov_update"] not in UPDATE_IDS:
            raise Exception("Unidentified update formula for covariance param")   # SSVI_TF_interface.py:612, 632

        c = _lib.SsviConfig()
        c.order, c.rank = self.N, self.D
        c.cov_kind = _lib.COV_DIAG if self.diag else _lib.COV_FULL
        c.model, c.likelihood = MODEL_IDS[cfg["model"]], LIK_IDS[cfg["likelihood"]]
        c.k1, c.k2 = int(cfg["k1"]), int(cfg["k2"])
        c.mean_update, c.cov_update = UPDATE_IDS[cfg["mean_update"]], UPDATE_IDS[cfg["cov_update"]]
        c.noise_dtype = _lib.NOISE_F32 if noise_dtype == torch.float32 else _lib.NOISE_F64
        c.kernel_path = int(kernel_path)      # 0 auto, 1 force the generic CUDA-core accumulate path
        c.tuning = int(tuning)                # SSVI_TUNE_* bits (include/ssvi_b200.h); never changes a result
        # sample table of the sweep path as floats (SSVI_TF_d, full covariance): the contract's fp32 option, tolerance 1e-4
        # table_dtype="f32_tc": the fp32 ARITHMETIC option, contraction on the tcgen05 tensor cores (accumulate_tc.cuh)
        c.table_dtype = (_lib.TABLE_F32_TC if table_dtype == "f32_tc"
                         else _lib.TABLE_F32 if table_dtype == torch.float32 else _lib.TABLE_F64)
        c.tc_scale_log2 = int(tc_scale_log2)
        self.table_dtype = table_dtype
        c.eta, c.cov_eta, c.sigma_eta = float(cfg["eta"]), float(cfg["cov_eta"]), float(cfg["sigma_eta"])
        c.poisson_eta = float(cfg.get("poisson_eta", 0.1))
        c.min_eigenval = float(cfg.get("min_eigenval", 0.01))
        c.cov_floor = float(cfg.get("epsilon", 1e-4))
        c.likelihood_param = float(cfg.get("likelihood_param", 1.0 if cfg["likelihood"] == "normal" else 0.0))
        for d, n in enumerate(self.dims):
            c.dims[d] = n
        self._c = c
        self.acc_len = int(self.lib.ssvi_acc_len(ctypes.byref(c)))
        h = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            rc = self.lib.ssvi_create(ctypes.byref(c), ctypes.byref(h))
        if rc != 0:
            raise _lib.SsviError("ssvi_create failed: %s" % self.lib.ssvi_last_error(None).decode())
        self.h = h
        self._quad_rule = None
        if cfg.get("quadrature"):                        # SSVI_TF_robust(quadrature=True): Gauss-Hermite degree
            self.set_quadrature(int(cfg["quadrature"]))

        # ---- data: COO + per-mode CSR -------------------------------------------------------
        if torch.is_tensor(coords) and coords.is_cuda:
            # device build of the index: stable sort by row == append order
            self.coords = coords.to(device=self.device, dtype=torch.int32).reshape(-1, self.N).contiguous()
            self.vals = vals.to(device=self.device, dtype=torch.float64).reshape(-1).contiguous()
            self.nnz = int(self.vals.shape[0])
            self.perm, self.rowptr_h = [], []
            for d, n in enumerate(self.dims):
                col = self.coords[:, d].long()
                if self.nnz and (int(col.min()) < 0 or int(col.max()) >= n):
                    raise ValueError("coordinate out of range in mode %d" % d)
                self.perm.append(torch.sort(col, stable=True).indices.contiguous())
                rp = torch.zeros(n + 1, dtype=torch.int64, device=self.device)
                rp[1:] = torch.cumsum(torch.bincount(col, minlength=n), 0)
                self.rowptr_h.append(rp.cpu().numpy())
                del col
            self._perm_h = None
        else:
            coords = np.ascontiguousarray(np.asarray(coords).reshape(-1, self.N))
            vals = np.ascontiguousarray(np.asarray(vals, dtype=np.float64).reshape(-1))
            self.nnz = int(vals.shape[0])
            self.rowptr_h, perm_h = build_csr(coords, self.dims)
            self.coords = torch.as_tensor(coords.astype(np.int32)).to(self.device)
            self.vals = torch.as_tensor(vals).to(self.device)
            self.perm = [torch.as_tensor(p).to(self.device) for p in perm_h]
            self._perm_h = perm_h
        self._call(self.lib.ssvi_bind_data, _ptr(self.coords), _ptr(self.vals), self.nnz)
        # index range this shard touches in every mode (a shard cut along index-space blocks touches a sub-range)
        self.row_window = []
        for d in range(self.N):
            ne = np.nonzero(np.diff(self.rowptr_h[d]))[0]
            self.row_window.append((int(ne[0]), int(ne[-1])) if len(ne) else (0, 0))
        # ... and the sweep path samples its table for those rows only (matters for sharded runs: at 8 ranks a rank touches
        # half of every mode of config 5, and sampling all 10k rows per mode was 5.5 % of its sweep)
        self._win_lo = np.array([w[0] for w in self.row_window], dtype=np.int64)
        self._win_n = np.array([w[1] - w[0] + 1 for w in self.row_window], dtype=np.int64)
        self._call(self.lib.ssvi_set_row_windows, ctypes.c_void_p(self._win_lo.ctypes.data),
                   ctypes.c_void_p(self._win_n.ctypes.data))

        # ---- state ------------------------------------------------------------------------------
        f64 = dict(device=self.device, dtype=torch.float64)
        rl = 2 * self.D if self.diag else (self.D + 1) * self.D
        self.row_len = rl
        self.params = [torch.zeros((n, rl // self.D, self.D), **f64) for n in self.dims]
        self.srow_len = int(self.lib.ssvi_sampler_row_len(ctypes.byref(c)))
        self.sampler = [torch.zeros((n, self.srow_len), **f64) for n in self.dims]
        self.ada_mean = [torch.ones((n, self.D), **f64) for n in self.dims]        # ada_acc_grad^T (:83)
        self.ada_cov = [torch.zeros((n, self.D), **f64) for n in self.dims]        # ada_acc_grad_cov^T (:41)
        self.norm_changes = [torch.zeros((n, 2), **f64) for n in self.dims]
        self.grad_norms = [torch.zeros((n, 2), **f64) for n in self.dims]
        self.pSigma = torch.ones((self.N,), **f64)
        self.w_state = torch.tensor([1.0, 0.0, 1.0], **f64)                        # w_sigma, w_ada_grad, w_tau
        self.hyper_out = torch.zeros((self.N, 3), **f64)
        self.time_step = [0] * self.N
        self._bind_state()
        self.noise = None
        self._full_plans = {}
        self._sweep_plans = {}
        self._sub_plans = {}
        self._rowptr_d = {}
        self.sweep_supported = bool(self.lib.ssvi_sweep_supported(self.h))
        self._scratch = {}
        self.launches = 0          # kernels of libssvi_b200 launched through this engine
        self.profile_events = None  # set to a list to time every ssvi_accumulate launch with CUDA events

    @property
    def perm_h(self):
        if self._perm_h is None:
            self._perm_h = [p.cpu().numpy() for p in self.perm]
        return self._perm_h

    # -------------------------------------------------------------------------------------------
    def _call(self, fn, *args):
        if torch.cuda.current_device() == self._dev_index:       # the per-row path makes ~10 calls per row update
            rc = fn(self.h, *args)
        else:
            with torch.cuda.device(self.device):
                rc = fn(self.h, *args)
        _lib.check(self.lib, self.h, rc)

    def _stream(self):
        """The caller's current CUDA stream on this device, as the void* the C ABI takes."""
        return ctypes.c_void_p(torch._C._cuda_getCurrentRawStream(self._dev_index))

    _PIN_SLOT, _PIN_SLOTS = 1 << 16, 16                    # pinned staging ring of _pack_to_device

    def _pack_to_device(self, arrays):
        """Host int32 / int64 arrays -> device tensors through ONE host-to-device copy (16-byte aligned sections of one
        buffer): a per-row plan is eight small tables, and eight synchronous copies were half of a row update's host
        time (tools/factorize_profile.py)."""
        arrays = [np.ascontiguousarray(a) for a in arrays]
        offs, total = [], 0
        for a in arrays:
            total = (total + 15) & ~15
            offs.append(total)
            total += a.nbytes
        total = max(total, 16)
        if total <= self._PIN_SLOT:
            # small plans (the per-row path) go through a ring of pinned slots with an asynchronous copy, so the host
            # builds the next row's plan while the device still works on this one; a slot is reused only after the
            # copy that read it has completed
            if self._pin_ring is None:
                self._pin_ring = torch.empty((self._PIN_SLOTS, self._PIN_SLOT), dtype=torch.uint8).pin_memory()
                self._pin_np = self._pin_ring.numpy()
                self._pin_events = [None] * self._PIN_SLOTS
                self._pin_next = 0
            i = self._pin_next
            self._pin_next = (i + 1) % self._PIN_SLOTS
            if self._pin_events[i] is not None:
                self._pin_events[i].synchronize()
            buf = self._pin_np[i]
            for a, o in zip(arrays, offs):
                buf[o:o + a.nbytes] = a.reshape(-1).view(np.uint8)
            d = torch.empty((total,), dtype=torch.uint8, device=self.device)
            d.copy_(self._pin_ring[i, :total], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self.device))
            self._pin_events[i] = ev
        else:
            buf = np.empty(total, dtype=np.uint8)
            for a, o in zip(arrays, offs):
                buf[o:o + a.nbytes] = a.reshape(-1).view(np.uint8)
            d = torch.from_numpy(buf).to(self.device)
        return [d[o:o + a.nbytes].view(torch.int32 if a.dtype == np.int32 else torch.int64)
                for a, o in zip(arrays, offs)]

    def _bind_state(self):
        def arr(lst):
            a = (ctypes.c_void_p * self.N)()
            for d, t in enumerate(lst):
                a[d] = t.data_ptr()
            return a
        self._call(self.lib.ssvi_bind_state, arr(self.params), arr(self.sampler), arr(self.ada_mean),
                   arr(self.ada_cov), arr(self.norm_changes), arr(self.grad_norms))

    def close(self):
        if getattr(self, "h", None) is not None:
            self.lib.ssvi_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -------------------------------------------------------------------------------------------
    def set_noise(self, pool):
        need2 = self.cfg["model"] != "d"
        blk = int(self.lib.ssvi_noise_block_len(ctypes.byref(self._c)))
        assert blk == noise_block_len(self.D, self.cfg["k1"], self.diag)
        assert pool.z_other.shape[1:] == (self.N - 1, blk), pool.z_other.shape
        assert pool.z_own.shape[1:] == (blk,)
        if need2:
            assert pool.r_w.shape[1:] == (self.cfg["k1"],) and pool.z_f.shape[1:] == (self.cfg["k2"], self.cfg["k1"])
        want = torch.float32 if self._c.noise_dtype == _lib.NOISE_F32 else torch.float64
        assert pool.dtype == want, "noise pool dtype %s does not match the engine's noise_dtype" % pool.dtype
        self.noise = pool
        self._call(self.lib.ssvi_bind_noise, _ptr(pool.z_other), _ptr(pool.z_own),
                   _ptr(pool.r_w if need2 else None), _ptr(pool.z_f if need2 else None), pool.slots)

    def fill_noise(self, pool, seed, draw):
        """Fresh standard variates into `pool` (ssvi_fill_noise: counter-based generator keyed by (seed, draw)); one
        launch on the caller's stream.  The stand-in for the reference's numpy RNG draws."""
        self._call(self.lib.ssvi_fill_noise, _ptr(pool.z_other), _ptr(pool.z_own), _ptr(pool.r_w), _ptr(pool.z_f),
                   int(pool.slots), int(seed) & 0xFFFFFFFFFFFFFFFF, int(draw) & 0xFFFFFFFFFFFFFFFF, self._stream())
        self.launches += 1

    # ---- factorize()'s loop enqueued from C (ssvi_factorize_iterations) ---------------------------------------------------
    def loop_begin(self, pool, batch_size, fixed_covariance, epsilon, noise_seed, noise_draws, minibatch_seed,
                   minibatch_draws, col):
        """Argument block for factorize_iterations(): the reference's one-row-per-mode schedule, single GPU."""
        lp = _lib.SsviLoop()
        self._loop_keep = []
        for d in range(self.N):
            if d not in self._rowptr_d:
                self._rowptr_d[d] = torch.as_tensor(np.asarray(self.rowptr_h[d], dtype=np.int64)).to(self.device)
            rp_h = np.ascontiguousarray(self.rowptr_h[d], dtype=np.int64)
            self._loop_keep.append(rp_h)
            lp.rowptr[d] = self._rowptr_d[d].data_ptr()
            lp.perm[d] = self.perm[d].data_ptr()
            lp.rowptr_host[d] = rp_h.ctypes.data
            lp.col[d] = int(col[d])
            lp.time_step[d] = int(self.time_step[d])
        lp.z_other, lp.z_own = pool.z_other.data_ptr(), pool.z_own.data_ptr()
        lp.r_w = pool.r_w.data_ptr() if pool.r_w is not None else None
        lp.z_f = pool.z_f.data_ptr() if pool.z_f is not None else None
        lp.pool_slots = int(pool.slots)
        lp.pSigma, lp.w_state, lp.hyper_out = self.pSigma.data_ptr(), self.w_state.data_ptr(), self.hyper_out.data_ptr()
        self._loop_stop = torch.zeros((2,), dtype=torch.int32, device=self.device)
        lp.stop = self._loop_stop.data_ptr()
        nbytes = int(self.lib.ssvi_loop_workspace_bytes(self.h, int(batch_size)))
        self._loop_ws = torch.empty((nbytes + 15,), dtype=torch.uint8, device=self.device)
        lp.workspace = (self._loop_ws.data_ptr() + 15) & ~15
        lp.workspace_bytes = nbytes
        lp.epsilon = float(epsilon)
        lp.noise_seed, lp.noise_draws = int(noise_seed) & (2 ** 64 - 1), int(noise_draws)
        lp.minibatch_seed, lp.minibatch_draws = int(minibatch_seed) & (2 ** 64 - 1), int(minibatch_draws)
        lp.batch_size, lp.chunk_entries = int(batch_size), int(self.chunk_entries)
        lp.fixed_covariance = int(bool(fixed_covariance))
        self.noise = pool
        return lp

    def factorize_iterations(self, lp, first, count):
        """Enqueue `count` iterations (no synchronisation); lp's counters, col and time_step advance in place."""
        before = int(lp.launches)
        self._call(self.lib.ssvi_factorize_iterations, ctypes.byref(lp), int(first), int(count), self._stream())
        self.launches += int(lp.launches) - before

    def loop_stop_state(self):
        """(fired, iteration) of the device stop word -- synchronises."""
        s = self._loop_stop.cpu().numpy()
        return bool(s[0]), int(s[1])

    def set_tuning(self, tuning):
        """SSVI_TUNE_* kernel options for the calls that follow (A/B timing inside one process)."""
        self._c.tuning = int(tuning)
        self._call(self.lib.ssvi_set_tuning, int(tuning))

    def set_quadrature(self, degree):
        """SSVI_TF_robust(quadrature=True), SSVI_TF_robust.py:22-25: Gauss-Hermite rule of `degree` nodes in place of
        the k2 second-layer draws (degree 0 / None: back to Monte Carlo)."""
        if not degree:
            self._quad_rule = None
            self._call(self.lib.ssvi_bind_quadrature, None, 0)
            return
        xs, ws = np.polynomial.hermite.hermgauss(int(degree))
        self._quad_rule = torch.as_tensor(np.concatenate([xs, ws]), dtype=torch.float64).to(self.device)
        self._call(self.lib.ssvi_bind_quadrature, _ptr(self._quad_rule), int(degree))

    def refresh_sampler(self, dim=None, rows=None):
        dims = range(self.N) if dim is None else [dim]
        for d in dims:
            n = 0 if rows is None else int(rows.numel())
            self._call(self.lib.ssvi_refresh_sampler, d, _ptr(rows), n, self._stream())
            self.launches += 1

    def check_flags(self):
        """Mirror of the reference's in-line asserts: raises AssertionError when a guard fired."""
        f = ctypes.c_int32(0)
        self._call(self.lib.ssvi_check_flags, ctypes.byref(f), self._stream())
        if f.value:
            raise AssertionError("; ".join(t for b, t in FLAG_TEXT.items() if f.value & b))

    # -------------------------------------------------------------------------------------------
    def _chunk_tables(self, dim, rows_h, seg_start, seg_len, slot_offset=None, extra=()):
        """rows_h: host int64 row ids; seg_start/seg_len: each row's minibatch as a run of entry_ids;
        slot_offset: per-row position of this shard's first entry inside the GLOBAL minibatch (multi-GPU).
        Returns the five device tables followed by the device copies of the host arrays in `extra` (one upload)."""
        # one warp works through one chunk: small plans (the per-row path of factorize: 128 entries) are cut finer so
        # that they spread over the SMs instead of queueing on eight warps (config 1 factorize loop: 1075 -> 1460
        # iterations/s at 4 entries per chunk); large plans keep chunk_entries
        total_entries = int(np.sum(seg_len))
        W = int(min(self.chunk_entries, max(2, -(-total_entries // (8 * self.sm_count)))))
        nch = (seg_len + W - 1) // W
        rcp = np.zeros(len(rows_h) + 1, dtype=np.int64)
        np.cumsum(nch, out=rcp[1:])
        total = int(rcp[-1])
        chunk_row = np.repeat(np.arange(len(rows_h), dtype=np.int64), nch)
        t = np.arange(total, dtype=np.int64) - rcp[chunk_row]
        chunk_start = seg_start[chunk_row] + t * W
        chunk_len = np.minimum(W, seg_len[chunk_row] - t * W)
        slot0 = t * W if slot_offset is None else t * W + np.asarray(slot_offset, dtype=np.int64)[chunk_row]
        return self._pack_to_device([chunk_row.astype(np.int32), chunk_start.astype(np.int64),
                                     chunk_len.astype(np.int32), slot0.astype(np.int32), rcp] +
                                    [np.asarray(a, dtype=np.int64) for a in extra])

    def plan_full(self, dim, rows=None, slot_offset=None):
        """Every observed entry of every listed row (scale = 1): the throughput sweep."""
        key = (dim, None if rows is None else tuple(int(r) for r in rows),
               None if slot_offset is None else hash(np.asarray(slot_offset).tobytes()))
        if key in self._full_plans:
            return self._full_plans[key]
        rows_h = np.arange(self.dims[dim], dtype=np.int64) if rows is None else np.asarray(rows, dtype=np.int64)
        rp = self.rowptr_h[dim]
        seg_start, seg_len = rp[rows_h], rp[rows_h + 1] - rp[rows_h]
        tabs = self._chunk_tables(dim, rows_h, seg_start, seg_len, slot_offset, extra=(rows_h, seg_len))
        plan = SweepPlan(dim, tabs[5], tabs[6], self.perm[dim], tabs[0], tabs[1], tabs[2], tabs[3], tabs[4], seg_len)
        self._full_plans[key] = plan
        return plan

    def plan_minibatch(self, dim, rows, batch_size, rng=None, sub_idx=None, slot_offset=None, device_seed=None):
        """The reference's minibatch (SSVI_TF_interface.py:170-174): rows holding more than batch_size
        entries are subsampled without replacement.  sub_idx: optional {row: index array} (noise-tape
        item #0) overriding the draw; rng: numpy Generator for a host draw; device_seed: draw on the device instead
        (ssvi_draw_minibatches keyed by this seed: nothing but the chunk tables crosses the bus, and no host copy of
        the mode's inverted index is needed)."""
        rows_h = np.asarray(rows, dtype=np.int64)
        rp = self.rowptr_h[dim]
        lens = rp[rows_h + 1] - rp[rows_h]
        scalar = np.ndim(batch_size) == 0          # an array gives every row its own size (a shard's share, dist.py)
        bs = np.broadcast_to(np.asarray(batch_size, dtype=np.int64), lens.shape)
        if not np.any(lens > bs):
            return self.plan_full(dim, rows_h, slot_offset)
        seg_len = np.minimum(lens, bs)
        seg_start = np.zeros(len(rows_h), dtype=np.int64)
        np.cumsum(seg_len[:-1], out=seg_start[1:])
        if device_seed is not None and sub_idx is None and scalar and batch_size <= 2048:
            tabs = self._chunk_tables(dim, rows_h, seg_start, seg_len, slot_offset, extra=(rows_h, lens, seg_start))
            if dim not in self._rowptr_d:
                self._rowptr_d[dim] = torch.as_tensor(np.asarray(rp, dtype=np.int64)).to(self.device)
            entry_ids = torch.empty((int(seg_len.sum()),), dtype=torch.int64, device=self.device)
            self._call(self.lib.ssvi_draw_minibatches, len(rows_h), _ptr(tabs[5]), _ptr(self._rowptr_d[dim]),
                       _ptr(self.perm[dim]), int(batch_size), int(device_seed) & 0xFFFFFFFFFFFFFFFF, _ptr(tabs[7]),
                       _ptr(entry_ids), self._stream())
            self.launches += 1
            return SweepPlan(dim, tabs[5], tabs[6], entry_ids, tabs[0], tabs[1], tabs[2], tabs[3], tabs[4], seg_len)
        ids = []
        for r, row in enumerate(rows_h):
            seg = self.perm_h[dim][rp[row]:rp[row + 1]]
            if lens[r] > bs[r]:
                if sub_idx is not None and int(row) in sub_idx:
                    pick = np.asarray(sub_idx[int(row)], dtype=np.int64)
                else:
                    pick = (rng or np.random).choice(int(lens[r]), int(bs[r]), replace=False)
                seg = seg[pick]
            ids.append(seg)
        entry_ids = np.concatenate(ids) if ids else np.zeros(0, np.int64)
        tabs = self._chunk_tables(dim, rows_h, seg_start, seg_len, slot_offset, extra=(rows_h, lens, entry_ids))
        return SweepPlan(dim, tabs[5], tabs[6], tabs[7], tabs[0], tabs[1], tabs[2], tabs[3], tabs[4], seg_len)

    def plan_subsampled_sweep(self, dim, batch_size, generator=None):
        """The reference's minibatch rule (SSVI_TF_interface.py:170-174) for EVERY row of mode `dim` at once, drawn on
        the device by ssvi_draw_minibatches: rows holding more than batch_size entries keep batch_size of them, a
        uniform random subset without replacement (one warp per row, no sort of the entry list); shorter rows keep
        all, in append order.  The chunk tables depend on (dim, batch_size) only and are cached; every call draws a
        fresh subsample (generator: its seed keys the draws).  Single shard."""
        B = int(batch_size)
        key = (dim, B)
        c = self._sub_plans.get(key)
        if c is None:
            rp = self.rowptr_h[dim]
            lens = rp[1:] - rp[:-1]
            seg_len = np.minimum(lens, B)
            seg_start = np.zeros(len(lens), dtype=np.int64)
            np.cumsum(seg_len[:-1], out=seg_start[1:])
            rows_h = np.arange(self.dims[dim], dtype=np.int64)
            tabs = self._chunk_tables(dim, rows_h, seg_start, seg_len)
            dev = self.device
            c = dict(tabs=tabs, seg_len=seg_len, rows=torch.as_tensor(rows_h).to(dev),
                     n_obs=torch.as_tensor(lens).to(dev), rowptr=torch.as_tensor(np.asarray(rp, dtype=np.int64)).to(dev),
                     seg_start=torch.as_tensor(seg_start).to(dev), total=int(seg_len.sum()), calls=0)
            self._sub_plans[key] = c
        c["calls"] += 1
        base = 319 if generator is None else int(generator.initial_seed())
        seed = (base * 0x9E3779B1 + dim * 0x85EBCA6B + c["calls"]) & 0xFFFFFFFFFFFFFFFF
        entry_ids = torch.empty((c["total"],), dtype=torch.int64, device=self.device)
        self._call(self.lib.ssvi_draw_minibatches, self.dims[dim], _ptr(c["rows"]), _ptr(c["rowptr"]),
                   _ptr(self.perm[dim]), B, seed, _ptr(c["seg_start"]), _ptr(entry_ids), self._stream())
        self.launches += 1
        t = c["tabs"]
        return SweepPlan(dim, c["rows"], c["n_obs"], entry_ids, t[0], t[1], t[2], t[3], t[4], c["seg_len"])

    def plan_sweep(self, dim, slot_offset=None):
        """Full sweep of mode `dim` (every observed entry of every row, scale = 1) for the sample-table
        path: entries regrouped by the noise slot they read (entry b of a row reads slot (b + offset) mod P),
        one chunk per (slot, row[, segment]).  Needs the noise pool (P) to be set."""
        assert self.noise is not None, "set_noise() first"
        assert self.sweep_supported, "configuration not supported by the sample-table path"
        P = int(self.noise.slots)
        key = (dim, P, None if slot_offset is None else hash(np.asarray(slot_offset).tobytes()))
        if key in self._sweep_plans:
            return self._sweep_plans[key]
        nb = self._sweep_tile_blocks(dim)
        if nb > 1:
            plan = self._plan_sweep_tiled(dim, slot_offset, nb)
            self._sweep_plans[key] = plan
            return plan
        n = self.dims[dim]
        dev = self.device
        rp = self.rowptr_h[dim]
        lens = rp[1:] - rp[:-1]
        off = np.zeros(n, dtype=np.int64) if slot_offset is None else np.asarray(slot_offset, dtype=np.int64) % P
        # ---- slot-major permutation of the mode's CSR entry list (device) ----
        lens_d = torch.as_tensor(lens).to(dev)
        row_of = torch.repeat_interleave(torch.arange(n, device=dev, dtype=torch.int64), lens_d)
        b = torch.arange(int(rp[-1]), device=dev, dtype=torch.int64) - torch.as_tensor(rp[:-1]).to(dev)[row_of]
        slot = (b + torch.as_tensor(off).to(dev)[row_of]) % P
        del b, row_of
        order = torch.sort(slot, stable=True).indices           # (slot, row, b): stable keeps (row, b)
        del slot
        entry_ids = self.perm[dim][order].contiguous()
        del order
        # ---- per (slot, row) counts and chunk tables (host) ----
        s_idx = np.arange(P, dtype=np.int64)[:, None]
        b0 = (s_idx - off[None, :]) % P                          # first position of row r that reads slot s
        cnt = np.where(lens[None, :] > b0, (lens[None, :] - b0 + P - 1) // P, 0)      # [P, n]
        W = self.sweep_chunk_entries
        if self.diag:
            # diagonal: a CTA stages N noise blocks (100 KB at config 4) per chunk, so chunks are as long as the launch
            # can afford: 16x the base length while that still leaves >= 8 waves of CTAs (2 per SM), else 8x, else 4x
            for mult in (16, 8, 4):
                if mult == 4 or int(((cnt + W * mult - 1) // (W * mult)).sum()) >= 8 * 2 * self.sm_count:
                    W *= mult
                    break
        nseg = (cnt + W - 1) // W                                # [P, n]
        flat_cnt, flat_seg = cnt.reshape(-1), nseg.reshape(-1)   # (slot, row) order == launch order
        pair_start = np.zeros(P * n + 1, dtype=np.int64)
        np.cumsum(flat_cnt, out=pair_start[1:])
        pair_chunk = np.zeros(P * n + 1, dtype=np.int64)
        np.cumsum(flat_seg, out=pair_chunk[1:])
        total = int(pair_chunk[-1])
        pair_of = np.repeat(np.arange(P * n, dtype=np.int64), flat_seg)
        seg = np.arange(total, dtype=np.int64) - pair_chunk[pair_of]
        c_cnt, c_nseg = flat_cnt[pair_of], flat_seg[pair_of]
        base, rem = c_cnt // c_nseg, c_cnt % c_nseg              # segments differ by at most one entry
        chunk_len = base + (seg < rem)
        chunk_start = pair_start[pair_of] + seg * base + np.minimum(seg, rem)
        chunk_slot = pair_of // n
        chunk_row = pair_of % n
        # ---- slot groups: as many slots per call as the table budget holds ----
        per_slot = int(self.lib.ssvi_sweep_table_bytes(self.h, dim, 1))
        gs = max(1, min(P, self.table_budget_bytes // max(per_slot, 1)))
        slot_chunk = pair_chunk[::n]                             # first chunk of each slot (+ total)
        groups = [(s0, min(gs, P - s0), int(slot_chunk[s0]), int(slot_chunk[min(s0 + gs, P)]))
                  for s0 in range(0, P, gs)]
        chunk_out, group_rcp, batch_of, batch_last, max_rec = self._group_records(chunk_slot // gs, chunk_row, n, groups)
        t = lambda a, dt: torch.as_tensor(np.ascontiguousarray(a.astype(dt))).to(dev)      # noqa: E731
        plan = SlotSweepPlan(dim, torch.arange(n, device=dev, dtype=torch.int64), lens_d, entry_ids,
                             t(chunk_row, np.int32), t(chunk_start, np.int64), t(chunk_len, np.int32),
                             t(chunk_slot, np.int32), t(chunk_out, np.int64), t(group_rcp, np.int64), lens,
                             groups, per_slot * gs)
        plan.chunk_start_h = np.concatenate([chunk_start, [int(pair_start[-1])]]).astype(np.int64)
        plan.batch_of, plan.batch_last, plan.max_batch_records = batch_of, batch_last, max_rec
        self._sweep_plans[key] = plan
        return plan

    def _group_records(self, chunk_group, chunk_row, n, groups):
        """Chunk records are folded into the row accumulators batch by batch: a batch = consecutive slot groups of ONE
        stream (groups alternate between the streams) whose records fit the record budget; its records live in that
        stream's buffer until the batch's last group has been launched.  For chunks listed in launch order returns
        chunk_out (index of the chunk's record inside its batch's buffer, row-major: a row's records contiguous, in
        launch order), batch_rcp [n_batches][n + 1] (each batch's row pointer into the buffer), batch_of [n_groups],
        batch_last [n_groups] (group closes its batch) and the largest batch's record count."""
        G = len(groups)
        ns = max(1, min(self.sweep_streams, G))
        cnt = np.array([g[3] - g[2] for g in groups], dtype=np.int64)
        cap = max(self.record_budget_bytes // (self.acc_len * 8), 1)
        batch_of = np.zeros(G, dtype=np.int64)
        batch_last = np.zeros(G, dtype=bool)
        nbatch = 0
        for i in range(ns):
            cur, prev = 0, None
            for k in range(i, G, ns):
                if prev is not None and cur + cnt[k] > cap:
                    batch_last[prev] = True
                    nbatch += 1
                    cur = 0
                batch_of[k] = nbatch
                cur += cnt[k]
                prev = k
            if prev is not None:
                batch_last[prev] = True
                nbatch += 1
        cb = batch_of[chunk_group]
        key = cb * n + chunk_row
        order = np.argsort(key, kind="stable")
        rank = np.empty(len(key), dtype=np.int64)
        rank[order] = np.arange(len(key), dtype=np.int64)
        per_batch = np.bincount(cb, minlength=nbatch)
        first = np.zeros(nbatch + 1, dtype=np.int64)
        np.cumsum(per_batch, out=first[1:])
        chunk_out = rank - first[cb]
        rcp = np.zeros((nbatch, n + 1), dtype=np.int64)
        np.cumsum(np.bincount(key, minlength=nbatch * n).reshape(nbatch, n), axis=1, out=rcp[:, 1:])
        return chunk_out, rcp, batch_of, batch_last, int(per_batch.max()) if nbatch else 0

    def _sweep_tile_blocks(self, dim):
        """Blocks per other mode for the L2-tiled launch order (1 = untiled).  When the sample table of ONE slot is
        larger than the table budget (config 5: 2 x 10k rows x 16 KB = 320 MB against a 126 MB L2), the other modes'
        index ranges are cut into nb blocks each and a slot's chunks are launched tile by tile, so the records the
        resident CTAs gather stay inside one tile's share of the table."""
        if self.diag:
            return 1
        if self.sweep_tile_blocks:
            return int(self.sweep_tile_blocks)
        per_slot = int(self.lib.ssvi_sweep_table_bytes(self.h, dim, 1))
        others = [d for d in range(self.N) if d != dim]
        touched = sum(self.row_window[d][1] - self.row_window[d][0] + 1 for d in others)
        # the records this shard gathers: the table covers the row windows only (ssvi_set_row_windows); the tensor-core
        # table also carries the swept mode's own samples, which are read once per chunk, not gathered
        own = (self.row_window[dim][1] - self.row_window[dim][0] + 1) if self.table_dtype == "f32_tc" else 0
        per_slot = per_slot * touched // max(touched + own, 1)
        budget = max(self.table_budget_bytes, 1)
        if self.table_dtype == "f32_tc":
            # the tensor-core kernel waits on its gathers, not on a math pipe: it wants the tile well inside the L2
            # (config 5 at full size, 2 / 3 / 4 / 5 blocks: 0.83 / 0.86 / 0.89 / 0.82 of the roofline, profiles/r3e, r3f)
            budget = max(budget // 2, 1)
        nb = -(-per_slot // budget)
        rp = self.rowptr_h[dim]
        rows_here = self.row_window[dim][1] - self.row_window[dim][0] + 1
        per_pair = float(rp[-1]) / max(rows_here * int(self.noise.slots), 1)          # entries per (row, slot)
        # short chunks cost more than the L2 misses they avoid (measured on config 5: 8 entries per chunk loses 13 %,
        # 31 per chunk gains 5 % even with tiles of 1.3 x L2): take the finest tiling that keeps chunks long
        while nb > 1 and per_pair / nb ** (self.N - 1) < self.sweep_tile_min_entries:
            nb -= 1
        return max(nb, 1)

    def _plan_sweep_tiled(self, dim, slot_offset, nb):
        """plan_sweep with the chunks of a slot launched tile by tile: one chunk per (slot, tile, row[, segment]),
        tile = the block coordinates of the entry's other-mode rows.  Same records, same reduction (each row's
        records are summed in (slot, tile, segment) order)."""
        P, n, dev, N = int(self.noise.slots), self.dims[dim], self.device, self.N
        rp = self.rowptr_h[dim]
        lens = rp[1:] - rp[:-1]
        off = np.zeros(n, dtype=np.int64) if slot_offset is None else np.asarray(slot_offset, dtype=np.int64) % P
        T = nb ** (N - 1)
        lens_d = torch.as_tensor(lens).to(dev)
        row_of = torch.repeat_interleave(torch.arange(n, device=dev, dtype=torch.int64), lens_d)
        b = torch.arange(int(rp[-1]), device=dev, dtype=torch.int64) - torch.as_tensor(rp[:-1]).to(dev)[row_of]
        key = (b + torch.as_tensor(off).to(dev)[row_of]) % P                  # slot
        del b
        ent = self.perm[dim]
        tile = torch.zeros_like(key)
        for d in range(N):
            if d == dim:
                continue
            lo, hi = self.row_window[d]
            bs = -(-(hi - lo + 1) // nb)
            tile = tile * nb + (self.coords[ent, d].long() - lo) // bs
        key = key * T + tile
        del tile
        key, order = torch.sort(key, stable=True)                # (slot, tile, row, b): stable keeps (row, b)
        entry_ids = ent[order].contiguous()
        triple = key * n + row_of[order]                         # non-decreasing
        del key, order, row_of
        uniq, counts = torch.unique_consecutive(triple, return_counts=True)
        del triple
        uniq, flat_cnt = uniq.cpu().numpy(), counts.cpu().numpy().astype(np.int64)
        W = self.sweep_chunk_entries
        flat_seg = (flat_cnt + W - 1) // W
        npair = len(uniq)
        pair_start = np.zeros(npair + 1, dtype=np.int64)
        np.cumsum(flat_cnt, out=pair_start[1:])
        pair_chunk = np.zeros(npair + 1, dtype=np.int64)
        np.cumsum(flat_seg, out=pair_chunk[1:])
        total = int(pair_chunk[-1])
        pair_of = np.repeat(np.arange(npair, dtype=np.int64), flat_seg)
        seg = np.arange(total, dtype=np.int64) - pair_chunk[pair_of]
        c_cnt, c_nseg = flat_cnt[pair_of], flat_seg[pair_of]
        base, rem = c_cnt // c_nseg, c_cnt % c_nseg
        chunk_len = base + (seg < rem)
        chunk_start = pair_start[pair_of] + seg * base + np.minimum(seg, rem)
        c_key = uniq[pair_of]
        chunk_row = c_key % n
        chunk_slot = c_key // (n * T)
        per_slot = int(self.lib.ssvi_sweep_table_bytes(self.h, dim, 1))
        slot_chunk = np.searchsorted(chunk_slot, np.arange(P + 1, dtype=np.int64))
        groups = [(s0, 1, int(slot_chunk[s0]), int(slot_chunk[s0 + 1])) for s0 in range(P)]
        chunk_out, group_rcp, batch_of, batch_last, max_rec = self._group_records(chunk_slot, chunk_row, n, groups)
        t = lambda a, dt: torch.as_tensor(np.ascontiguousarray(a.astype(dt))).to(dev)      # noqa: E731
        plan = SlotSweepPlan(dim, torch.arange(n, device=dev, dtype=torch.int64), lens_d, entry_ids,
                             t(chunk_row, np.int32), t(chunk_start, np.int64), t(chunk_len, np.int32),
                             t(chunk_slot, np.int32), t(chunk_out, np.int64), t(group_rcp, np.int64), lens,
                             groups, per_slot)
        plan.chunk_start_h = np.concatenate([chunk_start, [int(pair_start[-1])]]).astype(np.int64)
        plan.batch_of, plan.batch_last, plan.max_batch_records = batch_of, batch_last, max_rec
        plan.tile_blocks = nb
        return plan

    def _accumulate_sweep(self, plan, st):
        """Slot groups alternate between `sweep_streams` side streams, each with its own table, record buffer and row
        accumulator, so the tail of one group's launch overlaps the next group's sampling and ramp-up.  Each group's
        chunk records are folded into its stream's accumulator right after the launch (in group order per stream);
        the streams' accumulators are added in stream order after the join.  Deterministic for a fixed plan."""
        ns = max(1, min(self.sweep_streams, len(plan.groups)))
        n, L = plan.n_rows, self.acc_len
        tabs = [self._buf("sweep_table%d" % i, (plan.table_bytes + 7) // 8) for i in range(ns)]
        parts = [self._buf("sweep_partial%d" % i, plan.max_batch_records * L) for i in range(ns)]
        accs = [self._buf("acc" if i == 0 else "sweep_acc%d" % i, n * L) for i in range(ns)]
        # simple / robust + full covariance: per-group work buffers (F, weights, si) of the three-pass path
        e_pos = plan.chunk_start_h
        max_n = max([int(e_pos[c1] - e_pos[c0]) for (_, _, c0, c1) in plan.groups] + [0])
        work_bytes = int(self.lib.ssvi_sweep_work_bytes(self.h, max_n))
        works = [self._buf("sweep_work%d" % i, (work_bytes + 7) // 8) for i in range(ns)] if work_bytes else None
        per_call = 1 if self.diag else (2 if self.cfg["model"] == "d" else 4)

        def at(tensor, i):
            return ctypes.c_void_p(tensor.data_ptr() + i * tensor.element_size())
        main = torch.cuda.current_stream(self.device)
        if ns > 1:
            if len(self._side_streams) < ns:
                self._side_streams = [torch.cuda.Stream(device=self.device) for _ in range(ns)]
            fork = torch.cuda.Event()
            fork.record(main)
            for s_ in self._side_streams[:ns]:
                s_.wait_event(fork)
        seen = [False] * ns
        for k, (s0, nsl, c0, c1) in enumerate(plan.groups):
            i = k % ns
            stream = ctypes.c_void_p(self._side_streams[i].cuda_stream) if ns > 1 else st
            self._call(self.lib.ssvi_sweep_accumulate, plan.dim, s0, nsl, c1 - c0, at(plan.chunk_row, c0),
                       at(plan.chunk_start, c0), at(plan.chunk_len, c0), at(plan.chunk_slot, c0),
                       at(plan.chunk_out, c0), _ptr(plan.entry_ids), _ptr(plan.rows), _ptr(self.w_state), _ptr(tabs[i]),
                       tabs[i].numel() * 8, _ptr(works[i]) if works else ctypes.c_void_p(0),
                       works[i].numel() * 8 if works else 0, int(e_pos[c0]), int(e_pos[c1] - e_pos[c0]),
                       _ptr(parts[i]), stream)
            self.launches += per_call if c1 > c0 else (0 if self.diag else 1)
            if plan.batch_last[k]:                           # fold the batch's records into this stream's accumulator
                self._call(self.lib.ssvi_reduce_chunks_into, n, at(plan.group_rcp, int(plan.batch_of[k]) * (n + 1)),
                           _ptr(parts[i]), _ptr(accs[i]), 1 if seen[i] else 0, stream)
                seen[i] = True
                self.launches += 1
        if ns > 1:
            for s_ in self._side_streams[:ns]:
                join = torch.cuda.Event()
                join.record(s_)
                main.wait_event(join)
        acc = accs[0][:n * L]
        for i in range(1, ns):
            if seen[i]:
                acc.add_(accs[i][:n * L])
        if not seen[0]:
            acc.zero_()
        return acc

    # -------------------------------------------------------------------------------------------
    def _buf(self, name, numel):
        t = self._scratch.get(name)
        if t is None or t.numel() < numel:
            t = torch.empty((max(numel, 1),), device=self.device, dtype=torch.float64)
            self._scratch[name] = t
        return t

    def accumulate(self, plan):
        """Stages 1-3 + in-order chunk reduction -> acc [n_rows, acc_len] (a view of a scratch buffer)."""
        assert self.noise is not None, "set_noise() first"
        st = self._stream()
        ev = None
        if self.profile_events is not None:
            ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
            ev[0].record()
        if isinstance(plan, SlotSweepPlan):
            acc = self._accumulate_sweep(plan, st)           # reduces group by group (records stay in L2)
            if ev is not None:
                ev[1].record()
                self.profile_events.append((ev[0], ev[1], plan.n_entries))
            return acc.view(plan.n_rows, self.acc_len)
        partial = self._buf("partial", plan.n_chunks * self.acc_len)
        acc = self._buf("acc", plan.n_rows * self.acc_len)
        self._call(self.lib.ssvi_accumulate, plan.dim, plan.n_chunks, _ptr(plan.chunk_row),
                   _ptr(plan.chunk_start), _ptr(plan.chunk_len), _ptr(plan.chunk_slot0), _ptr(plan.entry_ids),
                   _ptr(plan.rows), _ptr(self.w_state), _ptr(partial), st)
        self.launches += 1 if plan.n_chunks else 0
        if ev is not None:
            ev[1].record()
            self.profile_events.append((ev[0], ev[1], plan.n_entries))
        self._call(self.lib.ssvi_reduce_chunks, plan.n_rows, _ptr(plan.row_chunk_ptr), _ptr(partial), _ptr(acc), st)
        self.launches += 1 if plan.n_chunks else 0
        return acc[:plan.n_rows * self.acc_len].view(plan.n_rows, self.acc_len)

    def row_update(self, plan, acc, fixed_covariance=False, n_obs=None):
        """Stage 4 + the w_sigma recurrence for the plan's rows."""
        si = self._buf("si", plan.n_rows)
        st = self._stream()
        n_obs = plan.n_obs if n_obs is None else n_obs
        self._call(self.lib.ssvi_row_update, plan.dim, plan.n_rows, _ptr(plan.rows), _ptr(acc), _ptr(n_obs),
                   _ptr(self.pSigma), _ptr(self.w_state), int(self.time_step[plan.dim]), int(bool(fixed_covariance)),
                   _ptr(si), st)
        self.launches += 1 if plan.n_rows else 0
        if self.cfg["model"] != "d":
            self._call(self.lib.ssvi_sigma_update, plan.n_rows, _ptr(si), _ptr(acc), _ptr(self.w_state), st)
            self.launches += 1 if plan.n_rows else 0

    def step(self, plan, fixed_covariance=False):
        """update_natural_param_batch for every row of the plan, from one state snapshot."""
        acc = self.accumulate(plan)
        self.row_update(plan, acc, fixed_covariance)

    def hyper_reduce(self, dim, update_psigma=True):
        """update_hyper_parameter(dim) (:146-160): refreshes pSigma[dim]; also leaves the mode's maximum
        mean / covariance change in hyper_out[dim, 1:3] for check_stop_cond (:967-984), which calls this with
        update_psigma=False (the reference's check_stop_cond does not touch pSigma)."""
        self._call(self.lib.ssvi_hyper_reduce, dim, _ptr(self.hyper_out[dim]), self._stream())
        if update_psigma:
            self.pSigma[dim:dim + 1].copy_(self.hyper_out[dim, 0:1])
        self.launches += 1

    def estimate_entries(self, what, entries, ys=None, slot0=0):
        """Per-entry Monte-Carlo estimator over the bound noise pool (include/ssvi_b200.h: ssvi_estimate_entries).
        entries: int32 device tensor [n, N]; ys: float64 device tensor [n] or None."""
        assert self.noise is not None, "set_noise() first"
        n = int(entries.shape[0])
        out = torch.empty((n,), device=self.device, dtype=torch.float64)
        self._call(self.lib.ssvi_estimate_entries, int(what), _ptr(entries), _ptr(ys) if ys is not None else None, n,
                   int(slot0), _ptr(self.w_state), _ptr(out), self._stream())
        self.launches += 1 if n else 0
        return out

    def predict_mean(self, entries):
        """entries: int32 device tensor [n, N] -> posterior-mean prediction per entry."""
        n = int(entries.shape[0])
        out = torch.empty((n,), device=self.device, dtype=torch.float64)
        self._call(self.lib.ssvi_predict_mean, _ptr(entries), n, _ptr(out), self._stream())
        self.launches += 1 if n else 0
        return out
