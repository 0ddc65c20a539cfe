"""SSVI_TF: the reference's factorizer surface (SSVI/SSVI_TF_interface.py) over the B200 engine.

Same constructor kwargs, same attribute names, same factorize() loop; the per-row stochastic
variational update (update_natural_param_batch, :163-219) runs as hand-written sm_100a kernels behind
the C ABI of include/ssvi_b200.h.  New, batched entry points: update_natural_param_rows(dim, rows)
(any set of rows of one mode from one state snapshot -- the equivalence class of DESIGN.md) and
sweep(dim) (every observed entry of every row, scale = 1).

There is no CPU fallback: constructing a model without a CUDA device raises.
"""
import time

import numpy as np
import torch

from ..engine import Engine, NoisePool
from ..Model.TF_Models import Posterior_Diag_Covariance, Posterior_Full_Covariance
from ..Probability import ProbFun as probs
from ..Probability.bernoulli import BernoulliDistribution
from ..Probability.normal import NormalDistribution
from ..Probability.poisson import PoissonDistribution
from .true_model import TrueModelMixin

EST_EXPECTED_OBS, EST_LOG_PDF, EST_PDF, EST_ROBUST_VLB = 0, 1, 2, 3      # include/ssvi_b200.h: SSVI_EST_*


class _RowBlock(object):
    """The rows of a sweep plan whose stage 4 this rank runs (Engine.row_update reads dim, rows, n_rows, n_obs)."""

    def __init__(self, dim, rows):
        self.dim, self.rows, self.n_rows, self.n_obs = dim, rows, int(rows.numel()), None


class _DeviceMirror(np.ndarray):
    """Host copy of a small device vector whose item assignment also writes the device tensor (the reference keeps
    pSigma as a numpy array that callers may modify in place, SSVI_TF_interface.py:56)."""

    def __new__(cls, dev):
        obj = np.asarray(dev.cpu().numpy()).view(cls)
        obj._dev = dev
        return obj

    def __array_finalize__(self, obj):
        self._dev = None                   # copies, slices and results of arithmetic are plain host arrays

    def __setitem__(self, key, value):
        np.ndarray.__setitem__(self, key, value)
        if self._dev is not None and self.shape == tuple(self._dev.shape):
            self._dev.copy_(torch.as_tensor(np.asarray(self, dtype=np.float64)))


class SSVI_TF(TrueModelMixin):
    _model = "d"          # "d" | "simple" | "robust"; set by the subclasses

    def __init__(self, tensor, rank, mean_update="S", cov_update="N", noise_update="N",
                 diag=False, mean0=None, cov0=None, sigma0=1,
                 unstable_cov=False, k1=64, k2=64, batch_size=128,
                 eta=1, cov_eta=1, sigma_eta=1, randstart=True,
                 device=None, noise_dtype=torch.float32, noise_slots=None, noise_seed=319,
                 chunk_entries=16, process_group=None, engine_kwargs=None):
        self.tensor = tensor
        self.dims = list(tensor.dims)
        self.datatype = tensor.datatype
        self.order = len(tensor.dims)
        self.randstart = randstart
        self.D = rank
        self.mean_update = mean_update
        self.cov_update = cov_update
        self.noise_update = noise_update
        self.diag = diag
        if mean_update not in ("S", "N") or cov_update not in ("S", "N"):
            raise Exception("Unidentified update formula for covariance param")     # :612, :632
        cls = Posterior_Diag_Covariance if diag else Posterior_Full_Covariance
        self.posterior = cls(self.dims, self.D, mean0, cov0, randstart)             # :37-41

        self.mean0 = mean0
        self.cov0 = cov0
        self.w_sigma0 = sigma0
        self.unstable_cov = unstable_cov
        self.cov_epsilon = 1e-7
        self.min_eigenval = 0.01
        self.k1 = k1
        self.k2 = k2
        self.batch_size = batch_size
        self.predict_num_samples = 32
        self.pmu = np.ones((self.D,))

        if self.datatype == "real":                                                # :59-77
            self.link_fun = lambda m: m
            self.likelihood_type = "normal"
            self.likelihood_param = 1
            self.likelihood = NormalDistribution()
        elif self.datatype == "binary":
            self.link_fun = lambda m: probs.sigmoid(m)
            self.likelihood_type = "bernoulli"
            self.likelihood_param = 0
            self.likelihood = BernoulliDistribution()
        elif self.datatype == "count":
            self.likelihood_type = "poisson"
            self.max_count = tensor.max_count
            self.min_count = tensor.min_count
            self.link_fun = lambda m: probs.poisson_link(m)
            self.likelihood_param = 0
            self.likelihood = PoissonDistribution()
        else:
            raise Exception("unsupported datatype " + str(self.datatype))

        self.epsilon = 0.0001
        self.eta = eta
        self.cov_eta = cov_eta
        self.poisson_eta = 0.1
        self.sigma_eta = sigma_eta
        self.noise_added = self._model != "d"
        self.d_mean = 1.
        self.d_cov = 1.
        self.w_changes = 0.
        self.missing = {}
        self.fixed_covariance = False
        self.rows_per_iteration = 1
        self.process_group = process_group
        self.shard_row_update = True       # multi-GPU full sweeps of SSVI_TF_d: stage 4 on row blocks + all-gather
        self.device_loop = True            # factorize(): enqueue the reference's schedule from C (ssvi_factorize_iterations)

        cfg = dict(model=self._model, likelihood=self.likelihood_type, D=rank, diag=diag, k1=k1, k2=k2,
                   batch_size=batch_size, mean_update=mean_update, cov_update=cov_update, eta=eta,
                   cov_eta=cov_eta, sigma_eta=sigma_eta, poisson_eta=self.poisson_eta,
                   min_eigenval=self.min_eigenval, epsilon=self.epsilon,
                   likelihood_param=float(self.likelihood_param))
        if device is None:
            device = "cuda:%d" % torch.cuda.current_device() if torch.cuda.is_available() else "cuda"
        self.engine = Engine(cfg, self.dims, tensor.train_entries, tensor.train_vals, device=device,
                             noise_dtype=noise_dtype, chunk_entries=chunk_entries, **(engine_kwargs or {}))
        eng = self.engine
        for d, p in enumerate(self.posterior.initial_values()):
            eng.params[d].copy_(torch.as_tensor(p))
        self.posterior.attach(eng.params, self._posterior_written)
        eng.refresh_sampler()
        eng.w_state.copy_(torch.tensor([1.0, 0.0, 1.0], dtype=torch.float64))       # SSVI_TF_simple.py:15-17
        self.norm_changes = eng.norm_changes                                        # :91  (n_d, 2) device
        self.grad_norms = eng.grad_norms                                            # :95
        self._noise_slots = int(noise_slots) if noise_slots else max(int(batch_size), 1)
        self._noise_gen_seed = int(noise_seed)
        self._noise_draws = 0
        self._own_pool = None
        self.keep_noise = False            # True: the bound pool is reused by every update (set by set_noise())
        self._minibatch_draws = 0
        self._np_rng = np.random.default_rng(noise_seed)
        self._sub_gen = torch.Generator(device=eng.device)
        self._sub_gen.manual_seed(int(noise_seed) + 7919)
        self._n_obs_global = {}

    # ---------------------------------------------------------------------------------------------
    # state the reference keeps as numpy arrays, exposed with the reference's names and shapes
    # ---------------------------------------------------------------------------------------------
    @property
    def ada_acc_grad(self):            # list of (D, n_d) views (:83)
        return [a.t() for a in self.engine.ada_mean]

    @property
    def ada_acc_grad_cov(self):        # (:41)
        return [a.t() for a in self.engine.ada_cov]

    @property
    def pSigma(self):                  # (:56); item assignment (model.pSigma[d] = v) writes through to the device
        return _DeviceMirror(self.engine.pSigma)

    @pSigma.setter
    def pSigma(self, v):
        self.engine.pSigma.copy_(torch.as_tensor(np.asarray(v, dtype=np.float64)))

    @property
    def time_step(self):               # (:79)
        return self.engine.time_step

    @time_step.setter
    def time_step(self, v):
        self.engine.time_step = [int(t) for t in v]

    def _w(self, i):
        return float(self.engine.w_state[i].item())

    w_sigma = property(lambda self: self._w(0), lambda self, v: self.engine.w_state[0:1].fill_(float(v)))
    w_ada_grad = property(lambda self: self._w(1), lambda self, v: self.engine.w_state[1:2].fill_(float(v)))
    w_tau = property(lambda self: self._w(2), lambda self, v: self.engine.w_state[2:3].fill_(float(v)))

    def _posterior_written(self, dim, row):
        rows = None if row is None else torch.tensor([int(row)], device=self.engine.device, dtype=torch.int64)
        self.engine.refresh_sampler(dim, rows)

    # ---------------------------------------------------------------------------------------------
    # noise
    # ---------------------------------------------------------------------------------------------
    def set_noise(self, pool):
        """Supply the Monte-Carlo variates explicitly (a NoisePool, e.g. NoisePool.from_tape).  The pool then stays
        bound -- every later update reads the same variates -- until fresh_noise() is called or keep_noise is
        cleared (parity tests, replays of a noise tape)."""
        self.engine.set_noise(pool)
        self.keep_noise = True

    def fresh_noise(self):
        """Draw a new pool of standard variates on the device (ssvi_fill_noise; the stand-in for the reference's
        global numpy RNG, which draws afresh in every update_natural_param_batch call): noise_slots slots
        (default batch_size, so every entry of a row's minibatch has its own draws).  Called by every update unless
        keep_noise is set; one kernel launch, nothing crosses the bus."""
        self._noise_draws += 1
        e = self.engine
        if self._own_pool is None:
            self._own_pool = NoisePool.empty(self._noise_slots, self.order, self.D, self.k1, self.k2,
                                             self.noise_added, e.device, e.noise_dtype, diag=self.diag)
        e.fill_noise(self._own_pool, self._noise_gen_seed, self._noise_draws)
        if e.noise is not self._own_pool:
            e.set_noise(self._own_pool)
        self.keep_noise = False

    # ---------------------------------------------------------------------------------------------
    # the hot path
    # ---------------------------------------------------------------------------------------------
    def _shard_info(self, dim, rows_h, lens):
        """Multi-GPU: (slot_offset, n_obs_global) of the listed rows; single GPU: (None, None)."""
        if self.process_group is None:
            return None, None
        from .. import dist as sdist
        key = (dim, rows_h.tobytes(), lens.tobytes())
        if key not in self._n_obs_global:
            local = torch.as_tensor(lens).to(self.engine.device)
            off, tot = sdist.row_slot_offsets(local, self.process_group)
            self._n_obs_global[key] = (off.cpu().numpy(), tot)
        return self._n_obs_global[key]

    def _run_plan(self, plan, n_obs_global=None, full_sweep=False):
        eng = self.engine
        if eng.noise is None or not self.keep_noise:
            self.fresh_noise()                                      # the reference draws afresh in every call
        if getattr(self, "_hyper_fresh", None):
            self._hyper_fresh[plan.dim] = False                     # check_stop_cond must re-reduce this mode
        acc = eng.accumulate(plan)
        if self.process_group is not None:
            from .. import dist as sdist
            sdist.allreduce_accumulators(acc, self.process_group)   # the exchange step: (di, Di, si, count)
            # row-block stage 4 only for sweep plans: their rows are arange(n), so block [lo, hi) of the plan IS rows
            # lo..hi-1 of the state arrays that allgather_rows exchanges
            own = sdist.owned_rows(plan.n_rows, self.process_group) if full_sweep else None
            if own is not None and self._model == "d" and self.shard_row_update:
                # full sweep: stage 4 on this rank's block of rows, then the updated rows of every state array
                lo, hi = own
                sub = _RowBlock(plan.dim, plan.rows[lo:hi])
                eng.row_update(sub, acc[lo:hi], self.fixed_covariance,
                               n_obs=(plan.n_obs if n_obs_global is None else n_obs_global)[lo:hi])
                d = plan.dim
                state = [eng.params[d], eng.sampler[d], eng.ada_mean[d], eng.norm_changes[d], eng.grad_norms[d]]
                if self.diag:
                    state.append(eng.ada_cov[d])
                sdist.allgather_rows(state, lo, hi, self.process_group)
                return
        eng.row_update(plan, acc, self.fixed_covariance, n_obs=n_obs_global)

    def update_natural_param_rows(self, dim, rows, noise=None, sub_idx=None):
        """update_natural_param_batch (:163-219) for every row in `rows` of mode `dim`, evaluated from one
        state snapshot (pSigma, time_step, w_sigma frozen; the w_sigma recurrence is replayed in row
        order afterwards).  Rows without observations are recorded in self.missing (:166-168)."""
        if noise is not None:
            self.set_noise(noise)
        rows = np.atleast_1d(np.asarray(rows, dtype=np.int64))
        rp = self.engine.rowptr_h[dim]
        full = rp[rows + 1] - rp[rows]
        B = int(self.batch_size)
        if self.process_group is None:
            total = full
        else:
            # sharded: this rank's share of every row's minibatch is proportional to its share of the row's entries
            # (largest-remainder rounding in rank order), so the global minibatch holds min(len, batch_size) entries
            # as the reference's does (:170-174)
            from .. import dist as sdist
            _, n_obs_t = self._shard_info(dim, rows, full)
            total = n_obs_t.cpu().numpy() if torch.is_tensor(n_obs_t) else np.asarray(n_obs_t)
            B = sdist.local_batch_sizes(full, self.batch_size, self.process_group)
        for r in rows[total == 0]:
            self.missing[(dim, int(r))] = True                      # :166-168 (no observation on any shard)
        lens = np.minimum(full, B)
        off, _ = self._shard_info(dim, rows, lens)
        _, n_obs = self._shard_info(dim, rows, full)
        self._minibatch_draws += 1                      # the draw itself happens on the device (:170-174)
        plan = self.engine.plan_minibatch(dim, rows, B, rng=self._np_rng, sub_idx=sub_idx,
                                          slot_offset=off,
                                          device_seed=self._noise_gen_seed * 0x9E3779B1 + self._minibatch_draws)
        self._run_plan(plan, n_obs)
        return plan

    def update_natural_param_batch(self, dim, i, noise=None, sub_idx=None):
        """The reference's per-row entry point (:163): the one-row case of update_natural_param_rows."""
        sub = None if sub_idx is None else {int(i): sub_idx}
        w_before = self.w_sigma if self.noise_added else 0.0
        self.update_natural_param_rows(dim, [i], noise=noise, sub_idx=sub)
        if self.noise_added:
            self.w_changes = abs(self.w_sigma - w_before)                            # :209

    def sweep_plan(self, dim):
        """(plan, n_obs) of a full sweep over mode `dim`: every observed entry of every row (scale = 1).
        SSVI_TF_d + full covariance and every diagonal-covariance model take the slot-major sweep kernels; the rest
        the direct kernels."""
        eng = self.engine
        rp = eng.rowptr_h[dim]
        rows = np.arange(self.dims[dim], dtype=np.int64)
        off, n_obs = self._shard_info(dim, rows, rp[1:] - rp[:-1])
        if eng.noise is None:
            self.fresh_noise()
        if not eng.sweep_supported and not getattr(self, "_warned_direct_sweep", False):
            self._warned_direct_sweep = True
            import warnings
            warnings.warn("sweep(): this configuration is outside the slot-major sweep kernels (full covariance needs "
                          "k1 == 64 and rank <= 32, or rank <= 64 for SSVI_TF_d; diagonal covariance rank <= 64 and k1 a "
                          "multiple of 16 up to 64; order <= 4); the direct / generic kernels run instead, several times slower (DESIGN.md 3)")
        plan = eng.plan_sweep(dim, off) if eng.sweep_supported else eng.plan_full(dim, None, off)
        return plan, n_obs

    def sweep(self, dim, batch_size=None):
        """One Jacobi sweep over mode `dim`: every row updated from one state snapshot.  batch_size=None takes every
        observed entry of every row (scale = 1); with a batch size each row takes the reference's minibatch
        (:170-174: at most batch_size entries, drawn without replacement, scale = len / batch_size), all rows
        subsampled in one device pass."""
        if batch_size is None:
            plan, n_obs = self.sweep_plan(dim)
        else:
            assert self.process_group is None, "subsampled sweeps are single-shard"
            if self.engine.noise is None:
                self.fresh_noise()
            plan, n_obs = self.engine.plan_subsampled_sweep(dim, batch_size, self._sub_gen), None
        self._run_plan(plan, n_obs, full_sweep=True)
        return plan

    def update_hyper_parameter(self, dim):
        """pSigma[dim] = sum_rows (tr S + m.m) / (n_dim D)  (:146-160), as a device reduction."""
        self.engine.hyper_reduce(dim)

    def check_stop_cond(self):
        """Largest per-row mean / covariance change over all modes (:967-984)."""
        eng = self.engine
        for d in range(self.order):
            if not getattr(self, "_hyper_fresh", {}).get(d, False):
                eng.hyper_reduce(d, update_psigma=False)
        mx = eng.hyper_out[:, 1:3].max(dim=0).values.cpu().numpy()
        self.d_mean, self.d_cov = float(mx[0]), float(mx[1])
        return self.d_mean, self.d_cov

    # ---------------------------------------------------------------------------------------------
    # factorize loop (:101-143)
    # ---------------------------------------------------------------------------------------------
    def factorize(self, report=100, max_iteration=2000, fixed_covariance=False, to_report=[], detailed_report=True,
                  output_folder="output", rows_per_iteration=None):
        """The reference's loop: per iteration and per mode, update the next row(s) in round-robin order,
        refresh pSigma, advance the pointers (time_step += 1 on wrap-around), test the stop condition.
        rows_per_iteration=1 (default) is the reference's schedule; "all" makes every iteration a full
        sweep of every mode."""
        self.report = report
        self.header_printed = False
        self.to_report = to_report
        self.max_iteration = max_iteration
        self.fixed_covariance = fixed_covariance
        self.detailed_report = detailed_report
        self.missing = {}
        self.max_changes = [[0, 0] for _ in self.dims]
        self.output_folder = output_folder
        rpi = self.rows_per_iteration if rows_per_iteration is None else rows_per_iteration
        print("Factorizing with max_iteration =", max_iteration, " fixing covariance?: ", fixed_covariance)

        update_column_pointer = [0 for _ in range(self.order)]
        start = time.time()
        if (rpi == 1 and self.process_group is None and not self.keep_noise and int(self.batch_size) <= 2048
                and self.device_loop):
            return self._factorize_device_loop(start)
        for iteration in range(self.max_iteration + 1):
            self._hyper_fresh = {}
            advance = []
            for dim in range(self.order):
                n = self.dims[dim]
                step = n if rpi == "all" else min(int(rpi), n)
                col = update_column_pointer[dim]
                rows = np.arange(col, min(col + step, n))
                self.update_natural_param_rows(dim, rows)                            # :124
                self.update_hyper_parameter(dim)                                     # :125
                self._hyper_fresh[dim] = True
                advance.append(len(rows))
            for dim in range(self.order):                                            # :128-133
                nxt = update_column_pointer[dim] + advance[dim]
                if nxt >= self.dims[dim]:
                    self.engine.time_step[dim] += 1
                    nxt = 0
                update_column_pointer[dim] = nxt
            mean_change, cov_change = self.check_stop_cond()                         # :135
            if (iteration != 0 and (iteration % self.report == 0)) or iteration in to_report:
                self.engine.check_flags()
                if self.detailed_report:
                    self.report_metrics(iteration, start, mean_change, cov_change)
                else:
                    self.report_metrics_mini(iteration, start, mean_change, cov_change, output_folder)
            if max(mean_change, cov_change) < self.epsilon:                          # :142
                break
        self._hyper_fresh = {}             # a later manual update + check_stop_cond() must reduce again
        self.engine.check_flags()

    def _factorize_device_loop(self, start):
        """The reference's schedule (one row per mode per iteration) with the loop body enqueued from C
        (ssvi_factorize_iterations): no per-row host work, no per-iteration read of the stop condition.  Blocks end at
        the iterations where a report is due; the device latches the stop condition and freezes the state at the
        iteration the reference would `break` at (:142), so the result equals the per-call loop's."""
        eng = self.engine
        total = self.max_iteration + 1
        if self._own_pool is None:
            self.fresh_noise()
            self._noise_draws -= 1                       # the pool only had to exist; the loop draws from here on
        lp = eng.loop_begin(self._own_pool, self.batch_size, self.fixed_covariance, self.epsilon, self._noise_gen_seed,
                            self._noise_draws, self._noise_gen_seed * 0x9E3779B1, self._minibatch_draws,
                            [0] * self.order)
        ts0 = list(eng.time_step)
        due = lambda i: (i != 0 and i % self.report == 0) or i in self.to_report          # noqa: E731
        it, done = 0, total
        while it < total:
            end = it
            while end < total - 1 and not due(end) and end - it < 255:
                end += 1
            eng.factorize_iterations(lp, it, end - it + 1)
            fired, at = eng.loop_stop_state()
            last = at if fired else end
            if due(last):
                mean_change, cov_change = self._loop_changes()
                eng.check_flags()
                if self.detailed_report:
                    self.report_metrics(last, start, mean_change, cov_change)
                else:
                    self.report_metrics_mini(last, start, mean_change, cov_change, self.output_folder)
            if fired:
                done = at + 1
                break
            it = end + 1
        # pointers / time steps after `done` iterations (:128-133); the counters keep what the loop consumed
        eng.time_step = [ts0[d] + done // self.dims[d] for d in range(self.order)]
        self._noise_draws, self._minibatch_draws = int(lp.noise_draws), int(lp.minibatch_draws)
        for d in range(self.order):                      # :166-168: visited rows without observations
            rp = eng.rowptr_h[d]
            for r in np.nonzero(rp[1:min(done, self.dims[d]) + 1] == rp[:min(done, self.dims[d])])[0]:
                self.missing[(d, int(r))] = True
        self.d_mean, self.d_cov = self._loop_changes()
        self._hyper_fresh = {}
        eng.check_flags()

    def _loop_changes(self):
        mx = self.engine.hyper_out[:, 1:3].max(dim=0).values.cpu().numpy()
        return float(mx[0]), float(mx[1])

    # ---------------------------------------------------------------------------------------------
    # prediction / evaluation (:987-1059)
    # ---------------------------------------------------------------------------------------------
    def _missing_mask(self, entries):
        keep = np.ones(len(entries), dtype=bool)
        if self.missing:
            for (dim, col) in self.missing:
                keep &= entries[:, dim] != col
        return keep

    def predict_entries(self, entries):
        """predict_entry (:1037-1059) for an (n, N) array of coordinates, evaluated on the device."""
        entries = np.asarray(entries, dtype=np.int64).reshape(-1, self.order)
        dev = torch.as_tensor(entries.astype(np.int32)).to(self.engine.device)
        if self.likelihood_type == "normal":
            return self.engine.predict_mean(dev).cpu().numpy()
        if self.likelihood_type == "bernoulli" and not self.noise_added:
            f = self.engine.predict_mean(dev).cpu().numpy()
            return np.where(f > 0, 1, -1)
        res = self._expected_observation_via_sampling(dev).cpu().numpy()
        if self.likelihood_type == "bernoulli":
            return np.where(res >= 0.5, 1, -1)
        return np.rint(res)

    def predict_entry(self, entry):
        return self.predict_entries(np.asarray(entry)[None, :])[0]

    def _expected_observation_via_sampling(self, entries_dev):
        """estimate_expected_observation_via_sampling (:1168-1189) for a batch of entries on the device
        (ssvi_estimate_entries): draws of every factor row from the noise pool, Hadamard product, (noise layer,)
        link, mean.  Uses the pool's k1 / k2 draws where the reference uses predict_num_samples."""
        if self.engine.noise is None:
            self.fresh_noise()
        return self.engine.estimate_entries(EST_EXPECTED_OBS, entries_dev.contiguous())

    # ---------------------------------------------------------------------------------------------
    # VLB / KL / NLL estimators of the detailed report (:874-964), SURVEY.md 8f rank 4
    # ---------------------------------------------------------------------------------------------
    def _estimate(self, what, entries, vals):
        entries = np.asarray(entries, dtype=np.int64).reshape(-1, self.order)
        if self.engine.noise is None:
            self.fresh_noise()
        dev = torch.as_tensor(entries.astype(np.int32)).to(self.engine.device)
        ys = torch.as_tensor(np.asarray(vals, dtype=np.float64).reshape(-1)).to(self.engine.device)
        return self.engine.estimate_entries(what, dev, ys)

    def estimate_expectation_term_vlb(self, entry, val):
        """(:889-908; SSVI_TF_robust.py:234-262 for the robust model) for one entry."""
        what = EST_ROBUST_VLB if self._model == "robust" else EST_LOG_PDF
        return float(self._estimate(what, np.asarray(entry)[None, :], [val]).item())

    def estimate_KL_divergence(self):
        """(:911-936).  The reference's loop overwrites KL instead of summing (:923-925): the value is the divergence
        of the LAST row of the LAST mode from N(pmu, pSigma[dim] I) -- kept as is."""
        dim = self.order - 1
        m, S = self.posterior.get_vector_distribution(dim, self.dims[dim] - 1)
        m, S = np.asarray(m, dtype=np.float64), np.asarray(S, dtype=np.float64)
        S0 = np.diag(S) if self.diag else S
        sigma = float(self.pSigma[dim])
        dm = self.pmu - m
        return float(0.5 * (np.trace(S0) / sigma + dm @ dm / sigma - self.D
                            + np.log(sigma ** self.D / np.linalg.det(S0))))

    def estimate_vlb(self, entries, values):
        """(:874-883): (sum over entries of the expectation term, KL)."""
        what = EST_ROBUST_VLB if self._model == "robust" else EST_LOG_PDF
        return float(self._estimate(what, entries, values).sum().item()), self.estimate_KL_divergence()

    def estimate_negative_log_likelihood(self, entries, vals):
        """(:939-964): -sum_e log mean_draws pdf(y_e, fs)."""
        return float(-torch.log(self._estimate(EST_PDF, entries, vals)).sum().item())

    def evaluate_error(self, entries, vals):
        """(rsme, error) over entries not touching an unlearned row (:997-1030)."""
        entries = np.asarray(entries, dtype=np.int64).reshape(-1, self.order)
        vals = np.asarray(vals, dtype=np.float64)
        keep = self._missing_mask(entries)
        num = int(keep.sum())
        if num == 0:
            return float("nan"), float("nan")
        pred = self.predict_entries(entries[keep])
        diff = pred - vals[keep]
        rsme = float(np.sqrt(np.sum(np.square(diff)) / num))
        if self.likelihood_type == "normal":
            error = float(np.sum(np.abs(diff)) / num)
        elif self.likelihood_type == "bernoulli":
            error = float(np.sum(pred != vals[keep]) / num)
        else:
            error = 0.0        # the reference compares self.likelihood (an object) with "poisson" (:1023): never true
        return rsme, error

    def evaluate_train_error(self):
        return self.evaluate_error(self.tensor.train_entries, self.tensor.train_vals)

    def evaluate_test_error(self):
        return self.evaluate_error(self.tensor.test_entries, self.tensor.test_vals)

    # ---------------------------------------------------------------------------------------------
    # reports (:764-871)
    # ---------------------------------------------------------------------------------------------
    def _print_header(self, detailed):
        self.header_printed = True
        print("Tensor dimensions: ", self.dims)
        print("Mean update scheme: ", self.mean_update)
        print("Covariance update : ", self.cov_update)
        print("Using diagonal covariance?: ", self.diag)
        print("Factorization rank: ", self.D)
        print("k1 samples = ", self.k1, " k2 samples = ", self.k2)
        print("eta = ", self.eta, " cov eta = ", self.cov_eta, " sigma eta = ", self.sigma_eta)
        if detailed:
            print("iteration |   time   | test_mae |rel-te-err|train_mae |rel-tr-err|   d_mean  |   d_cov  |avg_grad_m|avg_grad_c|"
                  " test_nll | train_nll |    VLB   |  E-term  |    KL    " + ("|   dw   |" if self.noise_added else ""))
        else:
            print("iteration |   time   |  test_err  |  test_mae  |  trainerr  |  trainmae |")

    def _header_due(self, iteration):
        return not self.header_printed and (iteration == self.report or
                                            (len(self.to_report) > 0 and iteration == self.to_report[0]))

    def report_metrics(self, iteration, start, mean_change, cov_change):
        if self._header_due(iteration):
            self._print_header(True)
        mae_train, error_train = self.evaluate_train_error()
        mae_test, error_test = self.evaluate_test_error()
        avg = [g.mean(dim=0).cpu().numpy() for g in self.grad_norms]                 # :797-805
        avg_grad_m = float(np.mean([a[0] for a in avg]))
        avg_grad_c = float(np.mean([a[1] for a in avg]))
        dec = 4
        cols = [iteration, np.around(time.time() - start, 2), np.around(mae_test, dec), np.around(error_test, dec),
                np.around(mae_train, dec), np.around(error_train, dec), np.around(mean_change, dec),
                np.around(cov_change, dec), np.around(avg_grad_m, dec), np.around(avg_grad_c, dec)]
        E_term, KL = self.estimate_vlb(self.tensor.train_entries, self.tensor.train_vals)       # :793
        test_nll = self.estimate_negative_log_likelihood(self.tensor.test_entries, self.tensor.test_vals)
        train_nll = self.estimate_negative_log_likelihood(self.tensor.train_entries, self.tensor.train_vals)
        cols += [np.around(test_nll, 2), np.around(train_nll, 2), np.around(E_term - KL), np.around(E_term, 3),
                 np.around(KL, 3)]                                                             # :821-826
        if self.noise_added:
            cols.append(np.around(self.w_changes, dec))
        print(" ".join("{:^10}".format(c) for c in cols))

    def report_metrics_mini(self, iteration, start, mean_change, cov_change, output_folder):
        if self._header_due(iteration):
            self._print_header(False)
        mae_train, error_train = self.evaluate_train_error()
        mae_test, error_test = self.evaluate_test_error()
        dec = 4
        print("{:^10} {:^10} {:^10} {:^10} {:^10} {:^10}".format(
            iteration, np.around(time.time() - start, 2), np.around(error_test, dec), np.around(mae_test, dec),
            np.around(error_train, dec), np.around(mae_train, dec)))                 # :856-858

    # ---------------------------------------------------------------------------------------------
    def reset(self):
        """reset() (:1195-1210), quirk Q9 included: ada_acc_grad restarts at ZEROS (the constructor uses
        ones), ada_acc_grad_cov is not reset."""
        eng = self.engine
        self.posterior.reset()
        eng.pSigma.fill_(1.0)
        eng.time_step = [0 for _ in range(self.order)]
        for d in range(self.order):
            eng.ada_mean[d].zero_()
            eng.norm_changes[d].zero_()
        if self.noise_added:
            eng.w_state.copy_(torch.tensor([1.0, 0.0, 1.0], dtype=torch.float64))
