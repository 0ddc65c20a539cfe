#!/usr/bin/env python
"""bench.py -- observed-entry SVI updates/s of the SSVI-TF hot path on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config c5] [--impl b200|reference]

Workload (every N): ONE fixed global problem -- BASELINE config 5, the configuration the metric's "at 1/2/4/8 B200" is
quoted on (real 10k^3, 1e9 observed entries, D = 32, SSVI_TF_d, full covariance) -- whose observed entries are sharded
over the N ranks along index-space blocks (strong scaling); per-row accumulators are all-reduced over NCCL.
A "step" is one pass of the hot path over the workload's train entries: a Jacobi sweep of ONE mode (stages 1-4 over every
observed train entry of every row of that mode, scale = 1, fresh Monte-Carlo variates) + the hyper-parameter reduction;
consecutive steps cycle through the modes.  One (entry, mode) contribution = one unit, so a step processes nnz_train
units.  At N = 1 the same process also measures the other BASELINE configs and the reference's own minibatch schedule
(`configs` in the JSON line).

Prints ONE JSON line (rank 0).  --impl reference times the reference's own CPU implementation of the path on all host
cores: the UNMODIFIED reference staged in baseline/_ref (tools/stage_reference.py; kind "reference") when present, else
the oracle port (oracle/ssvi_oracle.py, pinned to the reference by tests/golden; kind "port").
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "observed_entry_svi_updates_per_sec"
UNIT = "entry-updates/s"
REF_ROOT = os.path.join(ROOT, "baseline", "_ref")

# BASELINE.json configs.  nnz = observed entries, 80 % of which are TRAIN entries.
CONFIGS = {
    "c1": dict(dims=[50, 50, 50], nnz=125000, D=10, lik="normal", model="d", diag=False, k1=64, k2=128,
               desc="synthetic_test.py real 50x50x50 dense, D=10, normal, SSVI_TF_d full cov"),
    "c2": dict(dims=[1000, 1000, 1000], nnz=10_000_000, D=20, lik="bernoulli", model="d", diag=False, k1=64, k2=64,
               desc="binary 1000^3 at 1% density, D=20, bernoulli, SSVI_TF_d full cov"),
    "c3": dict(dims=[2000, 2000, 300], nnz=100_000_000, D=30, lik="poisson", model="robust", diag=False, k1=64,
               k2=64, cov_eta=0.001, desc="count 2000x2000x300, 1e8 nnz, D=30, poisson, SSVI_TF_robust full cov"),
    "c4": dict(dims=[500, 500, 500, 100], nnz=100_000_000, D=50, lik="normal", model="simple", diag=True, k1=64,
               k2=64, desc="real 500x500x500x100, 1e8 nnz, D=50, normal, SSVI_TF_simple diag cov"),
    "c5": dict(dims=[10000, 10000, 10000], nnz=1_000_000_000, D=32, lik="normal", model="d", diag=False, k1=64,
               k2=64, desc="real 10k^3, 1e9 nnz, D=32, normal, SSVI_TF_d full cov"),
    # diagnostic (not a BASELINE config): C5's rank with an index space whose one-slot sample table fits in L2
    "c5l2": dict(dims=[2500, 2500, 2500], nnz=50_000_000, D=32, lik="normal", model="d", diag=False, k1=64,
                 k2=64, desc="diagnostic: real 2500^3, 5e7 nnz, D=32, normal, SSVI_TF_d full cov"),
    # diagnostic (not a BASELINE config): C4's shape under SSVI_TF_d (the other diagonal-covariance sweep kernel)
    "c4d": dict(dims=[500, 500, 500, 100], nnz=100_000_000, D=50, lik="normal", model="d", diag=True, k1=64,
                k2=64, desc="diagnostic: real 500x500x500x100, 1e8 nnz, D=50, normal, SSVI_TF_d diag cov"),
}
DATATYPE = {"normal": "real", "bernoulli": "binary", "poisson": "count"}
FINE_BLOCKS = 8          # the global problem is generated as 8 index-space blocks, whatever N is (1, 2, 4 or 8)

# side measurements at N = 1 (`configs` in the JSON line): (key, config, scale of nnz, options)
SIDE_RUNS = [
    ("c1", "c1", 1.0, {}),
    ("c2", "c2", 1.0, {}),
    ("c2_reference_schedule_b128", "c2", 1.0, {"batch_size": 128}),
    ("c3", "c3", 0.25, {}),
    ("c3_quadrature", "c3", 0.25, {"quadrature": 30}),
    ("c4", "c4", 1.0, {}),
    # ranks above 32 with a full covariance (the reference's real-data default is rank 50, real_test.py:27): C2's tensor at
    # D = 50 through accumulate_wide.cuh, and the generic CUDA-core kernel it replaces there (on 5 % of the entries)
    ("c2_rank50", "c2", 1.0, {"rank": 50}),
    ("c2_rank50_generic_kernel", "c2", 0.05, {"rank": 50, "kernel_path": 1}),
    # (the float-TABLE option with fp64 arithmetic, SSVI_TABLE_F32, is slower than both -- DESIGN.md 3.5 -- and is
    # measured on request only: --only-configs c5_table_f32)
    # the fp32 ARITHMETIC option (SSVI_TABLE_F32_TC, tolerance 1e-4): V, f in fp32, contraction on the tcgen05 tensor cores
    ("c5_f32_tensor_core", "c5", 1.0, {"table_tc": True}),
    ("c2_f32_tensor_core", "c2", 1.0, {"table_tc": True}),
    ("c2_rank50_f32_tensor_core", "c2", 1.0, {"rank": 50, "table_tc": True}),
    # ... and on config 4 (diagonal covariance, closed-form weights): the fp32 instance of the entry-blocked kernel
    ("c4_f32", "c4", 1.0, {"table_tc": True}),
]


def release_device_memory():
    """Between workloads: a config-5-sized problem takes most of the 180 GB, and the caching allocator would keep it."""
    import gc
    import torch
    gc.collect()
    torch.cuda.synchronize()
    torch.cuda.empty_cache()


def algorithmic_bytes_per_unit(c, mean_row_len):
    """coords + value + gathered other-mode (mean, cov) rows + amortised own row."""
    N, D, w = len(c["dims"]), c["D"], 8
    r_row = (2 * D if c["diag"] else D + D * D) * w
    return 4 * N + w + (N - 1) * r_row + (2 * r_row + 2 * D * w) / max(mean_row_len, 1.0)


class ClockSampler(object):
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.rows, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
                for n, v in zip(names, r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


class DeviceTensorData(object):
    """Synthetic tensor of a BASELINE config, generated on the device: uniform distinct
    cells, ground-truth CP factors U_d ~ N(mu_d, 0.1 I) of rank D, values through the data link."""

    def __init__(self, c, device, seed, scale=1.0, window=None, keep_test=True):
        """window: [(lo, hi)] per mode -- a block of the index space (None: the whole space)."""
        import torch
        self.datatype = DATATYPE[c["lik"]]
        self.dims = list(c["dims"])
        N, D = len(self.dims), c["D"]
        nnz = max(int(c["nnz"] * scale), 1000)
        g = torch.Generator(device=device)
        g.manual_seed(seed)
        window = window or [(0, n) for n in self.dims]
        ext = [hi - lo for lo, hi in window]
        total = int(np.prod(ext))
        if nnz >= total:                      # dense (config 1): every cell, shuffled
            lin = torch.randperm(total, generator=g, device=device)
            nnz = total
        else:
            lin = torch.unique(torch.randint(0, total, (int(nnz * 1.03) + 64,), generator=g, device=device))
            lin = lin[torch.randperm(lin.numel(), generator=g, device=device)][:nnz]
            nnz = int(lin.numel())
        coords = torch.empty((nnz, N), dtype=torch.int32, device=device)
        rem = lin
        for d in range(N - 1, -1, -1):
            coords[:, d] = (rem % ext[d] + window[d][0]).to(torch.int32)
            rem = rem // ext[d]
        del lin, rem
        if c["lik"] == "bernoulli":
            mus = np.linspace(-1, 1, N)          # binary_tensor.py:12-19
        elif c["lik"] == "poisson":
            mus = np.linspace(0, 1.5, N)         # count_tensor.py:16-23
        else:
            mus = [0.5, 1.0, 0.2, 0.5][:N]
        gf = torch.Generator(device=device)              # ground-truth factors: the same on every rank / block
        gf.manual_seed(319 + N * 1000 + D)
        Us = [float(mus[d]) + math.sqrt(0.1) * torch.randn((self.dims[d], D), generator=gf, device=device)
              for d in range(N)]
        vals = torch.empty((nnz,), dtype=torch.float64, device=device)
        step = 1 << 24                                   # values in slabs: [nnz, D] at 1e9 entries would not fit
        for s0 in range(0, nnz, step):
            cc = coords[s0:s0 + step].long()
            f = Us[0][cc[:, 0]]
            for d in range(1, N):
                f = f * Us[d][cc[:, d]]
            f = f.sum(dim=1).double()
            if c["lik"] == "normal":
                v = f + 0.05 * f.abs() * torch.randn(f.shape, generator=g, device=device, dtype=torch.float64)
            elif c["lik"] == "bernoulli":
                u = torch.rand(f.shape, generator=g, device=device, dtype=torch.float64)
                v = torch.where(u < torch.sigmoid(f), 1.0, -1.0).double()
            else:
                lam = torch.clamp(torch.log1p(torch.exp(f)), min=1e-8)
                v = torch.clamp(torch.poisson(lam, generator=g), max=170.0)      # y! finite in fp64
            vals[s0:s0 + step] = v
        del Us
        n_train = int(nnz * 0.8)
        self.train_entries, self.train_vals = coords[:n_train].contiguous(), vals[:n_train].contiguous()
        if keep_test:
            self.test_entries = coords[n_train:].cpu().numpy().astype(np.int64)
            self.test_vals = vals[n_train:].cpu().numpy()
        else:
            self.test_entries, self.test_vals = np.zeros((0, N), np.int64), np.zeros((0,))
        self.max_count = int(vals.max().item()) if c["lik"] == "poisson" else 0
        self.min_count = int(vals.min().item()) if c["lik"] == "poisson" else 0

    @classmethod
    def from_blocks(cls, c, device, cfg_seed, scale, rank, world):
        """This rank's shard of the ONE global problem: the global tensor is the union of FINE_BLOCKS index-space
        blocks, block b drawn from seed (cfg_seed + 1000 b) at nnz * scale / FINE_BLOCKS entries, so its entries are
        the same whatever the rank count; a rank holds the blocks inside its window (shard_window)."""
        import torch
        blocks = owned_fine_blocks(c["dims"], rank, world)
        parts = [cls(c, device, cfg_seed + 1000 * b, scale / FINE_BLOCKS, window=win, keep_test=False)
                 for b, win in blocks]
        self = parts[0]
        if len(parts) > 1:
            self.train_entries = torch.cat([p.train_entries for p in parts])
            self.train_vals = torch.cat([p.train_vals for p in parts])
            self.max_count = max(p.max_count for p in parts)
            self.min_count = min(p.min_count for p in parts)
        return self


def shard_window(dims, rank, world):
    """Index-space block of `rank` when the observed entries are sharded along a grid of blocks (world = 2^k ranks:
    the modes are halved in turn, 8 ranks -> 2 x 2 x 2).  Every rank then gathers only its blocks' rows of the other
    modes and keeps whole (row, slot) chunks of the rows it owns."""
    grid = [1] * len(dims)
    w, d = world, 0
    while w > 1:
        assert w % 2 == 0, "block sharding needs a power-of-two world size"
        grid[d % len(dims)] *= 2
        w //= 2
        d += 1
    win, r = [], rank
    for d_, n in enumerate(dims):
        b = r % grid[d_]
        r //= grid[d_]
        step = -(-n // grid[d_])
        win.append((b * step, min(n, (b + 1) * step)))
    return win, grid


def owned_fine_blocks(dims, rank, world):
    """[(block id, window)] of the FINE_BLOCKS-grid blocks that lie inside `rank`'s window at `world` ranks."""
    assert FINE_BLOCKS % world == 0, "the global problem is cut into %d blocks: N must divide it" % FINE_BLOCKS
    mine, _ = shard_window(dims, rank, world)
    out = []
    for b in range(FINE_BLOCKS):
        win, _ = shard_window(dims, b, FINE_BLOCKS)
        if all(lo >= mlo and hi <= mhi for (lo, hi), (mlo, mhi) in zip(win, mine)):
            out.append((b, win))
    assert len(out) == FINE_BLOCKS // world
    return out


def build_model(c, data, device, group, chunk_entries, noise_slots, engine_kwargs=None, quadrature=None):
    """The reference's constructor call (synthetic_test.py:139) on a device-resident tensor."""
    import torch
    import ssvi_tf_b200 as pkg
    cls = {"d": pkg.SSVI_TF_d, "simple": pkg.SSVI_TF_simple, "robust": pkg.SSVI_TF_robust}[c["model"]]
    D = c["D"]
    mean0 = np.ones(D) if c["lik"] == "normal" else np.zeros(D)       # synthetic_test.py:25-35
    np.random.seed(319)
    extra = dict(quadrature=True) if quadrature else {}
    return cls(data, D, diag=c["diag"], mean0=mean0, cov0=np.eye(D) * 0.1, k1=c["k1"], k2=c["k2"], batch_size=128,
               cov_eta=c.get("cov_eta", 1), randstart=True, device=device, noise_dtype=torch.float32,
               noise_slots=noise_slots, chunk_entries=chunk_entries, process_group=group, engine_kwargs=engine_kwargs,
               **extra)


def fp64_pipe_stats(c, units_per_s):
    """Issued DMMA work of the sample-table path per unit: NP*16 contraction DMMAs per entry
    (+ the once-per-chunk own-row sample, ignored), 512 FLOP each, against the measured FP64-pipe peak."""
    if c["diag"] or c["model"] != "d" or c["D"] > 32 or c["k1"] != 64:
        return None
    nt = (c["D"] + 7) // 8
    dmma = nt * (nt + 1) // 2 * 2 * (c["k1"] // 8)
    peak = 37.1                           # TFLOP/s, profiles/r01_microbench_fp64_pipes.txt (mma.sync m8n8k4 f64)
    ach = units_per_s * dmma * 512 / 1e12
    return {"dmma_per_unit": dmma, "achieved_tflops": ach, "peak_tflops": peak, "frac": ach / peak,
            "note": "issued DMMA FLOP only; the other FP64 instructions of a unit add ~30 % pipe time"}


def hbm_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------------------------
# One measured workload: model + the two timed arms
# ------------------------------------------------------------------------------------------------
class Workload(object):
    def __init__(self, args, cfg_id, c, device, rank, world, group, scale=1.0, batch_size=None, quadrature=None,
                 sharded=False, table_f32=False, factor_rank=None, kernel_path=0, table_tc=False):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        if factor_rank:
            c = dict(c, D=int(factor_rank), desc=c["desc"].replace("D=%d" % c["D"], "D=%d" % int(factor_rank)))
        self.args, self.cfg_id, self.c = args, cfg_id, c
        self.device, self.rank, self.world, self.group = device, rank, world, group
        self.B = batch_size
        t0 = time.time()
        seed = 319 + int(cfg_id[1])
        if sharded:
            data = DeviceTensorData.from_blocks(c, device, seed, scale, rank, world)
            self.grid = shard_window(c["dims"], 0, world)[1]
        else:
            data = DeviceTensorData(c, device, seed, scale, keep_test=False)
            self.grid = None
        torch.cuda.synchronize()
        self.t_data = time.time() - t0
        t0 = time.time()
        self.model = build_model(c, data, device, group if sharded else None, args.chunk_entries, args.noise_slots,
                                 dict(table_budget_bytes=int(args.table_mb * (1 << 20)),
                                      sweep_chunk_entries=args.sweep_chunk, sweep_streams=args.sweep_streams,
                                      sweep_tile_blocks=args.tile_blocks, tuning=args.tuning,
                                      table_dtype="f32_tc" if table_tc else (torch.float32 if table_f32 else torch.float64),
                                      kernel_path=kernel_path),
                                 quadrature=quadrature)
        self.eng = self.model.engine
        self.arith32 = bool(table_tc)                                  # the fp32 arithmetic option was asked for
        self.table_tc = bool(table_tc) and self.eng.sweep_supported and not c["diag"] and c["D"] <= 64 and c["model"] == "d"
        self.N = self.model.order
        n = torch.tensor([self.eng.nnz], device=device, dtype=torch.int64)
        if sharded and world > 1 and group is not None:                # (group None: --shard-of, one rank's shard alone)
            dist.all_reduce(n)
        self.nnz_global = int(n.item())
        if batch_size is None:
            self.units = [self.nnz_global] * self.N                    # units of a sweep of mode d (all ranks)
        else:
            assert not sharded or world == 1, "subsampled sweeps are single-shard"
            self.units = [int(np.minimum(np.diff(self.eng.rowptr_h[d]), batch_size).sum()) for d in range(self.N)]
        self.model.fresh_noise()
        if batch_size is None:
            for d in range(self.N):
                self.model.sweep_plan(d)                               # plans are built once, outside the timed regions
        torch.cuda.synchronize()
        self.t_plan = time.time() - t0
        in_bytes = self.eng.coords.numel() * 4 + self.eng.vals.numel() * 8 + 8 * self.eng.nnz
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=device) if in_bytes < 256 << 20 else None
        self.l2_note = ("L2 flushed between steps (256 MB memset; a step's inputs are %.0f MB)" % (in_bytes / 1e6)) \
            if self.flush is not None else "a step's inputs are larger than L2 (%.0f MB per GPU)" % (in_bytes / 1e6)
        self.k = 0                                                     # steps done so far: step k sweeps mode k % N

    def barrier(self):
        if self.world > 1 and self.group is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def step(self):
        """One step: fresh variates (drawn inside model.sweep), one Jacobi sweep of mode k % N, its pSigma refresh."""
        if self.flush is not None:
            self.flush.zero_()                # inputs smaller than L2: evict them between steps
        dim = self.k % self.N
        self.model.sweep(dim, self.B)
        self.model.update_hyper_parameter(dim)
        self.k += 1
        return dim

    def max_over_ranks(self, ms):
        t = self.torch.tensor([ms], device=self.device, dtype=self.torch.float64)
        if self.world > 1 and self.group is not None:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def run_resident(self, steps, warmup, sample_clocks=False):
        torch, eng = self.torch, self.eng
        for _ in range(warmup):
            self.step()
        eng.check_flags()
        self.barrier()
        sampler = ClockSampler(torch.cuda.current_device()) if sample_clocks and self.rank == 0 else None
        if sampler:
            sampler.start()
        eng.profile_events = []
        launches0 = eng.launches
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.barrier()
        e0.record()
        units = 0
        for _ in range(steps):
            units += self.units[self.step()]
        e1.record()
        self.barrier()
        ms = self.max_over_ranks(e0.elapsed_time(e1))
        launches = eng.launches - launches0
        events, eng.profile_events = eng.profile_events, None
        clocks = sampler.stop() if sampler else None
        eng.check_flags()
        value = units / (ms * 1e-3)
        # roofline of the dominant kernel (stages 1-3 = Engine.accumulate), CUDA events inside the timed region
        acc_ms = [a.elapsed_time(b) for a, b, _ in events]
        acc_units = [n for _, _, n in events]                           # this rank's entries per launch
        c = self.c
        mean_row = float(np.mean(acc_units)) / (float(sum(c["dims"])) / self.N)   # entries per updated row, this rank
        bpu = algorithmic_bytes_per_unit(c, mean_row)
        peak, peak_src = hbm_peak()
        achieved = sum(acc_units) * bpu / (sum(acc_ms) * 1e-3) / 1e9
        roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                    "traffic": None, "kernel": self.dominant_kernel(), "peak_source": peak_src,
                    "algorithmic_bytes_per_unit": bpu, "kernel_ms_per_launch": float(np.mean(acc_ms)),
                    "units_per_launch": float(np.mean(acc_units)),
                    "kernel_share_of_step": sum(acc_ms) / (e0.elapsed_time(e1)),
                    "launch": "one call of Engine.accumulate = stages 1-3 of one mode sweep on this rank (CUDA events "
                              "around it; a sweep issues one sample-table + one accumulate launch per slot group on "
                              "side streams, profiles/*_launches_*)",
                    "fp64_pipe": fp64_pipe_stats(c, sum(acc_units) / (sum(acc_ms) * 1e-3)) if self.B is None and not self.table_tc else None,
                    "note": ("fp32 arithmetic option (1e-4): contraction on the tcgen05 tensor cores, the table gathers "
                             "(L2 -> SM) pace the kernel; DESIGN.md 3.8") if self.table_tc else
                            "fp64 contract: the FP64 pipe (37 TFLOP/s measured), not HBM, bounds this kernel; DESIGN.md 3.1"}
        tr_path = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tr_path):          # DRAM bytes per unit from the committed ncu capture of this kernel
            tr = json.load(open(tr_path)).get((self.cfg_id + ("_tc" if self.table_tc else "")) if self.B is None
                                              else self.cfg_id + "_b128")
            if tr:
                roofline["traffic"] = tr["dram_bytes_per_unit"] * float(np.mean(acc_units))
                roofline["traffic_source"] = tr["source"]
        return dict(value=value, ms_per_step=ms / steps, launches=launches, clocks=clocks, roofline=roofline)

    def dominant_kernel(self):
        c = self.c
        if self.B is not None or not self.eng.sweep_supported:
            return "accumulate_mma_kernel (direct evaluation, one warp per row-chunk)" if not c["diag"] and c["D"] <= 32 \
                else "accumulate_generic_kernel"
        if c["diag"]:
            if self.arith32 and c["model"] == "simple" and c["lik"] == "normal":
                return "accumulate_diag_lin32_kernel (fp32 arithmetic option)"
            return "accumulate_diag_lin_kernel" if c["model"] == "simple" and c["lik"] == "normal" else \
                ("accumulate_diag_d_kernel" if c["model"] == "d" else "accumulate_diag_kernel")
        if self.table_tc:
            return "accumulate_tc_kernel (tcgen05 contraction; + sample_table_kernel)" if c["D"] <= 32 else \
                "accumulate_tc_wide_kernel (tcgen05 contraction, two warps per entry; + sample_table_wide_kernel)"
        if c["D"] > 32:
            return "accumulate_wide_kernel (+ sample_table_wide_kernel)"
        return "accumulate_memo_kernel (+ sample_table_kernel)" if c["model"] == "d" else \
            "memo_weights_kernel (+ memo_f / memo_c / sample_table kernels)"

    # ---- end-to-end arm: host buffers in, host result out, every step --------------------------------------------
    # Every step's inputs (COO entries, values, the swept mode's entry list in the order the sweep kernels read it, the
    # noise pool) are copied from pinned host memory and the swept mode's posterior is copied back.  Two device input
    # sets: the copy of step k+1 runs on a copy stream while step k computes (all K copies are inside the timed
    # region).  The entry lists arrive PRE-PLANNED (slot-major order built once by Engine.plan_sweep, outside the timed
    # region), and the pool is one pinned pool re-sent every step (the resident arm draws fresh variates on the device).
    def run_e2e(self, steps):
        torch, eng, model, N = self.torch, self.eng, self.model, self.N
        from ssvi_tf_b200.engine import NoisePool, _ptr
        pin = lambda x: x.cpu().pin_memory()                                      # noqa: E731
        sweep = self.B is None
        plans = [model.sweep_plan(d)[0] for d in range(N)] if sweep else [None] * N
        lists0 = [pl.entry_ids for pl in plans] if sweep else list(eng.perm)      # per-mode entry lists the kernels read
        nz = eng.noise
        noise_names = [n for n in ("z_other", "z_own", "r_w", "z_f") if getattr(nz, n) is not None]
        noise0 = [getattr(nz, n) for n in noise_names]
        host = dict(coords=pin(eng.coords), vals=pin(eng.vals), lists=[pin(x) for x in lists0],
                    noise=[pin(x) for x in noise0])
        sets = [dict(coords=eng.coords if s == 0 else torch.empty_like(eng.coords),
                     vals=eng.vals if s == 0 else torch.empty_like(eng.vals),
                     lst=torch.empty_like(lists0[0]), noise=[torch.empty_like(x) for x in noise0]) for s in range(2)]
        h_out = [torch.empty(p.shape, dtype=p.dtype).pin_memory() for p in eng.params]
        nb = lambda x: x.numel() * x.element_size()                               # noqa: E731
        h2d = [nb(host["coords"]) + nb(host["vals"]) + nb(host["lists"][d]) + sum(nb(x) for x in host["noise"])
               for d in range(N)]
        d2h = [nb(h) for h in h_out]
        main_stream = torch.cuda.current_stream()
        copy_stream = torch.cuda.Stream()
        ready = [torch.cuda.Event(), torch.cuda.Event()]
        free = [torch.cuda.Event(), torch.cuda.Event()]
        k0 = self.k
        model.keep_noise = True

        def stage(k):                       # H2D of step k's inputs into set k % 2
            ds, dim = sets[k % 2], (k0 + k) % N
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(free[k % 2])
                ds["coords"].copy_(host["coords"], non_blocking=True)
                ds["vals"].copy_(host["vals"], non_blocking=True)
                ds["lst"][:host["lists"][dim].numel()].copy_(host["lists"][dim], non_blocking=True)
                for dn, hn in zip(ds["noise"], host["noise"]):
                    dn.copy_(hn, non_blocking=True)
                ready[k % 2].record(copy_stream)

        def bind(k):                        # point the library at set k % 2 (host-side pointer swaps only)
            ds, dim = sets[k % 2], (k0 + k) % N
            eng.coords, eng.vals = ds["coords"], ds["vals"]
            eng._call(eng.lib.ssvi_bind_data, _ptr(eng.coords), _ptr(eng.vals), eng.nnz)
            if sweep:
                plans[dim].entry_ids = ds["lst"]
            else:
                eng.perm[dim] = ds["lst"]
            kw = dict(zip(noise_names, ds["noise"]))
            eng.set_noise(NoisePool(kw.get("z_other"), kw.get("z_own"), kw.get("r_w"), kw.get("z_f")))

        def run(n_steps):
            units = 0
            for e in free:
                e.record(main_stream)
            stage(0)
            for k in range(n_steps):
                if k + 1 < n_steps:
                    stage(k + 1)
                main_stream.wait_event(ready[k % 2])
                bind(k)
                dim = self.step()
                units += self.units[dim]
                free[k % 2].record(main_stream)
                h_out[dim].copy_(eng.params[dim], non_blocking=True)
            return units

        nmax = max(x.numel() for x in lists0)
        for ds in sets:
            if ds["lst"].numel() < nmax:
                ds["lst"] = torch.empty((nmax,), dtype=lists0[0].dtype, device=self.device)
        run(min(2, steps))
        k0 = self.k
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        units = run(steps)
        e1.record()
        self.barrier()
        # restore the resident bindings
        eng.coords, eng.vals = sets[0]["coords"], sets[0]["vals"]
        eng._call(eng.lib.ssvi_bind_data, _ptr(eng.coords), _ptr(eng.vals), eng.nnz)
        for d in range(N):
            if sweep:
                plans[d].entry_ids = lists0[d]
            else:
                eng.perm[d] = lists0[d]
        model.fresh_noise()
        ms = self.max_over_ranks(e0.elapsed_time(e1))
        eng.check_flags()
        return {"value": units / (ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(np.mean(h2d)) * self.world,
                "d2h_bytes_per_step": int(np.mean(d2h)),
                "note": "per step, from pinned host memory: COO coords + values, the swept mode's entry list "
                        "(pre-planned slot-major order, built once outside the timed region), the noise pool; back: the "
                        "swept mode's posterior.  Bytes summed over ranks."}

    def close(self):
        self.eng.close()
        self.model = self.eng = None
        self.torch.cuda.empty_cache()


# ------------------------------------------------------------------------------------------------
# CPU baseline: the reference itself (baseline/_ref, under the SURVEY 8c shims) or the oracle port, on host cores
# ------------------------------------------------------------------------------------------------
def reference_kind():
    return "reference" if os.path.isdir(os.path.join(REF_ROOT, "SSVI")) else "port"


def _cpu_problem(c, rng, rows_per_mode=64):
    """Bounded sample of the workload for the CPU arms: same D / k1 / k2 / model / likelihood / order, index space shrunk
    to <= 64 per mode (the per-update cost does not depend on the tensor size), random distinct cells, values drawn as
    bench's generator does in spirit."""
    N = len(c["dims"])
    dims = [min(n, rows_per_mode) for n in c["dims"]]
    nnz = min(int(np.prod(dims)) // 2, 320 * max(dims))             # every row holds > 128 entries: minibatches of 128
    flat = rng.choice(int(np.prod(dims)), size=nnz, replace=False)
    coords = np.stack(np.unravel_index(flat, dims), axis=1).astype(np.int64)
    if c["lik"] == "normal":
        vals = rng.normal(0.5, 1.0, nnz)
    elif c["lik"] == "bernoulli":
        vals = rng.choice([-1.0, 1.0], nnz)
    else:
        vals = rng.poisson(2.0, nnz).astype(float)
    return dims, coords, vals


def _cpu_worker(args):
    c, seed, budget_s, max_rows, kind = args
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    rng = np.random.default_rng(seed)
    N, D, B = len(c["dims"]), c["D"], 128
    units, rows_done = 0, 0
    if kind == "reference":
        # the UNMODIFIED reference (baseline/_ref) through its own update_natural_param_batch, stock numpy RNG
        os.environ["SSVI_REFERENCE_ROOT"] = REF_ROOT
        from oracle import ref_shim
        np.random.seed(seed)
        dims, coords, vals = _cpu_problem(c, rng)
        obj = ref_shim.make_reference_model(c["model"], DATATYPE[c["lik"]], dims, coords, vals, D, c["diag"], c["k1"],
                                            c["k2"], B, cov_eta=c.get("cov_eta", 1))
        for d in range(N):                                          # start from a spread-out posterior, cov 0.1 I
            p = obj.posterior.params[d]
            p[:, 0, :] = rng.normal(1.0 if c["lik"] == "normal" else 0.0, 0.3, p[:, 0, :].shape)
            if c["diag"]:
                p[:, 1, :] = 0.1
            else:
                p[:, 1:, :] = np.eye(D) * 0.1
        t0 = time.perf_counter()
        while time.perf_counter() - t0 < budget_s and rows_done < max_rows:
            dim = rows_done % N
            row = (rows_done // N) % dims[dim]
            n_here = len(obj.tensor.find_observed_ui(dim, row))
            obj.update_natural_param_batch(dim, row)
            units += min(n_here, B)
            rows_done += 1
        return units, time.perf_counter() - t0
    from oracle import ssvi_oracle as so
    dims = [min(n, 64) for n in c["dims"]]
    cfg = so.make_config(c["model"], c["lik"], D, diag=c["diag"], k1=c["k1"], k2=c["k2"], batch_size=B,
                         cov_eta=c.get("cov_eta", 1.0))
    st = so.new_state(cfg, dims, rng=rng, mean0=np.ones(D) if c["lik"] == "normal" else np.zeros(D),
                      cov0=np.eye(D) * 0.1)
    t0 = time.perf_counter()
    while time.perf_counter() - t0 < budget_s and rows_done < max_rows:
        dim = rows_done % N
        row = (rows_done // N) % dims[dim]
        coords = rng.integers(0, dims, size=(B, N))
        coords[:, dim] = row
        if c["lik"] == "normal":
            ys = rng.normal(0.5, 1.0, B)
        elif c["lik"] == "bernoulli":
            ys = rng.choice([-1.0, 1.0], B)
        else:
            ys = rng.poisson(2.0, B).astype(float)
        so.row_step(st, cfg, dim, row, coords, ys, so.make_tape(rng, cfg, N, B))
        units += B
        rows_done += 1
    return units, time.perf_counter() - t0


def cpu_baseline(c, cores, budget_s, max_rows=10 ** 9, kind=None):
    """units/s of the CPU implementation over `cores` independent processes on disjoint synthetic problems (the
    reference has no intra-process parallelism; this mirrors its one-job-per-core Slurm usage)."""
    kind = kind or reference_kind()
    if cores == 1:
        res = [_cpu_worker((c, 1, budget_s, max_rows, kind))]
    else:
        import multiprocessing as mp
        with mp.get_context("spawn").Pool(cores) as pool:
            res = pool.map(_cpu_worker, [(c, 100 + i, budget_s, max_rows, kind) for i in range(cores)])
    units = sum(r[0] for r in res)
    wall = max(r[1] for r in res)
    return units / wall, units, wall


CPU_SAMPLE = ("%d row updates of B=128 entries: same D/k1/k2/model/likelihood/order as the workload, index space shrunk "
              "to <=64 per mode with random cells (per-unit cost does not depend on the tensor size), %s, %.1f s x %d "
              "process(es)")
CPU_IMPL = {"reference": "the UNMODIFIED reference's update_natural_param_batch (baseline/_ref, SURVEY 8c shims, stock "
                         "numpy RNG)",
            "port": "oracle/ssvi_oracle.py row_step (numpy port pinned to the reference by tests/golden)"}


def workload_config(args, c, world):
    """The `config` object of the JSON line: a function of the command line only, identical in both arms."""
    return {"workload": c["desc"] + (" at %.3g of its nnz" % args.scale if args.scale != 1.0 else "")
            + "; ONE global problem, observed entries sharded over the ranks along index-space blocks",
            "config_id": args.config, "nnz_observed_global": int(c["nnz"] * args.scale),
            "step": "one Jacobi sweep of one mode over every train entry (scale = 1) + its pSigma refresh; steps cycle "
                    "through the modes; fresh variates every step",
            "noise": "f32 pool, %d slots, redrawn on the device every step" % args.noise_slots,
            "entries_per_slot": "a row's entries share its %d slots: ~%.0f entries per (row, slot) at this size"
                                % (args.noise_slots, 0.8 * c["nnz"] * args.scale / (c["dims"][0] * args.noise_slots)),
            "reference_arm": "bounded sample: B=128 minibatch row updates on an index space shrunk to <=64 per mode "
                             "(same D/k1/k2/model/likelihood/order)"}


def run_reference_arm(args, c, rank, world):
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    kind = reference_kind()
    per_step = max(2.0, min(20.0, 90.0 / max(args.steps + args.warmup, 1)))
    for _ in range(args.warmup):
        cpu_baseline(c, cores, min(per_step, 2.0), kind=kind)
    tot_units, tot_wall = 0, 0.0
    for _ in range(args.steps):
        _, u, w = cpu_baseline(c, cores, per_step, kind=kind)
        tot_units += u
        tot_wall += w
    v = tot_units / tot_wall
    sample = CPU_SAMPLE % (tot_units // 128, CPU_IMPL[kind], per_step, cores)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * tot_wall / max(args.steps, 1), "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, c, world),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


# ------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--config", default="c5", choices=sorted(CONFIGS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--scale", type=float, default=1.0, help="fraction of the config's nnz")
    ap.add_argument("--chunk-entries", type=int, default=16)
    ap.add_argument("--noise-slots", type=int, default=128)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--table-mb", type=float, default=96.0, help="sample-table budget per call (L2-resident)")
    ap.add_argument("--sweep-chunk", type=int, default=64, help="max entries of one (row, slot) chunk")
    ap.add_argument("--sweep-streams", type=int, default=3, help="side streams the slot groups alternate between")
    ap.add_argument("--tile-blocks", type=int, default=0,
                    help="L2 tiling of the sweep launch order: blocks per other mode (0 = from the table budget, 1 = off)")
    ap.add_argument("--batch-size", type=int, default=0,
                    help="0: full-row sweeps (headline).  B > 0: reference-style subsampled sweeps, every row takes at "
                         "most B entries drawn afresh each sweep (SSVI_TF_interface.py:170-174), scale = len / B (N = 1)")
    ap.add_argument("--tuning", type=int, default=0, help="SSVI_TUNE_* bits for the headline run (include/ssvi_b200.h)")
    ap.add_argument("--ab-tunings", default="", help="comma-separated SSVI_TUNE_* values timed after the headline, same "
                                                     "process and plans (details.ab_tunings); diagnostic")
    ap.add_argument("--table-f32", action="store_true", help="headline with the sample table stored as floats (fp32 option)")
    ap.add_argument("--table-tc", action="store_true", help="headline with the fp32 arithmetic option: tcgen05 contraction (SSVI_TABLE_F32_TC)")
    ap.add_argument("--tc-scale-log2", type=int, default=0)
    ap.add_argument("--shard-of", type=int, default=0, help="diagnostic on ONE GPU: time rank 0's shard of the global problem as "
                    "it is cut for this many ranks (no collectives); for ncu captures of the per-rank kernel at N > 1")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="diagnostic runs only")
    ap.add_argument("--no-configs", action="store_true", help="skip the side measurements of the other configs (N = 1)")
    ap.add_argument("--only-configs", default="", help="comma-separated keys of SIDE_RUNS to measure (default: all)")
    args = ap.parse_args()
    c = CONFIGS[args.config]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return run_reference_arm(args, c, rank, world)

    import torch
    import torch.distributed as dist
    assert args.warmup >= 3 or args.scale != 1.0, "timing rules: W >= 3"
    t_start = time.time()
    torch.cuda.set_device(local)
    device = "cuda:%d" % local
    group, parity = None, "n/a (one rank)"
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device(device))
        group = dist.group.WORLD
        # in-run parity self-check of the multi-GPU path (ssvi_tf_b200/dist.py:selfcheck): raises on any mismatch
        from ssvi_tf_b200 import dist as sdist
        parity = sdist.selfcheck(group, device)

    shard_world = world
    if args.shard_of > 1:                                              # diagnostic: rank 0's shard of an N-rank job, alone
        assert world == 1, "--shard-of is a one-GPU diagnostic"
        shard_world = args.shard_of
    wl = Workload(args, args.config, c, device, rank, shard_world, group, scale=args.scale,
                  batch_size=args.batch_size or None, sharded=True, table_f32=args.table_f32, table_tc=args.table_tc)
    res = wl.run_resident(args.steps, args.warmup, sample_clocks=True)
    ab = {}
    for t in [int(x) for x in args.ab_tunings.split(",") if x]:
        wl.eng.set_tuning(t)
        r = wl.run_resident(max(wl.N, 3), wl.N)
        ab[str(t)] = {"value": r["value"], "ms_per_step": r["ms_per_step"], "roofline_frac": r["roofline"]["frac"]}
    wl.eng.set_tuning(args.tuning)
    e2e = None if args.no_e2e else wl.run_e2e(args.steps)
    eng = wl.eng
    plan0 = wl.model.sweep_plan(0)[0] if (eng.sweep_supported and wl.B is None) else None
    details = {"nnz_train_global": wl.nnz_global, "nnz_train_this_rank": eng.nnz,
               "units_per_step": wl.units[0], "chunk_entries": args.chunk_entries,
               "path": ("slot-major sweep kernels, %d slot(s)/call, %d L2 tile blocks per other mode, %d chunks per mode sweep"
                        % (plan0.groups[0][1], getattr(plan0, "tile_blocks", 1), plan0.n_chunks))
               if plan0 is not None else "direct (one warp per row-chunk)",
               "l2": wl.l2_note,
               "parallelism": "entries sharded x%d along index-space blocks %s; all-reduce of the row accumulators, stage 4 "
                              "on row blocks + all-gather" % (world, "x".join(map(str, wl.grid))) if world > 1 else "one GPU",
               "tuning": args.tuning, "ab_tunings": ab or None,
               "shard_of": ("DIAGNOSTIC: rank 0's shard of the problem as cut for %d ranks, alone on one GPU (no collectives); "
                            "value counts this shard's entries only" % args.shard_of) if args.shard_of > 1 else None,
               "setup_s": {"data": round(wl.t_data, 1), "model_and_plans": round(wl.t_plan, 1)}}
    wl.close()
    del wl
    release_device_memory()

    # ---- the other BASELINE configs and the reference's own schedule, same process, same method (N = 1 only) ----------
    side = None
    if world == 1 and not args.no_configs:
        side = {}
        only = [k for k in args.only_configs.split(",") if k]
        for key, cid, scale, opt in SIDE_RUNS + [("c5_table_f32", "c5", 1.0, {"table_f32": True})]:
            if (only and key not in only) or (not only and key == "c5_table_f32"):
                continue
            release_device_memory()
            cc = CONFIGS[cid]
            Nm = len(cc["dims"])
            t0 = time.time()
            try:
                w2 = Workload(args, cid, cc, device, 0, 1, None, scale=scale, batch_size=opt.get("batch_size"),
                              quadrature=opt.get("quadrature"), table_f32=opt.get("table_f32", False), table_tc=opt.get("table_tc", False),
                              factor_rank=opt.get("rank"), kernel_path=opt.get("kernel_path", 0))
                cc = w2.c
                r2 = w2.run_resident(2 * Nm, Nm)
                side[key] = {"workload": cc["desc"] + (" at %.3g of its nnz" % scale if scale != 1.0 else ""),
                             "schedule": ("reference minibatches: every row takes <= %d entries drawn on the device each "
                                          "sweep, scale = len/B" % opt["batch_size"]) if opt.get("batch_size") else
                             "full-row Jacobi sweeps (scale = 1)" + (", 30-node Gauss-Hermite second layer "
                                                                    "(SSVI_TF_robust quadrature=True)" if opt.get("quadrature") else "")
                             + (", sample table stored as float32 (SSVI_TABLE_F32), arithmetic fp64" if opt.get("table_f32") else "")
                             + (", fp32 arithmetic option (SSVI_TABLE_F32_TC): samples, f, weights and per-group sums in fp32 "
                                "(accumulate_diag32.cuh), chunk accumulators and stages 3-4 fp64" if opt.get("table_tc") and cc["diag"] else
                                ", fp32 arithmetic option (SSVI_TABLE_F32_TC): float table, V / f in fp32, contraction on the "
                                "tcgen05 tensor cores (fp16 hi/lo split, fp32 accumulation in tensor memory), likelihood and "
                                "stages 3-4 fp64" if opt.get("table_tc") else ""),
                             "dtype": ("f32 (tcgen05 contraction)" if not cc["diag"] else "f32") if opt.get("table_tc") else "f64",
                             "nnz_train": w2.eng.nnz, "steps": 2 * Nm, "warmup": Nm, "value": r2["value"], "unit": UNIT,
                             "ms_per_step": r2["ms_per_step"], "roofline_frac": r2["roofline"]["frac"],
                             "algorithmic_bytes_per_unit": r2["roofline"]["algorithmic_bytes_per_unit"],
                             "kernel": r2["roofline"]["kernel"], "kernel_share_of_step": r2["roofline"]["kernel_share_of_step"],
                             "gpu_launches": r2["launches"], "wall_s": round(time.time() - t0, 1)}
                w2.close()
                del w2
            except Exception as e:                                     # a side run must not take the headline down
                side[key] = {"error": "%s: %s" % (type(e).__name__, str(e)[:300])}

    if rank == 0:
        cpu = None
        if not args.no_cpu_baseline and world == 1:
            kind = reference_kind()
            v, u, w = cpu_baseline(c, 1, args.cpu_seconds, kind=kind)
            cpu = {"value": v, "unit": UNIT, "cores": 1, "kind": kind,
                   "sample": CPU_SAMPLE % (u // 128, CPU_IMPL[kind], w, 1)}
        print(json.dumps({
            "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": res["ms_per_step"], "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32 (tcgen05 contraction; likelihood and stage 4 fp64)" if args.table_tc else ("f64" if not args.table_f32 else "f64 (f32 sample table)"),
            "data": "synthetic",
            "config": workload_config(args, c, world), "details": details, "multi_gpu_parity": parity,
            "roofline": res["roofline"], "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": res["launches"],
            "clocks": res["clocks"], "configs": side, "wall_s": round(time.time() - t_start, 1)}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
