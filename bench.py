#!/usr/bin/env python
"""bench.py -- observed-entry SVI updates/s of the SSVI-TF hot path on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config c2] [--impl b200|reference]

A "step" is one full pass of the hot path over the workload: for every mode, stages 1-4 over every
observed TRAIN entry of every row (a Jacobi sweep, scale = 1) + the hyper-parameter reduction.  One
(entry, mode) contribution = one unit, so a step processes N * nnz_train units per GPU.
N > 1 (torchrun): observed entries are sharded by rank (weak scaling: every rank holds its own
nnz_train entries of the same index space), per-row accumulators are all-reduced over NCCL.

Prints ONE JSON line (rank 0).  --impl reference times the CPU oracle port (the reference is pure
Python and cannot travel to the GPU box; the oracle is its numpy restatement, pinned to the reference
by tests/golden) on all host cores.
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

# BASELINE.json configs (sizes per GPU).  nnz = observed entries, 80 % of which are TRAIN entries.
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

    def __init__(self, c, device, seed, scale=1.0, window=None):
        """window: [(lo, hi)] per mode -- this shard's block of the index space (None: the whole space)."""
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
        gf = torch.Generator(device=device)              # ground-truth factors: the same on every rank
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
        self.test_entries = coords[n_train:].cpu().numpy().astype(np.int64)
        self.test_vals = vals[n_train:].cpu().numpy()
        self.max_count = int(vals.max().item()) if c["lik"] == "poisson" else 0
        self.min_count = int(vals.min().item()) if c["lik"] == "poisson" else 0


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


def build_model(c, data, device, group, chunk_entries, noise_slots, engine_kwargs=None):
    import torch
    import ssvi_tf_b200 as pkg
    cls = {"d": pkg.SSVI_TF_d, "simple": pkg.SSVI_TF_simple, "robust": pkg.SSVI_TF_robust}[c["model"]]
    D = c["D"]
    mean0 = np.ones(D) if c["lik"] == "normal" else np.zeros(D)       # synthetic_test.py:25-35
    np.random.seed(319)
    return cls(data, D, diag=c["diag"], mean0=mean0, cov0=np.eye(D) * 0.1, k1=c["k1"], k2=c["k2"], batch_size=128,
               cov_eta=c.get("cov_eta", 1), randstart=True, device=device, noise_dtype=torch.float32,
               noise_slots=noise_slots, chunk_entries=chunk_entries, process_group=group, engine_kwargs=engine_kwargs)


def one_step(model, flush=None, batch_size=None):
    if flush is not None:
        flush.zero_()                     # inputs smaller than L2: evict them between iterations
    for dim in range(model.order):
        model.sweep(dim, batch_size)
        model.update_hyper_parameter(dim)


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


# ------------------------------------------------------------------------------------------------
# CPU baseline: the oracle port (numpy restatement of the reference step) on host cores
# ------------------------------------------------------------------------------------------------
def _cpu_worker(args):
    c, seed, budget_s, max_rows = args
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    from oracle import ssvi_oracle as so
    rng = np.random.default_rng(seed)
    N, D, B = len(c["dims"]), c["D"], 128
    dims = [min(n, 64) for n in c["dims"]]           # per-update cost does not depend on the tensor size
    cfg = so.make_config(c["model"], c["lik"], D, diag=c["diag"], k1=c["k1"], k2=c["k2"], batch_size=B,
                         cov_eta=c.get("cov_eta", 1.0))
    st = so.new_state(cfg, dims, rng=rng, mean0=np.ones(D) if c["lik"] == "normal" else np.zeros(D),
                      cov0=np.eye(D) * 0.1)
    units, rows_done = 0, 0
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


def cpu_baseline(c, cores, budget_s, max_rows=10 ** 9):
    """units/s of the oracle over `cores` independent processes on disjoint synthetic rows (the
    reference has no intra-process parallelism; this mirrors its one-job-per-core Slurm usage)."""
    if cores == 1:
        res = [_cpu_worker((c, 1, budget_s, max_rows))]
    else:
        import multiprocessing as mp
        with mp.get_context("spawn").Pool(cores) as pool:
            res = pool.map(_cpu_worker, [(c, 100 + i, budget_s, max_rows) for i in range(cores)])
    units = sum(r[0] for r in res)
    wall = max(r[1] for r in res)
    return units / wall, units, wall


def run_reference_arm(args, c, rank, world):
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    per_step = max(2.0, min(20.0, 90.0 / max(args.steps + args.warmup, 1)))
    for _ in range(args.warmup):
        cpu_baseline(c, cores, min(per_step, 2.0))
    tot_units, tot_wall = 0, 0.0
    for _ in range(args.steps):
        _, u, w = cpu_baseline(c, cores, per_step)
        tot_units += u
        tot_wall += w
    v = tot_units / tot_wall
    sample = ("%d row updates of B=128 entries (same D/k1/k2/model/likelihood as %s, dims shrunk to <=64), "
              "%.0f s per step x %d procs" % (tot_units // 128, args.config, per_step, cores))
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * tot_wall / max(args.steps, 1), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": c["desc"], "sweep": "reference-style minibatch rows (B=128)"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


# ------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--config", default="c2", choices=sorted(CONFIGS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--scale", type=float, default=1.0, help="fraction of the config's nnz (debug only)")
    ap.add_argument("--chunk-entries", type=int, default=16)
    ap.add_argument("--noise-slots", type=int, default=128)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--table-mb", type=float, default=96.0, help="sample-table budget per call (L2-resident)")
    ap.add_argument("--sweep-chunk", type=int, default=64, help="max entries of one (row, slot) chunk")
    ap.add_argument("--sweep-streams", type=int, default=3, help="side streams the slot groups alternate between")
    ap.add_argument("--shard", default="random", choices=["random", "blocks"],
                    help="N > 1: which entries a rank holds -- random: its own uniform sample of the whole index space; "
                         "blocks: a uniform sample of its block of the index space (shard_window)")
    ap.add_argument("--strong", action="store_true",
                    help="strong scaling: the config's nnz (x --scale) is divided over the ranks")
    ap.add_argument("--tile-blocks", type=int, default=0,
                    help="L2 tiling of the sweep launch order: blocks per other mode (0 = from the table budget, 1 = off)")
    ap.add_argument("--batch-size", type=int, default=0,
                    help="0: full-row sweeps (headline).  B > 0: reference-style subsampled sweeps, every row takes at "
                         "most B entries drawn afresh each sweep (SSVI_TF_interface.py:170-174), scale = len / B")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true",
                    help="diagnostic runs only (config 5 at full size pins 35 GB of host memory for the e2e arm)")
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
    torch.cuda.set_device(local)
    device = "cuda:%d" % local
    group = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device(device))
        group = dist.group.WORLD
    window, grid = (shard_window(c["dims"], rank, world) if args.shard == "blocks" and world > 1 else (None, None))
    data = DeviceTensorData(c, device, seed=319 + int(args.config[1]) + 1000 * rank,
                            scale=args.scale / (world if args.strong else 1), window=window)
    model = build_model(c, data, device, group, args.chunk_entries, args.noise_slots,
                        dict(table_budget_bytes=int(args.table_mb * (1 << 20)), sweep_chunk_entries=args.sweep_chunk,
                             sweep_streams=args.sweep_streams, sweep_tile_blocks=args.tile_blocks))
    eng = model.engine
    N = model.order
    nnz_train = eng.nnz
    B = args.batch_size or None
    if B is None:
        units_per_step = N * nnz_train                   # per rank
    else:
        assert world == 1, "subsampled sweeps are single-shard"
        units_per_step = int(sum(np.minimum(np.diff(eng.rowptr_h[d]), B).sum() for d in range(N)))
    model.fresh_noise()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    in_bytes = eng.coords.numel() * 4 + eng.vals.numel() * 8 + 8 * N * nnz_train
    flush = None
    if in_bytes < 256 << 20:
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=device)
    l2_note = ("L2 flushed between iterations (256 MB memset; inputs are %.0f MB)" % (in_bytes / 1e6)) if flush is not None \
        else "inputs larger than L2 (COO+CSR %.0f MB per GPU)" % (in_bytes / 1e6)
    # ---- device-resident arm: `value` ------------------------------------------------------------
    for _ in range(args.warmup):
        one_step(model, flush, B)
    eng.check_flags()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    eng.profile_events = []
    launches0 = eng.launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        one_step(model, flush, B)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    launches = eng.launches - launches0
    events, eng.profile_events = eng.profile_events, None
    clocks = sampler.stop() if rank == 0 else None
    eng.check_flags()
    t = torch.tensor([ms], device=device, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    value = world * units_per_step * args.steps / (ms * 1e-3)

    # roofline of the dominant kernel (ssvi_accumulate), from CUDA events inside the timed region
    acc_ms = [a.elapsed_time(b) for a, b, _ in events]
    acc_units = [n for _, _, n in events]
    mean_row = units_per_step / float(sum(c["dims"]))
    bpu = algorithmic_bytes_per_unit(c, mean_row)
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    achieved = sum(acc_units) * bpu / (sum(acc_ms) * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": None, "kernel": "accumulate (stages 1-3)", "peak_source": peak_src,
                "algorithmic_bytes_per_unit": bpu, "kernel_ms_per_launch": float(np.mean(acc_ms)),
                "kernel_share_of_step": sum(acc_ms) / ms,
                "launch": "one call of Engine.accumulate = stages 1-3 of one mode sweep (CUDA events around it; a sweep "
                          "issues one sample-table + one accumulate launch per slot group, profiles/*_launches_*)",
                "fp64_pipe": fp64_pipe_stats(c, sum(acc_units) / (sum(acc_ms) * 1e-3)) if B is None else None,
                "note": "fp64 contract: the FP64 pipe (37 TFLOP/s measured), not HBM, bounds this kernel; see DESIGN.md 3.1"}
    tr_path = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tr_path):          # DRAM bytes per unit from the committed ncu capture of this kernel
        tr = json.load(open(tr_path)).get(args.config)
        if tr:
            roofline["traffic"] = tr["dram_bytes_per_unit"] * float(np.mean(acc_units))
            roofline["traffic_source"] = tr["source"]

    # ---- end-to-end arm: host buffers in, host result out, every step ----------------------------
    # Every step's inputs (COO entries, values, the per-mode entry lists the sweep kernels read, the noise pool) are
    # copied from pinned host memory and the posterior is copied back.  Two device input sets: the copy of step k+1
    # runs on a copy stream while step k computes (all K copies are inside the timed region).
    def run_e2e():
        from ssvi_tf_b200.engine import NoisePool, _ptr
        pin = lambda x: x.cpu().pin_memory()                                      # noqa: E731
        plans = [model.sweep_plan(d)[0] for d in range(N)] if B is None else [None] * N
        nz = eng.noise
        noise_names = [n for n in ("z_other", "z_own", "r_w", "z_f") if getattr(nz, n) is not None]
        lists0 = [pl.entry_ids for pl in plans] if B is None else list(eng.perm)   # the per-mode entry lists the sweeps read
        dev_sets = [dict(coords=eng.coords, vals=eng.vals, perm=lists0,
                         noise=[getattr(nz, n) for n in noise_names])]
        dev_sets.append(dict(coords=torch.empty_like(eng.coords), vals=torch.empty_like(eng.vals),
                             perm=[torch.empty_like(x) for x in lists0],
                             noise=[torch.empty_like(x) for x in dev_sets[0]["noise"]]))
        host = dict(coords=pin(eng.coords), vals=pin(eng.vals), perm=[pin(x) for x in dev_sets[0]["perm"]],
                    noise=[pin(x) for x in dev_sets[0]["noise"]])
        h_out = [torch.empty(p.shape, dtype=p.dtype).pin_memory() for p in eng.params]
        h2d = (host["coords"].numel() * 4 + host["vals"].numel() * 8 + sum(x.numel() * 8 for x in host["perm"])
               + sum(x.numel() * x.element_size() for x in host["noise"]))
        d2h = sum(p.numel() * 8 for p in h_out)
        main_stream = torch.cuda.current_stream()
        copy_stream = torch.cuda.Stream()
        ready = [torch.cuda.Event(), torch.cuda.Event()]
        free = [torch.cuda.Event(), torch.cuda.Event()]

        def stage(k):                       # H2D of step k's inputs into set k % 2
            ds = dev_sets[k % 2]
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(free[k % 2])
                ds["coords"].copy_(host["coords"], non_blocking=True)
                ds["vals"].copy_(host["vals"], non_blocking=True)
                for dp, hp in zip(ds["perm"], host["perm"]):
                    dp.copy_(hp, non_blocking=True)
                for dn, hn in zip(ds["noise"], host["noise"]):
                    dn.copy_(hn, non_blocking=True)
                ready[k % 2].record(copy_stream)

        def bind(k):                        # point the library at set k % 2 (host-side pointer swaps only)
            ds = dev_sets[k % 2]
            eng.coords, eng.vals = ds["coords"], ds["vals"]
            eng._call(eng.lib.ssvi_bind_data, _ptr(eng.coords), _ptr(eng.vals), eng.nnz)
            for d_, (pl, t_) in enumerate(zip(plans, ds["perm"])):
                if pl is None:
                    eng.perm[d_] = t_
                else:
                    pl.entry_ids = t_
            kw = dict(zip(noise_names, ds["noise"]))
            eng.set_noise(NoisePool(kw.get("z_other"), kw.get("z_own"), kw.get("r_w"), kw.get("z_f")))

        def e2e_run(steps):
            for e in free:
                e.record(main_stream)
            stage(0)
            for k in range(steps):
                if k + 1 < steps:
                    stage(k + 1)
                main_stream.wait_event(ready[k % 2])
                bind(k)
                one_step(model, flush, B)
                free[k % 2].record(main_stream)
                for ho, p in zip(h_out, eng.params):
                    ho.copy_(p, non_blocking=True)

        e2e_run(2)
        barrier()
        e0.record()
        e2e_run(args.steps)
        e1.record()
        barrier()
        bind(0)
        t = torch.tensor([e0.elapsed_time(e1)], device=device, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_value = world * units_per_step * args.steps / (float(t.item()) * 1e-3)
        eng.check_flags()
        return e2e_value, h2d, d2h

    e2e = None
    if not args.no_e2e:
        e2e_value, h2d, d2h = run_e2e()
        e2e = {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h}

    if rank == 0:
        cpu = None
        if not args.no_cpu_baseline and world == 1:
            v, u, w = cpu_baseline(c, 1, args.cpu_seconds)
            cpu = {"value": v, "unit": UNIT, "cores": 1, "kind": "port",
                   "sample": "%d row updates of B=128 entries, same D/k1/k2/model/likelihood, dims shrunk to <=64, "
                             "%.1f s, 1 thread (oracle/ssvi_oracle.py)" % (u // 128, w)}
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "strong" if args.strong else "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": c["desc"], "config_id": args.config, "nnz_train_per_gpu": nnz_train,
                       "units_per_step_per_gpu": units_per_step,
                       "sweep": "full rows (scale=1), all modes" if B is None else
                       "reference-style minibatches: every row takes <= %d entries, drawn on the device each sweep "
                       "(scale = len/B), all rows of all modes" % B,
                       "noise": "f32 pool, %d slots" % eng.noise.slots, "chunk_entries": args.chunk_entries,
                       "path": ("slot-major sweep kernels, %d slots/call, %d L2 tile blocks per other mode"
                                % (model.sweep_plan(0)[0].groups[0][1], getattr(model.sweep_plan(0)[0], "tile_blocks", 1)))
                       if eng.sweep_supported and B is None else "direct (one warp per row-chunk)",
                       "l2": l2_note,
                       "parallelism": "entries sharded x%d (%s), all-reduce of row accumulators"
                                      % (world, "index-space blocks %s" % "x".join(map(str, grid)) if grid else
                                         "every rank a uniform sample of the whole index space")},
            "roofline": roofline, "cpu_baseline": cpu,
            "e2e": e2e, "gpu_launches": launches, "clocks": clocks}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
