#!/usr/bin/env python
"""Print the metrics that matter from an .ncu-rep (raw page) or from the raw page exported on the GPU box by
tools/ncu_export.sh: usage ncu_summary.py file.ncu-rep | file.raw.csv"""
import csv, subprocess, sys
if sys.argv[1].endswith(".csv"):
    out = open(sys.argv[1]).read()
else:
    out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]
keys = ["Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "launch__occupancy_limit_warps", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__ops_path_tensor_src_fp64.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.max", "smsp__inst_executed.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__warps_eligible.avg.per_cycle_active"]
for r in rows[2:]:
    d = dict(zip(hdr, r))
    print("=" * 100)
    for k in keys:
        if k in d:
            print("%-70s %-12s %s" % (k, units[hdr.index(k)], d[k][:110]))
    for k in hdr:
        if "issue_stalled" in k and k.endswith("per_issue_active.ratio") and float(d[k] or 0) > 0.05:
            print("%-70s %s" % (k.replace("smsp__average_warps_issue_stalled_", "stall:").replace("_per_issue_active.ratio", ""), d[k]))
    for k in hdr:
        if k.startswith("sm__inst_executed_pipe_") and k.endswith(".sum") and float((d[k] or "0").replace(",", "")) > 0:
            print("%-70s %s" % (k, d[k]))
