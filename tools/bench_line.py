#!/usr/bin/env python
"""One-screen digest of a bench.py JSON line: usage bench_line.py file.json"""
import json
import sys

try:
    d = json.loads([ln for ln in open(sys.argv[1]) if ln.startswith("{")][-1])      # NCCL prints its version first
except Exception as e:                                      # noqa: BLE001
    print("  no JSON line:", e)
    sys.exit(0)
r = d.get("roofline") or {}
print("  value %.4g %s  ms/step %.2f  frac %.3f  e2e %s  N=%s  parity=%s  wall %ss" % (
    d["value"], d["unit"], d["ms_per_step"], r.get("frac", float("nan")),
    "%.4g" % d["e2e"]["value"] if d.get("e2e") else None, d["n_gpus"], d.get("multi_gpu_parity"), d.get("wall_s")))
det = d.get("details") or {}
print("  path:", det.get("path"), "| setup", det.get("setup_s"), "| clocks", d.get("clocks"))
for t, v in (det.get("ab_tunings") or {}).items():
    print("  tuning %-3s value %.4g frac %.3f" % (t, v["value"], v["roofline_frac"]))
for k, v in (d.get("configs") or {}).items():
    print("  %-28s %s" % (k, v.get("error") or "value %.4g  ms/step %.3f  frac %.3f  %s" % (
        v["value"], v["ms_per_step"], v["roofline_frac"], v["kernel"][:48])))
