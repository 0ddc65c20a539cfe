#!/bin/bash
# quick GPU visit: sweep-path parity test + a few bench settings
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-q}
timeout 900 python -m pytest tests -m gpu -x -q -k "sweep_table or golden" > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 $OUT/${TAG}_pytest.log
for cfg in "48 64" "24 64" "96 64" "48 32" "48 128"; do
  set -- $cfg
  timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --table-mb $1 --sweep-chunk $2 > $OUT/${TAG}_bench_$1_$2.json 2> $OUT/${TAG}_bench_$1_$2.err
  echo "table-mb=$1 chunk=$2 rc=$?"; python - <<PY
import json
try:
    d=json.load(open("$OUT/${TAG}_bench_$1_$2.json"))
    print("  value %.4g  e2e %.4g  ms/step %.1f  frac %.3f  kernel_ms %.2f share %.3f launches %d" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d["roofline"]["frac"], d["roofline"]["kernel_ms_per_launch"], d["roofline"]["kernel_share_of_step"], d["gpu_launches"]))
except Exception as e:
    print("  failed", e); print(open("$OUT/${TAG}_bench_$1_$2.err").read()[-2000:])
PY
done
